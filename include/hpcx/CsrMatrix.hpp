// hpcx/CsrMatrix.hpp — sparse::CsrMatrix<T, Index> and sparse::spmv on the B200.
//
// The reference's lib/sparse is a hierarchical sparse *grid* (RootGrid / HashBlock /
// PointerBlock / DenseBlock, lib/sparse/include/SparseDS.hpp:4-7); it has no matrix and no
// y = A*x.  This header adds the CSR container next to it, in the same
// namespace and in the style of the reference's containers: a header-only C++17 template over
// flat aligned buffers (std::vector over hpc::AlignedAllocator<_, 64>,
// lib/hpc/include/AlignedAlloc.hpp:29), whose product is one C-ABI call into libhpcx.so.
//
//   row_ptr[nrows+1]  Index (int32_t or int64_t), row_ptr[0] == 0, non-decreasing
//   col_idx[nnz]      int32_t, ascending within a row
//   values[nnz]       T (float or double)
//
// Arithmetic contract of spmv (oracle/hpcx_oracle.c): per row, products values[p]*x[col_idx[p]]
// rounded separately and added in ascending p; the device reduction may reorder the adds of
// rows longer than 8 entries (tolerance 1e-12 fp64 / 1e-5 fp32 relative to sum |a*x|).
#pragma once
#ifndef HPCX_CSR_MATRIX_HPP_
#define HPCX_CSR_MATRIX_HPP_

#include <algorithm>
#include <cstdint>
#include <numeric>
#include <stdexcept>
#include <tuple>
#include <utility>
#include <mutex>
#include <vector>

#include "AlignedAlloc.hpp"
#include "CudaAllocator.hpp"
#include "Runtime.hpp"

namespace sparse {

template <typename T, typename Index = std::int32_t>
class CsrMatrix {
  static_assert(std::is_same_v<Index, std::int32_t> || std::is_same_v<Index, std::int64_t>,
                "row_ptr entries are int32_t or int64_t");

 public:
  using value_type = T;
  using index_type = Index;
  template <typename U>
  using buffer = std::vector<U, hpc::AlignedAllocator<U, 64>>;

  CsrMatrix() : row_ptr_(1, Index{0}) {}
  CsrMatrix(std::size_t nrows, std::size_t ncols, buffer<Index> row_ptr, buffer<std::int32_t> col_idx,
            buffer<T> values)
      : nrows_(nrows), ncols_(ncols), row_ptr_(std::move(row_ptr)), col_idx_(std::move(col_idx)),
        values_(std::move(values)) {
    validate();
  }

  // Coordinate triplets in any order; entries with equal (row, col) are merged by adding them
  // in their input order (the R-MAT generator of BASELINE configs[4] relies on that).
  static CsrMatrix from_triplets(std::size_t nrows, std::size_t ncols,
                                 std::vector<std::tuple<std::int64_t, std::int32_t, T>> entries) {
    std::stable_sort(entries.begin(), entries.end(), [](const auto &a, const auto &b) {
      return std::get<0>(a) != std::get<0>(b) ? std::get<0>(a) < std::get<0>(b)
                                              : std::get<1>(a) < std::get<1>(b);
    });
    CsrMatrix m;
    m.nrows_ = nrows;
    m.ncols_ = ncols;
    m.row_ptr_.assign(nrows + 1, Index{0});
    for (std::size_t i = 0; i < entries.size(); ++i) {
      const auto &[r, c, v] = entries[i];
      if (r < 0 || static_cast<std::size_t>(r) >= nrows || c < 0 || static_cast<std::size_t>(c) >= ncols)
        throw std::out_of_range("invalid index, out of boundary");
      if (i > 0 && std::get<0>(entries[i - 1]) == r && std::get<1>(entries[i - 1]) == c) {
        m.values_.back() = m.values_.back() + v;
        continue;
      }
      m.col_idx_.push_back(c);
      m.values_.push_back(v);
      ++m.row_ptr_[static_cast<std::size_t>(r) + 1];
    }
    std::partial_sum(m.row_ptr_.begin(), m.row_ptr_.end(), m.row_ptr_.begin());
    return m;
  }

  // 2-D 5-point Laplacian on an ny x nx grid (BASELINE configs[0], configs[1]): row i = y*nx+x,
  // columns ascending {i-nx, i-1, i, i+1, i+nx}, out-of-grid neighbours omitted,
  // values {-1,-1,4,-1,-1}; nnz = 5*nx*ny - 2*nx - 2*ny.
  static CsrMatrix laplace2d(std::size_t ny, std::size_t nx) {
    CsrMatrix m;
    m.nrows_ = m.ncols_ = ny * nx;
    const std::size_t nnz = 5 * nx * ny - 2 * nx - 2 * ny;
    m.row_ptr_.resize(m.nrows_ + 1);
    m.col_idx_.resize(nnz);
    m.values_.resize(nnz);
    std::size_t p = 0;
    for (std::size_t y = 0; y < ny; ++y)
      for (std::size_t x = 0; x < nx; ++x) {
        const std::size_t i = y * nx + x;
        m.row_ptr_[i] = static_cast<Index>(p);
        auto put = [&](std::size_t c, T v) {
          m.col_idx_[p] = static_cast<std::int32_t>(c);
          m.values_[p++] = v;
        };
        if (y > 0) put(i - nx, T(-1));
        if (x > 0) put(i - 1, T(-1));
        put(i, T(4));
        if (x + 1 < nx) put(i + 1, T(-1));
        if (y + 1 < ny) put(i + nx, T(-1));
      }
    m.row_ptr_[m.nrows_] = static_cast<Index>(p);
    return m;
  }

  // 3-D 27-point operator on n^3 (BASELINE configs[3]): 27 neighbours ascending, Dirichlet
  // truncation, centre 26, others -1; nnz = (3n-2)^3.
  static CsrMatrix op27(std::size_t n) {
    CsrMatrix m;
    m.nrows_ = m.ncols_ = n * n * n;
    const std::size_t nnz = (3 * n - 2) * (3 * n - 2) * (3 * n - 2);
    m.row_ptr_.resize(m.nrows_ + 1);
    m.col_idx_.resize(nnz);
    m.values_.resize(nnz);
    const std::int64_t N = static_cast<std::int64_t>(n);
    std::size_t p = 0;
    for (std::int64_t z = 0; z < N; ++z)
      for (std::int64_t y = 0; y < N; ++y)
        for (std::int64_t x = 0; x < N; ++x) {
          m.row_ptr_[static_cast<std::size_t>((z * N + y) * N + x)] = static_cast<Index>(p);
          for (std::int64_t dz = -1; dz <= 1; ++dz)
            for (std::int64_t dy = -1; dy <= 1; ++dy)
              for (std::int64_t dx = -1; dx <= 1; ++dx) {
                const std::int64_t zz = z + dz, yy = y + dy, xx = x + dx;
                if (zz < 0 || zz >= N || yy < 0 || yy >= N || xx < 0 || xx >= N) continue;
                m.col_idx_[p] = static_cast<std::int32_t>((zz * N + yy) * N + xx);
                m.values_[p++] = (dz == 0 && dy == 0 && dx == 0) ? T(26) : T(-1);
              }
        }
    m.row_ptr_[m.nrows_] = static_cast<Index>(p);
    return m;
  }

  std::size_t nrows() const noexcept { return nrows_; }
  std::size_t ncols() const noexcept { return ncols_; }
  std::size_t nnz() const noexcept { return col_idx_.size(); }
  const Index *row_ptr() const noexcept { return row_ptr_.data(); }
  const std::int32_t *col_idx() const noexcept { return col_idx_.data(); }
  const T *values() const noexcept { return values_.data(); }
  T *values() noexcept { return values_.data(); }

  // nnz*(sizeof T + 4) + nrows*(sizeof Index + 2*sizeof T) + sizeof Index.
  std::size_t algorithmic_bytes() const noexcept {
    return nnz() * (sizeof(T) + 4) + nrows_ * (sizeof(Index) + 2 * sizeof(T)) + sizeof(Index);
  }

  hpcx_csr_desc describe() const noexcept {
    hpcx_csr_desc d{};
    d.nrows = static_cast<std::int64_t>(nrows_);
    d.ncols = static_cast<std::int64_t>(ncols_);
    d.nnz = static_cast<std::int64_t>(nnz());
    d.row_ptr = row_ptr_.data();
    d.idx_bytes = static_cast<std::int32_t>(sizeof(Index));
    d.dtype = hpcx::dtypeOf<T>();
    d.col_idx = col_idx_.data();
    d.values = values_.data();
    d.mem = HPCX_MEM_HOST;
    return d;
  }

 private:
  void validate() const {
    if (row_ptr_.size() != nrows_ + 1 || row_ptr_.front() != 0 ||
        static_cast<std::size_t>(row_ptr_.back()) != col_idx_.size() || col_idx_.size() != values_.size())
      throw std::out_of_range("invalid index, out of boundary");
  }

  std::size_t nrows_ = 0, ncols_ = 0;
  buffer<Index> row_ptr_;
  buffer<std::int32_t> col_idx_;
  buffer<T> values_;
};

// Sparse-grid bridge (hpcx_csr_from_cells): the 5-point operator over the active
// cells of a sparse 2-D grid.  `Grid` is anything with the reference's
// sparse::RootGrid<T, Layout>::foreach(func(x, y, value&)) (lib/sparse/include/RootGrid.hpp:20-22,
// 62-75; the leaves call func concurrently, hence the lock).  Row r of the operator is the r-th
// distinct cell in (x, y) order; `cells` receives those coordinates, `x0` the cells' values.
struct ActiveCells {
  std::vector<std::int32_t> row_x, row_y;
};
template <typename T>
class DeviceCsr;
template <typename T, typename Grid>
DeviceCsr<T> csr_from_grid(hpcx::Context &ctx, Grid &grid, ActiveCells *cells = nullptr,
                           std::vector<T> *x0 = nullptr, const hpcx_csr_opts *opts = nullptr);

// Device-resident matrix (hpcx_csr_create): upload once, multiply many times.
template <typename T>
class DeviceCsr {
 public:
  template <typename Index>
  DeviceCsr(hpcx::Context &ctx, const CsrMatrix<T, Index> &A, const hpcx_csr_opts *opts = nullptr)
      : ctx_(ctx.handle()), nrows_(A.nrows()), ncols_(A.ncols()) {
    const hpcx_csr_desc d = A.describe();
    HPCX_CHECK(ctx_, hpcx_csr_create(ctx_, &d, opts, &A_));
  }
  ~DeviceCsr() {
    if (A_) hpcx_csr_destroy(A_);
  }
  DeviceCsr(const DeviceCsr &) = delete;
  DeviceCsr &operator=(const DeviceCsr &) = delete;

  // x, y in host memory (staged through the copy pipeline) ...
  void multiply(const T *x, T *y) const {
    HPCX_CHECK(ctx_, hpcx_spmv(A_, x, HPCX_MEM_HOST, y, HPCX_MEM_HOST, nullptr));
  }
  // ... or already on the device / in managed memory (cudaAllocator.hpp:39-83 vectors).
  void multiply_device(const T *x, T *y, void *stream = nullptr) const {
    HPCX_CHECK(ctx_, hpcx_spmv(A_, x, HPCX_MEM_DEVICE, y, HPCX_MEM_DEVICE, stream));
  }
  // ... or each where its allocator put it (HPCX_MEM_*; see hpcx::memKindOf).
  void multiply(const T *x, int32_t x_mem, T *y, int32_t y_mem, void *stream = nullptr) const {
    HPCX_CHECK(ctx_, hpcx_spmv(A_, x, x_mem, y, y_mem, stream));
  }

  // Device assembly from an edge (COO) list, duplicates merged by adding in list order
  // (hpcx_csr_from_coo): the arrays never visit the host.
  static DeviceCsr from_coo(hpcx::Context &ctx, std::size_t nrows, const std::vector<std::int32_t> &rows,
                            const std::vector<std::int32_t> &cols, const std::vector<T> &vals,
                            const hpcx_csr_opts *opts = nullptr) {
    if (rows.size() != cols.size() || rows.size() != vals.size())
      throw std::out_of_range("invalid index, out of boundary");
    hpcx_csr_arrays arr{};
    HPCX_CHECK(ctx.handle(), hpcx_csr_from_coo(ctx.handle(), static_cast<int64_t>(nrows),
                                               static_cast<int64_t>(rows.size()), rows.data(), cols.data(),
                                               vals.data(), hpcx::dtypeOf<T>(), HPCX_MEM_HOST, 4, &arr));
    hpcx_csr_desc d{};
    d.nrows = d.ncols = static_cast<int64_t>(nrows);
    d.nnz = arr.nnz;
    d.row_ptr = arr.row_ptr;
    d.idx_bytes = 4;
    d.dtype = hpcx::dtypeOf<T>();
    d.col_idx = arr.col_idx;
    d.values = arr.values;
    d.mem = HPCX_MEM_DEVICE;
    DeviceCsr out(ctx.handle(), nrows);
    const hpcx_status st = hpcx_csr_create(ctx.handle(), &d, opts, &out.A_);
    hpcx_device_free(ctx.handle(), arr.row_ptr);
    hpcx_device_free(ctx.handle(), arr.col_idx);
    hpcx_device_free(ctx.handle(), arr.values);
    HPCX_CHECK(ctx.handle(), st);
    return out;
  }
  DeviceCsr(DeviceCsr &&o) noexcept : ctx_(o.ctx_), A_(o.A_), nrows_(o.nrows_), ncols_(o.ncols_) { o.A_ = nullptr; }

  hpcx_csr handle() const noexcept { return A_; }
  hpcx_ctx context() const noexcept { return ctx_; }
  std::size_t nrows() const noexcept { return nrows_; }
  std::size_t ncols() const noexcept { return ncols_; }

 private:
  template <typename U, typename Grid>
  friend DeviceCsr<U> csr_from_grid(hpcx::Context &, Grid &, ActiveCells *, std::vector<U> *,
                                    const hpcx_csr_opts *);
  DeviceCsr(hpcx_ctx ctx, std::size_t n) : ctx_(ctx), nrows_(n), ncols_(n) {}
  hpcx_ctx ctx_ = nullptr;
  hpcx_csr A_ = nullptr;
  std::size_t nrows_ = 0, ncols_ = 0;
};

// Conjugate gradients around the product (hpcx_cg_*): q = A*p with <p, q> fused
// into the SpMV kernel, scalars resident on the device.  A must be symmetric positive definite.
struct CgResult {
  int iterations = 0;
  bool converged = false;
  double rr = 0, bb = 0;  // ||r||^2 (recurrence) and ||b||^2
};
template <typename T>
class CgSolver {
 public:
  explicit CgSolver(const DeviceCsr<T> &A) : ctx_(A.context()) { HPCX_CHECK(ctx_, hpcx_cg_create(A.handle(), &S_)); }
  ~CgSolver() {
    if (S_) hpcx_cg_destroy(S_);
  }
  CgSolver(const CgSolver &) = delete;
  CgSolver &operator=(const CgSolver &) = delete;

  // x is the start vector and receives the iterate (host containers).
  template <class AllocB, class AllocX>
  CgResult solve(const std::vector<T, AllocB> &b, std::vector<T, AllocX> &x, int max_iters, double rel_tol = 0.0,
                 int check_every = 0) {
    if (b.size() != x.size()) throw std::out_of_range("invalid index, out of boundary");
    hpcx_cg_result r{};
    HPCX_CHECK(ctx_, hpcx_cg_solve(S_, b.data(), HPCX_MEM_HOST, x.data(), HPCX_MEM_HOST, max_iters, rel_tol,
                                   check_every, &r, nullptr));
    return CgResult{r.iterations, r.converged == 1, r.rr, r.bb};
  }

 private:
  hpcx_ctx ctx_ = nullptr;
  hpcx_cg S_ = nullptr;
};

// y = A*x with host containers: the reference-facing free function.
template <typename T, typename Index, class AllocX, class AllocY>
void spmv(hpcx::Context &ctx, const CsrMatrix<T, Index> &A, const std::vector<T, AllocX> &x,
          std::vector<T, AllocY> &y) {
  if (x.size() != A.ncols()) throw std::out_of_range("invalid index, out of boundary");
  y.resize(A.nrows());
  DeviceCsr<T> D(ctx, A);
  D.multiply(x.data(), hpcx::memKindOf<AllocX>::value, y.data(), hpcx::memKindOf<AllocY>::value);
}

template <typename T, class AllocX, class AllocY>
void spmv(const DeviceCsr<T> &A, const std::vector<T, AllocX> &x, std::vector<T, AllocY> &y) {
  if (x.size() != A.ncols()) throw std::out_of_range("invalid index, out of boundary");
  y.resize(A.nrows());
  A.multiply(x.data(), hpcx::memKindOf<AllocX>::value, y.data(), hpcx::memKindOf<AllocY>::value);
}

template <typename T, typename Grid>
DeviceCsr<T> csr_from_grid(hpcx::Context &ctx, Grid &grid, ActiveCells *cells, std::vector<T> *x0,
                           const hpcx_csr_opts *opts) {
  std::vector<std::int32_t> xs, ys;
  std::vector<T> vs;
  std::mutex mu;
  grid.foreach ([&](auto x, auto y, auto &value) {
    std::lock_guard<std::mutex> lock(mu);
    xs.push_back(static_cast<std::int32_t>(x));
    ys.push_back(static_cast<std::int32_t>(y));
    vs.push_back(static_cast<T>(value));
  });
  hpcx_csr_arrays arr{};
  hpcx_cell_map map{};
  HPCX_CHECK(ctx.handle(), hpcx_csr_from_cells(ctx.handle(), static_cast<int64_t>(xs.size()), xs.data(), ys.data(),
                                               HPCX_MEM_HOST, hpcx::dtypeOf<T>(), 4, &arr, &map));
  const std::size_t m = static_cast<std::size_t>(map.nrows);
  std::vector<std::int32_t> cell_row(xs.size());
  hpcx_status st = HPCX_OK;
  if (!xs.empty()) st = hpcx_memcpy(ctx.handle(), cell_row.data(), map.cell_row, xs.size() * 4);
  if (cells && st == HPCX_OK) {
    cells->row_x.resize(m);
    cells->row_y.resize(m);
    if (m) st = hpcx_memcpy(ctx.handle(), cells->row_x.data(), map.row_x, m * 4);
    if (m && st == HPCX_OK) st = hpcx_memcpy(ctx.handle(), cells->row_y.data(), map.row_y, m * 4);
  }
  if (x0 && st == HPCX_OK) {
    x0->assign(m, T{});
    for (std::size_t i = 0; i < xs.size(); ++i) (*x0)[cell_row[i]] = vs[i];
  }
  hpcx_csr_desc d{};
  d.nrows = d.ncols = map.nrows;
  d.nnz = arr.nnz;
  d.row_ptr = arr.row_ptr;
  d.idx_bytes = 4;
  d.dtype = hpcx::dtypeOf<T>();
  d.col_idx = arr.col_idx;
  d.values = arr.values;
  d.mem = HPCX_MEM_DEVICE;
  DeviceCsr<T> out(ctx.handle(), m);
  if (st == HPCX_OK) st = hpcx_csr_create(ctx.handle(), &d, opts, &out.A_);
  for (void *p : {arr.row_ptr, (void *)arr.col_idx, arr.values, (void *)map.cell_row, (void *)map.row_x,
                  (void *)map.row_y})
    if (p) hpcx_device_free(ctx.handle(), p);
  HPCX_CHECK(ctx.handle(), st);
  return out;
}

}  // namespace sparse
#endif  // HPCX_CSR_MATRIX_HPP_
