// hpcx/Multi.hpp — several GPUs from ONE process and ONE host thread.
//
// The reference's callers are single-process OpenMP programs that index their containers from
// their own loops (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:356); the one multi-stream
// idiom it has is driven by a single thread as well (lib/gpu/transfer_overlap_testsuite/src/
// cuda_tut_transfer_overlap.cu:92-103).  The overloads below take the same host containers as the
// single-GPU entry points of Stencil.hpp / CsrMatrix.hpp, plus a hpcx::MultiContext instead of a
// hpcx::Context, and spread the work over every device of it:
//
//   hpcx::MultiContext gpus;                               // all visible devices
//   sparse::MultiDeviceCsr<double> dA(gpus, A);            // row blocks + halo tables, built here
//   sparse::spmv(dA, x, y);                                // host vectors in, host vector out
//   hpc::stencil_sweep(gpus, a, b, hpc::Jacobi7{100});     // z-slabs, ghost planes by peer stores
//   hpc::stencil_sweep(gpus, a, b, hpc::Box1D{1, 8});      // y-slabs of the reference's blur sweeps
//
// Partitioning, halo index maps and the exchange are the library's (hpcx_multi_*, hpcx.h): the
// same kernels and the same fused peer-store transport as the one-process-per-GPU path, with
// plain peer pointers; no MPI, no NCCL, no torchrun.  Results are bit-identical to the
// single-device sweeps, and to the single-device product for rows short enough to keep the
// CPU summation order.
#pragma once
#ifndef HPCX_MULTI_HPP_
#define HPCX_MULTI_HPP_

#include <vector>

#include "CsrMatrix.hpp"
#include "Stencil.hpp"

namespace hpcx {

class MultiContext {
 public:
  // devices.empty(): devices 0 .. count-1 of the process (count <= 0: every visible device)
  explicit MultiContext(const std::vector<int> &devices = {}, int count = 0) {
    std::vector<int32_t> ids(devices.begin(), devices.end());
    int32_t n = static_cast<int32_t>(ids.size());
    if (ids.empty()) {
      n = count;
      if (n <= 0) {
        n = hpcx_device_count();
        if (n <= 0) throwStatus(HPCX_CUDA_ERROR, "no CUDA device is visible", __FILE__, __LINE__);
      }
    }
    const hpcx_status st = hpcx_init_multi(n, ids.empty() ? nullptr : ids.data(), &m_);
    if (st != HPCX_OK) {
      const std::string msg = m_ ? hpcx_multi_last_error_string(m_) : "hpcx_init_multi failed";
      if (m_) hpcx_multi_finalize(m_);
      m_ = nullptr;
      throwStatus(st, msg.c_str(), __FILE__, __LINE__);
    }
  }
  ~MultiContext() {
    if (m_) hpcx_multi_finalize(m_);
  }
  MultiContext(const MultiContext &) = delete;
  MultiContext &operator=(const MultiContext &) = delete;

  hpcx_multi handle() const noexcept { return m_; }
  int size() const noexcept { return hpcx_multi_size(m_); }
  void check(hpcx_status st, const char *file, int line) const {
    if (st != HPCX_OK) throwStatus(st, hpcx_multi_last_error_string(m_), file, line);
  }

 private:
  hpcx_multi m_ = nullptr;
};
#define HPCX_MCHECK(mc, expr) (mc).check((expr), __FILE__, __LINE__)

}  // namespace hpcx

namespace sparse {

// A host CsrMatrix spread over the devices of a MultiContext (contiguous row blocks; by rows, or by
// nonzeros for power-law matrices).
template <typename T>
class MultiDeviceCsr {
 public:
  template <typename Index>
  MultiDeviceCsr(hpcx::MultiContext &gpus, const CsrMatrix<T, Index> &A, bool balance_by_nnz = false,
                 const hpcx_csr_opts *opts = nullptr)
      : gpus_(gpus), nrows_(A.nrows()), ncols_(A.ncols()) {
    const hpcx_csr_desc d = A.describe();
    HPCX_MCHECK(gpus_, hpcx_multi_csr_create(gpus.handle(), &d, opts, balance_by_nnz ? 1 : 0, &A_));
  }
  ~MultiDeviceCsr() {
    if (A_) hpcx_multi_csr_destroy(A_);
  }
  MultiDeviceCsr(const MultiDeviceCsr &) = delete;
  MultiDeviceCsr &operator=(const MultiDeviceCsr &) = delete;

  void multiply(const T *x, T *y) const { HPCX_MCHECK(gpus_, hpcx_multi_spmv(A_, x, y)); }
  std::vector<int64_t> bounds() const {
    std::vector<int64_t> b(static_cast<std::size_t>(gpus_.size()) + 1);
    HPCX_MCHECK(gpus_, hpcx_multi_csr_bounds(A_, b.data()));
    return b;
  }
  std::size_t nrows() const noexcept { return nrows_; }
  std::size_t ncols() const noexcept { return ncols_; }

 private:
  hpcx::MultiContext &gpus_;
  hpcx_multi_csr A_ = nullptr;
  std::size_t nrows_ = 0, ncols_ = 0;
};

template <typename T, class AllocX, class AllocY>
void spmv(const MultiDeviceCsr<T> &A, const std::vector<T, AllocX> &x, std::vector<T, AllocY> &y) {
  if (x.size() != A.ncols()) throw std::out_of_range("invalid index, out of boundary");
  y.resize(A.nrows());
  A.multiply(x.data(), y.data());
}

}  // namespace sparse

namespace hpc {

// Host containers in, host containers out, all devices of `gpus`.
template <typename T, std::size_t Lo, std::size_t Hi, std::size_t A, class Al, std::size_t Lo2, std::size_t Hi2,
          std::size_t A2, class Al2>
void stencil_sweep(hpcx::MultiContext &gpus, const HPCHighDimensionFlatArray<2, T, Lo, Hi, A, Al> &a,
                   HPCHighDimensionFlatArray<2, T, Lo2, Hi2, A2, Al2> &b, Box1D box) {
  hpcx_grid_desc da{2, hpcx::dtypeOf<T>(), (int32_t)Lo, (int32_t)Hi, {(int64_t)a.dim()[0], (int64_t)a.dim()[1], 0}};
  hpcx_grid_desc db{2, hpcx::dtypeOf<T>(), (int32_t)Lo2, (int32_t)Hi2, {(int64_t)b.dim()[0], (int64_t)b.dim()[1], 0}};
  hpcx_multi_grid ga = nullptr, gb = nullptr;
  HPCX_MCHECK(gpus, hpcx_multi_grid_create(gpus.handle(), &da, &ga));
  hpcx_status st = hpcx_multi_grid_create(gpus.handle(), &db, &gb);
  hpcx_stencil_desc sd{};
  sd.kind = box.axis == 1 ? HPCX_STENCIL_BOX_X : HPCX_STENCIL_BOX_Y;
  sd.radius = box.radius;
  if (st == HPCX_OK) st = hpcx_multi_grid_upload(ga, a.data());
  if (st == HPCX_OK) st = hpcx_multi_stencil_sweep(ga, gb, &sd, 1);
  if (st == HPCX_OK) st = hpcx_multi_grid_download(gb, b.data());
  std::string msg = st == HPCX_OK ? "" : hpcx_multi_last_error_string(gpus.handle());
  hpcx_multi_grid_destroy(ga);
  if (gb) hpcx_multi_grid_destroy(gb);
  if (st != HPCX_OK) hpcx::throwStatus(st, msg.c_str(), __FILE__, __LINE__);
}

template <typename T, std::size_t Lo, std::size_t Hi, std::size_t A, class Al>
void stencil_sweep(hpcx::MultiContext &gpus, HPCHighDimensionFlatArray<3, T, Lo, Hi, A, Al> &a,
                   HPCHighDimensionFlatArray<3, T, Lo, Hi, A, Al> &b, Jacobi7 j) {
  static_assert(Lo >= 1 && Hi >= 1, "Jacobi7 needs at least one ghost cell per side");
  hpcx_grid_desc d{3, hpcx::dtypeOf<T>(), (int32_t)Lo, (int32_t)Hi,
                   {(int64_t)a.dim()[0], (int64_t)a.dim()[1], (int64_t)a.dim()[2]}};
  hpcx_multi_grid ga = nullptr, gb = nullptr;
  HPCX_MCHECK(gpus, hpcx_multi_grid_create(gpus.handle(), &d, &ga));
  hpcx_status st = hpcx_multi_grid_create(gpus.handle(), &d, &gb);
  hpcx_stencil_desc sd{};
  sd.kind = HPCX_STENCIL_JACOBI7;
  sd.variant = j.variant;
  if (st == HPCX_OK) st = hpcx_multi_grid_upload(ga, a.data());
  if (st == HPCX_OK) st = hpcx_multi_grid_upload(gb, b.data());
  if (st == HPCX_OK) st = hpcx_multi_stencil_sweep(ga, gb, &sd, j.steps);
  if (st == HPCX_OK) st = hpcx_multi_grid_download(gb, b.data());
  if (st == HPCX_OK && j.steps > 1) st = hpcx_multi_grid_download(ga, a.data());
  std::string msg = st == HPCX_OK ? "" : hpcx_multi_last_error_string(gpus.handle());
  hpcx_multi_grid_destroy(ga);
  if (gb) hpcx_multi_grid_destroy(gb);
  if (st != HPCX_OK) hpcx::throwStatus(st, msg.c_str(), __FILE__, __LINE__);
}

}  // namespace hpc
#endif  // HPCX_MULTI_HPP_
