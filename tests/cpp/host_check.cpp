// host_check.cpp — exercises the C++ face of libhpcx (include/hpcx/*.hpp).
//
//   host_check index            prints strides / totals / offsets / at() verdicts of
//                               hpc::HPCHighDimensionFlatArray for tests/test_cpp_host.py to compare
//                               with the reference header (oracle/_ref) and the golden file
//   host_check csr              prints FNV-1a digests of the CsrMatrix builders' arrays
//   host_check multi            one process, one thread, every visible GPU (hpcx::MultiContext): the
//                               reference's sweep loop nests and a CPU SpMV loop against the multi-device
//                               overloads of hpc::stencil_sweep / sparse::spmv
//   host_check gpu              runs the sweeps and the product on cuda:0 through the C++ entry
//                               points and compares with reference-style loop nests written here
//
// Built by tests/test_cpp_host.py with g++ -std=c++17 -fopenmp -Iinclude -Llibhpc_b200 -lhpcx.
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <tuple>
#include <vector>

#include <hpcx/CsrMatrix.hpp>
#include <hpcx/Distributed.hpp>
#include <hpcx/HPCHighDimensionFlatArray.hpp>
#include <hpcx/Multi.hpp>
#include <hpcx/Stencil.hpp>

static uint64_t fnv1a(const void *p, size_t n, uint64_t h = 1469598103934665603ull) {
  const unsigned char *b = static_cast<const unsigned char *>(p);
  for (size_t i = 0; i < n; ++i) h = (h ^ b[i]) * 1099511628211ull;
  return h;
}

// Exposes the protected helpers the reference header also keeps protected.
template <std::size_t D, typename T, std::size_t Lo, std::size_t Hi>
struct Probe : hpc::HPCHighDimensionFlatArray<D, T, Lo, Hi> {
  using Base = hpc::HPCHighDimensionFlatArray<D, T, Lo, Hi>;
  using Base::Base;
  static void layout(const std::array<std::size_t, D> &dims) {
    const auto st = Base::compute_stride_and_total(dims);
    std::printf("layout D=%zu lo=%zu hi=%zu dims", D, Lo, Hi);
    for (auto d : dims) std::printf(" %zu", d);
    std::printf(" stride");
    for (auto s : st.first) std::printf(" %zu", s);
    std::printf(" total %zu\n", st.second);
  }
};

template <std::size_t Lo, std::size_t Hi>
static void probe2(std::size_t ny, std::size_t nx) {
  Probe<2, float, Lo, Hi>::layout({ny, nx});
  Probe<2, float, Lo, Hi> a(ny, nx);
  const float *base = a.data();
  for (std::intptr_t y = -(std::intptr_t)Lo - 1; y <= (std::intptr_t)(ny + Hi); ++y)
    for (std::intptr_t x = -(std::intptr_t)Lo - 1; x <= (std::intptr_t)(nx + Hi); ++x) {
      long long off = -1;
      try {
        off = &a.at({y, x}) - base;
      } catch (const std::out_of_range &) {
        off = -1;  // the reference throws std::out_of_range here
      }
      std::printf("at2 lo=%zu hi=%zu dims %zu %zu idx %td %td off %lld\n", Lo, Hi, ny, nx, y, x, off);
    }
}

template <std::size_t Lo, std::size_t Hi>
static void probe3(std::size_t nz, std::size_t ny, std::size_t nx) {
  Probe<3, double, Lo, Hi>::layout({nz, ny, nx});
  Probe<3, double, Lo, Hi> a(nz, ny, nx);
  const double *base = a.data();
  for (std::intptr_t z = -(std::intptr_t)Lo; z < (std::intptr_t)(nz + Hi); ++z)
    for (std::intptr_t y = -(std::intptr_t)Lo; y < (std::intptr_t)(ny + Hi); ++y)
      for (std::intptr_t x = -(std::intptr_t)Lo; x < (std::intptr_t)(nx + Hi); ++x)
        std::printf("op3 lo=%zu hi=%zu dims %zu %zu %zu idx %td %td %td off %lld\n", Lo, Hi, nz, ny, nx, z,
                    y, x, (long long)(&a(z, y, x) - base));
}

static int run_index() {
  probe2<0, 0>(3, 4);
  probe2<8, 8>(4, 5);
  probe2<1, 3>(2, 7);
  probe3<1, 1>(2, 3, 4);
  probe3<2, 1>(3, 2, 2);
  Probe<1, float, 2, 2>::layout({9});
  Probe<2, float, 8, 8>::layout({8192, 8192});
  Probe<3, float, 1, 1>::layout({512, 512, 512});
  hpc::HPCHighDimensionFlatArray<2, float, 8> z(4, 4);
  bool zero = true;
  for (std::size_t i = 0; i < z.size(); ++i) zero = zero && z.data()[i] == 0.0f;
  std::printf("value_initialised %d\n", zero ? 1 : 0);
  return 0;
}

template <typename M>
static void digest(const char *name, const M &m) {
  std::printf("csr %s nrows %zu nnz %zu row_ptr %016" PRIx64 " col %016" PRIx64 " val %016" PRIx64 " bytes %zu\n",
              name, m.nrows(), m.nnz(), fnv1a(m.row_ptr(), (m.nrows() + 1) * sizeof(*m.row_ptr())),
              fnv1a(m.col_idx(), m.nnz() * 4), fnv1a(m.values(), m.nnz() * sizeof(*m.values())),
              m.algorithmic_bytes());
}

static int run_csr() {
  digest("laplace2d_64x48_f64_i32", sparse::CsrMatrix<double, int32_t>::laplace2d(64, 48));
  digest("laplace2d_1024_f64_i32", sparse::CsrMatrix<double, int32_t>::laplace2d(1024, 1024));
  digest("op27_12_f64_i64", sparse::CsrMatrix<double, int64_t>::op27(12));
  using Trip = std::tuple<std::int64_t, std::int32_t, float>;
  std::vector<Trip> t = {{2, 1, 1.f}, {0, 3, 2.f}, {2, 1, 0.5f}, {1, 0, -1.f}, {0, 0, 4.f}, {2, 1, 0.25f}};
  auto m = sparse::CsrMatrix<float, int32_t>::from_triplets(3, 4, t);
  std::printf("triplets nnz %zu rp %d %d %d %d col %d %d %d %d val %g %g %g %g\n", m.nnz(), m.row_ptr()[0],
              m.row_ptr()[1], m.row_ptr()[2], m.row_ptr()[3], m.col_idx()[0], m.col_idx()[1], m.col_idx()[2],
              m.col_idx()[3], m.values()[0], m.values()[1], m.values()[2], m.values()[3]);
  try {
    std::vector<Trip> bad = {{5, 0, 1.f}};
    (void)sparse::CsrMatrix<float, int32_t>::from_triplets(3, 4, bad);
    std::printf("triplets_oob no_throw\n");
  } catch (const std::out_of_range &) {
    std::printf("triplets_oob out_of_range\n");
  }
  // host-side partitioners of the multi-GPU face (static chunking, test_hpc_benchmark.cpp:1356)
  {
    const auto A = sparse::CsrMatrix<double, int32_t>::laplace2d(10, 10);
    const auto rows = sparse::partition_rows(A, 3);
    const auto by_nnz = sparse::partition_rows(A, 3, sparse::RowSplit::EqualNnz);
    const auto slabs = hpc::partition_slabs(13, 8);
    std::printf("partition_rows");
    for (auto b : rows) std::printf(" %lld", (long long)b);
    std::printf(" | by_nnz");
    for (auto b : by_nnz) std::printf(" %lld", (long long)b);
    std::printf(" | slabs");
    for (auto b : slabs) std::printf(" %lld", (long long)b);
    std::printf("\n");
  }
  return 0;
}

static float frand(uint64_t &s) {
  s = s * 6364136223846793005ull + 1442695040888963407ull;
  return (float)((s >> 40) & 0xFFFF) / 65536.0f - 0.5f;
}

// Stand-in with sparse::RootGrid's foreach(func(x, y, value&)) signature (RootGrid.hpp:20-22).
struct LineGrid {
  std::vector<std::tuple<std::intptr_t, std::intptr_t, float>> cells;
  template <typename F>
  void foreach (const F &f) {
    for (auto &c : cells) f(std::get<0>(c), std::get<1>(c), std::get<2>(c));
  }
};

static int run_gpu() {
  int bad = 0;
  hpcx::Context ctx(0);
  uint64_t seed = 7;
  {  // the reference's sweeps: BM_x_blur / BM_y_blur loop nests (test_hpc_benchmark.cpp:354-368, 444-457)
    constexpr int nblur = 8;
    const int ny = 96, nx = 256;
    hpc::HPCHighDimensionFlatArray<2, float, nblur> a(ny, nx);
    hpc::HPCHighDimensionFlatArray<2, float> b(ny, nx), want(ny, nx);
    for (std::size_t i = 0; i < a.size(); ++i) a.data()[i] = frand(seed);
    for (int axis = 0; axis < 2; ++axis) {
#pragma omp parallel for collapse(2)
      for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x) {
          float res = 0.0f;
          for (int t = -nblur; t <= nblur; ++t) res += axis == 1 ? a(y, x + t) : a(y + t, x);
          want(y, x) = res;
        }
      hpc::stencil_sweep(ctx, a, b, hpc::Box1D{axis, nblur});
      const int same = std::memcmp(b.data(), want.data(), b.size() * sizeof(float)) == 0;
      std::printf("box axis %d bit_exact %d\n", axis, same);
      bad += !same;
    }
    try {
      hpc::stencil_sweep(ctx, a, b, hpc::Box1D{1, nblur + 1});  // radius beyond the ghost cells
      std::printf("box_radius no_throw\n");
      ++bad;
    } catch (const std::system_error &e) {
      std::printf("box_radius system_error %s\n", e.what());
    }
  }
  {  // 7-point Jacobi, two steps
    const int nz = 20, ny = 18, nx = 130;
    hpc::HPCHighDimensionFlatArray<3, float, 1> u(nz, ny, nx), v(nz, ny, nx), w0(nz, ny, nx), w1(nz, ny, nx);
    for (std::size_t i = 0; i < u.size(); ++i) u.data()[i] = v.data()[i] = w0.data()[i] = w1.data()[i] = frand(seed);
    auto sweep = [&](auto &src, auto &dst) {
      for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
          for (int x = 0; x < nx; ++x) {
            float s = src(z - 1, y, x) + src(z, y - 1, x);
            s = s + src(z, y, x - 1);
            s = s + src(z, y, x + 1);
            s = s + src(z, y + 1, x);
            s = s + src(z + 1, y, x);
            dst(z, y, x) = s * (1.0f / 6.0f);
          }
    };
    sweep(w0, w1);
    sweep(w1, w0);  // newest iterate in w0
    hpc::stencil_sweep(ctx, u, v, hpc::Jacobi7{2});
    const int same = std::memcmp(v.data(), w0.data(), v.size() * sizeof(float)) == 0 &&
                     std::memcmp(u.data(), w1.data(), u.size() * sizeof(float)) == 0;
    std::printf("jacobi7 bit_exact %d\n", same);
    bad += !same;
    // two sweeps per pass over HBM (HPCX_VARIANT_TEMPORAL): the newest iterate is the same, bit for bit
    hpc::HPCHighDimensionFlatArray<3, float, 1> u2(nz, ny, nx), v2(nz, ny, nx);
    std::memcpy(u2.data(), w1.data(), u2.size() * sizeof(float));  // iterate 1
    std::memcpy(v2.data(), w1.data(), v2.size() * sizeof(float));
    sweep(w0, w1);
    sweep(w1, w0);
    sweep(w0, w1);  // iterate 5 in w1; u2 holds iterate 1: four more sweeps
    hpc::stencil_sweep(ctx, u2, v2, hpc::Jacobi7{4, HPCX_VARIANT_TEMPORAL});
    const int same_tb = std::memcmp(v2.data(), w1.data(), v2.size() * sizeof(float)) == 0;
    std::printf("jacobi7_temporal bit_exact %d\n", same_tb);
    bad += !same_tb;
  }
  {  // SpMV, 5-point Laplacian (rows of <= 5 entries: the device order is the CPU order)
    auto A = sparse::CsrMatrix<double>::laplace2d(200, 300);
    std::vector<double> x(A.ncols()), y, want(A.nrows());
    for (auto &v : x) v = frand(seed);
    for (std::size_t i = 0; i < A.nrows(); ++i) {
      double s = 0.0;
      for (auto p = A.row_ptr()[i]; p < A.row_ptr()[i + 1]; ++p) {
        const double prod = A.values()[p] * x[A.col_idx()[p]];
        s = s + prod;
      }
      want[i] = s;
    }
    sparse::spmv(ctx, A, x, y);
    double err = 0, mag = 0;
    for (std::size_t i = 0; i < y.size(); ++i) {
      err = std::fmax(err, std::fabs(y[i] - want[i]));
      mag = std::fmax(mag, std::fabs(want[i]));
    }
    std::printf("spmv rel_err %.3e\n", err / mag);
    bad += !(err <= 1e-12 * mag);
    sparse::DeviceCsr<double> dA(ctx, A);
    std::vector<double> y2;
    sparse::spmv(dA, x, y2);
    bad += std::memcmp(y.data(), y2.data(), y.size() * 8) != 0;
    {  // row a6: managed-memory vectors, the way the reference hands buffers to kernels
       // (std::vector<T, CudaAllocator<T, CudaMemManaged>>, radix_sort_gpu.h:22-23)
      std::vector<double, CudaAllocator<double, CudaMemManaged>> mx(A.ncols()), my(A.nrows());
      std::memcpy(mx.data(), x.data(), x.size() * 8);
      sparse::spmv(dA, mx, my);  // goes through the ABI as HPCX_MEM_MANAGED: no staging copy
      const int same = std::memcmp(my.data(), y.data(), y.size() * 8) == 0;
      static_assert(hpcx::memKindOf<decltype(mx)::allocator_type>::value == HPCX_MEM_MANAGED, "managed vectors");
      std::printf("managed_vectors bit_exact %d\n", same);
      bad += !same;
    }
    try {
      std::vector<double> shortx(3);
      sparse::spmv(dA, shortx, y2);
      ++bad;
    } catch (const std::out_of_range &) {
      std::printf("spmv_short_x out_of_range\n");
    }
  }
  {  // BM_conv's loop nest (test_hpc_benchmark.cpp:793-806) and BM_transpose's (:672-682)
    constexpr int nkern = 16;
    const int n = 96;
    hpc::HPCHighDimensionFlatArray<2, float> ka(n, n), want(n, n), kc(nkern, nkern);
    hpc::HPCHighDimensionFlatArray<2, float, nkern> kb(n, n);
    for (std::size_t i = 0; i < ka.size(); ++i) ka.data()[i] = want.data()[i] = frand(seed);
    for (std::size_t i = 0; i < kb.size(); ++i) kb.data()[i] = frand(seed);
    for (std::size_t i = 0; i < kc.size(); ++i) kc.data()[i] = frand(seed);
    for (int y = 0; y < n; y++)
      for (int x = 0; x < n; ++x)
        for (int l = 0; l < nkern; ++l)
          for (int k = 0; k < nkern; ++k) want(y, x) += kb(y + l, x + k) * kc(l, k);
    hpc::conv2d(ctx, ka, kb, kc);
    int same = std::memcmp(ka.data(), want.data(), ka.size() * sizeof(float)) == 0;
    std::printf("conv2d bit_exact %d\n", same);
    bad += !same;
    const int ny = 70, nx = 200;
    hpc::HPCHighDimensionFlatArray<2, float> a_t(ny, nx), b_t(nx, ny), bw(nx, ny);
    for (std::size_t i = 0; i < a_t.size(); ++i) a_t.data()[i] = frand(seed);
    for (int y = 0; y < ny; ++y)
      for (int x = 0; x < nx; ++x) bw(x, y) = a_t(y, x);
    hpc::transpose(ctx, a_t, b_t);
    same = std::memcmp(b_t.data(), bw.data(), bw.size() * sizeof(float)) == 0;
    std::printf("transpose bit_exact %d\n", same);
    bad += !same;
  }
  {  // COO assembly on the device + CG around the product
    const int n = 40;
    auto A = sparse::CsrMatrix<double>::laplace2d(n, n);
    std::vector<std::int32_t> rows, cols;
    std::vector<double> vals;
    for (std::size_t i = 0; i < A.nrows(); ++i)
      for (auto p = A.row_ptr()[i]; p < A.row_ptr()[i + 1]; ++p) {
        // every entry split in two halves listed far apart: the merge must add them back
        rows.push_back((std::int32_t)i);
        cols.push_back(A.col_idx()[p]);
        vals.push_back(0.5 * A.values()[p]);
      }
    const std::size_t half = rows.size();
    for (std::size_t e = half; e-- > 0;) {
      rows.push_back(rows[e]);
      cols.push_back(cols[e]);
      vals.push_back(vals[e]);
    }
    auto dA = sparse::DeviceCsr<double>::from_coo(ctx, A.nrows(), rows, cols, vals);
    sparse::DeviceCsr<double> dRef(ctx, A);
    std::vector<double> x(A.ncols()), y1, y2;
    for (auto &v : x) v = frand(seed);
    sparse::spmv(dA, x, y1);
    sparse::spmv(dRef, x, y2);
    const int same = std::memcmp(y1.data(), y2.data(), y1.size() * 8) == 0;
    std::printf("from_coo product bit_exact %d\n", same);
    bad += !same;
    std::vector<double> xs(A.nrows()), b, sol(A.nrows(), 0.0);
    for (auto &v : xs) v = frand(seed);
    sparse::spmv(dRef, xs, b);
    sparse::CgSolver<double> cg(dA);
    const sparse::CgResult res = cg.solve(b, sol, 1000, 1e-11, 16);
    double err = 0, mag = 0;
    for (std::size_t i = 0; i < sol.size(); ++i) {
      err = std::fmax(err, std::fabs(sol[i] - xs[i]));
      mag = std::fmax(mag, std::fabs(xs[i]));
    }
    std::printf("cg iterations %d converged %d rel_err %.3e\n", res.iterations, (int)res.converged, err / mag);
    bad += !(res.converged && err <= 1e-8 * mag);
  }
  {  // sparse-grid bridge: a stand-in with RootGrid's foreach(func(x, y, value&)) signature
     // (lib/sparse/include/RootGrid.hpp:20-22; the reference tree itself needs oneTBB)
    LineGrid g;
    const int nx = 9, ny = 6;
    for (int y = ny - 1; y >= 0; --y)
      for (int x = 0; x < nx; ++x) g.cells.emplace_back(x - 4, y + 10, frand(seed));
    sparse::ActiveCells ac;
    std::vector<float> x0, y1, y2;
    auto dG = sparse::csr_from_grid<float>(ctx, g, &ac, &x0);
    // the same cells as a full rectangle: rows are x-major, i.e. laplace2d(nx, ny)
    auto L = sparse::CsrMatrix<float>::laplace2d(nx, ny);
    sparse::DeviceCsr<float> dL(ctx, L);
    sparse::spmv(dG, x0, y1);
    sparse::spmv(dL, x0, y2);
    int ok = dG.nrows() == (std::size_t)(nx * ny) && ac.row_x.front() == -4 && ac.row_y.front() == 10 &&
             ac.row_x.back() == 4 && ac.row_y.back() == 15 && std::memcmp(y1.data(), y2.data(), y1.size() * 4) == 0;
    for (auto &c : g.cells)  // x0 carries each cell's value at its row
      ok = ok && x0[(std::get<0>(c) + 4) * ny + (std::get<1>(c) - 10)] == std::get<2>(c);
    std::printf("grid_bridge ok %d\n", ok);
    bad += !ok;
  }
  {  // the multi-GPU face with a job of one process: same answers as the single-GPU calls
    hpcx::Communicator comm(ctx);
    auto A = sparse::CsrMatrix<double>::laplace2d(64, 96);
    const auto bounds = sparse::partition_rows(A, comm.size());
    sparse::DistributedCsr<double> D(ctx, comm, A, bounds);
    std::vector<double> x(A.ncols()), y(A.nrows(), -1.0), want;
    for (auto &v : x) v = frand(seed);
    sparse::spmv(ctx, A, x, want);
    void *dx = nullptr, *dy = nullptr;
    hpcx_device_alloc(ctx.handle(), x.size() * 8, &dx);
    hpcx_device_alloc(ctx.handle(), y.size() * 8, &dy);
    hpcx_memcpy(ctx.handle(), dx, x.data(), x.size() * 8);
    D.multiply(static_cast<const double *>(dx), static_cast<double *>(dy));
    hpcx_memcpy(ctx.handle(), y.data(), dy, y.size() * 8);
    hpcx_device_free(ctx.handle(), dx);
    hpcx_device_free(ctx.handle(), dy);
    int ok = bounds.front() == 0 && bounds.back() == (std::int64_t)A.nrows() && D.owned() == (std::int64_t)A.nrows() &&
             D.halo_columns().empty() && std::memcmp(y.data(), want.data(), y.size() * 8) == 0;

    const int nz = 12, ny = 10, nx = 70;
    const auto slabs = hpc::partition_slabs(nz, comm.size());
    hpc::HPCHighDimensionFlatArray<3, float, 2> u(nz, ny, nx), v(nz, ny, nx), w(nz, ny, nx);
    for (std::size_t i = 0; i < u.size(); ++i) u.data()[i] = v.data()[i] = w.data()[i] = frand(seed);
    hpc::stencil_sweep(ctx, w, v, hpc::Jacobi7{5});  // reference answer: five single sweeps (w -> v)
    hpc::HPCHighDimensionFlatArray<3, float, 2> r(nz, ny, nx);
    std::memcpy(r.data(), u.data(), u.size() * sizeof(float));
    {
      hpc::DeviceGrid<3, float, 2> du(ctx, u), dr(ctx, r);
      hpc::HaloPlan<float, 2, 2> plan(du, dr, hpc::Jacobi7{1, HPCX_VARIANT_TEMPORAL});
      plan.sweep(5);
      dr.download(r);
    }
    ok = ok && slabs.back() == nz && std::memcmp(r.data(), v.data(), r.size() * sizeof(float)) == 0;
    {  // the block-split face of the box sweeps on a 1 x 1 block grid: same answer as hpc::stencil_sweep
      const int by = 40, bx = 128;
      hpc::HPCHighDimensionFlatArray<2, float, 8> a(by, bx);
      hpc::HPCHighDimensionFlatArray<2, float> b0(by, bx), b1(by, bx);
      for (std::size_t i = 0; i < a.size(); ++i) a.data()[i] = frand(seed);
      for (int axis = 0; axis < 2; ++axis) {
        hpc::stencil_sweep(ctx, a, b0, hpc::Box1D{axis, 8});
        hpc::DeviceGrid<2, float, 8> da(ctx, a);
        hpc::DeviceGrid<2, float> db(ctx, b1);
        hpc::BoxHaloPlan<float, 8, 8, 0, 0> plan(da, db, hpc::Box1D{axis, 8}, comm.size(), 1);
        plan.sweep();
        db.download(b1);
        ok = ok && std::memcmp(b0.data(), b1.data(), b0.size() * sizeof(float)) == 0;
      }
    }
    std::printf("distributed_face ok %d\n", ok);
    bad += !ok;
  }
  std::printf("launches %lld\n", (long long)ctx.launchCount());
  std::printf(bad ? "CPP GPU FAIL\n" : "CPP GPU OK\n");
  return bad;
}

// One process, one host thread, every visible GPU.
static int run_multi() {
  int bad = 0;
  hpcx::MultiContext gpus;
  std::printf("multi devices %d\n", gpus.size());
  uint64_t seed = 11;
  {  // BM_x_blur / BM_y_blur loop nests (test_hpc_benchmark.cpp:354-368, 444-457) over y-slabs
    constexpr int nblur = 8;
    const int ny = 64 * gpus.size() + 10, nx = 256;
    hpc::HPCHighDimensionFlatArray<2, float, nblur> a(ny, nx);
    hpc::HPCHighDimensionFlatArray<2, float> b(ny, nx), want(ny, nx);
    for (std::size_t i = 0; i < a.size(); ++i) a.data()[i] = frand(seed);
    for (int axis = 0; axis < 2; ++axis) {
#pragma omp parallel for collapse(2)
      for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x) {
          float res = 0.0f;
          for (int t = -nblur; t <= nblur; ++t) res += axis == 1 ? a(y, x + t) : a(y + t, x);
          want(y, x) = res;
        }
      hpc::stencil_sweep(gpus, a, b, hpc::Box1D{axis, nblur});
      const int same = std::memcmp(b.data(), want.data(), b.size() * sizeof(float)) == 0;
      std::printf("multi box axis %d bit_exact %d\n", axis, same);
      bad += !same;
    }
  }
  {  // 7-point Jacobi over z-slabs, ghost planes traded by peer stores inside the sweep kernel
    const int nz = 12 * gpus.size() + 5, ny = 18, nx = 130, steps = 6;
    hpc::HPCHighDimensionFlatArray<3, float, 1> u(nz, ny, nx), v(nz, ny, nx), w0(nz, ny, nx), w1(nz, ny, nx);
    for (std::size_t i = 0; i < u.size(); ++i) u.data()[i] = v.data()[i] = w0.data()[i] = w1.data()[i] = frand(seed);
    auto sweep = [&](auto &src, auto &dst) {
#pragma omp parallel for collapse(2)
      for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
          for (int x = 0; x < nx; ++x) {
            float s = src(z - 1, y, x) + src(z, y - 1, x);
            s = s + src(z, y, x - 1);
            s = s + src(z, y, x + 1);
            s = s + src(z, y + 1, x);
            s = s + src(z + 1, y, x);
            dst(z, y, x) = s * (1.0f / 6.0f);
          }
    };
    for (int t = 0; t < steps; t += 2) {
      sweep(w0, w1);
      sweep(w1, w0);
    }  // newest iterate (steps) in w0, the one before it in w1
    hpc::stencil_sweep(gpus, u, v, hpc::Jacobi7{steps});
    const int same = std::memcmp(v.data(), w0.data(), v.size() * sizeof(float)) == 0 &&
                     std::memcmp(u.data(), w1.data(), u.size() * sizeof(float)) == 0;
    std::printf("multi jacobi7 bit_exact %d\n", same);
    bad += !same;
  }
  {  // SpMV: 27-point operator, row blocks over the devices, x halo by peer stores
    auto A = sparse::CsrMatrix<double>::op27(14);
    std::vector<double> x(A.ncols()), y, want(A.nrows());
    for (auto &v : x) v = frand(seed);
    for (std::size_t i = 0; i < A.nrows(); ++i) {
      double s = 0.0;
      for (auto p = A.row_ptr()[i]; p < A.row_ptr()[i + 1]; ++p) {
        const double prod = A.values()[p] * x[A.col_idx()[p]];
        s = s + prod;
      }
      want[i] = s;
    }
    sparse::MultiDeviceCsr<double> dA(gpus, A);
    for (int rep = 0; rep < 3; ++rep) sparse::spmv(dA, x, y);
    const int same = y.size() == want.size() && std::memcmp(y.data(), want.data(), y.size() * 8) == 0;
    const auto b = dA.bounds();
    std::printf("multi spmv bit_exact %d blocks %zu last %lld\n", same, b.size() - 1, (long long)b.back());
    bad += !same;
  }
  std::printf(bad ? "CPP MULTI FAILED %d\n" : "CPP MULTI OK\n", bad);
  return bad ? 3 : 0;
}

int main(int argc, char **argv) {
  const std::string mode = argc > 1 ? argv[1] : "index";
  try {
    if (mode == "index") return run_index();
    if (mode == "csr") return run_csr();
    if (mode == "gpu") return run_gpu();
    if (mode == "multi") return run_multi();
  } catch (const std::exception &e) {
    std::printf("exception %s\n", e.what());
    return 2;
  }
  return 1;
}
