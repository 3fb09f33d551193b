// hpcx/Stencil.hpp — stencil sweeps over hpc::HPCHighDimensionFlatArray on the B200.
//
// The reference has no sweep function: each sweep is a doubly nested OpenMP loop in the
// benchmark TU that indexes the container's operator().  These entry points
// replace the loop nests; their bodies are one C-ABI call into libhpcx.so (CUDA, sm_100a):
//
//   hpc::stencil_sweep(ctx, a, b, hpc::Box1D{axis, r})
//       axis == Dimension-1:  b(y,x) = sum_{k=-r..r} a(y, x+k)   BM_x_blur,
//           BM_x_blur_tiling_simd_prefetch (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368, 425-441)
//       axis == Dimension-2:  b(y,x) = sum_{k=-r..r} a(y+k, x)   BM_y_blur,
//           BM_YXx_blur_tiling_prefetch_streamed_IPL (:444-457, 575-601)
//       res starts at 0 and taps are added for k ascending: bit-identical to the reference.
//   hpc::stencil_sweep(ctx, a, b, hpc::Jacobi7{steps})
//       3-D 7-point Jacobi (new); a and b hold the same Dirichlet ghosts;
//       afterwards b is the newest iterate and a the one before it.
//
//   hpc::conv2d(ka, kb, kc)      ka(y,x) += sum_{l,k} kb(y+l, x+k) * kc(l,k): BM_conv / BM_conv_block
//       (:786-818; kb is <2,T,nkern>, kc an nkern x nkern host array); bit-identical for 16 x 16.
//   hpc::transpose(a_t, b_t)     b_t(x,y) = a_t(y,x): BM_transpose / BM_transpose_tiling (:669-698).
//
// hpc::DeviceGrid is the device-resident mirror for callers that iterate without going back
// to the host (the reference's analogue is a std::vector over CudaAllocator,
// lib/gpu/util/include/cudaAllocator.hpp:39-83).
#pragma once
#ifndef HPCX_STENCIL_HPP_
#define HPCX_STENCIL_HPP_

#include <utility>

#include "HPCHighDimensionFlatArray.hpp"
#include "Runtime.hpp"

namespace hpc {

struct Box1D {
  int axis;        // 0 = slowest dimension
  int radius = 8;  // the reference's nblur (test_hpc_benchmark.cpp:29)
};
struct Jacobi7 {
  int steps = 1;
  int variant = HPCX_VARIANT_AUTO;
};

template <std::size_t Dimension, typename T, std::size_t Low_Bound = 0,
          std::size_t High_Bound = Low_Bound>
class DeviceGrid {
  static_assert(Dimension >= 1 && Dimension <= 3, "libhpcx grids are 1-D to 3-D");

 public:
  template <std::size_t A, class Al>
  using Host = HPCHighDimensionFlatArray<Dimension, T, Low_Bound, High_Bound, A, Al>;

  DeviceGrid(hpcx::Context &ctx, const std::array<std::size_t, Dimension> &dims) : ctx_(ctx.handle()) {
    desc_.ndim = static_cast<int32_t>(Dimension);
    desc_.dtype = hpcx::dtypeOf<T>();
    desc_.lo = static_cast<int32_t>(Low_Bound);
    desc_.hi = static_cast<int32_t>(High_Bound);
    for (std::size_t i = 0; i < 3; ++i) desc_.dim[i] = i < Dimension ? static_cast<int64_t>(dims[i]) : 0;
    HPCX_CHECK(ctx_, hpcx_grid_create(ctx_, &desc_, &grid_));
  }
  template <std::size_t A, class Al>
  DeviceGrid(hpcx::Context &ctx, const Host<A, Al> &a) : DeviceGrid(ctx, a.dim()) {
    upload(a);
  }
  ~DeviceGrid() {
    if (grid_) hpcx_grid_destroy(grid_);
  }
  DeviceGrid(const DeviceGrid &) = delete;
  DeviceGrid &operator=(const DeviceGrid &) = delete;
  DeviceGrid(DeviceGrid &&o) noexcept : ctx_(o.ctx_), grid_(std::exchange(o.grid_, nullptr)), desc_(o.desc_) {}

  // Whole padded array (ghosts included), reference layout on the host side.
  template <std::size_t A, class Al>
  void upload(const Host<A, Al> &a, void *stream = nullptr) {
    HPCX_CHECK(ctx_, hpcx_grid_upload(grid_, a.data(), stream));
  }
  template <std::size_t A, class Al>
  void download(Host<A, Al> &a, void *stream = nullptr) const {
    HPCX_CHECK(ctx_, hpcx_grid_download(grid_, a.data(), stream));
  }

  hpcx_grid handle() const noexcept { return grid_; }
  hpcx_ctx context() const noexcept { return ctx_; }

 private:
  hpcx_ctx ctx_ = nullptr;
  hpcx_grid grid_ = nullptr;
  hpcx_grid_desc desc_{};
};

// ---- device-resident sweeps
template <std::size_t D, typename T, std::size_t Lo, std::size_t Hi, std::size_t Lo2, std::size_t Hi2>
void stencil_sweep(DeviceGrid<D, T, Lo, Hi> &in, DeviceGrid<D, T, Lo2, Hi2> &out, Box1D box,
                   void *stream = nullptr) {
  static_assert(D == 2, "box sweeps take 2-D grids");
  hpcx_stencil_desc st{};
  if (box.axis != 0 && box.axis != 1)
    hpcx::throwStatus(HPCX_INVALID_ARG, "Box1D axis must be 0 or 1", __FILE__, __LINE__);
  st.kind = box.axis == 1 ? HPCX_STENCIL_BOX_X : HPCX_STENCIL_BOX_Y;
  st.radius = box.radius;
  HPCX_CHECK(in.context(), hpcx_stencil_sweep(in.handle(), out.handle(), &st, 1, stream));
}

template <typename T, std::size_t Lo, std::size_t Hi>
void stencil_sweep(DeviceGrid<3, T, Lo, Hi> &in, DeviceGrid<3, T, Lo, Hi> &out, Jacobi7 j,
                   void *stream = nullptr) {
  static_assert(Lo >= 1 && Hi >= 1, "Jacobi7 needs at least one ghost cell per side");
  hpcx_stencil_desc st{};
  st.kind = HPCX_STENCIL_JACOBI7;
  st.variant = j.variant;
  HPCX_CHECK(in.context(), hpcx_stencil_sweep(in.handle(), out.handle(), &st, j.steps, stream));
}

// ka(y,x) += sum_{l<KH} sum_{k<KW} kb(y+l, x+k) * kc(l,k)   (BM_conv, test_hpc_benchmark.cpp:793-806)
template <typename T, std::size_t Lo, std::size_t Hi, std::size_t Lo2, std::size_t Hi2, std::size_t A,
          class Al>
void conv2d(DeviceGrid<2, T, Lo, Hi> &ka, DeviceGrid<2, T, Lo2, Hi2> &kb,
            const HPCHighDimensionFlatArray<2, T, 0, 0, A, Al> &kc, void *stream = nullptr,
            int variant = HPCX_VARIANT_AUTO) {
  const auto d = kc.dim();
  HPCX_CHECK(ka.context(), hpcx_conv2d(kb.handle(), ka.handle(), kc.data(), static_cast<int32_t>(d[0]),
                                       static_cast<int32_t>(d[1]), variant, stream));
}

// b_t(x,y) = a_t(y,x)   (BM_transpose, test_hpc_benchmark.cpp:672-682)
template <typename T, std::size_t Lo, std::size_t Hi, std::size_t Lo2, std::size_t Hi2>
void transpose(DeviceGrid<2, T, Lo, Hi> &a_t, DeviceGrid<2, T, Lo2, Hi2> &b_t, void *stream = nullptr) {
  HPCX_CHECK(a_t.context(), hpcx_grid_transpose(a_t.handle(), b_t.handle(), stream));
}

// ---- host containers in, host containers out (the reference-facing call)
template <typename T, std::size_t Hi, std::size_t Lo2, std::size_t Hi2, std::size_t A, class Al, std::size_t A2,
          class Al2, std::size_t A3, class Al3>
void conv2d(hpcx::Context &ctx, HPCHighDimensionFlatArray<2, T, 0, Hi, A, Al> &ka,
            const HPCHighDimensionFlatArray<2, T, Lo2, Hi2, A2, Al2> &kb,
            const HPCHighDimensionFlatArray<2, T, 0, 0, A3, Al3> &kc) {
  DeviceGrid<2, T, 0, Hi> da(ctx, ka);
  DeviceGrid<2, T, Lo2, Hi2> db(ctx, kb);
  conv2d(da, db, kc);
  da.download(ka);
}

template <typename T, std::size_t Lo, std::size_t Hi, std::size_t A, class Al, std::size_t Lo2, std::size_t Hi2,
          std::size_t A2, class Al2>
void transpose(hpcx::Context &ctx, const HPCHighDimensionFlatArray<2, T, Lo, Hi, A, Al> &a_t,
               HPCHighDimensionFlatArray<2, T, Lo2, Hi2, A2, Al2> &b_t) {
  DeviceGrid<2, T, Lo, Hi> da(ctx, a_t);
  DeviceGrid<2, T, Lo2, Hi2> db(ctx, b_t.dim());
  transpose(da, db);
  db.download(b_t);
}

template <typename T, std::size_t Lo, std::size_t Hi, std::size_t A, class Al, std::size_t Lo2,
          std::size_t Hi2, std::size_t A2, class Al2>
void stencil_sweep(hpcx::Context &ctx, const HPCHighDimensionFlatArray<2, T, Lo, Hi, A, Al> &a,
                   HPCHighDimensionFlatArray<2, T, Lo2, Hi2, A2, Al2> &b, Box1D box) {
  DeviceGrid<2, T, Lo, Hi> da(ctx, a);
  DeviceGrid<2, T, Lo2, Hi2> db(ctx, b.dim());
  stencil_sweep(da, db, box);
  db.download(b);
}

template <typename T, std::size_t Lo, std::size_t Hi, std::size_t A, class Al>
void stencil_sweep(hpcx::Context &ctx, HPCHighDimensionFlatArray<3, T, Lo, Hi, A, Al> &a,
                   HPCHighDimensionFlatArray<3, T, Lo, Hi, A, Al> &b, Jacobi7 j) {
  DeviceGrid<3, T, Lo, Hi> da(ctx, a), db(ctx, b);
  stencil_sweep(da, db, j);
  db.download(b);
  if (j.steps > 1) da.download(a);
}

}  // namespace hpc
#endif  // HPCX_STENCIL_HPP_
