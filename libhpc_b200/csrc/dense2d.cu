// dense2d.cu — the two other grid sweeps of the reference's benchmark family (SURVEY.md §8f
// ranks 2 and 3), behind the same ghost-padded grid container:
//
//   CONV2D     out(y,x) += sum_{l<KH} sum_{k<KW} in(y+l, x+k) * w(l,k), l outer, k inner, one
//              rounding per multiply and per add, accumulating onto out's previous value:
//              BM_conv / BM_conv_block (reference tests/test_hpc_benchmark/
//              test_hpc_benchmark.cpp:786-818; nkern = 16, `kb` has ghost 16, `ka` none).
//   TRANSPOSE  out(x,y) = in(y,x): BM_transpose / BM_transpose_tiling (…:669-700).
//
// conv2d_tile_kernel<T, 16, 16>: a CTA stages a (32+15)-wide, (64+15)-tall input tile in shared memory;
// each thread owns one column x and RY = 8 consecutive output rows, walks the 23 input rows of
// its column window in ascending order and, for every row r and tap k, feeds in(r, x+k) to the
// outputs j with 0 <= r-j < 16.  Each output therefore sees its 256 products in exactly the
// reference's (l, k) order, so the result is bit-identical, while one shared-memory load
// (conflict-free: a warp reads 32 consecutive floats) feeds up to 8 multiply-add pairs.  The
// weights travel as a __grid_constant__ kernel parameter (parameters live in the constant bank,
// and nothing is shared between calls, streams or contexts); the r/k/j loops are fully unrolled
// so every weight is an immediate constant-bank operand.  The sweep is FP32-issue bound (512 separately rounded
// operations per cell), not HBM bound: 8 B/cell of traffic against 512 flops.
//
// transpose_tile_kernel<T>: 64 x 64 tiles through padded shared memory, 16-byte global accesses
// on both sides; 2 * sizeof(T) bytes per cell, HBM bound.
#include "grid.cuh"

#include <algorithm>
#include <atomic>
#include <utility>

using namespace hpcx;

namespace {

struct Geom2 {
  int64_t ny, nx;       // logical extents
  int64_t glo1;         // ghost rows below y = 0
  int64_t pitch, xoff;  // device row pitch / offset of x = 0
};
__host__ Geom2 geom2(const hpcx_grid_s *g) { return Geom2{g->n[1], g->n[2], g->glo[1], g->pitch, g->xoff}; }
__device__ __forceinline__ int64_t at2(const Geom2 &g, int64_t y, int64_t x) {
  return (y + g.glo1) * g.pitch + g.xoff + x;
}

__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

constexpr int kMaxTaps = 32 * 32;

// Weights of the tiled kernel: a by-value kernel parameter (1 KB fp32 / 2 KB fp64 for 16 x 16).
template <typename T, int N>
struct ConvWeights {
  T v[N];
};

constexpr int kConvTX = 32, kConvRY = 8, kConvGY = 8;  // 32 columns x (8 groups of 8 rows) per CTA

// Compile-time row / tap loops (a pack expansion instead of `#pragma unroll`, which the compiler
// only honours partially at this size): every weight index is a constant expression, so the
// multiplies take their weight straight from the constant bank.
template <typename T, int KH, int KW, int SP, int R, int K>
__device__ __forceinline__ void conv_tap(const T *__restrict__ col, T (&acc)[kConvRY],
                                         const ConvWeights<T, KH * KW> &w) {
  const T v = col[R * SP + K];
#pragma unroll
  for (int j = 0; j < kConvRY; ++j) {
    const int l = R - j;
    if (l >= 0 && l < KH) acc[j] = add_rn(acc[j], mul_rn(v, w.v[l * KW + K]));
  }
}
template <typename T, int KH, int KW, int SP, int R, int... K>
__device__ __forceinline__ void conv_row(const T *__restrict__ col, T (&acc)[kConvRY],
                                         const ConvWeights<T, KH * KW> &w, std::integer_sequence<int, K...>) {
  (conv_tap<T, KH, KW, SP, R, K>(col, acc, w), ...);
}
template <typename T, int KH, int KW, int SP, int... R>
__device__ __forceinline__ void conv_rows(const T *__restrict__ col, T (&acc)[kConvRY],
                                          const ConvWeights<T, KH * KW> &w, std::integer_sequence<int, R...>) {
  (conv_row<T, KH, KW, SP, R>(col, acc, w, std::make_integer_sequence<int, KW>{}), ...);
}

template <typename T, int KH, int KW>
__global__ void __launch_bounds__(kConvTX *kConvGY)
conv2d_tile_kernel(const T *__restrict__ in, T *__restrict__ out, Geom2 gi, Geom2 go,
                   const __grid_constant__ ConvWeights<T, KH * KW> w) {
  constexpr int TY = kConvRY * kConvGY;           // 64 output rows per CTA
  constexpr int SW = kConvTX + KW - 1;            // staged tile width
  constexpr int SH = TY + KH - 1;                 // staged tile height
  constexpr int SP = SW | 1;                      // odd pitch
  __shared__ T tile[SH * SP];
  const int tx = threadIdx.x % kConvTX, gy = threadIdx.x / kConvTX;
  const int64_t x0 = (int64_t)blockIdx.x * kConvTX, y0 = (int64_t)blockIdx.y * TY;
  // stage: rows y0 .. y0+SH-1, columns x0 .. x0+SW-1, clipped to the rows/columns any output of
  // the grid can touch (y < ny + KH - 1, x < nx + KW - 1: inside the ghost region by contract)
  for (int i = threadIdx.x; i < SH * SW; i += kConvTX * kConvGY) {
    const int r = i / SW, c = i % SW;
    const int64_t y = y0 + r, x = x0 + c;
    tile[r * SP + c] = (y < gi.ny + KH - 1 && x < gi.nx + KW - 1) ? __ldg(in + at2(gi, y, x)) : T(0);
  }
  __syncthreads();
  const int64_t x = x0 + tx, yb = y0 + (int64_t)gy * kConvRY;
  T acc[kConvRY];
#pragma unroll
  for (int j = 0; j < kConvRY; ++j)
    acc[j] = (x < go.nx && yb + j < go.ny) ? out[at2(go, yb + j, x)] : T(0);
  const T *col = tile + (gy * kConvRY) * SP + tx;
  conv_rows<T, KH, KW, SP>(col, acc, w, std::make_integer_sequence<int, kConvRY + KH - 1>{});
#pragma unroll
  for (int j = 0; j < kConvRY; ++j)
    if (x < go.nx && yb + j < go.ny) out[at2(go, yb + j, x)] = acc[j];
}

// Any kernel extent: one thread per output, taps through L1, weights from a per-call device buffer.
template <typename T>
__global__ void __launch_bounds__(256)
conv2d_direct_kernel(const T *__restrict__ in, T *__restrict__ out, Geom2 gi, Geom2 go, int kh, int kw,
                     const T *__restrict__ w) {
  const int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t y = blockIdx.y * (int64_t)blockDim.y + threadIdx.y;
  if (x >= go.nx || y >= go.ny) return;
  T s = out[at2(go, y, x)];
  for (int l = 0; l < kh; ++l) {
    const T *p = in + at2(gi, y + l, x);
    for (int k = 0; k < kw; ++k) s = add_rn(s, mul_rn(__ldg(p + k), __ldg(w + l * kw + k)));
  }
  out[at2(go, y, x)] = s;
}

// out(x, y) = in(y, x).  64 x 64 tile, 256 threads; thread t moves 4-element vectors: reads row
// ty + 16 i of the tile at columns 4 tx .. 4 tx + 3, writes transposed rows the same way.
template <typename T>
struct alignas(4 * sizeof(T)) Vec4 {
  T v[4];
};

template <typename T, bool VEC>
__global__ void __launch_bounds__(256)
transpose_tile_kernel(const T *__restrict__ in, T *__restrict__ out, Geom2 gi, Geom2 go) {
  constexpr int TS = 64;
  __shared__ T tile[TS][TS + 1];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int64_t x0 = (int64_t)blockIdx.x * TS, y0 = (int64_t)blockIdx.y * TS;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = ty + 16 * i;
    const int64_t y = y0 + r, x = x0 + 4 * tx;
    if (y < gi.ny) {
      if (VEC && x + 3 < gi.nx) {
        const Vec4<T> q = *reinterpret_cast<const Vec4<T> *>(in + at2(gi, y, x));
#pragma unroll
        for (int e = 0; e < 4; ++e) tile[r][4 * tx + e] = q.v[e];
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (x + e < gi.nx) tile[r][4 * tx + e] = in[at2(gi, y, x + e)];
      }
    }
  }
  __syncthreads();
  // output rows are input columns: out row (x0 + r), out columns y0 + 4 tx ..
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = ty + 16 * i;
    const int64_t oy = x0 + r, ox = y0 + 4 * tx;
    if (oy < go.ny) {
      if (VEC && ox + 3 < go.nx) {
        Vec4<T> q;
#pragma unroll
        for (int e = 0; e < 4; ++e) q.v[e] = tile[4 * tx + e][r];
        *reinterpret_cast<Vec4<T> *>(out + at2(go, oy, ox)) = q;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (ox + e < go.nx) out[at2(go, oy, ox + e)] = tile[4 * tx + e][r];
      }
    }
  }
}

// ---- weights scratch held by the library itself (hpcx_conv2d_static_scratch)
// One copy per GPU that loads libhpcx.so, resident until the process exits: 1024 taps of the
// widest dtype = 8 KB.  In host code `conv_static_weights` is only a handle for the runtime's
// symbol table — never a device address — so the host side asks cudaGetSymbolAddress for the
// address on each device and hands THAT to the kernel.
constexpr size_t kStaticScratchBytes = (size_t)kMaxTaps * sizeof(double);
constexpr int kStaticScratchMaxDevices = 64;
}  // namespace

static __device__ __align__(256) unsigned char conv_static_weights[kStaticScratchBytes];

namespace {

std::atomic<void *> g_static_weights_addr[kStaticScratchMaxDevices];

// Device address of conv_static_weights on the CURRENT device, resolved once per device.
// Idempotent, so two threads resolving the same device at once both store the same value.
hpcx_status static_weights_addr(hpcx_ctx ctx, void **out) {
  if (ctx->device < 0 || ctx->device >= kStaticScratchMaxDevices)
    return fail(ctx, HPCX_UNSUPPORTED, "hpcx_conv2d_static_scratch: device ordinal beyond 63");
  void *p = g_static_weights_addr[ctx->device].load(std::memory_order_acquire);
  if (!p) {
    HPCX_CUDA(ctx, cudaGetSymbolAddress(&p, conv_static_weights));
    if (!p) return fail(ctx, HPCX_CUDA_ERROR, "cudaGetSymbolAddress gave no address for the weights scratch");
    g_static_weights_addr[ctx->device].store(p, std::memory_order_release);
  }
  *out = p;
  return HPCX_OK;
}

// The direct kernel with its weights staged in the library-held array instead of a per-call
// allocation.  No cudaMalloc / cudaFree / pool call on this path.
template <typename T>
hpcx_status conv2d_static_t(hpcx_grid in, hpcx_grid out, const T *w, int kh, int kw, void *scratch, cudaStream_t s) {
  hpcx_ctx ctx = in->ctx;
  const Geom2 gi = geom2(in), go = geom2(out);
  T *dw = static_cast<T *>(scratch);
  // every tap the kernel reads is overwritten here, so nothing left by an earlier call is used
  HPCX_CUDA(ctx, cudaMemcpyAsync(dw, w, (size_t)kh * kw * sizeof(T), cudaMemcpyHostToDevice, s));
  dim3 block(64, 4), grid((unsigned)((go.nx + 63) / 64), (unsigned)((go.ny + 3) / 4));
  conv2d_direct_kernel<T><<<grid, block, 0, s>>>((const T *)in->dev, (T *)out->dev, gi, go, kh, kw, dw);
  return after_launch(ctx, "conv2d_direct_kernel (static scratch)");
}

template <typename T>
hpcx_status conv2d_t(hpcx_grid in, hpcx_grid out, const T *w, int kh, int kw, int variant, cudaStream_t s) {
  hpcx_ctx ctx = in->ctx;
  const Geom2 gi = geom2(in), go = geom2(out);
  if (kh == 16 && kw == 16 && variant != HPCX_VARIANT_DIRECT) {
    ConvWeights<T, 256> wk;
    memcpy(wk.v, w, sizeof wk.v);  // captured by the launch: no state outlives the call
    dim3 grid((unsigned)((go.nx + kConvTX - 1) / kConvTX),
              (unsigned)((go.ny + kConvRY * kConvGY - 1) / (kConvRY * kConvGY)));
    conv2d_tile_kernel<T, 16, 16><<<grid, kConvTX * kConvGY, 0, s>>>((const T *)in->dev, (T *)out->dev, gi, go, wk);
    return after_launch(ctx, "conv2d_tile_kernel");
  }
  // stream-ordered scratch for the weights (the host array may be reused as soon as this returns:
  // a pageable cudaMemcpyAsync has staged it by then)
  const size_t wb = (size_t)kh * kw * sizeof(T);
  T *dw = nullptr;
  HPCX_CUDA(ctx, cudaMallocAsync((void **)&dw, wb, s));
  HPCX_CUDA(ctx, cudaMemcpyAsync(dw, w, wb, cudaMemcpyHostToDevice, s));
  dim3 block(64, 4), grid((unsigned)((go.nx + 63) / 64), (unsigned)((go.ny + 3) / 4));
  conv2d_direct_kernel<T><<<grid, block, 0, s>>>((const T *)in->dev, (T *)out->dev, gi, go, kh, kw, dw);
  hpcx_status st = after_launch(ctx, "conv2d_direct_kernel");
  HPCX_CUDA(ctx, cudaFreeAsync(dw, s));
  return st;
}

template <typename T>
hpcx_status transpose_t(hpcx_grid in, hpcx_grid out, cudaStream_t s) {
  const Geom2 gi = geom2(in), go = geom2(out);
  dim3 grid((unsigned)((gi.nx + 63) / 64), (unsigned)((gi.ny + 63) / 64));
  // 16-byte vectors need 4-element aligned rows on both sides (device pitches are multiples of
  // 32 bytes; the x offset of interior 0 is 32-byte aligned by construction)
  const bool vec = (gi.pitch % 4 == 0) && (go.pitch % 4 == 0) && (gi.xoff % 4 == 0) && (go.xoff % 4 == 0) &&
                   ((gi.glo1 * gi.pitch) % 4 == 0) && ((go.glo1 * go.pitch) % 4 == 0);
  if (vec)
    transpose_tile_kernel<T, true><<<grid, 256, 0, s>>>((const T *)in->dev, (T *)out->dev, gi, go);
  else
    transpose_tile_kernel<T, false><<<grid, 256, 0, s>>>((const T *)in->dev, (T *)out->dev, gi, go);
  return after_launch(in->ctx, "transpose_tile_kernel");
}

}  // namespace

extern "C" {

hpcx_status hpcx_conv2d(hpcx_grid in, hpcx_grid out, const void *weights, int32_t kh, int32_t kw,
                        int32_t variant, void *stream) {
  if (!in || !out || !weights) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = in->ctx;
  HPCX_REQUIRE(ctx, in->ctx == out->ctx && in != out && in->dev != out->dev, "in and out must be two grids of one context");
  if (in->desc.ndim != 2 || out->desc.ndim != 2) return fail(ctx, HPCX_UNSUPPORTED, "conv2d takes 2-D grids");
  HPCX_REQUIRE(ctx, in->desc.dtype == out->desc.dtype, "grids differ in dtype");
  HPCX_REQUIRE(ctx, in->n[1] == out->n[1] && in->n[2] == out->n[2], "grids differ in logical extent");
  HPCX_REQUIRE(ctx, kh >= 1 && kw >= 1 && kh * kw <= kMaxTaps, "kernel extent must be within 1..1024 taps");
  HPCX_REQUIRE(ctx, in->ghi[1] >= kh - 1 && in->ghi[2] >= kw - 1,
               "the input grid needs High_Bound >= kernel extent - 1 (taps reach y+kh-1, x+kw-1)");
  HPCX_REQUIRE(ctx, in->n[1] <= 65535LL * 4, "grid too tall for the conv2d sweep");
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  hpcx_status rc = in->desc.dtype == HPCX_F64
                       ? conv2d_t<double>(in, out, (const double *)weights, kh, kw, variant, s)
                       : conv2d_t<float>(in, out, (const float *)weights, kh, kw, variant, s);
  if (rc != HPCX_OK) return rc;
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}

// hpcx_conv2d's direct sweep for callers that must not allocate while they run: a solver that has
// already handed every free byte of HBM to its own arena (any later cudaMalloc fails), a time-step
// loop that cannot afford cudaMalloc's device-wide synchronisation, or a process whose driver set-up
// has no stream-ordered memory pool for cudaMallocAsync.  The grids are made once at set-up like
// any others; this call itself only copies the weights into the library's own 8 KB device array
// and launches.  The first call on a device resolves that array's address (cudaGetSymbolAddress,
// with the device made current); make it before starting any stream capture.  All calls on a
// device share the array, so two calls in flight at once race on the weights.
hpcx_status hpcx_conv2d_static_scratch(hpcx_grid in, hpcx_grid out, const void *weights, int32_t kh,
                                       int32_t kw, void *stream) {
  if (!in || !out || !weights) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = in->ctx;
  HPCX_REQUIRE(ctx, in->ctx == out->ctx && in != out && in->dev != out->dev, "in and out must be two grids of one context");
  if (in->desc.ndim != 2 || out->desc.ndim != 2) return fail(ctx, HPCX_UNSUPPORTED, "conv2d takes 2-D grids");
  HPCX_REQUIRE(ctx, in->desc.dtype == out->desc.dtype, "grids differ in dtype");
  HPCX_REQUIRE(ctx, in->n[1] == out->n[1] && in->n[2] == out->n[2], "grids differ in logical extent");
  HPCX_REQUIRE(ctx, kh >= 1 && kw >= 1, "kernel extent must be at least 1 x 1");
  const size_t es = dtype_size(in->desc.dtype);
  if ((size_t)kh * (size_t)kw * es > kStaticScratchBytes) {
    char msg[200];
    snprintf(msg, sizeof msg,
             "hpcx_conv2d_static_scratch: %d x %d taps of %zu bytes need %zu bytes, the library-held "
             "scratch is %zu bytes; use hpcx_conv2d",
             kh, kw, es, (size_t)kh * (size_t)kw * es, kStaticScratchBytes);
    return fail(ctx, HPCX_INVALID_ARG, msg);
  }
  HPCX_REQUIRE(ctx, in->ghi[1] >= kh - 1 && in->ghi[2] >= kw - 1,
               "the input grid needs High_Bound >= kernel extent - 1 (taps reach y+kh-1, x+kw-1)");
  HPCX_REQUIRE(ctx, in->n[1] <= 65535LL * 4, "grid too tall for the conv2d sweep");
  int prev = -1;
  HPCX_CUDA(ctx, cudaGetDevice(&prev));
  if (prev != ctx->device) HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  void *scratch = nullptr;
  hpcx_status rc = static_weights_addr(ctx, &scratch);
  if (rc == HPCX_OK)
    rc = in->desc.dtype == HPCX_F64
             ? conv2d_static_t<double>(in, out, (const double *)weights, kh, kw, scratch, s)
             : conv2d_static_t<float>(in, out, (const float *)weights, kh, kw, scratch, s);
  if (rc == HPCX_OK && !stream) {
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) rc = fail(ctx, HPCX_CUDA_ERROR, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
  }
  if (prev != ctx->device) cudaSetDevice(prev);
  return rc;
}

hpcx_status hpcx_grid_transpose(hpcx_grid in, hpcx_grid out, void *stream) {
  if (!in || !out) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = in->ctx;
  HPCX_REQUIRE(ctx, in->ctx == out->ctx && in != out && in->dev != out->dev, "in and out must be two grids of one context");
  if (in->desc.ndim != 2 || out->desc.ndim != 2) return fail(ctx, HPCX_UNSUPPORTED, "transpose takes 2-D grids");
  HPCX_REQUIRE(ctx, in->desc.dtype == out->desc.dtype, "grids differ in dtype");
  HPCX_REQUIRE(ctx, in->n[1] == out->n[2] && in->n[2] == out->n[1], "out must be (nx, ny) for an (ny, nx) input");
  HPCX_REQUIRE(ctx, in->n[1] <= 65535LL * 64 && in->n[2] <= 0x7fffffffLL * 64, "grid too large for the transpose");
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  hpcx_status rc = in->desc.dtype == HPCX_F64 ? transpose_t<double>(in, out, s) : transpose_t<float>(in, out, s);
  if (rc != HPCX_OK) return rc;
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}

}  // extern "C"
