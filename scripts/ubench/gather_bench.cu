// gather_bench.cu — microbenchmark of x-gather strategies for power-law SpMV (C5), not product code.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o scripts/ubench/libgather.so scripts/ubench/gather_bench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); return -1.f; } } while (0)

template <int MODE>  // 0 stream only, 1 ldg gather (4 consecutive per thread), 2 tex gather
__global__ void __launch_bounds__(256) k_vec4(const int4 *__restrict__ col, const float4 *__restrict__ val,
                                              const float *__restrict__ x, cudaTextureObject_t tx, int64_t n4,
                                              float *__restrict__ out) {
  float s = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int4 c = __ldg(col + i);
    const float4 v = __ldg(val + i);
    float a, b, cc, d;
    if (MODE == 0) { a = (float)c.x; b = (float)c.y; cc = (float)c.z; d = (float)c.w; }
    else if (MODE == 1) { a = __ldg(x + c.x); b = __ldg(x + c.y); cc = __ldg(x + c.z); d = __ldg(x + c.w); }
    else { a = tex1Dfetch<float>(tx, c.x); b = tex1Dfetch<float>(tx, c.y); cc = tex1Dfetch<float>(tx, c.z); d = tex1Dfetch<float>(tx, c.w); }
    s += v.x * a + v.y * b + v.z * cc + v.w * d;
  }
  out[blockIdx.x * (int64_t)blockDim.x + threadIdx.x] = s;
}

// one nonzero per lane per load (a warp request covers 32 consecutive nonzeros), U independent loads
template <int MODE, int U>
__global__ void __launch_bounds__(256) k_lane(const int *__restrict__ col, const float *__restrict__ val,
                                              const float *__restrict__ x, cudaTextureObject_t tx, int64_t n,
                                              float *__restrict__ out) {
  float s = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride * U) {
    int c[U]; float v[U], a[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int64_t j = i + u * stride; c[u] = j < n ? __ldg(col + j) : 0; v[u] = j < n ? __ldg(val + j) : 0.f; }
#pragma unroll
    for (int u = 0; u < U; ++u) a[u] = MODE == 1 ? __ldg(x + c[u]) : tex1Dfetch<float>(tx, c[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) s += v[u] * a[u];
  }
  out[blockIdx.x * (int64_t)blockDim.x + threadIdx.x] = s;
}

// persistent CTA, hub slice of x in shared memory; colh: hub entries encoded as (0x80000000 | slot)
template <int MODE, int U, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_hub(const int *__restrict__ colh, const float *__restrict__ val,
                                                    const float *__restrict__ x, cudaTextureObject_t tx,
                                                    const int *__restrict__ hub_col, int K, int64_t n,
                                                    float *__restrict__ out) {
  extern __shared__ float sx[];
  for (int i = threadIdx.x; i < K; i += THREADS) sx[i] = __ldg(x + hub_col[i]);
  __syncthreads();
  float s = 0.f;
  const int64_t stride = (int64_t)gridDim.x * THREADS;
  for (int64_t i = blockIdx.x * (int64_t)THREADS + threadIdx.x; i < n; i += stride * U) {
    int c[U]; float v[U], a[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int64_t j = i + u * stride; c[u] = j < n ? __ldg(colh + j) : 0; v[u] = j < n ? __ldg(val + j) : 0.f; }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (c[u] < 0) a[u] = sx[c[u] & 0x7fffffff];
      else a[u] = MODE == 1 ? __ldg(x + c[u]) : tex1Dfetch<float>(tx, c[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) s += v[u] * a[u];
  }
  out[blockIdx.x * (int64_t)THREADS + threadIdx.x] = s;
}

// same hub kernel, but every warp streams its OWN contiguous range of nonzeros (what the product
// kernel does) instead of the grid-wide sliding window: 4736 concurrent streams per array
template <int U, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_hub_contig(const int *__restrict__ colh, const float *__restrict__ val,
                                                           const float *__restrict__ x, const int *__restrict__ hub_col, int K,
                                                           int64_t n, float *__restrict__ out, int vec) {
  extern __shared__ float sx[];
  for (int i = threadIdx.x; i < K; i += THREADS) sx[i] = __ldg(x + hub_col[i]);
  __syncthreads();
  float s = 0.f;
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * (THREADS / 32) + (threadIdx.x >> 5), nw = (int64_t)gridDim.x * (THREADS / 32);
  const int64_t ntiles = n / (32 * U);
  for (int64_t t = ntiles * w / nw; t < ntiles * (w + 1) / nw; ++t) {
    int c[U]; float v[U], a[U];
    if (vec) {  // lane owns 4 consecutive entries per half tile: 128-bit loads
#pragma unroll
      for (int h = 0; h < U / 4; ++h) {
        const int4 cc = __ldcs(reinterpret_cast<const int4 *>(colh + t * 32 * U + h * 128 + lane * 4));
        const float4 vv = __ldcs(reinterpret_cast<const float4 *>(val + t * 32 * U + h * 128 + lane * 4));
        c[4 * h] = cc.x; c[4 * h + 1] = cc.y; c[4 * h + 2] = cc.z; c[4 * h + 3] = cc.w;
        v[4 * h] = vv.x; v[4 * h + 1] = vv.y; v[4 * h + 2] = vv.z; v[4 * h + 3] = vv.w;
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) { const int64_t j = t * 32 * U + u * 32 + lane; c[u] = __ldg(colh + j); v[u] = __ldg(val + j); }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) a[u] = c[u] < 0 ? sx[c[u] & 0x7fffffff] : __ldg(x + c[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) s += v[u] * a[u];
  }
  out[blockIdx.x * (int64_t)THREADS + threadIdx.x] = s;
}

static cudaTextureObject_t make_tex(const float *x, int64_t n) {
  cudaResourceDesc rd{}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = (void *)x;
  rd.res.linear.desc = cudaCreateChannelDesc<float>(); rd.res.linear.sizeInBytes = (size_t)n * 4;
  cudaTextureDesc td{}; td.readMode = cudaReadModeElementType;
  cudaTextureObject_t t = 0; cudaCreateTextureObject(&t, &rd, &td, nullptr); return t;
}

template <typename F> static float time_it(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) f();
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) f();
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) { printf("kernel error %s\n", cudaGetErrorString(e)); return -1.f; }
  return ms / 10;
}

extern "C" float gb_run(int which, const int *col, const int *colh, const float *val, const float *x, int64_t nx,
                        const int *hub_col, int K, int64_t n, float *out, int sms) {
  static cudaTextureObject_t tx = 0; static const float *tx_for = nullptr;
  if (tx_for != x) { tx = make_tex(x, nx); tx_for = x; }
  const int64_t n4 = n / 4;
  const int g = sms * 8;
  switch (which) {
    case 0: return time_it([&] { k_vec4<0><<<g, 256>>>((const int4 *)col, (const float4 *)val, x, tx, n4, out); });
    case 1: return time_it([&] { k_vec4<1><<<g, 256>>>((const int4 *)col, (const float4 *)val, x, tx, n4, out); });
    case 2: return time_it([&] { k_vec4<2><<<g, 256>>>((const int4 *)col, (const float4 *)val, x, tx, n4, out); });
    case 3: return time_it([&] { k_lane<1, 4><<<g, 256>>>(col, val, x, tx, n, out); });
    case 4: return time_it([&] { k_lane<1, 8><<<g, 256>>>(col, val, x, tx, n, out); });
    case 5: return time_it([&] { k_lane<2, 4><<<g, 256>>>(col, val, x, tx, n, out); });
    case 6: return time_it([&] { k_lane<2, 8><<<g, 256>>>(col, val, x, tx, n, out); });
    case 7: { CK(cudaFuncSetAttribute(k_hub<1, 8, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub<1, 8, 1024><<<sms, 1024, K * 4>>>(colh, val, x, tx, hub_col, K, n, out); }); }
    case 8: { CK(cudaFuncSetAttribute(k_hub<2, 8, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub<2, 8, 1024><<<sms, 1024, K * 4>>>(colh, val, x, tx, hub_col, K, n, out); }); }
    case 9: { CK(cudaFuncSetAttribute(k_hub<1, 16, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub<1, 16, 1024><<<sms, 1024, K * 4>>>(colh, val, x, tx, hub_col, K, n, out); }); }
    case 10: { CK(cudaFuncSetAttribute(k_hub<1, 8, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub<1, 8, 512><<<sms, 512, K * 4>>>(colh, val, x, tx, hub_col, K, n, out); }); }
    // hub with small K and several CTAs per SM is covered by passing K and sms * ctas as `sms`
    case 11: { CK(cudaFuncSetAttribute(k_hub<1, 8, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub<1, 8, 512><<<sms, 512, K * 4>>>(colh, val, x, tx, hub_col, K, n, out); }); }
    case 12: { CK(cudaFuncSetAttribute(k_hub_contig<8, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub_contig<8, 1024><<<sms, 1024, K * 4>>>(colh, val, x, hub_col, K, n, out, 0); }); }
    case 13: { CK(cudaFuncSetAttribute(k_hub_contig<8, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, K * 4));
      return time_it([&] { k_hub_contig<8, 1024><<<sms, 1024, K * 4>>>(colh, val, x, hub_col, K, n, out, 1); }); }
  }
  return -2.f;
}
