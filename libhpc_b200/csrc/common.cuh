// common.cuh — shared internals of libhpcx (context, error plumbing, PTX helpers).
// Error convention follows the reference's checkCuda/throwCudaError
// (lib/gpu/util/include/cudaHelper.cuh:10-27,96-101) but nothing is thrown across the
// C ABI: failures become an hpcx_status plus a message stored in the context.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/hpcx.h"

struct hpcx_ctx_s {
  int device = 0;
  int sm_count = 148;
  int64_t l2_bytes = 126 << 20;
  int cc_major = 0, cc_minor = 0;
  cudaStream_t compute = nullptr;  // default stream of the context
  cudaStream_t comm = nullptr;     // halo exchange / copy pipeline
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev[8] = {};
  int64_t launches = 0;
  bool debug_sync = false;  // HPCX_DEBUG_SYNC=1: sync + check after every launch
                            // (reference idiom, cuda_radix_sort_v4.cu:104-107)
  std::string last_error;
  // kernels whose dynamic shared-memory limit has been raised on this context's device
  // (the attribute is per device, so the record lives here and not in a function-local static)
  std::vector<const void *> smem_configured;
  // multi-GPU
  int nranks = 1, rank = 0;
  void *nccl_lib = nullptr;
  void *nccl_comm = nullptr;
  float *barrier_token = nullptr;  // 16 bytes of device memory for hpcx_comm_barrier
  // Peer-flag waits (fused halo transports): a kernel that gives up on a neighbour stores 1 here
  // and traps.  The word lives in mapped pinned host memory, so the host can still read it after
  // the trap has killed the CUDA context.  halo_timeout_ns == 0: wait without a bound (NCCL's
  // behaviour); hpcx_comm_set_timeout / HPCX_HALO_TIMEOUT_MS set a bound.
  uint32_t *err_host = nullptr, *err_dev = nullptr;
  uint64_t halo_timeout_ns = 0;
  // driver entry points
  void *encode_tiled = nullptr;  // cuTensorMapEncodeTiled
};

namespace hpcx {

inline hpcx_status fail(hpcx_ctx ctx, hpcx_status st, const std::string &msg) {
  if (ctx) ctx->last_error = msg;
  return st;
}

#define HPCX_CUDA(ctx, expr)                                                          \
  do {                                                                                \
    cudaError_t e__ = (expr);                                                         \
    if (e__ != cudaSuccess) {                                                         \
      char b__[512];                                                                  \
      snprintf(b__, sizeof b__, "%s:%d: %s: %s", __FILE__, __LINE__, #expr,           \
               cudaGetErrorString(e__));                                              \
      return hpcx::fail((ctx),                                                        \
                        e__ == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR, \
                        b__);                                                         \
    }                                                                                 \
  } while (0)

#define HPCX_REQUIRE(ctx, cond, msg)                                            \
  do {                                                                          \
    if (!(cond)) return hpcx::fail((ctx), HPCX_INVALID_ARG, std::string(msg)); \
  } while (0)

// cudaFuncAttributeMaxDynamicSharedMemorySize, once per context (= per device).
inline hpcx_status ensure_dyn_smem(hpcx_ctx ctx, const void *func, int bytes) {
  for (const void *f : ctx->smem_configured)
    if (f == func) return HPCX_OK;
  cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess)
    return fail(ctx, HPCX_CUDA_ERROR, std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
  ctx->smem_configured.push_back(func);
  return HPCX_OK;
}

// Has a peer-flag wait of this context timed out?  Reports (and clears) the error word.
inline hpcx_status comm_check(hpcx_ctx ctx) {
  if (ctx->err_host && *reinterpret_cast<volatile uint32_t *>(ctx->err_host)) {
    *reinterpret_cast<volatile uint32_t *>(ctx->err_host) = 0;
    return fail(ctx, HPCX_NCCL_ERROR, "timed out waiting for a neighbour's halo flag (peer transport)");
  }
  return HPCX_OK;
}

// Synchronising exit of the peer-transport paths: a timed-out wait traps, so the stream
// synchronise fails; the error word says why.
inline hpcx_status sync_and_check(hpcx_ctx ctx, cudaStream_t s) {
  cudaError_t e = cudaStreamSynchronize(s);
  if (hpcx_status st = comm_check(ctx)) return st;
  if (e != cudaSuccess)
    return fail(ctx, HPCX_CUDA_ERROR, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
  return HPCX_OK;
}

// Called after every kernel launch of the library.
inline hpcx_status after_launch(hpcx_ctx ctx, const char *what, int n = 1) {
  ctx->launches += n;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && ctx->debug_sync) e = cudaDeviceSynchronize();
  if (e != cudaSuccess)
    return fail(ctx, HPCX_CUDA_ERROR, std::string(what) + ": " + cudaGetErrorString(e));
  return HPCX_OK;
}

// NVTX range per phase (reference idiom: lib/gpu/transfer_overlap_testsuite/src/
// cuda_tut_transfer_overlap.cu:89,104; lib/gpu/radix_gpu/src/cuda_radix_sort_v4.cu:99-102).
// Header-only NVTX3: a no-op unless a profiler has injected itself.
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange &) = delete;
  NvtxRange &operator=(const NvtxRange &) = delete;
};
#define HPCX_RANGE_CAT2(a, b) a##b
#define HPCX_RANGE_CAT(a, b) HPCX_RANGE_CAT2(a, b)
#define HPCX_RANGE(name) hpcx::NvtxRange HPCX_RANGE_CAT(nvtx_range_, __LINE__)(name)

inline size_t dtype_size(int dtype) { return dtype == HPCX_F64 ? 8 : 4; }
inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

// What a kernel needs to wait on a neighbour's flag: where to report a timeout and the bound.
struct PeerWait {
  uint32_t *err = nullptr;   // mapped host word (hpcx_ctx_s::err_dev)
  uint64_t timeout_ns = 0;   // 0 = wait without a bound
};
inline PeerWait peer_wait_of(hpcx_ctx ctx) { return PeerWait{ctx->err_dev, ctx->halo_timeout_ns}; }

// ----------------------------------------------------------------- device PTX helpers
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D bulk copy global -> shared (TMA engine, SASS UBLKCP). 16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes,
                                         uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// Tiled tensor loads (TMA, SASS UTMALDG).
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1,
                                            uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, int c0, int c1,
                                            int c2, uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}
// ---- cross-GPU flags (system scope), used by the peer-store halo paths
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Spin until *flag >= want (wrap-safe).  Never returns with the condition unmet: without a bound
// it waits like an NCCL receive would; with one (hpcx_comm_set_timeout) it records the timeout in
// the context's host-visible error word and traps, so that no kernel goes on to read a stale
// ghost plane or x-halo — the caller sees HPCX_NCCL_ERROR at its next synchronising call.
__device__ __forceinline__ void wait_flag(const uint32_t *flag, uint32_t want, const PeerWait &pw) {
  if ((int32_t)(ld_acquire_sys(flag) - want) < 0) {
    const uint64_t t0 = pw.timeout_ns ? global_timer_ns() : 0;
    for (uint32_t it = 1; (int32_t)(ld_acquire_sys(flag) - want) < 0; ++it) {
      __nanosleep(100);
      if (pw.timeout_ns && (it & 255u) == 0 && global_timer_ns() - t0 > pw.timeout_ns) {
        *reinterpret_cast<volatile uint32_t *>(pw.err) = 1u;
        __threadfence_system();
        __trap();
      }
    }
  }
  // the ghost plane was written by the neighbour's generic stores; TMA reads it through the
  // async proxy
  asm volatile("fence.proxy.async;" ::: "memory");
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
#endif

// Host: build a tiled tensor map through the driver entry point held by the context.
hpcx_status make_tensor_map(hpcx_ctx ctx, CUtensorMap *map, int dtype, int rank, void *base,
                            const uint64_t *dims, const uint64_t *strides_bytes,
                            const uint32_t *box);

}  // namespace hpcx
