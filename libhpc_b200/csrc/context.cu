// context.cu — context lifetime, memory helpers, tensor-map construction.
#include "common.cuh"

namespace hpcx {

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

hpcx_status make_tensor_map(hpcx_ctx ctx, CUtensorMap *map, int dtype, int rank, void *base,
                            const uint64_t *dims, const uint64_t *strides_bytes,
                            const uint32_t *box) {
  if (!ctx->encode_tiled)
    return fail(ctx, HPCX_CUDA_ERROR, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];  // stride of dim i+1
  CUresult r = reinterpret_cast<encode_tiled_fn>(ctx->encode_tiled)(
      map, dtype == HPCX_F64 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
      (cuuint32_t)rank, base, gdim, gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char b[128];
    snprintf(b, sizeof b, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return fail(ctx, HPCX_CUDA_ERROR, b);
  }
  return HPCX_OK;
}

}  // namespace hpcx

using namespace hpcx;

extern "C" {

const char *hpcx_version(void) { return "hpcx 0.1 (sm_100a)"; }

hpcx_status hpcx_init(int32_t device, hpcx_ctx *out) {
  if (!out) return HPCX_INVALID_ARG;
  *out = nullptr;
  hpcx_ctx ctx = new (std::nothrow) hpcx_ctx_s();
  if (!ctx) return HPCX_OOM;
  *out = ctx;  // handed back even on failure so the caller can read the message
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(ctx, HPCX_CUDA_ERROR,
                std::string("no CUDA device: ") + cudaGetErrorString(e) +
                    " (libhpcx has no CPU fallback)");
  HPCX_REQUIRE(ctx, device >= 0 && device < ndev, "device index out of range");
  HPCX_CUDA(ctx, cudaSetDevice(device));
  ctx->device = device;
  cudaDeviceProp prop;
  HPCX_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
  ctx->sm_count = prop.multiProcessorCount;
  ctx->l2_bytes = prop.l2CacheSize > 0 ? prop.l2CacheSize : ctx->l2_bytes;
  ctx->cc_major = prop.major;
  ctx->cc_minor = prop.minor;
  if (prop.major != 10)
    return fail(ctx, HPCX_UNSUPPORTED,
                "libhpcx is built for sm_100a only; device is sm_" + std::to_string(prop.major) +
                    std::to_string(prop.minor));
  HPCX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->compute, cudaStreamNonBlocking));
  // halo exchange stream: highest priority, so that the boundary strips and the NCCL kernels of a
  // distributed sweep get SM slots ahead of the interior kernel they run beside (dist.cu)
  int prio_lo = 0, prio_hi = 0;
  HPCX_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  HPCX_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->comm, cudaStreamNonBlocking, prio_hi));
  HPCX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
  HPCX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
  for (auto &ev : ctx->ev) HPCX_CUDA(ctx, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  const char *dbg = getenv("HPCX_DEBUG_SYNC");
  ctx->debug_sync = dbg && dbg[0] == '1';
  // error word of the peer-flag waits: mapped pinned host memory (readable after a device trap)
  HPCX_CUDA(ctx, cudaHostAlloc((void **)&ctx->err_host, 64, cudaHostAllocMapped | cudaHostAllocPortable));
  memset(ctx->err_host, 0, 64);
  HPCX_CUDA(ctx, cudaHostGetDevicePointer((void **)&ctx->err_dev, ctx->err_host, 0));
  if (const char *to = getenv("HPCX_HALO_TIMEOUT_MS")) ctx->halo_timeout_ns = (uint64_t)atoll(to) * 1000000ull;
  cudaDriverEntryPointQueryResult qres;
  void *fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) ctx->encode_tiled = fn;
  (void)cudaGetLastError();
  return HPCX_OK;
}

hpcx_status hpcx_finalize(hpcx_ctx ctx) {
  if (!ctx) return HPCX_INVALID_ARG;
  if (ctx->compute) {
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (auto &ev : ctx->ev)
      if (ev) cudaEventDestroy(ev);
    cudaStreamDestroy(ctx->compute);
    cudaStreamDestroy(ctx->comm);
    cudaStreamDestroy(ctx->copy_in);
    cudaStreamDestroy(ctx->copy_out);
    cudaFree(ctx->barrier_token);
    if (ctx->err_host) cudaFreeHost(ctx->err_host);
  }
  delete ctx;
  return HPCX_OK;
}

const char *hpcx_last_error_string(hpcx_ctx ctx) {
  return ctx ? ctx->last_error.c_str() : "null context";
}

hpcx_status hpcx_synchronize(hpcx_ctx ctx) {
  if (!ctx) return HPCX_INVALID_ARG;
  cudaError_t e = cudaDeviceSynchronize();
  if (hpcx_status st = comm_check(ctx)) return st;
  HPCX_CUDA(ctx, e);
  return HPCX_OK;
}

int64_t hpcx_launch_count(hpcx_ctx ctx) { return ctx ? ctx->launches : -1; }

int32_t hpcx_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return n;
}

hpcx_status hpcx_host_alloc(hpcx_ctx ctx, size_t bytes, void **out) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  HPCX_CUDA(ctx, cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
  return HPCX_OK;
}
hpcx_status hpcx_host_free(hpcx_ctx ctx, void *p) {
  if (!ctx) return HPCX_INVALID_ARG;
  HPCX_CUDA(ctx, cudaFreeHost(p));
  return HPCX_OK;
}
hpcx_status hpcx_device_alloc(hpcx_ctx ctx, size_t bytes, void **out) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  HPCX_CUDA(ctx, cudaMalloc(out, bytes ? bytes : 1));
  return HPCX_OK;
}
hpcx_status hpcx_device_free(hpcx_ctx ctx, void *p) {
  if (!ctx) return HPCX_INVALID_ARG;
  HPCX_CUDA(ctx, cudaFree(p));
  return HPCX_OK;
}
hpcx_status hpcx_mem_alloc(int32_t mem, size_t bytes, void **out) {
  if (!out) return HPCX_INVALID_ARG;
  *out = nullptr;
  cudaError_t e;
  switch (mem) {
    case HPCX_MEM_DEVICE: e = cudaMalloc(out, bytes ? bytes : 1); break;
    case HPCX_MEM_MANAGED: e = cudaMallocManaged(out, bytes ? bytes : 1); break;
    case HPCX_MEM_HOST: e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault); break;
    default: return HPCX_INVALID_ARG;
  }
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? HPCX_OK : e == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR;
}
hpcx_status hpcx_mem_free(int32_t mem, void *p) {
  if (mem != HPCX_MEM_DEVICE && mem != HPCX_MEM_MANAGED && mem != HPCX_MEM_HOST) return HPCX_INVALID_ARG;
  const cudaError_t e = mem == HPCX_MEM_HOST ? cudaFreeHost(p) : cudaFree(p);
  return e == cudaSuccess ? HPCX_OK : HPCX_CUDA_ERROR;
}
hpcx_status hpcx_memcpy(hpcx_ctx ctx, void *dst, const void *src, size_t bytes) {
  if (!ctx) return HPCX_INVALID_ARG;
  // Contract: ordered after ALL work already issued to the device (the caller's own streams
  // included — they are unknown here, hence one device-wide drain) and complete on return.  The
  // copy itself runs on the compute stream, so one stream synchronise ends the call.
  HPCX_CUDA(ctx, cudaDeviceSynchronize());
  HPCX_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ctx->compute));
  HPCX_CUDA(ctx, cudaStreamSynchronize(ctx->compute));
  return HPCX_OK;
}

}  // extern "C"
