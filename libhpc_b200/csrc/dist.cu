// dist.cu — multi-GPU transport: one process per GPU, NCCL
// send/recv between neighbouring ranks over NVLink, overlapped with interior compute on a
// second stream; the event-chained two-stream shape generalises the only overlap idiom the
// reference has (lib/gpu/transfer_overlap_testsuite/src/cuda_tut_transfer_overlap.cu:92-103).
//
// NCCL is bound at run time (dlopen of libnccl.so.2 — the copy torch already mapped when the
// caller is a torch.distributed program) so that libhpcx.so loads on machines without it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <functional>

#include "dist.cuh"

using namespace hpcx;

namespace hpcx {
hpcx_status jacobi7_launch(hpcx_grid_s *g, void *src, void *dst, int variant, int zsplit,
                           int64_t zbeg, int64_t zend, cudaStream_t s, const JacobiPeer *peer);
hpcx_status allreduce_sum(hpcx_ctx ctx, void *buf, size_t count, int dtype, cudaStream_t s);
bool jacobi7_tb2_supported(hpcx_grid_s *g);
hpcx_status box_launch_rows(hpcx_grid_s *in, hpcx_grid_s *out, const hpcx_stencil_desc *st, int64_t y0, int64_t y1,
                            cudaStream_t s);
hpcx_status box_launch_tile(hpcx_grid_s *in, hpcx_grid_s *out, const hpcx_stencil_desc *st, int64_t y0, int64_t y1,
                            int64_t x0, int64_t x1, cudaStream_t s);
hpcx_status jacobi7_tb2_launch(hpcx_grid_s *g, void *src, void *dst, int zsplit, int64_t zbeg, int64_t zend,
                               bool nb_lo, bool nb_hi, cudaStream_t s, int64_t zbeg2 = 0, int64_t zend2 = 0);
hpcx_status jacobi7_tb2_peer_launch(hpcx_grid_s *g, void *src, void *dst, int zsplit, const JacobiPeer &peer,
                                    cudaStream_t s);
hpcx_status spmv_launch_blocks(hpcx_csr A, const RowBlock *d_blocks, int64_t nblocks, const void *x,
                               void *y, cudaStream_t s, void *dot_part, const SpmvDistArgs *da);
hpcx_status spmv_launch_fixup_tables(hpcx_csr A, const int32_t *long_row, const int32_t *long_off,
                                     int64_t nlong, void *y, cudaStream_t s, const void *x,
                                     void *dot_part);
hpcx_status spmv_dist_host_pipelined(hpcx_csr A, const void *x, void *y, cudaStream_t s,
                                     const std::function<hpcx_status()> &push_and_boundary);
hpcx_status csr_plan_rowblocks(hpcx_csr A, const uint8_t *d_kind, int64_t maxlen, RowBlock **d_blocks,
                               std::vector<RowBlock> *host, int64_t *n0, int32_t **d_long_row,
                               int32_t **d_long_off, int64_t *nlong, int64_t *nchunks);
}  // namespace hpcx

namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId *);
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
  ncclResult_t (*CommInitRankConfig)(ncclComm_t *, int, ncclUniqueId, int, ncclConfig_t *);  // optional (NCCL >= 2.14)
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t);
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  const char *(*GetErrorString)(ncclResult_t);
};
NcclApi g_nccl{};

hpcx_status load_nccl(hpcx_ctx ctx) {
  if (ctx->nccl_lib) return HPCX_OK;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(ctx, HPCX_NCCL_ERROR, std::string("cannot load libnccl: ") + dlerror());
#define HPCX_SYM(field, name)                                                        \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));           \
  if (!g_nccl.field) return fail(ctx, HPCX_NCCL_ERROR, "libnccl lacks " name)
  HPCX_SYM(GetUniqueId, "ncclGetUniqueId");
  HPCX_SYM(CommInitRank, "ncclCommInitRank");
  g_nccl.CommInitRankConfig =
      reinterpret_cast<decltype(g_nccl.CommInitRankConfig)>(dlsym(h, "ncclCommInitRankConfig"));
  HPCX_SYM(CommDestroy, "ncclCommDestroy");
  HPCX_SYM(Send, "ncclSend");
  HPCX_SYM(Recv, "ncclRecv");
  HPCX_SYM(GroupStart, "ncclGroupStart");
  HPCX_SYM(GroupEnd, "ncclGroupEnd");
  HPCX_SYM(AllReduce, "ncclAllReduce");
  HPCX_SYM(AllGather, "ncclAllGather");
  HPCX_SYM(GetErrorString, "ncclGetErrorString");
#undef HPCX_SYM
  ctx->nccl_lib = h;
  return HPCX_OK;
}

#define HPCX_NCCL(ctx, expr)                                                             \
  do {                                                                                   \
    ncclResult_t r__ = (expr);                                                           \
    if (r__ != ncclSuccess)                                                              \
      return hpcx::fail((ctx), HPCX_NCCL_ERROR,                                          \
                        std::string(#expr) + ": " + g_nccl.GetErrorString(r__));         \
  } while (0)

inline ncclComm_t comm_of(hpcx_ctx ctx) { return static_cast<ncclComm_t>(ctx->nccl_comm); }
inline ncclDataType_t nccl_type(int dtype) { return dtype == HPCX_F64 ? ncclFloat64 : ncclFloat32; }

}  // namespace

// ------------------------------------------------------------------ distributed stencil
// (struct hpcx_halo_plan_s: dist.cuh)
namespace {

// Exchange w ghost planes per side of buffer `buf` (device layout of grid g) with the z
// neighbours.  Planes are contiguous in the device layout, so nothing is packed.
hpcx_status exchange_planes(hpcx_grid g, void *buf, cudaStream_t s, int64_t w = 1, int stride = 1) {
  hpcx_ctx ctx = g->ctx;
  if (ctx->nranks <= 1) return HPCX_OK;
  // slices of the split axis: z planes of a 3-D grid, rows of a 2-D grid (both contiguous)
  const bool rows2d = g->desc.ndim == 2;
  const int64_t plane = rows2d ? g->pitch : g->ext[1] * g->pitch;
  const int64_t nown = rows2d ? g->n[1] : g->n[0], glo = rows2d ? g->glo[1] : g->glo[0];
  char *base = static_cast<char *>(buf);
  auto at = [&](int64_t z) { return base + (z + glo) * plane * g->es; };
  const ncclDataType_t ty = nccl_type(g->desc.dtype);
  // stride > 1: rows of a py x px block grid, neighbours are `stride` = px ranks away
  const int lo = ctx->rank - stride, hi = ctx->rank + stride;
  HPCX_NCCL(ctx, g_nccl.GroupStart());
  if (lo >= 0) {
    HPCX_NCCL(ctx, g_nccl.Send(at(0), (size_t)(w * plane), ty, lo, comm_of(ctx), s));
    HPCX_NCCL(ctx, g_nccl.Recv(at(-w), (size_t)(w * plane), ty, lo, comm_of(ctx), s));
  }
  if (hi < ctx->nranks) {
    HPCX_NCCL(ctx, g_nccl.Send(at(nown - w), (size_t)(w * plane), ty, hi, comm_of(ctx), s));
    HPCX_NCCL(ctx, g_nccl.Recv(at(nown), (size_t)(w * plane), ty, hi, comm_of(ctx), s));
  }
  HPCX_NCCL(ctx, g_nccl.GroupEnd());
  return HPCX_OK;
}

// x faces of a block (2-D grids, block split): the `r` owned columns next to each x neighbour are
// packed (face_copy_kernel, a11), exchanged with rank -1 / +1 of the same block row, and unpacked
// into the ghost columns.  All on stream s.
hpcx_status exchange_cols(hpcx_halo_plan p, cudaStream_t s) {
  hpcx_ctx ctx = p->ctx;
  hpcx_grid g = p->in;
  const int r = p->st.radius, rx = ctx->rank % p->px;
  const bool has[2] = {rx > 0, rx + 1 < p->px};
  const int peer[2] = {ctx->rank - 1, ctx->rank + 1};
  int64_t elems = 0;
  if (hpcx_status st = hpcx_grid_face_elems(g, 1, r, &elems)) return st;
  for (int side = 0; side < 2; ++side)
    if (has[side])
      if (hpcx_status st = hpcx_grid_pack_face(g, 1, side, r, p->face_send[side], HPCX_MEM_DEVICE, s)) return st;
  const ncclDataType_t ty = nccl_type(g->desc.dtype);
  HPCX_NCCL(ctx, g_nccl.GroupStart());
  for (int side = 0; side < 2; ++side)
    if (has[side]) {
      HPCX_NCCL(ctx, g_nccl.Send(p->face_send[side], (size_t)elems, ty, peer[side], comm_of(ctx), s));
      HPCX_NCCL(ctx, g_nccl.Recv(p->face_recv[side], (size_t)elems, ty, peer[side], comm_of(ctx), s));
    }
  HPCX_NCCL(ctx, g_nccl.GroupEnd());
  for (int side = 0; side < 2; ++side)
    if (has[side])
      if (hpcx_status st = hpcx_grid_unpack_face(g, 1, side, r, p->face_recv[side], HPCX_MEM_DEVICE, s)) return st;
  return HPCX_OK;
}

// ---- transport 1: peer windows over CUDA IPC ------------------------------------------
// What every rank publishes about itself (gathered with ncclAllGather at plan time).
struct PeerInfo {
  cudaIpcMemHandle_t handle[3];  // allocations holding grid buffer A, grid buffer B, flag block
  int64_t offset[3];             // byte offset of each inside its allocation
  int64_t n0, n1, n2, pitch, ext1;
};

// cudaIpcGetMemHandle names the whole allocation a pointer lives in; carry the offset too.
hpcx_status ipc_export(hpcx_ctx ctx, void *ptr, cudaIpcMemHandle_t *h, int64_t *offset) {
  typedef CUresult (*range_fn)(CUdeviceptr *, size_t *, CUdeviceptr);
  static range_fn get_range = nullptr;
  if (!get_range) {
    cudaDriverEntryPointQueryResult q;
    void *fn = nullptr;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn)
      return fail(ctx, HPCX_CUDA_ERROR, "cuMemGetAddressRange entry point unavailable");
    get_range = reinterpret_cast<range_fn>(fn);
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  if (get_range(&base, &size, (CUdeviceptr)ptr) != CUDA_SUCCESS)
    return fail(ctx, HPCX_CUDA_ERROR, "cuMemGetAddressRange failed");
  *offset = (int64_t)((CUdeviceptr)ptr - base);
  HPCX_CUDA(ctx, cudaIpcGetMemHandle(h, (void *)base));
  return HPCX_OK;
}

hpcx_status allgather_bytes(hpcx_ctx ctx, const void *mine, void *all, size_t bytes) {
  char *d = nullptr;
  HPCX_CUDA(ctx, cudaMalloc((void **)&d, bytes * (size_t)(ctx->nranks + 1)));
  HPCX_CUDA(ctx, cudaMemcpy(d, mine, bytes, cudaMemcpyHostToDevice));
  ncclResult_t r = g_nccl.AllGather(d, d + bytes, bytes, ncclChar, comm_of(ctx), ctx->comm);
  cudaError_t e = cudaStreamSynchronize(ctx->comm);
  if (r == ncclSuccess && e == cudaSuccess)
    e = cudaMemcpy(all, d + bytes, bytes * (size_t)ctx->nranks, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (r != ncclSuccess) return fail(ctx, HPCX_NCCL_ERROR, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r));
  if (e != cudaSuccess) return fail(ctx, HPCX_CUDA_ERROR, std::string("allgather: ") + cudaGetErrorString(e));
  return HPCX_OK;
}

void close_peers(hpcx_halo_plan p) {
  for (int side = 0; side < 2; ++side)
    for (int k = 0; k < 3; ++k)
      if (p->peer_open[side][k]) {
        cudaIpcCloseMemHandle(p->peer_open[side][k]);
        p->peer_open[side][k] = nullptr;
      }
  cudaFree(p->flags);
  p->flags = nullptr;
}

}  // namespace
namespace hpcx {
// The local flag block of a peer-store plan, and the two buffers the plan was made for.
hpcx_status plan_prepare(hpcx_halo_plan p) {
  hpcx_ctx ctx = p->ctx;
  p->buf[0] = p->in->dev;
  p->buf[1] = p->out->dev;
  // A flag block of its own allocation (2 MiB keeps cudaMalloc from sub-allocating it).
  HPCX_CUDA(ctx, cudaMalloc((void **)&p->flags, 2u << 20));
  HPCX_CUDA(ctx, cudaMemset(p->flags, 0, 256));
  HPCX_CUDA(ctx, cudaDeviceSynchronize());
  return HPCX_OK;
}

}  // namespace hpcx
namespace {
// Opens the neighbours' grid buffers and flag blocks.  Collective over the communicator.
hpcx_status open_peers(hpcx_halo_plan p) {
  hpcx_ctx ctx = p->ctx;
  hpcx_grid g = p->in;
  hpcx_status st = plan_prepare(p);
  if (st != HPCX_OK) return st;
  PeerInfo mine{};
  void *ptrs[3] = {p->buf[0], p->buf[1], p->flags};
  for (int k = 0; k < 3; ++k) {
    st = ipc_export(ctx, ptrs[k], &mine.handle[k], &mine.offset[k]);
    if (st != HPCX_OK) return st;
  }
  mine.n0 = g->n[0];
  mine.n1 = g->n[1];
  mine.n2 = g->n[2];
  mine.pitch = g->pitch;
  mine.ext1 = g->ext[1];
  std::vector<PeerInfo> all(ctx->nranks);
  st = allgather_bytes(ctx, &mine, all.data(), sizeof(PeerInfo));
  if (st != HPCX_OK) return st;
  for (int side = 0; side < 2; ++side) {
    const int q = ctx->rank + (side ? 1 : -1);
    if (q < 0 || q >= ctx->nranks) continue;
    const PeerInfo &o = all[q];
    if (o.n1 != mine.n1 || o.n2 != mine.n2 || o.pitch != mine.pitch || o.ext1 != mine.ext1)
      return fail(ctx, HPCX_INVALID_ARG, "neighbouring slabs differ in their (y, x) extents");
    void *base[3];
    for (int k = 0; k < 3; ++k) {
      HPCX_CUDA(ctx, cudaIpcOpenMemHandle(&base[k], o.handle[k], cudaIpcMemLazyEnablePeerAccess));
      p->peer_open[side][k] = base[k];
    }
    p->peer_buf[side][0] = reinterpret_cast<float *>((char *)base[0] + o.offset[0]);
    p->peer_buf[side][1] = reinterpret_cast<float *>((char *)base[1] + o.offset[1]);
    p->peer_flags[side] = reinterpret_cast<uint32_t *>((char *)base[2] + o.offset[2]);
    p->peer_n0[side] = o.n0;
  }
  return HPCX_OK;
}

// Last launch of a peer-store sweep sequence: the stream does not complete before both
// neighbours have delivered the ghost planes of the newest iterate.
__global__ void halo_wait_kernel(const uint32_t *flags, PeerWait pw, uint32_t want, int has_lo,
                                 int has_hi) {
  if (has_lo) wait_flag(flags + 0, want, pw);
  if (has_hi) wait_flag(flags + 1, want, pw);
}

// Two-sweep passes: after the pass kernel, copy this slab's two boundary planes per side straight
// into the neighbours' ghost planes (peer stores over NVLink; the planes are contiguous), then
// the last CTA publishes the pass number in the neighbours' flag words and waits for theirs, so
// that the next pass kernel may read its ghost planes as soon as this kernel has ended.
__global__ void __launch_bounds__(256)
halo_push_planes_kernel(const float4 *__restrict__ lo_src, float4 *__restrict__ lo_dst,
                        const float4 *__restrict__ hi_src, float4 *__restrict__ hi_dst, int64_t n4,
                        uint32_t *counter, uint32_t *sig_lo, uint32_t *sig_hi, uint32_t sig_val,
                        const uint32_t *flags, PeerWait pw) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
    if (lo_dst) lo_dst[i] = __ldg(lo_src + i);
    if (hi_dst) hi_dst[i] = __ldg(hi_src + i);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0 && atomicAdd(counter, 1u) + 1 == gridDim.x) {
    atomicExch(counter, 0u);
    __threadfence_system();
    if (sig_lo) st_release_sys(sig_lo, sig_val);
    if (sig_hi) st_release_sys(sig_hi, sig_val);
    if (sig_lo) wait_flag(flags + 0, sig_val, pw);
    if (sig_hi) wait_flag(flags + 1, sig_val, pw);
  }
}

}  // namespace

extern "C" {

hpcx_status hpcx_comm_unique_id(hpcx_ctx ctx, void *id128) {
  if (!ctx || !id128) return HPCX_INVALID_ARG;
  hpcx_status st = load_nccl(ctx);
  if (st != HPCX_OK) return st;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  HPCX_NCCL(ctx, g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof id);
  return HPCX_OK;
}

hpcx_status hpcx_comm_init(hpcx_ctx ctx, int32_t nranks, int32_t rank, const void *id128) {
  if (!ctx || !id128) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(ctx, nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
  HPCX_REQUIRE(ctx, ctx->nccl_comm == nullptr, "communicator already initialised");
  hpcx_status st = load_nccl(ctx);
  if (st != HPCX_OK) return st;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof id);
  ncclComm_t comm;
  HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  // NCCL's own choice of CTAs is kept by default: capping it at 4 for the few-MiB halo messages was
  // measured SLOWER at 2 GPUs (one sweep 219 vs 199 us, two sweeps 130 vs 105 us per sweep: the
  // exchange then outlasts the interior sweep it hides behind).  HPCX_NCCL_MAX_CTAS > 0 sets a cap.
  int max_ctas = 0;
  if (const char *e = getenv("HPCX_NCCL_MAX_CTAS")) max_ctas = atoi(e);
  if (g_nccl.CommInitRankConfig && max_ctas > 0) {
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    cfg.maxCTAs = max_ctas;
    cfg.minCTAs = 1;
    HPCX_NCCL(ctx, g_nccl.CommInitRankConfig(&comm, nranks, id, rank, &cfg));
  } else {
    HPCX_NCCL(ctx, g_nccl.CommInitRank(&comm, nranks, id, rank));
  }
  ctx->nccl_comm = comm;
  ctx->nranks = nranks;
  ctx->rank = rank;
  return HPCX_OK;
}

hpcx_status hpcx_comm_finalize(hpcx_ctx ctx) {
  if (!ctx) return HPCX_INVALID_ARG;
  if (ctx->nccl_comm) {
    cudaDeviceSynchronize();
    g_nccl.CommDestroy(comm_of(ctx));
    ctx->nccl_comm = nullptr;
  }
  ctx->nranks = 1;
  ctx->rank = 0;
  return HPCX_OK;
}

hpcx_status hpcx_comm_barrier(hpcx_ctx ctx) {
  if (!ctx) return HPCX_INVALID_ARG;
  if (ctx->nranks <= 1 || !ctx->nccl_comm) {
    HPCX_CUDA(ctx, cudaDeviceSynchronize());
    return HPCX_OK;
  }
  if (!ctx->barrier_token) {
    HPCX_CUDA(ctx, cudaMalloc((void **)&ctx->barrier_token, 16));
    HPCX_CUDA(ctx, cudaMemsetAsync(ctx->barrier_token, 0, 16, ctx->comm));
  }
  float *token = ctx->barrier_token;
  HPCX_CUDA(ctx, cudaDeviceSynchronize());
  HPCX_NCCL(ctx, g_nccl.AllReduce(token, token, 1, ncclFloat32, ncclSum, comm_of(ctx), ctx->comm));
  HPCX_CUDA(ctx, cudaStreamSynchronize(ctx->comm));
  return HPCX_OK;
}

hpcx_status hpcx_comm_set_timeout(hpcx_ctx ctx, int64_t milliseconds) {
  if (!ctx || milliseconds < 0) return HPCX_INVALID_ARG;
  ctx->halo_timeout_ns = (uint64_t)milliseconds * 1000000ull;
  return HPCX_OK;
}

hpcx_status hpcx_comm_check(hpcx_ctx ctx) {
  if (!ctx) return HPCX_INVALID_ARG;
  return comm_check(ctx);
}

hpcx_status hpcx_comm_ipc_export(hpcx_ctx ctx, const void *dev_ptr, void *handle64) {
  if (!ctx || !dev_ptr || !handle64) return HPCX_INVALID_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaIpcMemHandle_t h;
  HPCX_CUDA(ctx, cudaIpcGetMemHandle(&h, const_cast<void *>(dev_ptr)));
  memcpy(handle64, &h, sizeof h);
  return HPCX_OK;
}

hpcx_status hpcx_comm_ipc_open(hpcx_ctx ctx, const void *handle64, void **peer_ptr) {
  if (!ctx || !handle64 || !peer_ptr) return HPCX_INVALID_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof h);
  HPCX_CUDA(ctx, cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return HPCX_OK;
}

hpcx_status hpcx_comm_ipc_close(hpcx_ctx ctx, void *peer_ptr) {
  if (!ctx) return HPCX_INVALID_ARG;
  HPCX_CUDA(ctx, cudaIpcCloseMemHandle(peer_ptr));
  return HPCX_OK;
}

// Plan of a distributed box sweep over a py x px grid of blocks (px == 1: y-slabs).
static hpcx_status box_plan_create(hpcx_ctx ctx, hpcx_grid in, hpcx_grid out, const hpcx_stencil_desc *st, int py,
                                   int px, hpcx_halo_plan *out_plan) {
  HPCX_REQUIRE(ctx, in->desc.ndim == 2 && out->desc.ndim == 2, "box sweeps take 2-D grids");
  HPCX_REQUIRE(ctx, in->desc.dtype == out->desc.dtype, "grids differ in dtype");
  HPCX_REQUIRE(ctx, in->desc.dim[0] == out->desc.dim[0] && in->desc.dim[1] == out->desc.dim[1],
               "grids differ in logical extent");
  HPCX_REQUIRE(ctx, st->radius >= 0 && st->radius <= in->desc.lo && st->radius <= in->desc.hi,
               "radius exceeds the ghost width of the input grid");
  HPCX_REQUIRE(ctx, in->n[1] >= std::max(st->radius, 1), "a block needs at least `radius` rows");
  HPCX_REQUIRE(ctx, px == 1 || in->n[2] >= std::max(st->radius, 1), "a block needs at least `radius` columns");
  HPCX_REQUIRE(ctx, py >= 1 && px >= 1 && py * px == ctx->nranks, "py * px must equal the number of ranks");
  if (ctx->nranks > 1 && !ctx->nccl_comm) return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  hpcx_halo_plan p = new (std::nothrow) hpcx_halo_plan_s();
  if (!p) return fail(ctx, HPCX_OOM, "host allocation failed");
  p->ctx = ctx;
  p->in = in;
  p->out = out;
  p->st = *st;
  p->py = py;
  p->px = px;
  cudaError_t e = p->create_events();
  if (e == cudaSuccess && px > 1 && st->kind == HPCX_STENCIL_BOX_X && st->radius > 0) {
    int64_t elems = 0;
    hpcx_grid_face_elems(in, 1, st->radius, &elems);
    for (int side = 0; side < 2 && e == cudaSuccess; ++side) {
      e = cudaMalloc(&p->face_send[side], (size_t)elems * in->es);
      if (e == cudaSuccess) e = cudaMalloc(&p->face_recv[side], (size_t)elems * in->es);
    }
  }
  if (e != cudaSuccess) {
    hpcx_stencil_dist_destroy(p);
    return fail(ctx, e == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR,
                std::string("box plan: ") + cudaGetErrorString(e));
  }
  *out_plan = p;
  return HPCX_OK;
}

hpcx_status hpcx_stencil_dist_plan_blocks(hpcx_ctx ctx, hpcx_grid in, hpcx_grid out, const hpcx_stencil_desc *st,
                                          int32_t py, int32_t px, hpcx_halo_plan *out_plan) {
  if (!ctx || !in || !out || !st || !out_plan) return HPCX_INVALID_ARG;
  *out_plan = nullptr;
  if (st->kind != HPCX_STENCIL_BOX_X && st->kind != HPCX_STENCIL_BOX_Y)
    return fail(ctx, HPCX_UNSUPPORTED, "the block split serves the 2-D box sweeps; JACOBI7 is split into z-slabs");
  return box_plan_create(ctx, in, out, st, py, px, out_plan);
}

hpcx_status hpcx_stencil_dist_plan(hpcx_ctx ctx, hpcx_grid in, hpcx_grid out,
                                   const hpcx_stencil_desc *st, int32_t transport,
                                   hpcx_halo_plan *out_plan) {
  if (!ctx || !in || !out || !st || !out_plan) return HPCX_INVALID_ARG;
  *out_plan = nullptr;
  if (st->kind == HPCX_STENCIL_BOX_X || st->kind == HPCX_STENCIL_BOX_Y) {
    // the reference's 17-point sweeps over y-slabs (rows split like its static chunking,
    // test_hpc_benchmark.cpp:1356,1371-1372): BOX_X needs no exchange, BOX_Y trades `radius` rows
    if (transport != 0)
      return fail(ctx, HPCX_UNSUPPORTED, "distributed box sweeps use the NCCL transport (0)");
    return box_plan_create(ctx, in, out, st, ctx->nranks, 1, out_plan);
  }
  HPCX_REQUIRE(ctx, st->kind == HPCX_STENCIL_JACOBI7, "unknown stencil kind");
  HPCX_REQUIRE(ctx, in->desc.ndim == 3 && out->desc.ndim == 3, "JACOBI7 takes 3-D grids");
  HPCX_REQUIRE(ctx, in->desc.dtype == out->desc.dtype, "grids differ in dtype");
  for (int i = 0; i < 3; ++i)
    HPCX_REQUIRE(ctx, in->desc.dim[i] == out->desc.dim[i], "grids differ in logical extent");
  HPCX_REQUIRE(ctx, in->desc.lo >= 1 && in->desc.hi >= 1 && in->desc.lo == out->desc.lo &&
                        in->desc.hi == out->desc.hi,
               "JACOBI7 needs identical ghost widths >= 1 on both grids");
  HPCX_REQUIRE(ctx, in->n[0] >= 1, "a slab needs at least one plane");
  if (transport != 0 && transport != 1)
    return fail(ctx, HPCX_INVALID_ARG, "transport must be 0 (NCCL send/recv) or 1 (peer stores)");
  if (transport == 1 && (in->desc.dtype != HPCX_F32 || st->variant == HPCX_VARIANT_DIRECT ||
                         !ctx->encode_tiled))
    return fail(ctx, HPCX_UNSUPPORTED, "the peer-store transport needs the fp32 TMA sweep");
  if (ctx->nranks > 1 && !ctx->nccl_comm)
    return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  if (st->variant == HPCX_VARIANT_TEMPORAL) {
    // Two sweeps per pass: a slab trades TWO planes per side and pass, so every slab needs ghost
    // width >= 2 and at least two planes.  The ranks agree on the verdict (sum of per-rank
    // violations) so that nobody is left waiting in a receive.
    if (!jacobi7_tb2_supported(in))
      return fail(ctx, HPCX_UNSUPPORTED, "the temporally blocked Jacobi needs fp32 and a tensor-map driver");
    if (ctx->nranks > 1) {
      float bad = (in->desc.lo < 2 || in->desc.hi < 2 || in->n[0] < 2) ? 1.f : 0.f, *d_bad = nullptr;
      HPCX_CUDA(ctx, cudaMalloc(&d_bad, sizeof(float)));
      cudaMemcpyAsync(d_bad, &bad, sizeof(float), cudaMemcpyHostToDevice, ctx->comm);
      hpcx_status rc = hpcx::allreduce_sum(ctx, d_bad, 1, HPCX_F32, ctx->comm);
      if (rc == HPCX_OK && cudaMemcpyAsync(&bad, d_bad, sizeof(float), cudaMemcpyDeviceToHost, ctx->comm) == cudaSuccess)
        cudaStreamSynchronize(ctx->comm);
      cudaFree(d_bad);
      if (rc != HPCX_OK) return rc;
      if (bad != 0.f)
        return fail(ctx, HPCX_INVALID_ARG, "TEMPORAL over slabs needs ghost widths >= 2 and >= 2 planes on every rank");
    }
  }
  hpcx_halo_plan p = new (std::nothrow) hpcx_halo_plan_s();
  if (!p) return fail(ctx, HPCX_OOM, "host allocation failed");
  p->ctx = ctx;
  p->in = in;
  p->out = out;
  p->st = *st;
  p->transport = transport;
  if (const char *g = getenv("HPCX_NCCL_GRAPHS")) p->graphs_ok = atoi(g) != 0;  // 0: replay nothing, launch every call eagerly
  cudaError_t e = p->create_events();
  if (e != cudaSuccess) {
    p->destroy_events();
    delete p;
    return fail(ctx, HPCX_CUDA_ERROR, std::string("cudaEventCreate: ") + cudaGetErrorString(e));
  }
  if (transport == 1 && ctx->nranks > 1) {
    hpcx_status rc = open_peers(p);
    if (rc != HPCX_OK) {
      close_peers(p);
      hpcx_stencil_dist_destroy(p);
      return rc;
    }
  }
  *out_plan = p;
  return HPCX_OK;
}

// Transport 1: ONE kernel per sweep.  Its boundary CTAs run in the first wave (the short top z
// segment is scheduled second), write their plane into this rank's dst AND into the neighbour's
// ghost plane with peer stores over NVLink, and release a sweep-number flag in the neighbour; the
// next sweep's CTAs acquire that flag before their first TMA load of a ghost plane.  No NCCL call,
// no second stream and no pack kernel on the data path; the exchange overlaps the interior
// planes of the same kernel.  Two-sweep passes (HPCX_VARIANT_TEMPORAL) run the plain pass kernel
// and push their two boundary planes per side with halo_push_planes_kernel right behind it.
}  // extern "C"
namespace hpcx {
// Slabs of at most this many planes push their boundary planes from inside the two-sweep pass
// kernel (stencil.cu, jacobi7_tb2_peer_kernel); thicker ones through halo_push_planes_kernel.
constexpr int64_t kThinSlabPlanes = 128;
hpcx_status stencil_peer_sweep(hpcx_halo_plan p, int32_t steps, cudaStream_t cs, bool sync) {
  hpcx_ctx ctx = p->ctx;
  hpcx_grid g = p->in;
  const bool has_lo = ctx->rank > 0, has_hi = ctx->rank + 1 < ctx->nranks;
  const int64_t plane = g->ext[1] * g->pitch;
  int cur = p->in->dev == p->buf[0] ? 0 : 1;  // which of (A, B) is the source
  if (p->in->dev != p->buf[cur] || p->out->dev != p->buf[cur ^ 1])
    return fail(ctx, HPCX_INVALID_ARG, "grid buffers changed since the plan was made");
  hpcx_status rc = HPCX_OK;
  // HPCX_VARIANT_TEMPORAL: steps/2 two-sweep passes (two planes per side copied into the
  // neighbours by a push kernel behind each pass), then one single sweep when steps is odd.  `epoch` counts
  // passes: flags carry the number of the pass whose planes have landed.
  const bool temporal = p->st.variant == HPCX_VARIANT_TEMPORAL;
  const int passes = temporal ? steps / 2 + (steps & 1) : steps;
  for (int t = 0; t < passes; ++t) {
    const int dst = cur ^ 1;
    JacobiPeer pr;
    if (has_lo) {
      pr.ghost_lo = p->peer_buf[0][dst] + (p->peer_n0[0] + g->glo[0]) * plane;
      pr.sig_lo = p->peer_flags[0] + 1;  // this rank is the low neighbour's HIGH side
    }
    if (has_hi) {
      pr.ghost_hi = p->peer_buf[1][dst] + (g->glo[0] - 1) * plane;
      pr.sig_hi = p->peer_flags[1] + 0;
    }
    pr.flags = p->flags;
    pr.counters = p->flags + 4;
    pr.pw = peer_wait_of(ctx);
    pr.wait_val = t == 0 ? 0 : p->epoch + (uint32_t)t;  // pass 0 reads NCCL-exchanged ghosts
    pr.sig_val = p->epoch + (uint32_t)t + 1;
    if (temporal && t < steps / 2 && g->n[0] >= 2 && g->n[0] <= kThinSlabPlanes) {
      // thin slab: ONE launch per pass — the pass kernel awaits its ghost planes itself and the CTAs
      // that own planes 0, 1 / n0-2, n0-1 push them into the neighbours' ghosts n, n+1 / -2, -1
      JacobiPeer pp = pr;
      if (has_hi) pp.ghost_hi = pr.ghost_hi - plane;  // first of the two ghost planes
      pp.counters = p->flags + 6;                     // tile counts differ from the one-sweep kernel's
      rc = jacobi7_tb2_peer_launch(g, p->buf[cur], p->buf[dst], p->st.zsplit, pp, cs);
    } else if (temporal && t < steps / 2) {
      // pass kernel ; push kernel (planes 0, 1 -> low neighbour's ghosts n, n+1; planes n0-2, n0-1
      // -> high neighbour's ghosts -2, -1; then waits until the neighbours' planes have landed here)
      rc = jacobi7_tb2_launch(g, p->buf[cur], p->buf[dst], p->st.zsplit, 0, g->n[0], has_lo, has_hi, cs);
      if (rc != HPCX_OK) return rc;
      const float *mine = static_cast<const float *>(p->buf[dst]);
      halo_push_planes_kernel<<<2 * ctx->sm_count, 256, 0, cs>>>(
          reinterpret_cast<const float4 *>(mine + g->glo[0] * plane),
          reinterpret_cast<float4 *>(has_lo ? pr.ghost_lo : nullptr),
          reinterpret_cast<const float4 *>(mine + (g->glo[0] + g->n[0] - 2) * plane),
          reinterpret_cast<float4 *>(has_hi ? pr.ghost_hi - plane : nullptr), 2 * plane / 4, p->flags + 8,
          pr.sig_lo, pr.sig_hi, pr.sig_val, p->flags, pr.pw);
      rc = after_launch(ctx, "halo_push_planes_kernel");
    } else {
      rc = jacobi7_launch(g, p->buf[cur], p->buf[dst], HPCX_VARIANT_TMA, temporal ? 0 : p->st.zsplit, 0, g->n[0],
                          cs, &pr);
    }
    if (rc != HPCX_OK) return rc;
    cur = dst;
  }
  p->epoch += (uint32_t)passes;
  steps = passes;  // parity of the buffer swap below
  halo_wait_kernel<<<1, 1, 0, cs>>>(p->flags, peer_wait_of(ctx), p->epoch, has_lo, has_hi);
  rc = after_launch(ctx, "halo_wait_kernel");
  if (rc != HPCX_OK) return rc;
  if ((steps & 1) == 0) std::swap(p->in->dev, p->out->dev);  // newest iterate is `out`
  if (sync) return sync_and_check(ctx, cs);
  return HPCX_OK;
}
}  // namespace hpcx
// Everything one NCCL-transport call puts on the two streams (no host synchronisation inside, so the
// sequence can be captured into a graph): the initial ghost exchange of `in`, then per pass
//   comm stream   : boundary planes z = 0, n-1, then send dst boundary planes / receive dst ghost planes
//   compute stream: interior planes 1 .. n-2 (beside both)
// HPCX_VARIANT_TEMPORAL runs steps/2 two-sweep passes the same way with TWO planes per side (the
// neighbour's boundary plane is advanced redundantly on this rank), then one single sweep when
// steps is odd: the bytes exchanged per sweep are unchanged, the messages per sweep halve.
static hpcx_status nccl_sweeps_enqueue(hpcx_halo_plan p, int32_t steps, cudaStream_t cs) {
  hpcx_ctx ctx = p->ctx;
  hpcx_grid g = p->in;
  cudaStream_t ms = ctx->comm;
  const int64_t nz = g->n[0];
  hpcx_status rc = HPCX_OK;
  // Make the ghosts of the first source iterate consistent with the neighbours.
  HPCX_CUDA(ctx, cudaEventRecord(p->ev_start, cs));
  HPCX_CUDA(ctx, cudaStreamWaitEvent(ms, p->ev_start, 0));
  rc = exchange_planes(g, p->in->dev, ms, (p->st.variant == HPCX_VARIANT_TEMPORAL && steps >= 2) ? 2 : 1);
  if (rc != HPCX_OK) return rc;
  // A pass advances the slab by w sweeps (w = 2: the temporally blocked kernel) and needs w valid
  // ghost planes per side of its source; the exchange after a pass carries the width the NEXT pass
  // reads.  Boundary = the w planes next to each z neighbour.
  const bool temporal = p->st.variant == HPCX_VARIANT_TEMPORAL;
  const bool nb_lo = ctx->rank > 0, nb_hi = ctx->rank + 1 < ctx->nranks;
  const int passes = temporal ? steps / 2 + (steps & 1) : steps;
  void *src = p->in->dev, *dst = p->out->dev;
  // The boundary strips run on the COMM stream (high priority), ahead of the exchange they feed,
  // while the compute stream goes straight from one interior launch to the next: the interior of
  // pass t needs the boundary strips of pass t-1 (ev_bnd[(t-1) & 1]) but neither the strips nor the
  // exchange of pass t, and the strips of pass t need the interior of pass t-1 (ev_int) and the
  // exchange of pass t-1 (stream order).  With the strips in front of the interior on one stream
  // (round 1) every sweep paid for their launch, their 128-CTA tail and the hand-over twice.
  for (int t = 0; t < passes; ++t) {
    const bool two = temporal && t < steps / 2;
    const int64_t w = two ? 2 : 1;
    const int64_t lo_end = std::min<int64_t>(w, nz), hi_beg = std::max<int64_t>(lo_end, nz - w);
    auto launch = [&](int64_t zb, int64_t ze, int zsplit, cudaStream_t st) -> hpcx_status {
      if (ze <= zb) return HPCX_OK;
      if (two) return jacobi7_tb2_launch(g, src, dst, zsplit, zb, ze, nb_lo, nb_hi, st);
      return jacobi7_launch(g, src, dst, temporal ? HPCX_VARIANT_TMA : p->st.variant, zsplit, zb, ze, st, nullptr);
    };
    if (t > 0) HPCX_CUDA(ctx, cudaStreamWaitEvent(ms, p->ev_int, 0));
    nvtxRangePushA("sweep: boundary planes");
    if (two) {  // both boundary strips in one launch
      rc = jacobi7_tb2_launch(g, src, dst, 1, 0, lo_end, nb_lo, nb_hi, ms, hi_beg, nz);
    } else {
      rc = launch(0, lo_end, 1, ms);
      if (rc == HPCX_OK) rc = launch(hi_beg, nz, 1, ms);
    }
    nvtxRangePop();
    if (rc != HPCX_OK) return rc;
    HPCX_CUDA(ctx, cudaEventRecord(p->ev_bnd[t & 1], ms));
    if (t + 1 < passes) {
      HPCX_RANGE("sweep: ghost-plane exchange (NCCL send/recv)");
      const int64_t wnext = (temporal && t + 1 < steps / 2) ? 2 : 1;
      rc = exchange_planes(g, dst, ms, wnext);
      if (rc != HPCX_OK) return rc;
    }
    {
      HPCX_RANGE("sweep: interior planes");
      if (t > 0) HPCX_CUDA(ctx, cudaStreamWaitEvent(cs, p->ev_bnd[(t - 1) & 1], 0));
      rc = launch(lo_end, hi_beg, temporal ? (two ? p->st.zsplit : 0) : p->st.zsplit, cs);
      if (rc != HPCX_OK) return rc;
      if (t + 1 < passes) HPCX_CUDA(ctx, cudaEventRecord(p->ev_int, cs));
    }
    std::swap(src, dst);
  }
  // the caller's stream ends behind the strips of the last pass
  HPCX_CUDA(ctx, cudaStreamWaitEvent(cs, p->ev_bnd[(passes - 1) & 1], 0));
  return HPCX_OK;
}

// One distributed box sweep in -> out over this rank's block (a y-slab when px == 1).  The `radius`
// ghost rows (BOX_Y; contiguous, nothing is packed) or ghost columns (BOX_X with px > 1; packed
// faces) of `in` come from the neighbours on the comm stream while the part of the block that needs
// none of them is swept; the two boundary strips follow the receive.
static hpcx_status box_sweep_enqueue(hpcx_halo_plan p, int32_t, cudaStream_t cs) {
  hpcx_ctx ctx = p->ctx;
  hpcx_grid g = p->in;
  cudaStream_t ms = ctx->comm;
  const int64_t ny = g->n[1], nx = g->n[2];
  const bool along_y = p->st.kind == HPCX_STENCIL_BOX_Y;
  const int64_t r = p->st.radius, n = along_y ? ny : nx;
  hpcx_status rc = HPCX_OK;
  HPCX_CUDA(ctx, cudaEventRecord(p->ev_start, cs));
  HPCX_CUDA(ctx, cudaStreamWaitEvent(ms, p->ev_start, 0));
  {
    HPCX_RANGE("box sweep: ghost exchange (NCCL send/recv)");
    rc = along_y ? exchange_planes(g, p->in->dev, ms, r, p->px) : exchange_cols(p, ms);
  }
  if (rc != HPCX_OK) return rc;
  HPCX_CUDA(ctx, cudaEventRecord(p->ev_recv, ms));
  const int64_t lo_end = std::min<int64_t>(r, n), hi_beg = std::max<int64_t>(lo_end, n - r);
  auto strip = [&](int64_t b0, int64_t b1) {
    return along_y ? box_launch_rows(p->in, p->out, &p->st, b0, b1, cs)
                   : box_launch_tile(p->in, p->out, &p->st, 0, ny, b0, b1, cs);
  };
  {
    HPCX_RANGE("box sweep: interior");
    rc = strip(lo_end, hi_beg);
  }
  if (rc != HPCX_OK) return rc;
  HPCX_CUDA(ctx, cudaStreamWaitEvent(cs, p->ev_recv, 0));
  HPCX_RANGE("box sweep: boundary strips");
  rc = strip(0, lo_end);
  if (rc == HPCX_OK) rc = strip(hi_beg, n);
  return rc;
}

// Runs `enqueue(p, steps, cs)` — a sequence of launches, NCCL groups and event hand-overs between
// the caller's stream and the comm stream, with no host synchronisation inside — as ONE CUDA graph
// replay: captured the first time a (source buffer, destination buffer, steps) triple is seen,
// replayed afterwards.  A Jacobi sweep is three launches plus an NCCL group: at 90-170 us per sweep
// and 100 sweeps per call their launch gaps and the hand-overs between the two streams are what
// the replay removes.  Falls back to eager launches
// when the stream is already being captured, HPCX_NCCL_GRAPHS=0, or a capture fails.
typedef hpcx_status (*enqueue_fn)(hpcx_halo_plan, int32_t, cudaStream_t);
static hpcx_status replay_or_enqueue(hpcx_halo_plan p, int32_t steps, cudaStream_t cs, enqueue_fn enqueue) {
  hpcx_ctx ctx = p->ctx;
  hpcx_status rc = HPCX_OK;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (ctx->nranks > 1 && p->graphs_ok && cudaStreamIsCapturing(cs, &cap) == cudaSuccess &&
      cap == cudaStreamCaptureStatusNone) {
    const hpcx_halo_plan_s::GraphEntry *hit = nullptr;
    for (auto &e : p->graphs)
      if (e.in == p->in->dev && e.out == p->out->dev && e.steps == steps) hit = &e;
    if (!hit) {
      HPCX_RANGE("sweep: capture into a CUDA graph");
      if (hpcx_status st = preload_stencil_kernels(ctx)) return st;  // no attribute calls inside the capture
      cudaGraph_t graph = nullptr;
      cudaGraphExec_t exec = nullptr;
      const int64_t before = ctx->launches;
      cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed);
      if (e == cudaSuccess) {
        rc = enqueue(p, steps, cs);
        e = cudaStreamEndCapture(cs, &graph);
        if (rc == HPCX_OK && e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
        if (graph) cudaGraphDestroy(graph);
      }
      const int64_t per_replay = ctx->launches - before;
      ctx->launches = before;  // nothing has run yet
      if (rc != HPCX_OK || e != cudaSuccess || !exec) {
        (void)cudaGetLastError();
        p->graphs_ok = false;  // every rank takes the same decision: same code, same library
        rc = HPCX_OK;
      } else {
        if (p->graphs.size() >= 8) {
          for (auto &old : p->graphs) cudaGraphExecDestroy(old.exec);
          p->graphs.clear();
        }
        p->graphs.push_back({p->in->dev, p->out->dev, steps, exec, per_replay});
        hit = &p->graphs.back();
      }
    }
    if (hit) {
      HPCX_CUDA(ctx, cudaGraphLaunch(hit->exec, cs));
      ctx->launches += hit->launches;
      return HPCX_OK;
    }
  }
  return enqueue(p, steps, cs);
}

extern "C" {

// Per step t (src -> dst):
//   comm stream   : [wait interior of t-1] boundary planes z = 0, n-1 -> ev_bnd[t & 1];
//                   send dst boundary planes / receive dst ghost planes
//   compute stream: [wait boundary planes of t-1] interior planes 1 .. n-2 -> ev_int
// The ghost planes a step needs were produced by the previous step's exchange (stream order).
// HPCX_VARIANT_TEMPORAL runs steps/2 two-sweep passes the same way with TWO planes per side (the
// neighbour's boundary plane is advanced redundantly on this rank), then one single sweep when
// steps is odd: the bytes exchanged per sweep are unchanged, the messages per sweep halve.
hpcx_status hpcx_stencil_dist_sweep(hpcx_halo_plan p, int32_t steps, void *stream) {
  if (!p) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_stencil_dist_sweep");
  hpcx_ctx ctx = p->ctx;
  HPCX_REQUIRE(ctx, steps >= 1, "steps must be >= 1");
  hpcx_grid g = p->in;
  cudaStream_t cs = stream ? (cudaStream_t)stream : ctx->compute;
  cudaStream_t ms = ctx->comm;
  const int64_t nz = g->n[0];
  hpcx_status rc = HPCX_OK;

  if (p->st.kind == HPCX_STENCIL_BOX_X || p->st.kind == HPCX_STENCIL_BOX_Y) {
    HPCX_REQUIRE(ctx, steps == 1, "box sweeps are single-step");
    const bool along_y = p->st.kind == HPCX_STENCIL_BOX_Y;
    const bool exchange = p->st.radius > 0 && ctx->nranks > 1 && (along_y ? p->py > 1 : p->px > 1);
    // eager: replaying the sweep as a CUDA graph was measured and lost (x sweep over x-slabs at
    // N = 2: 1046 against 1175 Gcell/s; a single 100 us sweep per call leaves a replay nothing to hide)
    rc = exchange ? box_sweep_enqueue(p, 1, cs) : box_launch_rows(p->in, p->out, &p->st, 0, g->n[1], cs);
    if (rc != HPCX_OK) return rc;
    if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(cs));
    return HPCX_OK;
  }

  if (p->transport == 1 && ctx->nranks > 1) {
    // Make the ghosts of the first source iterate consistent with the neighbours, then the fused sweeps.
    HPCX_CUDA(ctx, cudaEventRecord(p->ev_start, cs));
    HPCX_CUDA(ctx, cudaStreamWaitEvent(ms, p->ev_start, 0));
    rc = exchange_planes(g, p->in->dev, ms, (p->st.variant == HPCX_VARIANT_TEMPORAL && steps >= 2) ? 2 : 1);
    if (rc != HPCX_OK) return rc;
    HPCX_CUDA(ctx, cudaEventRecord(p->ev_recv, ms));
    HPCX_CUDA(ctx, cudaStreamWaitEvent(cs, p->ev_recv, 0));
    return hpcx::stencil_peer_sweep(p, steps, cs, stream == nullptr);
  }

  // NCCL transport: the whole call as one graph replay (replay_or_enqueue).
  rc = replay_or_enqueue(p, steps, cs, nccl_sweeps_enqueue);
  if (rc != HPCX_OK) return rc;
  const bool temporal = p->st.variant == HPCX_VARIANT_TEMPORAL;
  const int passes = temporal ? steps / 2 + (steps & 1) : steps;
  if ((passes & 1) == 0) std::swap(p->in->dev, p->out->dev);
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(cs));
  return HPCX_OK;
}

hpcx_status hpcx_stencil_dist_destroy(hpcx_halo_plan p) {
  if (!p) return HPCX_INVALID_ARG;
  cudaDeviceSynchronize();
  const hpcx_status pending = comm_check(p->ctx);  // a timed-out wait of an asynchronous sweep surfaces here
  close_peers(p);
  p->destroy_events();
  for (int side = 0; side < 2; ++side) {
    cudaFree(p->face_send[side]);
    cudaFree(p->face_recv[side]);
  }
  for (auto &e : p->graphs) cudaGraphExecDestroy(e.exec);
  delete p;
  return pending;
}

}  // extern "C"

// ------------------------------------------------------------------ distributed SpMV
namespace hpcx {
void csr_close_peers(hpcx_csr A);
}
namespace {

template <typename T>
__global__ void __launch_bounds__(256)
gather_boundary_x_kernel(const T *__restrict__ x, const int32_t *__restrict__ send_idx,
                         T *__restrict__ sendbuf, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    sendbuf[i] = __ldg(x + send_idx[i]);
}

// ---- transport 1: the push (gather + peer stores + flags), the flag waits and the ack all live
// in the SpMV kernel itself (spmv.cu, DIST instantiation; SpmvDistArgs in spmv.cuh).
constexpr int kSyncData = 0, kSyncAck = 64, kMaxPeers = 64;  // words of the sync block; counters follow at 128, 129

struct CsrPeerInfo {
  cudaIpcMemHandle_t handle[2];  // allocations holding [owned | halo] x and the sync block
  int64_t offset[2];
  int64_t nloc;
  int64_t recv_off[kMaxPeers];   // element offset inside the halo of the segment from rank q
  int64_t recv_cnt[kMaxPeers];
};

void close_csr_peers(hpcx_csr A) {
  for (void *b : A->peer_open) cudaIpcCloseMemHandle(b);
  A->peer_open.clear();
  cudaFree(A->d_sync);
  cudaFree(A->d_seg_start);
  cudaFree(A->d_seg_rank);
  cudaFree(A->d_seg_dst);
  cudaFree(A->d_seg_flag);
  cudaFree(A->d_src_rank);
  cudaFree(A->d_src_ack);
  A->d_sync = nullptr;
  A->d_seg_start = A->d_seg_rank = A->d_src_rank = nullptr;
  A->d_seg_dst = nullptr;
  A->d_seg_flag = A->d_src_ack = nullptr;
  A->nseg = A->nsrc = 0;
  A->transport = 0;
}

}  // namespace

namespace hpcx {
void csr_close_peers(hpcx_csr A) { close_csr_peers(A); }

// see preload_stencil_kernels (stencil.cu)
hpcx_status preload_dist_kernels(hpcx_ctx ctx) {
  cudaFuncAttributes a;
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)halo_wait_kernel));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)halo_push_planes_kernel));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)gather_boundary_x_kernel<float>));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)gather_boundary_x_kernel<double>));
  return HPCX_OK;
}

// The sync block of the fused transport: flags by source rank, acks by destination rank, counters.
hpcx_status csr_peer_prepare(hpcx_csr A) {
  hpcx_ctx ctx = A->ctx;
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_sync, 2u << 20));  // an allocation of its own (IPC exports whole allocations)
  HPCX_CUDA(ctx, cudaMemset(A->d_sync, 0, 1024));
  HPCX_CUDA(ctx, cudaDeviceSynchronize());
  return HPCX_OK;
}

// Builds the push / wait tables of the fused transport from what this rank can see of its peers:
// views[q] = rank q's [owned | halo] x buffer and sync block as pointers valid on THIS device
// (CUDA IPC windows between processes, plain peer pointers inside one process) plus where this
// rank's segment lands in q's halo.
hpcx_status csr_peer_wire(hpcx_csr A, const std::vector<CsrPeerView> &views) {
  hpcx_ctx ctx = A->ctx;
  const int P = ctx->nranks, me = ctx->rank;
  std::vector<int32_t> seg_start{0}, seg_rank, src_rank;
  std::vector<void *> seg_dst;
  std::vector<uint32_t *> seg_flag, src_ack;
  int64_t soff = 0;
  for (int q = 0; q < P; ++q) {
    const bool sends = A->send_counts[q] > 0, recvs = A->recv_counts[q] > 0;
    if (q == me || (!sends && !recvs)) {
      soff += A->send_counts[q];
      continue;
    }
    const CsrPeerView &v = views[q];
    if (!v.xfull || !v.sync) return fail(ctx, HPCX_INVALID_ARG, "no window into a rank this one trades with");
    if (v.recv_cnt_me != A->send_counts[q])
      return fail(ctx, HPCX_INVALID_ARG, "send/recv counts of two ranks do not match");
    if (sends) {
      seg_rank.push_back(q);
      seg_dst.push_back(v.xfull + (v.nloc + v.recv_off_me) * A->es);
      seg_flag.push_back(v.sync + kSyncData + me);
      soff += A->send_counts[q];
      seg_start.push_back((int32_t)soff);
    }
    if (recvs) {
      src_rank.push_back(q);
      src_ack.push_back(v.sync + kSyncAck + me);
    }
  }
  A->nseg = (int32_t)seg_rank.size();
  A->nsrc = (int32_t)src_rank.size();
  auto up = [&](const void *h, size_t bytes, void **d) -> cudaError_t {
    cudaError_t e = cudaMalloc(d, std::max<size_t>(bytes, 16));
    if (e == cudaSuccess && bytes) e = cudaMemcpy(*d, h, bytes, cudaMemcpyHostToDevice);
    return e;
  };
  HPCX_CUDA(ctx, up(seg_start.data(), seg_start.size() * 4, (void **)&A->d_seg_start));
  HPCX_CUDA(ctx, up(seg_rank.data(), seg_rank.size() * 4, (void **)&A->d_seg_rank));
  HPCX_CUDA(ctx, up(seg_dst.data(), seg_dst.size() * sizeof(void *), (void **)&A->d_seg_dst));
  HPCX_CUDA(ctx, up(seg_flag.data(), seg_flag.size() * sizeof(void *), (void **)&A->d_seg_flag));
  HPCX_CUDA(ctx, up(src_rank.data(), src_rank.size() * 4, (void **)&A->d_src_rank));
  HPCX_CUDA(ctx, up(src_ack.data(), src_ack.size() * sizeof(void *), (void **)&A->d_src_ack));
  A->epoch = 0;
  A->transport = 1;
  return HPCX_OK;
}
}  // namespace hpcx

extern "C" {

hpcx_status hpcx_csr_set_halo(hpcx_csr A, int64_t nloc, int64_t nhalo, const int64_t *recv_counts,
                              const int64_t *send_counts, const int32_t *send_idx,
                              const uint8_t *row_kind) {
  if (!A) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = A->ctx;
  HPCX_REQUIRE(ctx, nloc == A->nrows && nloc + nhalo == A->ncols,
               "local matrix must have nloc rows and nloc + nhalo columns");
  HPCX_REQUIRE(ctx, recv_counts && send_counts && row_kind, "null halo table");
  const int P = ctx->nranks;
  close_csr_peers(A);  // peer windows point into the buffers replaced below
  // the distributed product, its host-operand pipeline and the CG step all run the STREAM kernel
  // (its tables always exist); HUB tables built at create time would only hold memory from here on
  if (A->hub.ready) hub_free(A);
  if (A->kernel == HPCX_SPMV_HUB) A->kernel = HPCX_SPMV_STREAM;
  A->nloc = nloc;
  A->nhalo = nhalo;
  A->recv_counts.assign(recv_counts, recv_counts + P);
  A->send_counts.assign(send_counts, send_counts + P);
  int64_t nsend = 0, nrecv = 0;
  for (int q = 0; q < P; ++q) {
    HPCX_REQUIRE(ctx, send_counts[q] >= 0 && recv_counts[q] >= 0, "negative halo count");
    nsend += send_counts[q];
    nrecv += recv_counts[q];
  }
  HPCX_REQUIRE(ctx, nrecv == nhalo, "recv_counts do not add up to nhalo");
  HPCX_REQUIRE(ctx, nsend == 0 || send_idx, "null send_idx");
  for (int64_t i = 0; i < nsend; ++i)
    if (send_idx[i] < 0 || send_idx[i] >= nloc)
      return fail(ctx, HPCX_OUT_OF_RANGE, "send_idx entry outside [0, nloc)");
  A->nsend = nsend;
  cudaFree(A->d_send_idx);
  cudaFree(A->d_sendbuf);
  cudaFree(A->d_xfull);
  cudaFree(A->d_blocks_dist);
  cudaFree(A->d_long_row_dist);
  cudaFree(A->d_long_off_dist);
  A->d_send_idx = nullptr;
  A->d_sendbuf = A->d_xfull = nullptr;
  A->d_blocks_dist = A->d_blocks_int = A->d_blocks_bnd = nullptr;
  A->d_long_row_dist = A->d_long_off_dist = nullptr;
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_send_idx, std::max<size_t>((size_t)nsend, 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc(&A->d_sendbuf, std::max<size_t>((size_t)nsend, 1) * A->es));
  HPCX_CUDA(ctx, cudaMalloc(&A->d_xfull, std::max<size_t>((size_t)A->ncols, 1) * A->es));
  if (nsend)
    HPCX_CUDA(ctx, cudaMemcpy(A->d_send_idx, send_idx, (size_t)nsend * 4, cudaMemcpyHostToDevice));

  // One row-block table [interior blocks | boundary blocks] over the device-resident row_ptr
  // (csr_plan.cu), cut wherever the row kind changes.
  // Interior rows may run before the halo lands, so the table is cut wherever the kind changes.  On a
  // power-law matrix the two kinds alternate row by row (nearly every row of some length references
  // a foreign column): cutting there shreds the table into millions of blocks of a few rows (C5 over
  // 2 GPUs: 7.1 ms per product).  When the kinds are that fragmented every row is treated as a
  // boundary row instead — one contiguous table, all blocks start after the flags, nothing overlaps
  // the exchange, which for such matrices is the larger part of x anyway.
  std::vector<uint8_t> all_boundary;
  {
    int64_t changes = 0;
    for (int64_t i = 1; i < nloc; ++i) changes += row_kind[i] != row_kind[i - 1];
    if (changes > 64 && changes > nloc / 256) {
      all_boundary.assign((size_t)nloc, (uint8_t)1);
      row_kind = all_boundary.data();
    }
  }
  uint8_t *d_kind = nullptr;
  HPCX_CUDA(ctx, cudaMalloc((void **)&d_kind, std::max<size_t>((size_t)nloc, 1)));
  cudaError_t ke = cudaMemcpy(d_kind, row_kind, (size_t)nloc, cudaMemcpyHostToDevice);
  std::vector<RowBlock> all;
  int64_t n0 = 0, nchunks = 0;
  hpcx_status st = ke == cudaSuccess
                       ? csr_plan_rowblocks(A, d_kind, A->max_row_len, &A->d_blocks_dist, &all, &n0, &A->d_long_row_dist,
                                            &A->d_long_off_dist, &A->nlong_dist, &nchunks)
                       : fail(ctx, HPCX_CUDA_ERROR, std::string("row_kind upload: ") + cudaGetErrorString(ke));
  cudaFree(d_kind);
  if (st != HPCX_OK) return st;
  A->h_blocks_int.assign(all.begin(), all.begin() + n0);
  A->h_blocks_bnd.assign(all.begin() + n0, all.end());
  for (cudaEvent_t e : A->pipe_dist.ev) cudaEventDestroy(e);
  A->pipe_dist = PipePlan{};  // the chunking of the host-operand pipeline follows the new tables
  A->d_blocks_int = A->d_blocks_dist;
  A->d_blocks_bnd = A->d_blocks_dist + n0;
  return HPCX_OK;
}

hpcx_status hpcx_csr_halo_transport(hpcx_csr A, int32_t transport) {
  if (!A) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = A->ctx;
  HPCX_REQUIRE(ctx, transport == 0 || transport == 1, "transport must be 0 or 1");
  HPCX_REQUIRE(ctx, A->d_xfull != nullptr, "hpcx_csr_set_halo has not been called");
  HPCX_CUDA(ctx, cudaDeviceSynchronize());
  close_csr_peers(A);
  if (transport == 0 || ctx->nranks <= 1) return HPCX_OK;
  if (!ctx->nccl_comm) return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  const int P = ctx->nranks, me = ctx->rank;
  HPCX_REQUIRE(ctx, P <= kMaxPeers, "the peer-store transport supports up to 64 ranks");

  hpcx_status st = hpcx::csr_peer_prepare(A);
  if (st != HPCX_OK) return st;
  CsrPeerInfo mine{};
  st = ipc_export(ctx, A->d_xfull, &mine.handle[0], &mine.offset[0]);
  if (st == HPCX_OK) st = ipc_export(ctx, A->d_sync, &mine.handle[1], &mine.offset[1]);
  if (st != HPCX_OK) return st;
  mine.nloc = A->nloc;
  int64_t roff = 0;
  for (int q = 0; q < P; ++q) {
    mine.recv_off[q] = roff;
    mine.recv_cnt[q] = A->recv_counts[q];
    roff += A->recv_counts[q];
  }
  std::vector<CsrPeerInfo> all(P);
  st = allgather_bytes(ctx, &mine, all.data(), sizeof(CsrPeerInfo));
  if (st != HPCX_OK) return st;
  // open the windows of the ranks this one trades with (CUDA IPC) and wire the tables
  std::vector<hpcx::CsrPeerView> views(P);
  for (int q = 0; q < P; ++q) {
    const bool trades = q != me && (A->send_counts[q] > 0 || A->recv_counts[q] > 0);
    if (!trades) continue;
    void *bx = nullptr, *bs = nullptr;
    HPCX_CUDA(ctx, cudaIpcOpenMemHandle(&bx, all[q].handle[0], cudaIpcMemLazyEnablePeerAccess));
    A->peer_open.push_back(bx);
    HPCX_CUDA(ctx, cudaIpcOpenMemHandle(&bs, all[q].handle[1], cudaIpcMemLazyEnablePeerAccess));
    A->peer_open.push_back(bs);
    views[q].xfull = (char *)bx + all[q].offset[0];
    views[q].sync = reinterpret_cast<uint32_t *>((char *)bs + all[q].offset[1]);
    views[q].nloc = all[q].nloc;
    views[q].recv_off_me = all[q].recv_off[me];
    views[q].recv_cnt_me = all[q].recv_cnt[me];
  }
  st = hpcx::csr_peer_wire(A, views);
  if (st != HPCX_OK) return st;
  // Nobody may push before every rank's sync block is zeroed and its windows are open.
  return hpcx_comm_barrier(ctx);
}

}  // extern "C"

namespace hpcx {
// In-place sum over the ranks of `count` scalars on stream s (the CG solver's dot products).
hpcx_status allreduce_sum(hpcx_ctx ctx, void *buf, size_t count, int dtype, cudaStream_t s) {
  if (ctx->nranks <= 1) return HPCX_OK;
  if (!ctx->nccl_comm) return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  HPCX_NCCL(ctx, g_nccl.AllReduce(buf, buf, count, nccl_type(dtype), ncclSum, comm_of(ctx), s));
  return HPCX_OK;
}
}  // namespace hpcx

extern "C" {

// Where the fused <x, y> partials of the distributed product go when a CG solver has attached a
// partial array (A->dot_part): [interior blocks | boundary blocks | long rows].
static inline void *dot_slot(hpcx_csr A, int64_t first) {
  return A->dot_part ? static_cast<char *>(A->dot_part) + first * A->es : nullptr;
}

// Transport 1: ONE launch per product on the caller's stream.  Grid = [push CTAs | interior row
// blocks | boundary row blocks]: the push CTAs run first (gather + peer stores + flags), the
// interior blocks overlap the neighbours' pushes, the boundary blocks acquire the flags themselves
// and the last of them acks the source ranks (round 1 needed five launches for the same).
static hpcx_status spmv_dist_peer(hpcx_csr A, char *xfull, void *y_local, cudaStream_t cs, bool boundary_only = false) {
  hpcx_ctx ctx = A->ctx;
  HPCX_RANGE("spmv_dist: one launch (push | interior | boundary + ack)");
  SpmvDistArgs da;
  da.epoch = ++A->epoch;
  da.nseg = A->nseg;
  da.nsrc = A->nsrc;
  // boundary_only (host-operand pipeline): the interior rows run as separate chunked launches
  da.n_int = boundary_only ? 0 : (int)A->h_blocks_int.size();
  da.n_bnd = (int)A->h_blocks_bnd.size();
  int64_t longest = 0;
  for (int q = 0; q < ctx->nranks; ++q) longest = std::max(longest, A->send_counts[q]);
  // about 2048 entries per push CTA, at most two CTAs per SM and segment: a banded operator pushes one
  // grid row or plane with a handful of CTAs, a power-law one its millions of entries with hundreds
  // (16 CTAs for 4.65 M entries kept every other CTA of C5 waiting for a millisecond)
  da.push_per_seg = (int)std::min<int64_t>(std::max<int64_t>((longest + 2047) / 2048, 1), 2 * (int64_t)ctx->sm_count);
  da.npush = A->nseg * da.push_per_seg;
  da.send_idx = A->d_send_idx;
  da.seg_start = A->d_seg_start;
  da.seg_rank = A->d_seg_rank;
  da.seg_dst = A->d_seg_dst;
  da.seg_flag = A->d_seg_flag;
  da.src_rank = A->d_src_rank;
  da.src_ack = A->d_src_ack;
  da.sync = A->d_sync;
  da.pw = peer_wait_of(ctx);
  if (A->nsrc > 0 && da.n_bnd == 0)
    return fail(ctx, HPCX_INVALID_ARG, "halo entries are received but no row is marked as a boundary row");
  hpcx_status rc = spmv_launch_blocks(A, boundary_only ? A->d_blocks_bnd : A->d_blocks_dist,
                                      (int64_t)da.n_int + da.n_bnd, xfull, y_local, cs, dot_slot(A, 0), &da);
  if (rc != HPCX_OK || boundary_only) return rc;
  return spmv_launch_fixup_tables(A, A->d_long_row_dist, A->d_long_off_dist, A->nlong_dist, y_local, cs, xfull,
                                  dot_slot(A, (int64_t)da.n_int + da.n_bnd));
}

// Host operands: x_host / y_host hold this rank's nloc owned entries.  With the fused transport
// the call is pipelined against PCIe like hpcx_spmv with host operands: the last piece of x goes
// up first, then the others; as soon as both ends of the owned block have landed one launch pushes
// the boundary entries to the neighbours and computes the boundary rows, and the interior rows
// follow in chunks whose rows of y travel back while later chunks still compute.  Otherwise:
// copy, product, copy.  Synchronous (the result is in y_host on return).
hpcx_status hpcx_spmv_dist_host(hpcx_csr A, const void *x_host, void *y_host, void *stream) {
  if (!A || !x_host || !y_host) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_spmv_dist_host");
  hpcx_ctx ctx = A->ctx;
  HPCX_REQUIRE(ctx, A->d_xfull != nullptr, "hpcx_csr_set_halo has not been called");
  cudaStream_t cs = stream ? (cudaStream_t)stream : ctx->compute;
  char *xfull = static_cast<char *>(A->d_xfull);
  const size_t nb = (size_t)A->nloc * A->es;
  if (A->transport == 1 && ctx->nranks > 1 && !A->dot_part) {
    hpcx_status st = spmv_dist_host_pipelined(A, x_host, y_host, cs, [&]() {
      return spmv_dist_peer(A, xfull, A->d_y, cs, true);
    });
    if (st == HPCX_OK) return sync_and_check(ctx, cs);
    if (st != HPCX_UNSUPPORTED) return st;
  }
  if (!A->d_y) HPCX_CUDA(ctx, cudaMalloc(&A->d_y, std::max<size_t>(nb, 1)));
  if (nb) HPCX_CUDA(ctx, cudaMemcpyAsync(xfull, x_host, nb, cudaMemcpyHostToDevice, cs));
  if (hpcx_status st = hpcx_spmv_dist(A, xfull, A->d_y, cs)) return st;
  if (nb) HPCX_CUDA(ctx, cudaMemcpyAsync(y_host, A->d_y, nb, cudaMemcpyDeviceToHost, cs));
  return sync_and_check(ctx, cs);
}

hpcx_status hpcx_csr_x_buffer(hpcx_csr A, void **x_full) {
  if (!A || !x_full) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(A->ctx, A->d_xfull != nullptr, "hpcx_csr_set_halo has not been called");
  *x_full = A->d_xfull;
  return HPCX_OK;
}

// Row-block product for matrices whose rows reference most of x anyway (power-law graphs: C5):
// no [owned | halo] renumbering and no per-entry push — every rank's slice of x is gathered into a
// full-length x on each rank, then the block's own kernel — HUB for power-law blocks — runs on the
// local rows with GLOBAL column ids.  SURVEY.md 8(e): "for C5 fall back to ncclAllGather of x".
// The slices of x are the caller's choice (bounds) and need not follow the row blocks: equal slices
// go through ncclAllGather (every rank sends and receives the same amount, all channels busy);
// unequal ones — x split like the nnz-balanced rows — through one grouped send/recv, where the rank
// that owns the longest slice sends it P - 1 times and sets the pace (measured on C5 over 8 GPUs:
// 0.68 ms against 0.17 ms for the block kernel).
hpcx_status hpcx_spmv_allgather(hpcx_csr A, const int64_t *bounds, const void *x_local, void *x_full,
                                void *y_local, void *stream) {
  if (!A || !bounds || !x_local || !x_full || !y_local) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_spmv_allgather");
  hpcx_ctx ctx = A->ctx;
  const int P = ctx->nranks, me = ctx->rank;
  HPCX_REQUIRE(ctx, A->d_xfull == nullptr, "hpcx_spmv_allgather takes a row block with global column ids (no hpcx_csr_set_halo)");
  bool even = true;
  for (int q = 0; q < P; ++q) {
    HPCX_REQUIRE(ctx, bounds[q] <= bounds[q + 1], "bounds must not decrease");
    even = even && bounds[q + 1] - bounds[q] == bounds[1] - bounds[0];
  }
  HPCX_REQUIRE(ctx, bounds[0] == 0 && bounds[P] == A->ncols, "bounds must split the ncols entries of x into one slice per rank");
  if (P > 1 && !ctx->nccl_comm) return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  cudaStream_t cs = stream ? (cudaStream_t)stream : ctx->compute;
  const size_t es = (size_t)A->es;
  const int64_t nmine = bounds[me + 1] - bounds[me];
  char *xf = static_cast<char *>(x_full);
  char *mine = xf + (size_t)bounds[me] * es;
  if (P > 1 && even) {
    if (nmine > 0)  // in place when x_local already is this rank's slice of x_full
      HPCX_NCCL(ctx, g_nccl.AllGather(x_local, x_full, (size_t)nmine, nccl_type(A->dtype), comm_of(ctx), cs));
  } else {
    if (x_local != mine && nmine > 0)
      HPCX_CUDA(ctx, cudaMemcpyAsync(mine, x_local, (size_t)nmine * es, cudaMemcpyDeviceToDevice, cs));
    if (P > 1) {
      const ncclDataType_t ty = nccl_type(A->dtype);
      HPCX_NCCL(ctx, g_nccl.GroupStart());
      for (int d = 1; d < P; ++d) {
        const int to = (me + d) % P, from = (me - d + P) % P;
        if (nmine > 0) HPCX_NCCL(ctx, g_nccl.Send(mine, (size_t)nmine, ty, to, comm_of(ctx), cs));
        if (bounds[from + 1] > bounds[from])
          HPCX_NCCL(ctx, g_nccl.Recv(xf + (size_t)bounds[from] * es, (size_t)(bounds[from + 1] - bounds[from]), ty, from,
                                     comm_of(ctx), cs));
      }
      HPCX_NCCL(ctx, g_nccl.GroupEnd());
    }
  }
  if (hpcx_status st = hpcx_spmv(A, x_full, HPCX_MEM_DEVICE, y_local, HPCX_MEM_DEVICE, cs)) return st;
  if (!stream) return sync_and_check(ctx, cs);
  return HPCX_OK;
}

// compute stream: gather boundary x -> ev ; interior rows ; [wait recv] boundary rows
// comm stream   : [wait ev] grouped send/recv of the x halo segments -> ev_recv
hpcx_status hpcx_spmv_dist(hpcx_csr A, const void *x_local, void *y_local, void *stream) {
  if (!A || !x_local || !y_local) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_spmv_dist");
  hpcx_ctx ctx = A->ctx;
  HPCX_REQUIRE(ctx, A->d_xfull != nullptr, "hpcx_csr_set_halo has not been called");
  if (ctx->nranks > 1 && A->transport == 0 && !ctx->nccl_comm)  // peer stores need no communicator
    return fail(ctx, HPCX_NCCL_ERROR, "hpcx_comm_init has not been called");
  cudaStream_t cs = stream ? (cudaStream_t)stream : ctx->compute;
  cudaStream_t ms = ctx->comm;
  const int P = ctx->nranks;
  char *xfull = static_cast<char *>(A->d_xfull);
  if (x_local != A->d_xfull)
    HPCX_CUDA(ctx, cudaMemcpyAsync(xfull, x_local, (size_t)A->nloc * A->es,
                                   cudaMemcpyDeviceToDevice, cs));
  hpcx_status rc;
  if (A->transport == 1 && P > 1) {
    rc = spmv_dist_peer(A, xfull, y_local, cs);
    if (rc != HPCX_OK) return rc;
    if (!stream) return sync_and_check(ctx, cs);
    return HPCX_OK;
  }
  if (A->nsend) {
    const int blocks = (int)std::min<int64_t>((A->nsend + 255) / 256, (int64_t)ctx->sm_count * 8);
    if (A->dtype == HPCX_F64)
      gather_boundary_x_kernel<double><<<blocks, 256, 0, cs>>>(
          (const double *)xfull, A->d_send_idx, (double *)A->d_sendbuf, A->nsend);
    else
      gather_boundary_x_kernel<float><<<blocks, 256, 0, cs>>>(
          (const float *)xfull, A->d_send_idx, (float *)A->d_sendbuf, A->nsend);
    rc = after_launch(ctx, "gather_boundary_x_kernel");
    if (rc != HPCX_OK) return rc;
  }
  const bool exchange = P > 1 && (A->nsend || A->nhalo);
  if (exchange) {
    HPCX_CUDA(ctx, cudaEventRecord(ctx->ev[0], cs));
    HPCX_CUDA(ctx, cudaStreamWaitEvent(ms, ctx->ev[0], 0));
    const ncclDataType_t ty = nccl_type(A->dtype);
    HPCX_NCCL(ctx, g_nccl.GroupStart());
    int64_t soff = 0, roff = 0;
    for (int q = 0; q < P; ++q) {
      if (A->send_counts[q])
        HPCX_NCCL(ctx, g_nccl.Send((char *)A->d_sendbuf + soff * A->es, (size_t)A->send_counts[q],
                                   ty, q, comm_of(ctx), ms));
      if (A->recv_counts[q])
        HPCX_NCCL(ctx, g_nccl.Recv(xfull + (A->nloc + roff) * A->es, (size_t)A->recv_counts[q], ty,
                                   q, comm_of(ctx), ms));
      soff += A->send_counts[q];
      roff += A->recv_counts[q];
    }
    HPCX_NCCL(ctx, g_nccl.GroupEnd());
    HPCX_CUDA(ctx, cudaEventRecord(ctx->ev[1], ms));
  }
  rc = spmv_launch_blocks(A, A->d_blocks_int, (int64_t)A->h_blocks_int.size(), xfull, y_local, cs,
                          dot_slot(A, 0), nullptr);
  if (rc != HPCX_OK) return rc;
  if (exchange) HPCX_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev[1], 0));
  rc = spmv_launch_blocks(A, A->d_blocks_bnd, (int64_t)A->h_blocks_bnd.size(), xfull, y_local, cs,
                          dot_slot(A, (int64_t)A->h_blocks_int.size()), nullptr);
  if (rc != HPCX_OK) return rc;
  rc = spmv_launch_fixup_tables(A, A->d_long_row_dist, A->d_long_off_dist, A->nlong_dist, y_local, cs, xfull,
                                dot_slot(A, (int64_t)(A->h_blocks_int.size() + A->h_blocks_bnd.size())));
  if (rc != HPCX_OK) return rc;
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(cs));
  return HPCX_OK;
}

}  // extern "C"
