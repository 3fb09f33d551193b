// multi.cu — single-process, multi-GPU driver ("multi-GPU driven from ONE host
// thread").  Every caller the reference has is a single-process OpenMP program that indexes its
// containers from its own loops (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:356); the only
// multi-stream idiom it contains runs from one thread too (lib/gpu/transfer_overlap_testsuite/src/
// cuda_tut_transfer_overlap.cu:92-103).  Such a program reaches N GPUs through this file without
// adopting MPI or torchrun: hpcx_init_multi makes one context per device inside the process and
// enables peer access between them; a host CSR matrix or a host grid is partitioned here (the same
// partitioner and localiser as the one-process-per-GPU path, partition.cpp), and the products /
// sweeps run through the SAME kernels and the same fused peer-store transport as dist.cu — the
// neighbours' buffers are plain peer pointers instead of CUDA IPC windows, and nothing needs NCCL.
#include <algorithm>

#include "dist.cuh"

using namespace hpcx;

struct hpcx_multi_s {
  std::vector<hpcx_ctx> ctx;
  std::string last_error;
};

struct hpcx_multi_csr_s {
  hpcx_multi m = nullptr;
  std::vector<hpcx_csr> A;
  std::vector<int64_t> bounds;
  int es = 8;
};

struct hpcx_multi_grid_s {
  hpcx_multi m = nullptr;
  hpcx_grid_desc desc{};
  std::vector<hpcx_grid> g;
  std::vector<int64_t> bounds;   // split of dim[0]
  int64_t plane_host = 0;        // elements of one padded slice of axis 0 in the host layout
  // cached plan of the last Jacobi sweep (per rank), valid for (partner, variant)
  std::vector<hpcx_halo_plan> plans;
  hpcx_multi_grid_s *plan_partner = nullptr;
  int plan_variant = -1, plan_zsplit = -1;
};

namespace {

struct DeviceGuard {
  int prev = 0;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    cudaSetDevice(dev);
  }
  ~DeviceGuard() { cudaSetDevice(prev); }
};

hpcx_status mfail(hpcx_multi m, hpcx_status st, const std::string &msg) {
  if (m) m->last_error = msg;
  return st;
}
// status of a per-rank call -> status of the multi call, message carried over
hpcx_status lift(hpcx_multi m, int rank, hpcx_status st) {
  if (st != HPCX_OK) m->last_error = "rank " + std::to_string(rank) + ": " + m->ctx[rank]->last_error;
  return st;
}
#define HPCX_MTRY(m, rank, expr)                          \
  do {                                                    \
    hpcx_status st__ = lift((m), (rank), (expr));         \
    if (st__ != HPCX_OK) return st__;                     \
  } while (0)

hpcx_status sync_all(hpcx_multi m) {
  hpcx_status out = HPCX_OK;
  for (size_t i = 0; i < m->ctx.size(); ++i) {
    DeviceGuard g(m->ctx[i]->device);
    hpcx_status st = lift(m, (int)i, hpcx_synchronize(m->ctx[i]));
    if (st != HPCX_OK && out == HPCX_OK) out = st;
  }
  return out;
}

void free_plans(hpcx_multi_grid G) {
  for (size_t i = 0; i < G->plans.size(); ++i)
    if (G->plans[i]) {
      DeviceGuard g(G->m->ctx[i]->device);
      hpcx_halo_plan p = G->plans[i];
      cudaDeviceSynchronize();
      cudaFree(p->flags);  // peer pointers are not IPC windows here: nothing to close
      p->destroy_events();
      for (auto &e : p->graphs) cudaGraphExecDestroy(e.exec);
      delete p;
    }
  G->plans.clear();
  G->plan_partner = nullptr;
  G->plan_variant = -1;
  G->plan_zsplit = -1;
}

// Copies the padded slices [s0, s1) of axis 0 between a rank's device grid and the global host
// array (host layout of the reference container: dense, slice = plane_host elements).
hpcx_status copy_slices(hpcx_grid g, char *host_slab, int64_t s0, int64_t s1, bool to_device, cudaStream_t s) {
  if (s1 <= s0) return HPCX_OK;
  hpcx_ctx ctx = g->ctx;
  if (g->desc.ndim == 1) return fail(ctx, HPCX_UNSUPPORTED, "1-D grids are not split");
  // device rows per slice of the descriptor's axis 0: a z plane of a 3-D grid, one row of a 2-D grid
  const int64_t rows = g->desc.ndim == 3 ? g->ext[1] : 1;
  const size_t width = (size_t)g->ext[2] * g->es;
  char *dev = (char *)g->dev + ((size_t)(s0 * rows) * g->pitch + (g->xoff - g->glo[2])) * g->es;
  char *host = host_slab + (size_t)(s0 * rows) * width;
  const size_t nrows = (size_t)((s1 - s0) * rows);
  if (to_device)
    HPCX_CUDA(ctx, cudaMemcpy2DAsync(dev, (size_t)g->pitch * g->es, host, width, width, nrows, cudaMemcpyHostToDevice, s));
  else
    HPCX_CUDA(ctx, cudaMemcpy2DAsync(host, width, dev, (size_t)g->pitch * g->es, width, nrows, cudaMemcpyDeviceToHost, s));
  return HPCX_OK;
}

}  // namespace

extern "C" {

hpcx_status hpcx_init_multi(int32_t ndev, const int32_t *dev_ids, hpcx_multi *out) {
  if (!out || ndev < 1) return HPCX_INVALID_ARG;
  *out = nullptr;
  hpcx_multi m = new (std::nothrow) hpcx_multi_s();
  if (!m) return HPCX_OOM;
  *out = m;  // handed back even on failure so that the caller can read the message
  int prev = 0;
  cudaGetDevice(&prev);
  for (int i = 0; i < ndev; ++i) {
    const int dev = dev_ids ? dev_ids[i] : i;
    for (int j = 0; j < i; ++j)
      if (m->ctx[j]->device == dev) return mfail(m, HPCX_INVALID_ARG, "a device is listed twice");
    hpcx_ctx c = nullptr;
    hpcx_status st = hpcx_init(dev, &c);
    if (c) m->ctx.push_back(c);
    if (st != HPCX_OK) {
      m->last_error = c ? c->last_error : "hpcx_init failed";
      cudaSetDevice(prev);
      return st;
    }
    c->nranks = ndev;
    c->rank = i;
    // One host thread queues every device's kernels.  If a launch fails half way through a round
    // the kernels already queued wait for neighbours that will never signal, and the thread's next
    // synchronise would hang with them: here (unlike the one-process-per-GPU case, where a late
    // rank is legitimate) the flag waits get a bound unless the caller chose one.
    if (c->halo_timeout_ns == 0) c->halo_timeout_ns = 120ull * 1000000000ull;
    // one thread launches every device's kernels, and kernels of different devices wait on each
    // other's flags: no kernel may be loaded lazily in the middle of such a sequence
    cudaSetDevice(dev);
    st = preload_stencil_kernels(c);
    if (st == HPCX_OK) st = preload_dist_kernels(c);
    if (st == HPCX_OK) st = preload_spmv_kernels(c);
    if (st != HPCX_OK) {
      m->last_error = c->last_error;
      cudaSetDevice(prev);
      return st;
    }
  }
  // every device may load from and store to every other one (the fused transport's peer stores)
  for (int i = 0; i < ndev; ++i)
    for (int j = 0; j < ndev; ++j) {
      if (i == j) continue;
      cudaSetDevice(m->ctx[i]->device);
      int can = 0;
      cudaDeviceCanAccessPeer(&can, m->ctx[i]->device, m->ctx[j]->device);
      if (!can) {
        cudaSetDevice(prev);
        return mfail(m, HPCX_UNSUPPORTED, "devices " + std::to_string(m->ctx[i]->device) + " and " +
                                              std::to_string(m->ctx[j]->device) + " cannot access each other's memory");
      }
      cudaError_t e = cudaDeviceEnablePeerAccess(m->ctx[j]->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        cudaSetDevice(prev);
        return mfail(m, HPCX_CUDA_ERROR, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
      }
      (void)cudaGetLastError();
    }
  cudaSetDevice(prev);
  return HPCX_OK;
}

hpcx_status hpcx_multi_finalize(hpcx_multi m) {
  if (!m) return HPCX_INVALID_ARG;
  for (hpcx_ctx c : m->ctx) {
    DeviceGuard g(c->device);
    hpcx_finalize(c);
  }
  delete m;
  return HPCX_OK;
}

int32_t hpcx_multi_size(hpcx_multi m) { return m ? (int32_t)m->ctx.size() : 0; }
hpcx_ctx hpcx_multi_ctx(hpcx_multi m, int32_t rank) {
  return m && rank >= 0 && rank < (int32_t)m->ctx.size() ? m->ctx[rank] : nullptr;
}
const char *hpcx_multi_last_error_string(hpcx_multi m) { return m ? m->last_error.c_str() : "null multi-context"; }

// ------------------------------------------------------------------ CSR SpMV over all devices
hpcx_status hpcx_multi_csr_create(hpcx_multi m, const hpcx_csr_desc *d, const hpcx_csr_opts *opts,
                                  int32_t partition_mode, hpcx_multi_csr *out) {
  if (!m || !d || !out) return HPCX_INVALID_ARG;
  *out = nullptr;
  if (d->mem != HPCX_MEM_HOST && d->mem != HPCX_MEM_MANAGED)
    return mfail(m, HPCX_INVALID_ARG, "hpcx_multi_csr_create partitions a HOST (or managed) matrix");
  if (d->nrows != d->ncols) return mfail(m, HPCX_INVALID_ARG, "the row partition needs a square matrix (x split like y)");
  if (d->idx_bytes != 4 && d->idx_bytes != 8) return mfail(m, HPCX_INVALID_ARG, "idx_bytes must be 4 or 8");
  const int P = (int)m->ctx.size();
  if (P > 64) return mfail(m, HPCX_UNSUPPORTED, "the peer-store transport supports up to 64 devices (sync-block slots)");
  const int es = (int)dtype_size(d->dtype);
  hpcx_multi_csr M = new (std::nothrow) hpcx_multi_csr_s();
  if (!M) return mfail(m, HPCX_OOM, "host allocation failed");
  M->m = m;
  M->es = es;
  M->A.assign(P, nullptr);
  M->bounds.assign(P + 1, 0);
  auto bail = [&](hpcx_status st) {
    hpcx_multi_csr_destroy(M);
    return st;
  };
  hpcx_status st = hpcx_partition_rows(d->nrows, d->row_ptr, d->idx_bytes, P, partition_mode, M->bounds.data());
  if (st != HPCX_OK) return bail(mfail(m, st, "hpcx_partition_rows failed"));
  auto rp_at = [&](int64_t i) {
    return d->idx_bytes == 4 ? (int64_t)((const int32_t *)d->row_ptr)[i] : ((const int64_t *)d->row_ptr)[i];
  };
  for (int r = 0; r < P; ++r) {
    hpcx_ctx ctx = m->ctx[r];
    DeviceGuard guard(ctx->device);
    hpcx_halo_sizes sz{};
    st = hpcx_csr_localize(d->nrows, d->row_ptr, d->idx_bytes, d->col_idx, M->bounds.data(), P, r, &sz, nullptr, nullptr,
                           nullptr, nullptr, nullptr, nullptr, nullptr);
    if (st != HPCX_OK) return bail(mfail(m, st, "hpcx_csr_localize failed"));
    std::vector<int64_t> lrp(sz.nloc + 1), halo(std::max<int64_t>(sz.nhalo, 1)), rc(P), sc(P);
    std::vector<int32_t> lcol(std::max<int64_t>(sz.nnz_loc, 1)), sidx(std::max<int64_t>(sz.nsend, 1));
    std::vector<uint8_t> kind(std::max<int64_t>(sz.nloc, 1));
    st = hpcx_csr_localize(d->nrows, d->row_ptr, d->idx_bytes, d->col_idx, M->bounds.data(), P, r, &sz, lrp.data(),
                           lcol.data(), halo.data(), rc.data(), sc.data(), sidx.data(), kind.data());
    if (st != HPCX_OK) return bail(mfail(m, st, "hpcx_csr_localize failed"));
    hpcx_csr_desc ld{};
    ld.nrows = sz.nloc;
    ld.ncols = sz.nloc + sz.nhalo;
    ld.nnz = sz.nnz_loc;
    ld.row_ptr = lrp.data();
    ld.idx_bytes = 8;
    ld.dtype = d->dtype;
    ld.col_idx = lcol.data();
    ld.values = (const char *)d->values + (size_t)rp_at(M->bounds[r]) * es;
    ld.mem = HPCX_MEM_HOST;
    hpcx_csr_opts o{};
    if (opts) o = *opts;
    if (o.kernel == HPCX_SPMV_AUTO) o.kernel = HPCX_SPMV_STREAM;  // the distributed product is the STREAM kernel's
    st = lift(m, r, hpcx_csr_create(ctx, &ld, &o, &M->A[r]));
    if (st != HPCX_OK) return bail(st);
    st = lift(m, r, hpcx_csr_set_halo(M->A[r], sz.nloc, sz.nhalo, rc.data(), sc.data(), sidx.data(), kind.data()));
    if (st != HPCX_OK) return bail(st);
    if (cudaError_t e = cudaMalloc(&M->A[r]->d_y, std::max<size_t>((size_t)sz.nloc * es, 1)); e != cudaSuccess)
      return bail(lift(m, r, fail(ctx, e == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR,
                                  std::string("hpcx_multi_csr_create: y block: ") + cudaGetErrorString(e))));
  }
  if (P > 1) {
    // fused transport: every rank's sync block, then the tables from plain peer pointers
    for (int r = 0; r < P; ++r) {
      DeviceGuard guard(m->ctx[r]->device);
      st = lift(m, r, csr_peer_prepare(M->A[r]));
      if (st != HPCX_OK) return bail(st);
    }
    for (int r = 0; r < P; ++r) {
      DeviceGuard guard(m->ctx[r]->device);
      std::vector<CsrPeerView> views(P);
      for (int q = 0; q < P; ++q) {
        if (q == r) continue;
        hpcx_csr Q = M->A[q];
        views[q].xfull = (char *)Q->d_xfull;
        views[q].sync = Q->d_sync;
        views[q].nloc = Q->nloc;
        int64_t off = 0;
        for (int p = 0; p < r; ++p) off += Q->recv_counts[p];
        views[q].recv_off_me = off;
        views[q].recv_cnt_me = Q->recv_counts[r];
      }
      st = lift(m, r, csr_peer_wire(M->A[r], views));
      if (st != HPCX_OK) return bail(st);
    }
    st = sync_all(m);
    if (st != HPCX_OK) return bail(st);
  }
  *out = M;
  return HPCX_OK;
}

hpcx_status hpcx_multi_csr_destroy(hpcx_multi_csr M) {
  if (!M) return HPCX_INVALID_ARG;
  sync_all(M->m);
  for (size_t r = 0; r < M->A.size(); ++r)
    if (M->A[r]) {
      DeviceGuard g(M->m->ctx[r]->device);
      hpcx_csr_destroy(M->A[r]);
    }
  delete M;
  return HPCX_OK;
}

hpcx_status hpcx_multi_csr_bounds(hpcx_multi_csr M, int64_t *bounds) {
  if (!M || !bounds) return HPCX_INVALID_ARG;
  std::copy(M->bounds.begin(), M->bounds.end(), bounds);
  return HPCX_OK;
}

// y = A*x, x and y HOST arrays of the global size.  One host thread: the x blocks go up, every
// device's one-launch product (push | interior | boundary + ack) is queued, the y blocks come
// back; the devices run concurrently and trade their boundary entries by peer stores.
hpcx_status hpcx_multi_spmv(hpcx_multi_csr M, const void *x_host, void *y_host) {
  if (!M || !x_host || !y_host) return HPCX_INVALID_ARG;
  hpcx_multi m = M->m;
  HPCX_RANGE("hpcx_multi_spmv");
  const int P = (int)M->A.size();
  for (int r = 0; r < P; ++r) {
    hpcx_csr A = M->A[r];
    DeviceGuard g(m->ctx[r]->device);
    const size_t off = (size_t)M->bounds[r] * M->es, nb = (size_t)A->nloc * M->es;
    if (nb) HPCX_CUDA(A->ctx, cudaMemcpyAsync(A->d_xfull, (const char *)x_host + off, nb, cudaMemcpyHostToDevice, A->ctx->compute));
  }
  // every device's product is queued BEFORE any result is fetched: a device-to-host copy into
  // pageable memory blocks this thread until that device is done, and it is not done before its
  // neighbours — whose kernels this same thread has yet to launch — have pushed their entries
  for (int r = 0; r < P; ++r) {
    hpcx_csr A = M->A[r];
    DeviceGuard g(m->ctx[r]->device);
    HPCX_MTRY(m, r, hpcx_spmv_dist(A, A->d_xfull, A->d_y, A->ctx->compute));
  }
  for (int r = 0; r < P; ++r) {
    hpcx_csr A = M->A[r];
    DeviceGuard g(m->ctx[r]->device);
    const size_t off = (size_t)M->bounds[r] * M->es, nb = (size_t)A->nloc * M->es;
    if (nb) HPCX_CUDA(A->ctx, cudaMemcpyAsync((char *)y_host + off, A->d_y, nb, cudaMemcpyDeviceToHost, A->ctx->compute));
  }
  return sync_all(m);
}

// ------------------------------------------------------------------ grids split along axis 0
hpcx_status hpcx_multi_grid_create(hpcx_multi m, const hpcx_grid_desc *d, hpcx_multi_grid *out) {
  if (!m || !d || !out) return HPCX_INVALID_ARG;
  *out = nullptr;
  if (d->ndim < 2 || d->ndim > 3) return mfail(m, HPCX_UNSUPPORTED, "grids of 2 or 3 dimensions are split along their slowest axis");
  const int P = (int)m->ctx.size();
  if (d->dim[0] < P) return mfail(m, HPCX_INVALID_ARG, "fewer slices along the slowest axis than devices");
  hpcx_multi_grid G = new (std::nothrow) hpcx_multi_grid_s();
  if (!G) return mfail(m, HPCX_OOM, "host allocation failed");
  G->m = m;
  G->desc = *d;
  G->g.assign(P, nullptr);
  G->bounds.assign(P + 1, 0);
  // even split (every device gets at least one slice; the static chunking of hpcx_partition_slabs
  // can leave trailing parts empty, which a slab transport cannot have)
  for (int r = 0; r <= P; ++r) G->bounds[r] = d->dim[0] * r / P;
  G->plane_host = 1;
  for (int i = 1; i < d->ndim; ++i) G->plane_host *= d->dim[i] + d->lo + d->hi;
  for (int r = 0; r < P; ++r) {
    DeviceGuard guard(m->ctx[r]->device);
    hpcx_grid_desc ld = *d;
    ld.dim[0] = G->bounds[r + 1] - G->bounds[r];
    hpcx_status st = lift(m, r, hpcx_grid_create(m->ctx[r], &ld, &G->g[r]));
    if (st != HPCX_OK) {
      hpcx_multi_grid_destroy(G);
      return st;
    }
  }
  *out = G;
  return HPCX_OK;
}

hpcx_status hpcx_multi_grid_destroy(hpcx_multi_grid G) {
  if (!G) return HPCX_INVALID_ARG;
  sync_all(G->m);
  free_plans(G);
  for (size_t r = 0; r < G->g.size(); ++r)
    if (G->g[r]) {
      DeviceGuard g(G->m->ctx[r]->device);
      hpcx_grid_destroy(G->g[r]);
    }
  delete G;
  return HPCX_OK;
}

// Every slab takes its own slices AND its ghost slices from the global array: a slab's ghosts are
// its neighbours' boundary slices, so the devices start consistent without any exchange.
hpcx_status hpcx_multi_grid_upload(hpcx_multi_grid G, const void *host_flat) {
  if (!G || !host_flat) return HPCX_INVALID_ARG;
  hpcx_multi m = G->m;
  const int es = (int)dtype_size(G->desc.dtype);
  for (size_t r = 0; r < G->g.size(); ++r) {
    DeviceGuard guard(m->ctx[r]->device);
    char *slab = (char *)const_cast<void *>(host_flat) + (size_t)G->bounds[r] * G->plane_host * es;
    const int64_t n = G->bounds[r + 1] - G->bounds[r] + G->desc.lo + G->desc.hi;
    HPCX_MTRY(m, (int)r, copy_slices(G->g[r], slab, 0, n, true, m->ctx[r]->compute));
  }
  return sync_all(m);
}

// Owned slices only (the outermost ghost slices are Dirichlet values the host already holds).
hpcx_status hpcx_multi_grid_download(hpcx_multi_grid G, void *host_flat) {
  if (!G || !host_flat) return HPCX_INVALID_ARG;
  hpcx_multi m = G->m;
  const int es = (int)dtype_size(G->desc.dtype);
  for (size_t r = 0; r < G->g.size(); ++r) {
    DeviceGuard guard(m->ctx[r]->device);
    char *slab = (char *)host_flat + (size_t)G->bounds[r] * G->plane_host * es;
    const int64_t lo = G->desc.lo, n = G->bounds[r + 1] - G->bounds[r];
    HPCX_MTRY(m, (int)r, copy_slices(G->g[r], slab, lo, lo + n, false, m->ctx[r]->compute));
  }
  return sync_all(m);
}

hpcx_status hpcx_multi_stencil_sweep(hpcx_multi_grid in, hpcx_multi_grid out, const hpcx_stencil_desc *st, int32_t steps) {
  if (!in || !out || !st) return HPCX_INVALID_ARG;
  hpcx_multi m = in->m;
  if (out->m != m || in == out) return mfail(m, HPCX_INVALID_ARG, "in and out must be two grids of one multi-context");
  HPCX_RANGE("hpcx_multi_stencil_sweep");
  const int P = (int)m->ctx.size();
  if (st->kind == HPCX_STENCIL_BOX_X || st->kind == HPCX_STENCIL_BOX_Y) {
    // one sweep, input ghosts uploaded with the slab: the slabs are independent
    for (int r = 0; r < P; ++r) {
      DeviceGuard guard(m->ctx[r]->device);
      HPCX_MTRY(m, r, hpcx_stencil_sweep(in->g[r], out->g[r], st, steps, m->ctx[r]->compute));
    }
    return sync_all(m);
  }
  if (st->kind != HPCX_STENCIL_JACOBI7) return mfail(m, HPCX_INVALID_ARG, "unknown stencil kind");
  if (P == 1) {
    DeviceGuard guard(m->ctx[0]->device);
    HPCX_MTRY(m, 0, hpcx_stencil_sweep(in->g[0], out->g[0], st, steps, nullptr));
    return HPCX_OK;
  }
  if (steps < 1) return mfail(m, HPCX_INVALID_ARG, "steps must be >= 1");
  if (in->desc.ndim != 3 || out->desc.ndim != 3 || in->desc.dtype != HPCX_F32 || out->desc.dtype != HPCX_F32)
    return mfail(m, HPCX_UNSUPPORTED, "the multi-device Jacobi sweep is the fp32 3-D TMA kernel");
  for (int i = 0; i < 3; ++i)
    if (in->desc.dim[i] != out->desc.dim[i]) return mfail(m, HPCX_INVALID_ARG, "grids differ in logical extent");
  if (in->desc.lo != out->desc.lo || in->desc.hi != out->desc.hi || in->desc.lo < 1 || in->desc.hi < 1)
    return mfail(m, HPCX_INVALID_ARG, "JACOBI7 needs identical ghost widths >= 1 on both grids");
  const bool temporal = st->variant == HPCX_VARIANT_TEMPORAL;
  const int64_t w = temporal && steps >= 2 ? 2 : 1;  // ghost planes the first pass reads
  if (temporal)
    for (int r = 0; r < P; ++r)
      if (in->desc.lo < 2 || in->desc.hi < 2 || in->g[r]->n[0] < 2)
        return mfail(m, HPCX_INVALID_ARG, "TEMPORAL over slabs needs ghost widths >= 2 and >= 2 planes on every device");
  // ---- plans: one per rank, wired with plain peer pointers; cached while (in, out, variant) repeat
  bool reuse = in->plan_partner == out && in->plan_variant == st->variant && in->plan_zsplit == st->zsplit &&
               (int)in->plans.size() == P;
  for (int r = 0; reuse && r < P; ++r) {
    hpcx_halo_plan p = in->plans[r];
    const bool fwd = in->g[r]->dev == p->buf[0] && out->g[r]->dev == p->buf[1];
    const bool bwd = in->g[r]->dev == p->buf[1] && out->g[r]->dev == p->buf[0];
    reuse = fwd || bwd;
  }
  if (!reuse) {
    free_plans(in);
    in->plans.assign(P, nullptr);
    for (int r = 0; r < P; ++r) {
      DeviceGuard guard(m->ctx[r]->device);
      hpcx_halo_plan p = new (std::nothrow) hpcx_halo_plan_s();
      if (!p) return mfail(m, HPCX_OOM, "host allocation failed");
      in->plans[r] = p;
      p->ctx = m->ctx[r];
      p->in = in->g[r];
      p->out = out->g[r];
      p->st = *st;
      p->transport = 1;
      p->create_events();
      HPCX_MTRY(m, r, plan_prepare(p));
    }
    for (int r = 0; r < P; ++r)
      for (int side = 0; side < 2; ++side) {
        const int q = r + (side ? 1 : -1);
        if (q < 0 || q >= P) continue;
        hpcx_halo_plan p = in->plans[r], o = in->plans[q];
        if (o->in->n[1] != p->in->n[1] || o->in->n[2] != p->in->n[2] || o->in->pitch != p->in->pitch)
          return mfail(m, HPCX_INVALID_ARG, "neighbouring slabs differ in their (y, x) extents");
        p->peer_buf[side][0] = static_cast<float *>(o->buf[0]);
        p->peer_buf[side][1] = static_cast<float *>(o->buf[1]);
        p->peer_flags[side] = o->flags;
        p->peer_n0[side] = o->in->n[0];
      }
    in->plan_partner = out;
    in->plan_variant = st->variant;
    in->plan_zsplit = st->zsplit;
  }
  // ---- ghosts of the source iterate: w boundary planes of every slab into its neighbours (peer copies)
  if (hpcx_status rc = sync_all(m)) return rc;
  for (int r = 0; r < P; ++r) {
    hpcx_grid g = in->g[r];
    const size_t plane = (size_t)g->ext[1] * g->pitch * g->es;
    auto at = [&](hpcx_grid h, int64_t z) { return (char *)h->dev + (size_t)(z + h->glo[0]) * plane; };
    if (r + 1 < P) {  // my top w planes -> the upper neighbour's low ghost planes
      hpcx_grid h = in->g[r + 1];
      HPCX_CUDA(m->ctx[r], cudaMemcpyPeer(at(h, -w), h->ctx->device, at(g, g->n[0] - w), g->ctx->device, (size_t)w * plane));
    }
    if (r > 0) {  // my bottom w planes -> the lower neighbour's high ghost planes
      hpcx_grid h = in->g[r - 1];
      HPCX_CUDA(m->ctx[r], cudaMemcpyPeer(at(h, h->n[0]), h->ctx->device, at(g, 0), g->ctx->device, (size_t)w * plane));
    }
  }
  if (hpcx_status rc = sync_all(m)) return rc;
  // ---- the sweeps: every rank's launches are queued from this one thread; the kernels of
  // neighbouring devices synchronise among themselves through the flags.  Long runs go out in
  // chunks of 64 sweeps per device, round robin: a device's launch queue is finite, and a thread
  // blocked on device 0's full queue could never launch the device-1 kernels device 0 waits for.
  // A chunk sweeps in -> out, so the roles of the two grids alternate from chunk to chunk.
  constexpr int kChunk = 64;
  bool flipped = false;  // true: the newest iterate sits in the `in` handles
  for (int done = 0; done < steps; done += kChunk) {
    const int k = std::min(kChunk, steps - done);
    for (int r = 0; r < P; ++r) {
      DeviceGuard guard(m->ctx[r]->device);
      hpcx_halo_plan p = in->plans[r];
      p->in = flipped ? out->g[r] : in->g[r];
      p->out = flipped ? in->g[r] : out->g[r];
      HPCX_MTRY(m, r, stencil_peer_sweep(p, k, m->ctx[r]->compute, false));
    }
    if (done + k < steps) flipped = !flipped;  // the next chunk starts from this chunk's newest iterate
  }
  if (hpcx_status rc = sync_all(m)) return rc;
  if (flipped)  // hand the newest iterate to `out`, as hpcx_stencil_sweep does
    for (int r = 0; r < P; ++r) std::swap(in->g[r]->dev, out->g[r]->dev);
  return HPCX_OK;
}

}  // extern "C"
