// csr_plan.cu — create-time analysis of a CSR matrix on the device:
// row_ptr validation and narrowing, the row-length histogram, and the row-block table of the
// STREAM kernel (spmv.cu).  Round 1 did all of this in serial host loops over the rows (three
// passes over 67 M rows for configs[1]: 1.1 s); here every step is a kernel over the device copy
// of row_ptr, so creating a matrix costs its upload plus a few milliseconds.
//
// Row blocks without a serial scan.  A greedy cut ("add rows until the tile is full") makes every
// block boundary depend on the previous one.  Instead the NONZEROS are cut at multiples of a
// quantum Q and each cut is snapped to the next row start: quantum k owns the rows whose first
// nonzero lies in [kQ, (k+1)Q), found by two binary searches in row_ptr — independent for every
// k.  Such a run of rows holds fewer than Q + len(last row) nonzeros, so with
//     L = min(longest row, tile/2),  Q = tile - L
// it fits a tile unless its LAST row is longer than L; that row is then peeled off into a block of
// its own (whole-CTA reduction) or, when it exceeds a tile, into tile-sized chunks whose partial
// sums the fix-up kernel adds in order.  For matrices of short rows (configs[1]: 5 per row) L is
// the longest row and every block is filled to within L of the tile.  Runs are further cut where
// the row kind changes (interior / boundary rows of a distributed matrix: the two kinds go to two
// consecutive segments of ONE table, interior first) and every 4096 rows (long runs of empty rows).
// Two passes — count, exclusive scan, write — keep the table in ascending row order.
#include "spmv.cuh"

#include <cub/cub.cuh>

using namespace hpcx;

namespace {

constexpr int kRowCap = 4096;  // rows per block at most

struct ScanOut {
  unsigned long long bins[34];
  unsigned long long nnz_long_rows;  // nonzeros in rows longer than 256 (HPCX_SPMV_AUTO)
  int maxlen;
  int bad;  // 1: row_ptr decreases somewhere, 2: an entry outside [0, nnz]
};

// One pass over row_ptr (int32 or int64 source): monotonicity, narrowing to int32, histogram by
// ceil-log2 bin (bin 0 = empty, bin k = 2^(k-1) <= len < 2^k), longest row.
template <typename I>
__global__ void __launch_bounds__(256)
rowptr_scan_kernel(const I *__restrict__ src, int64_t nrows, int64_t nnz, int32_t *__restrict__ rp32,
                   ScanOut *__restrict__ out) {
  __shared__ unsigned int sbins[34];
  __shared__ int smax, sbad;
  __shared__ unsigned long long slong;
  if (threadIdx.x < 34) sbins[threadIdx.x] = 0;
  if (threadIdx.x == 0) smax = 0, sbad = 0, slong = 0;
  __syncthreads();
  int mymax = 0, bad = 0;
  unsigned long long mylong = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i <= nrows; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t a = (int64_t)src[i];
    if (a < 0 || a > nnz) bad |= 2;
    rp32[i] = (int32_t)a;
    if (i < nrows) {
      const int64_t b = (int64_t)src[i + 1];
      if (b < a) {
        bad |= 1;
      } else {
        int64_t len = b - a;
        if (len > 256) mylong += (unsigned long long)len;
        mymax = max(mymax, len > (int64_t)INT32_MAX ? INT32_MAX : (int)len);
        int bin = 0;
        while (len) ++bin, len >>= 1;
        atomicAdd(&sbins[bin], 1u);
      }
    }
  }
  if (mymax) atomicMax(&smax, mymax);
  if (bad) atomicOr(&sbad, bad);
  if (mylong) atomicAdd(&slong, mylong);
  __syncthreads();
  if (threadIdx.x < 34 && sbins[threadIdx.x]) atomicAdd(&out->bins[threadIdx.x], (unsigned long long)sbins[threadIdx.x]);
  if (threadIdx.x == 0) {
    if (smax) atomicMax(&out->maxlen, smax);
    if (sbad) atomicOr(&out->bad, sbad);
    if (slong) atomicAdd(&out->nnz_long_rows, slong);
  }
}

// Threads per row of a block (row-length binning at block granularity): rows no longer than
// short_row_max take the one-thread-per-row path (lanes = 1, the oracle's summation order);
// otherwise lanes grow with the mean row length and are raised until the block's longest row
// costs a lane at most 64 entries; 0 = the whole CTA works on a single long row.
__device__ __forceinline__ int lanes_for(int64_t nnz_blk, int64_t nrows, int64_t max_len, int short_row_max) {
  if (nrows <= 0) return 1;
  const int64_t avg = (nnz_blk + nrows - 1) / nrows;
  if (nrows == 1 && nnz_blk > 256) return 0;
  if (max_len <= short_row_max) return 1;
  int lanes = avg <= 16 ? 2 : avg <= 32 ? 4 : avg <= 64 ? 8 : avg <= 128 ? 16 : 32;
  while (lanes < 32 && max_len > 64 * (int64_t)lanes) lanes <<= 1;
  return lanes;
}

__device__ __forceinline__ int64_t lower_bound_rp(const int32_t *__restrict__ rp, int64_t nrows, int64_t v) {
  int64_t lo = 0, hi = nrows;  // first i in [0, nrows] with rp[i] >= v; nrows when there is none below it
  while (lo < hi) {
    const int64_t mid = lo + (hi - lo) / 2;
    if ((int64_t)rp[mid] >= v) hi = mid; else lo = mid + 1;
  }
  return lo;
}

struct PlanParams {
  const int32_t *rp;
  const uint8_t *kind;  // may be null: every row is kind 0
  int64_t nrows, nnz, nq;
  int tile, Q, L, short_row_max;
};

// WRITE = false: counts per quantum (blocks of kind 0 / kind 1, chunk blocks, long rows).
// WRITE = true : emits the blocks at the offsets the scans produced.
template <bool WRITE>
__global__ void __launch_bounds__(128)
rowblock_plan_kernel(PlanParams P, int32_t *__restrict__ c0, int32_t *__restrict__ c1, int32_t *__restrict__ cch,
                     int32_t *__restrict__ clg, int64_t n0_total, RowBlock *__restrict__ blocks,
                     int32_t *__restrict__ long_row, int32_t *__restrict__ long_off) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= P.nq) return;
  const int64_t r0 = lower_bound_rp(P.rp, P.nrows, k * (int64_t)P.Q);
  const int64_t r1 = lower_bound_rp(P.rp, P.nrows, (k + 1) * (int64_t)P.Q);
  int n[2] = {0, 0}, nch = 0, nlg = 0;
  int64_t o[2] = {0, 0}, och = 0, olg = 0;
  if (WRITE) o[0] = c0[k], o[1] = n0_total + c1[k], och = cch[k], olg = clg[k];
  auto emit = [&](int kd, const RowBlock &b) {
    if (WRITE) blocks[o[kd] + n[kd]] = b;
    ++n[kd];
  };
  int64_t rs = r1;  // rows [r0, rs) are blocked together; a long last row is peeled off
  bool peel = false;
  if (r1 > r0 && (int64_t)P.rp[r1] - P.rp[r1 - 1] > P.L) rs = r1 - 1, peel = true;
  // runs of equal kind, at most kRowCap rows each
  int64_t a = r0;
  while (a < rs) {
    const int kd = P.kind ? P.kind[a] : 0;
    int64_t b = a, maxlen = 0;
    while (b < rs && b - a < kRowCap && (P.kind ? P.kind[b] : 0) == kd) {
      maxlen = max(maxlen, (int64_t)P.rp[b + 1] - P.rp[b]);
      ++b;
    }
    const int64_t p0 = P.rp[a], p1 = P.rp[b];
    emit(kd, RowBlock{(int32_t)a, (int32_t)(b - a), -1, lanes_for(p1 - p0, b - a, maxlen, P.short_row_max), p0, p1});
    a = b;
  }
  if (peel) {
    const int64_t r = r1 - 1, p0 = P.rp[r], p1 = P.rp[r + 1];
    const int kd = P.kind ? P.kind[r] : 0;
    if (p1 - p0 <= P.tile) {
      emit(kd, RowBlock{(int32_t)r, 1, -1, lanes_for(p1 - p0, 1, p1 - p0, P.short_row_max), p0, p1});
    } else {
      if (WRITE) long_row[olg] = (int32_t)r, long_off[olg] = (int32_t)och;
      ++nlg;
      for (int64_t p = p0; p < p1; p += P.tile) {
        emit(kd, RowBlock{(int32_t)r, 1, (int32_t)(och + nch), 0, p, min(p + (int64_t)P.tile, p1)});
        ++nch;
      }
    }
  }
  if (!WRITE) c0[k] = n[0], c1[k] = n[1], cch[k] = nch, clg[k] = nlg;
}

struct Scratch {
  void *p = nullptr;
  ~Scratch() { cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

}  // namespace

namespace hpcx {

// row_ptr -> A->d_row_ptr (int32), A->bins, longest row.  src lives in `mem`.
hpcx_status csr_scan_rowptr(hpcx_csr A, const void *src, int idx_bytes, int mem, int64_t *maxlen,
                            int64_t *nnz_long_rows) {
  hpcx_ctx ctx = A->ctx;
  cudaStream_t s = ctx->compute;
  const int64_t n = A->nrows;
  Scratch raw, out;
  const void *dsrc = src;
  if (mem == HPCX_MEM_HOST) {
    HPCX_CUDA(ctx, raw.alloc((size_t)(n + 1) * idx_bytes));
    HPCX_CUDA(ctx, cudaMemcpyAsync(raw.p, src, (size_t)(n + 1) * idx_bytes, cudaMemcpyHostToDevice, s));
    dsrc = raw.p;
  }
  HPCX_CUDA(ctx, out.alloc(sizeof(ScanOut)));
  HPCX_CUDA(ctx, cudaMemsetAsync(out.p, 0, sizeof(ScanOut), s));
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 256) / 256, (int64_t)ctx->sm_count * 8));
  if (idx_bytes == 4)
    rowptr_scan_kernel<int32_t><<<grid, 256, 0, s>>>((const int32_t *)dsrc, n, A->nnz, A->d_row_ptr, (ScanOut *)out.p);
  else
    rowptr_scan_kernel<int64_t><<<grid, 256, 0, s>>>((const int64_t *)dsrc, n, A->nnz, A->d_row_ptr, (ScanOut *)out.p);
  if (hpcx_status st = after_launch(ctx, "rowptr_scan_kernel")) return st;
  ScanOut h;
  int32_t ends[2] = {0, 0};
  HPCX_CUDA(ctx, cudaMemcpyAsync(&h, out.p, sizeof h, cudaMemcpyDeviceToHost, s));
  HPCX_CUDA(ctx, cudaMemcpyAsync(&ends[0], A->d_row_ptr, 4, cudaMemcpyDeviceToHost, s));
  HPCX_CUDA(ctx, cudaMemcpyAsync(&ends[1], A->d_row_ptr + n, 4, cudaMemcpyDeviceToHost, s));
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  if (h.bad & 2) return fail(ctx, HPCX_INVALID_ARG, "row_ptr entry outside [0, nnz]");
  if (ends[0] != 0 || (int64_t)ends[1] != A->nnz)
    return fail(ctx, HPCX_INVALID_ARG, "row_ptr[0] != 0 or row_ptr[nrows] != nnz");
  if (h.bad & 1) return fail(ctx, HPCX_INVALID_ARG, "row_ptr is not non-decreasing");
  for (int b = 0; b < 34; ++b) A->bins[b] = (int64_t)h.bins[b];
  *maxlen = h.maxlen;
  *nnz_long_rows = (int64_t)h.nnz_long_rows;
  return HPCX_OK;
}

// Row-block table over the device row_ptr.  kind (device, may be null) splits the table into
// [kind 0 blocks | kind 1 blocks]; *n0 receives the size of the first segment.  The table, the
// long-row list and the chunk offsets are freshly allocated device arrays (cudaFree by the caller);
// `host` receives a copy of the table.
hpcx_status csr_plan_rowblocks(hpcx_csr A, const uint8_t *d_kind, int64_t maxlen, RowBlock **d_blocks,
                               std::vector<RowBlock> *host, int64_t *n0, int32_t **d_long_row,
                               int32_t **d_long_off, int64_t *nlong, int64_t *nchunks) {
  hpcx_ctx ctx = A->ctx;
  cudaStream_t s = ctx->compute;
  PlanParams P;
  P.rp = A->d_row_ptr;
  P.kind = d_kind;
  P.nrows = A->nrows;
  P.nnz = A->nnz;
  P.tile = A->tile;
  P.L = (int)std::min<int64_t>(std::max<int64_t>(maxlen, 1), A->tile / 2);
  P.Q = A->tile - P.L;
  P.nq = A->nnz / P.Q + 1;
  P.short_row_max = A->short_row_max;
  *d_blocks = nullptr;
  *d_long_row = *d_long_off = nullptr;
  Scratch c[4], tmp;
  const size_t nq = (size_t)P.nq;
  for (auto &x : c) HPCX_CUDA(ctx, x.alloc((nq + 1) * 4));
  // one spare entry so that an exclusive scan over nq + 1 entries leaves the totals at [nq]
  for (auto &x : c) HPCX_CUDA(ctx, cudaMemsetAsync((int32_t *)x.p + nq, 0, 4, s));
  const unsigned grid = (unsigned)((P.nq + 127) / 128);
  rowblock_plan_kernel<false><<<grid, 128, 0, s>>>(P, (int32_t *)c[0].p, (int32_t *)c[1].p, (int32_t *)c[2].p,
                                                   (int32_t *)c[3].p, 0, nullptr, nullptr, nullptr);
  if (hpcx_status st = after_launch(ctx, "rowblock_plan_kernel<count>")) return st;
  size_t tb = 0;
  HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, (int32_t *)c[0].p, (int32_t *)c[0].p, (int)(nq + 1), s));
  HPCX_CUDA(ctx, tmp.alloc(tb));
  int32_t tot[4];
  for (int i = 0; i < 4; ++i) {
    HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, tb, (int32_t *)c[i].p, (int32_t *)c[i].p, (int)(nq + 1), s));
    HPCX_CUDA(ctx, cudaMemcpyAsync(&tot[i], (int32_t *)c[i].p + nq, 4, cudaMemcpyDeviceToHost, s));
  }
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  const int64_t nb = (int64_t)tot[0] + tot[1];
  *n0 = tot[0];
  *nchunks = tot[2];
  *nlong = tot[3];
  HPCX_CUDA(ctx, cudaMalloc((void **)d_blocks, std::max<size_t>((size_t)nb, 1) * sizeof(RowBlock)));
  HPCX_CUDA(ctx, cudaMalloc((void **)d_long_row, std::max<size_t>((size_t)tot[3], 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc((void **)d_long_off, (size_t)(tot[3] + 1) * 4));
  HPCX_CUDA(ctx, cudaMemcpyAsync(*d_long_off + tot[3], &tot[2], 4, cudaMemcpyHostToDevice, s));
  rowblock_plan_kernel<true><<<grid, 128, 0, s>>>(P, (int32_t *)c[0].p, (int32_t *)c[1].p, (int32_t *)c[2].p,
                                                  (int32_t *)c[3].p, tot[0], *d_blocks, *d_long_row, *d_long_off);
  if (hpcx_status st = after_launch(ctx, "rowblock_plan_kernel<write>")) return st;
  if (host) {
    host->resize((size_t)nb);
    if (nb) HPCX_CUDA(ctx, cudaMemcpyAsync(host->data(), *d_blocks, (size_t)nb * sizeof(RowBlock), cudaMemcpyDeviceToHost, s));
  }
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}

}  // namespace hpcx
