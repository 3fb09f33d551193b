// spmv.cu — CSR sparse matrix-vector product y = A*x for sm_100a.
//
// No reference implementation exists; the arithmetic contract is
// the CPU restatement in oracle/hpcx_oracle.c (hpcx_oracle_spmv_*): per row, products
// values[p]*x[col_idx[p]] rounded separately and added in ascending p.
//
// STREAM kernel (default).  The matrix is cut at create time into row blocks whose
// nonzeros fit one shared-memory tile.  One CTA per block:
//   1. one thread issues two 1-D bulk copies (TMA engine, cp.async.bulk -> SASS UBLKCP) that
//      bring the block's contiguous values[] and col_idx[] ranges into shared memory and
//      complete on an mbarrier — perfectly coalesced, no registers held while in flight;
//      several CTAs are resident per SM, so copies of one overlap the arithmetic of others;
//   2. all threads gather x[col] through the read-only path (ld.global.nc) and turn the
//      staged values into products in place;
//   3. rows are reduced out of shared memory with a per-block choice of 1, 2..32 lanes or
//      the whole CTA per row (row-length binning at block granularity); with 1 lane per row
//      the summation order is exactly the oracle's, so short-row matrices are bit-exact.
// Rows longer than a tile are split into tile-sized chunks whose partial sums are added in
// chunk order by a small fix-up kernel (deterministic, no atomics).
#include "spmv.cuh"

#include <algorithm>
#include <functional>

using namespace hpcx;

namespace hpcx {
void csr_close_peers(hpcx_csr A);  // dist.cu
// csr_plan.cu
hpcx_status csr_scan_rowptr(hpcx_csr A, const void *src, int idx_bytes, int mem, int64_t *maxlen,
                            int64_t *nnz_long_rows);
hpcx_status csr_plan_rowblocks(hpcx_csr A, const uint8_t *d_kind, int64_t maxlen, RowBlock **d_blocks,
                               std::vector<RowBlock> *host, int64_t *n0, int32_t **d_long_row,
                               int32_t **d_long_off, int64_t *nlong, int64_t *nchunks);
}

namespace {

constexpr int kThreads = 256;

template <typename T>
__device__ __forceinline__ T ldg_nc(const T *p) {
  return __ldg(p);
}

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

template <typename T>
__device__ __forceinline__ T shfl_xor(T v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}

// Dynamic shared memory layout: [values TILE+8][cols TILE+8][mbarrier][warp sums].
template <typename T, int TILE>
struct StreamSmem {
  static constexpr int kVal = (TILE + 8) * (int)sizeof(T);
  static constexpr int kCol = (TILE + 8) * 4;
  static constexpr int kBar = kVal + kCol;
  static constexpr int kWsum = kBar + 16;
  static constexpr int kBytes = kWsum + (kThreads / 32) * (int)sizeof(T);
};

constexpr int kSyncData = 0, kSyncAck = 64, kSyncCount = 128, kSyncBnd = 129;  // words of the sync block (dist.cu)

// x gather: the read-only path for owned entries; boundary row blocks of the one-launch distributed
// product (L2X) read halo slots a peer GPU wrote DURING this kernel, so they go to L2
// (ld.global.cg) — an L1 line fetched earlier by an interior CTA of the same SM could be stale.
template <typename T, bool L2X>
__device__ __forceinline__ T gather_x(const T *__restrict__ x, int c) {
  if constexpr (L2X) return __ldcg(x + c);
  else return __ldg(x + c);
}

// One row block: stage the tile, gather, reduce the rows.
// DOT: additionally emits dot_part[block] = sum over the block's rows of y_i * x_i (square
// matrices: the fused q = A*p, <p, q> of the CG step, cg.cu), reduced in a fixed order.
template <typename T, int TILE, bool DOT, bool L2X>
__device__ __forceinline__ void stream_block(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                             const T *__restrict__ val, const T *__restrict__ x, T *__restrict__ y,
                                             const RowBlock *__restrict__ blocks, T *__restrict__ partials,
                                             T *__restrict__ dot_part, int bidx, unsigned char *smem) {
  using L = StreamSmem<T, TILE>;
  T *sval = reinterpret_cast<T *>(smem);
  int32_t *scol = reinterpret_cast<int32_t *>(smem + L::kVal);
  uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L::kBar);
  T *wsum = reinterpret_cast<T *>(smem + L::kWsum);
  const int tid = threadIdx.x;
  // Block descriptor: 32 bytes, same for every thread (two broadcast 16-byte loads).
  const int4 d0 = __ldg(reinterpret_cast<const int4 *>(blocks + bidx));
  const longlong2 d1 = __ldg(reinterpret_cast<const longlong2 *>(blocks + bidx) + 1);
  const int row0 = d0.x, nrows = d0.y, part = d0.z, lanes = d0.w;
  const int64_t p0 = d1.x, p1 = d1.y;
  const int64_t p0a = p0 & ~int64_t(3);          // 16-byte aligned start in col_idx[]
  const int lead = (int)(p0 - p0a);              // staged entries before the block's first
  const int cnt = (int)(p1 - p0a);               // staged entries up to the block's last
  const int cnt_a = (cnt + 3) & ~3;

  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (tid == 0 && cnt_a > 0) {
    mbar_expect_tx(bar, (uint32_t)cnt_a * (uint32_t)(sizeof(T) + 4));
    bulk_g2s(sval, val + p0a, (uint32_t)cnt_a * (uint32_t)sizeof(T), bar);
    bulk_g2s(scol, col + p0a, (uint32_t)cnt_a * 4u, bar);
  }
  if (cnt_a > 0) mbar_wait(bar, 0);

  T dot = T(0);
  (void)dot;
  if (lanes == 1) {
    // Short rows: one thread per row straight out of the staged tile, ascending p — the oracle's
    // order, so the result is bit-identical.  A warp's 32 lanes hold 32 consecutive rows, so for
    // banded matrices the k-th gathers of a warp fall into one or two 128-byte lines of x
    // (columns i+off_k for consecutive i), and the staged col/val reads are stride-(row length)
    // accesses to shared memory (conflict-free for odd lengths such as 5 and 27).  No product is
    // written back and no second barrier is needed.
    for (int r = tid; r < nrows; r += kThreads) {
      const int a = (int)((int64_t)row_ptr[row0 + r] - p0a);
      const int e = (int)((int64_t)row_ptr[row0 + r + 1] - p0a);
      T s = T(0);
      for (int i = a; i < e; i += 4) {
        T xv[4], vv[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool on = i + k < e;
          xv[k] = on ? gather_x<T, L2X>(x, scol[i + k]) : T(0);
          vv[k] = on ? sval[i + k] : T(0);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (i + k < e) s = add_rn(s, mul_rn(vv[k], xv[k]));
      }
      y[row0 + r] = s;
      if constexpr (DOT) dot = add_rn(dot, mul_rn(s, ldg_nc(x + row0 + r)));
    }
    if constexpr (!DOT) return;
  } else {

  // Products in place. Four gathers in flight per thread per round.
  for (int base = lead + tid; base < cnt; base += 4 * kThreads) {
    T xv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = base + k * kThreads;
      xv[k] = i < cnt ? gather_x<T, L2X>(x, scol[i]) : T(0);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = base + k * kThreads;
      if (i < cnt) sval[i] = mul_rn(sval[i], xv[k]);
    }
  }
  __syncthreads();

  if (lanes > 1) {
    // `lanes` threads per row: strided partial sums, xor-shuffle tree.
    const int g = tid / lanes, l = tid % lanes;
    const int groups = kThreads / lanes;
    for (int rb = 0; rb < nrows; rb += groups) {
      const int r = rb + g;
      T s = T(0);
      if (r < nrows) {
        const int a = (int)((int64_t)row_ptr[row0 + r] - p0a);
        const int e = (int)((int64_t)row_ptr[row0 + r + 1] - p0a);
        for (int i = a + l; i < e; i += lanes) s = add_rn(s, sval[i]);
      }
      for (int m = lanes >> 1; m > 0; m >>= 1) s = add_rn(s, shfl_xor(s, m));
      if (r < nrows && l == 0) {
        y[row0 + r] = s;
        if constexpr (DOT) dot = add_rn(dot, mul_rn(s, ldg_nc(x + row0 + r)));
      }
    }
  } else {
    // Whole CTA on one row (or one chunk of a long row).
    T s = T(0);
    for (int i = lead + tid; i < cnt; i += kThreads) s = add_rn(s, sval[i]);
    for (int m = 16; m > 0; m >>= 1) s = add_rn(s, shfl_xor(s, m));
    if ((tid & 31) == 0) wsum[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
      T t = wsum[0];
#pragma unroll
      for (int w = 1; w < kThreads / 32; ++w) t = add_rn(t, wsum[w]);
      if (part >= 0) {
        partials[part] = t;  // the row's <y, x> term is added by the fix-up kernel
      } else {
        y[row0] = t;
        if constexpr (DOT) dot = mul_rn(t, ldg_nc(x + row0));
      }
    }
    if constexpr (DOT) __syncthreads();  // wsum is reused below
  }
  }
  if constexpr (DOT) {
    for (int m = 16; m > 0; m >>= 1) dot = add_rn(dot, shfl_xor(dot, m));
    if ((tid & 31) == 0) wsum[tid >> 5] = dot;
    __syncthreads();
    if (tid == 0) {
      T t = wsum[0];
#pragma unroll
      for (int w = 1; w < kThreads / 32; ++w) t = add_rn(t, wsum[w]);
      dot_part[bidx] = t;
    }
  }
}

// DIST: the one-launch distributed product (SpmvDistArgs, spmv.cuh): grid = [push CTAs | interior
// row blocks | boundary row blocks].
template <typename T, int TILE, bool DOT, bool DIST>
__global__ void __launch_bounds__(kThreads, 8)  // 32 registers: 8 CTAs per SM, as many as the tiles allow
spmv_stream_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                   const T *__restrict__ val, const T *__restrict__ x, T *__restrict__ y,
                   const RowBlock *__restrict__ blocks, T *__restrict__ partials,
                   T *__restrict__ dot_part, const __grid_constant__ SpmvDistArgs da) {
  extern __shared__ __align__(128) unsigned char smem[];
  if constexpr (!DIST) {
    stream_block<T, TILE, DOT, false>(row_ptr, col, val, x, y, blocks, partials, dot_part, blockIdx.x, smem);
  } else {
    const int tid = threadIdx.x;
    if ((int)blockIdx.x < da.npush) {
      // ---- push role: segment s of the send list goes into rank seg_rank[s]'s halo slots
      const int sgm = blockIdx.x / da.push_per_seg, j = blockIdx.x % da.push_per_seg;
      // the destination must have finished reading the previous product's halo (its ack)
      if (tid == 0) wait_flag(da.sync + kSyncAck + da.seg_rank[sgm], da.epoch - 1, da.pw);
      __syncthreads();
      const int a = da.seg_start[sgm], e = da.seg_start[sgm + 1];
      T *dst = static_cast<T *>(da.seg_dst[sgm]);
      for (int i = a + j * kThreads + tid; i < e; i += da.push_per_seg * kThreads) dst[i - a] = __ldg(x + da.send_idx[i]);
      __threadfence_system();
      __syncthreads();
      if (tid == 0 && (atomicAdd(da.sync + kSyncCount, 1u) + 1) % (uint32_t)da.npush == 0)
        for (int k = 0; k < da.nseg; ++k) st_release_sys(da.seg_flag[k], da.epoch);
      return;
    }
    const int bidx = blockIdx.x - da.npush;
    if (bidx < da.n_int) {  // interior rows: owned x only, exactly the single-GPU block
      stream_block<T, TILE, DOT, false>(row_ptr, col, val, x, y, blocks, partials, dot_part, bidx, smem);
      return;
    }
    // boundary rows: every source rank's entries of this product must have landed
    if (tid < da.nsrc) wait_flag(da.sync + kSyncData + da.src_rank[tid], da.epoch, da.pw);
    __syncthreads();
    stream_block<T, TILE, DOT, true>(row_ptr, col, val, x, y, blocks, partials, dot_part, bidx, smem);
    // the last boundary block to finish tells the source ranks that their entries have been read
    __syncthreads();
    if (tid == 0 && (atomicAdd(da.sync + kSyncBnd, 1u) + 1) % (uint32_t)da.n_bnd == 0)
      for (int k = 0; k < da.nsrc; ++k) st_release_sys(da.src_ack[k], da.epoch);
  }
}

// Adds the chunk partial sums of every long row in chunk order.
// dot_part (may be null): dot_part[j] = y_row * x_row of long row j.
template <typename T>
__global__ void spmv_longrow_fixup_kernel(const int32_t *__restrict__ long_row,
                                          const int32_t *__restrict__ long_off,
                                          const T *__restrict__ partials, T *__restrict__ y,
                                          int nlong, const T *__restrict__ x,
                                          T *__restrict__ dot_part) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nlong) return;
  T s = T(0);
  for (int k = long_off[j]; k < long_off[j + 1]; ++k) s = add_rn(s, partials[k]);
  y[long_row[j]] = s;
  if (dot_part) dot_part[j] = mul_rn(s, x[long_row[j]]);
}

// Classic vector-per-row kernel: G lanes per row straight from global memory.
// Kept as the A/B baseline for the STREAM kernel (ncu: sectors/request, achieved GB/s).
template <typename T, int G>
__global__ void __launch_bounds__(kThreads)
spmv_vector_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                   const T *__restrict__ val, const T *__restrict__ x, T *__restrict__ y,
                   int64_t nrows) {
  const int64_t t = blockIdx.x * (int64_t)kThreads + threadIdx.x;
  const int64_t row = t / G;
  const int l = (int)(t % G);
  T s = T(0);
  if (row < nrows) {
    const int a = row_ptr[row], e = row_ptr[row + 1];
    for (int i = a + l; i < e; i += G) s = add_rn(s, mul_rn(__ldg(val + i), ldg_nc(x + __ldg(col + i))));
  }
#pragma unroll
  for (int m = G >> 1; m > 0; m >>= 1) s = add_rn(s, shfl_xor(s, m));
  if (row < nrows && l == 0) y[row] = s;
}

// SELL-C-sigma: one thread per row of a slice, slices stored column-major (C consecutive
// rows interleaved) so a warp reads C consecutive values per step.
template <typename T>
__global__ void __launch_bounds__(kThreads)
spmv_sell_kernel(const int64_t *__restrict__ slice_ptr, const int32_t *__restrict__ slice_len,
                 const int32_t *__restrict__ perm, const int32_t *__restrict__ col,
                 const T *__restrict__ val, const T *__restrict__ x, T *__restrict__ y,
                 int64_t nrows, int C) {
  const int64_t t = blockIdx.x * (int64_t)kThreads + threadIdx.x;
  const int64_t slice = t / C;
  const int lane = (int)(t % C);
  const int64_t srow = slice * C + lane;  // position in the sorted order
  if (srow >= nrows) return;
  const int64_t base = slice_ptr[slice];
  const int len = slice_len[slice];
  T s = T(0);
  for (int k = 0; k < len; ++k) {
    const int64_t i = base + (int64_t)k * C + lane;
    s = add_rn(s, mul_rn(__ldg(val + i), ldg_nc(x + __ldg(col + i))));
  }
  y[perm[srow]] = s;
}

template <typename T, int TILE, bool DOT, bool DIST>
hpcx_status launch_stream_t(hpcx_csr A, const RowBlock *d_blocks, int64_t nblocks, const T *x, T *y,
                            T *dot_part, cudaStream_t s, const SpmvDistArgs &da) {
  const int64_t grid = nblocks + (DIST ? da.npush : 0);
  if (grid == 0) return HPCX_OK;
  using L = StreamSmem<T, TILE>;
  if (hpcx_status st = ensure_dyn_smem(A->ctx, (const void *)spmv_stream_kernel<T, TILE, DOT, DIST>, L::kBytes)) return st;
  spmv_stream_kernel<T, TILE, DOT, DIST><<<(unsigned)grid, kThreads, L::kBytes, s>>>(
      A->d_row_ptr, A->d_col, (const T *)A->d_val, x, y, d_blocks, (T *)A->d_partials, dot_part, da);
  return after_launch(A->ctx, "spmv_stream_kernel");
}

template <typename T, bool DOT, bool DIST>
hpcx_status launch_stream_tile(hpcx_csr A, const RowBlock *d_blocks, int64_t nblocks, const T *x, T *y, T *dot_part,
                               cudaStream_t s, const SpmvDistArgs &da) {
  switch (A->tile) {
    case 1024: return launch_stream_t<T, 1024, DOT, DIST>(A, d_blocks, nblocks, x, y, dot_part, s, da);
    case 2048: return launch_stream_t<T, 2048, DOT, DIST>(A, d_blocks, nblocks, x, y, dot_part, s, da);
    case 4096: return launch_stream_t<T, 4096, DOT, DIST>(A, d_blocks, nblocks, x, y, dot_part, s, da);
    default: return fail(A->ctx, HPCX_INVALID_ARG, "tile_nnz must be 1024, 2048 or 4096");
  }
}

// da != nullptr: the one-launch distributed product (push CTAs + flag waits + ack in the kernel)
template <typename T>
hpcx_status launch_stream(hpcx_csr A, const RowBlock *d_blocks, int64_t nblocks, const T *x, T *y,
                          cudaStream_t s, T *dot_part = nullptr, const SpmvDistArgs *da = nullptr) {
  static const SpmvDistArgs none{};
  if (da)
    return dot_part ? launch_stream_tile<T, true, true>(A, d_blocks, nblocks, x, y, dot_part, s, *da)
                    : launch_stream_tile<T, false, true>(A, d_blocks, nblocks, x, y, nullptr, s, *da);
  return dot_part ? launch_stream_tile<T, true, false>(A, d_blocks, nblocks, x, y, dot_part, s, none)
                  : launch_stream_tile<T, false, false>(A, d_blocks, nblocks, x, y, nullptr, s, none);
}

template <typename T>
hpcx_status launch_fixup(hpcx_csr A, const int32_t *long_row, const int32_t *long_off,
                         int64_t nlong, T *y, cudaStream_t s, const T *x = nullptr,
                         T *dot_part = nullptr) {
  if (nlong == 0) return HPCX_OK;
  spmv_longrow_fixup_kernel<T><<<(unsigned)((nlong + 127) / 128), 128, 0, s>>>(
      long_row, long_off, (const T *)A->d_partials, y, (int)nlong, x, dot_part);
  return after_launch(A->ctx, "spmv_longrow_fixup_kernel");
}

template <typename T>
hpcx_status launch_vector(hpcx_csr A, const T *x, T *y, cudaStream_t s) {
  int G = A->lanes_per_row;
  if (G <= 0) {
    const int64_t avg = A->nrows ? (A->nnz + A->nrows - 1) / A->nrows : 1;
    G = avg <= 2 ? 1 : avg <= 4 ? 2 : avg <= 8 ? 4 : avg <= 16 ? 8 : avg <= 32 ? 16 : 32;
  }
  const int64_t threads = A->nrows * G;
  const unsigned blocks = (unsigned)((threads + kThreads - 1) / kThreads);
  if (blocks == 0) return HPCX_OK;
#define HPCX_VEC(GG)                                                                       \
  case GG:                                                                                 \
    spmv_vector_kernel<T, GG><<<blocks, kThreads, 0, s>>>(A->d_row_ptr, A->d_col,          \
                                                          (const T *)A->d_val, x, y, A->nrows); \
    break;
  switch (G) {
    HPCX_VEC(1) HPCX_VEC(2) HPCX_VEC(4) HPCX_VEC(8) HPCX_VEC(16) HPCX_VEC(32)
    default: return fail(A->ctx, HPCX_INVALID_ARG, "lanes_per_row must be a power of two <= 32");
  }
#undef HPCX_VEC
  return after_launch(A->ctx, "spmv_vector_kernel");
}

template <typename T>
hpcx_status launch_sell(hpcx_csr A, const T *x, T *y, cudaStream_t s) {
  const int64_t threads = A->sell_nslices * A->sell_c;
  const unsigned blocks = (unsigned)((threads + kThreads - 1) / kThreads);
  if (blocks == 0) return HPCX_OK;
  spmv_sell_kernel<T><<<blocks, kThreads, 0, s>>>(A->d_slice_ptr, A->d_slice_len, A->d_perm,
                                                  A->d_sell_col, (const T *)A->d_sell_val, x, y,
                                                  A->nrows, A->sell_c);
  return after_launch(A->ctx, "spmv_sell_kernel");
}

template <typename T>
hpcx_status spmv_device(hpcx_csr A, const T *x, T *y, cudaStream_t s) {
  hpcx_status st;
  switch (A->kernel) {
    case HPCX_SPMV_STREAM:
      // A->dot_part (set by the CG solver): [row blocks | long rows] partials of <x, y>
      st = launch_stream<T>(A, A->d_blocks, (int64_t)A->h_blocks.size(), x, y, s, (T *)A->dot_part);
      if (st == HPCX_OK)
        st = launch_fixup<T>(A, A->d_long_row, A->d_long_off, A->nlong, y, s, x,
                             A->dot_part ? (T *)A->dot_part + A->h_blocks.size() : nullptr);
      return st;
    case HPCX_SPMV_HUB:
      if (A->dot_part)  // the fused <x, y> of the CG step lives in the STREAM kernel (its tables always exist)
        return fail(A->ctx, HPCX_UNSUPPORTED, "fused dot product needs the STREAM kernel");
      return hub_launch(A, x, y, s);
    case HPCX_SPMV_VECTOR: return launch_vector<T>(A, x, y, s);
    case HPCX_SPMV_SELL: return launch_sell<T>(A, x, y, s);
    default: return fail(A->ctx, HPCX_INVALID_ARG, "unknown SpMV kernel");
  }
}

void free_csr(hpcx_csr A) {
  hpcx::csr_close_peers(A);
  hpcx::hub_free(A);
  cudaFree(A->d_row_ptr);
  cudaFree(A->d_col);
  cudaFree(A->d_val);
  cudaFree(A->d_blocks);
  cudaFree(A->d_long_row);
  cudaFree(A->d_long_off);
  cudaFree(A->d_partials);
  cudaFree(A->d_perm);
  cudaFree(A->d_slice_len);
  cudaFree(A->d_sell_col);
  cudaFree(A->d_slice_ptr);
  cudaFree(A->d_sell_val);
  cudaFree(A->d_x);
  cudaFree(A->d_y);
  for (cudaEvent_t e : A->pipe.ev) cudaEventDestroy(e);
  for (cudaEvent_t e : A->pipe_dist.ev) cudaEventDestroy(e);
  cudaFree(A->d_send_idx);
  cudaFree(A->d_sendbuf);
  cudaFree(A->d_xfull);
  cudaFree(A->d_blocks_dist);
  cudaFree(A->d_long_row_dist);
  cudaFree(A->d_long_off_dist);
  delete A;
}

}  // namespace

namespace hpcx {
// see preload_stencil_kernels (stencil.cu): the kernels of the distributed product, all tiles
template <typename T, int TILE>
static hpcx_status preload_stream(hpcx_ctx ctx) {
  using L = StreamSmem<T, TILE>;
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)spmv_stream_kernel<T, TILE, false, true>, L::kBytes)) return st;
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)spmv_stream_kernel<T, TILE, true, true>, L::kBytes)) return st;
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)spmv_stream_kernel<T, TILE, false, false>, L::kBytes)) return st;
  return ensure_dyn_smem(ctx, (const void *)spmv_stream_kernel<T, TILE, true, false>, L::kBytes);
}
hpcx_status preload_spmv_kernels(hpcx_ctx ctx) {
  cudaFuncAttributes a;
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)spmv_longrow_fixup_kernel<float>));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)spmv_longrow_fixup_kernel<double>));
  if (hpcx_status st = preload_stream<float, 1024>(ctx)) return st;
  if (hpcx_status st = preload_stream<float, 2048>(ctx)) return st;
  if (hpcx_status st = preload_stream<float, 4096>(ctx)) return st;
  if (hpcx_status st = preload_stream<double, 1024>(ctx)) return st;
  if (hpcx_status st = preload_stream<double, 2048>(ctx)) return st;
  return preload_stream<double, 4096>(ctx);
}

// Shared with dist.cu.
hpcx_status spmv_launch_blocks(hpcx_csr A, const RowBlock *d_blocks, int64_t nblocks, const void *x,
                               void *y, cudaStream_t s, void *dot_part, const SpmvDistArgs *da) {
  if (A->dtype == HPCX_F64)
    return launch_stream<double>(A, d_blocks, nblocks, (const double *)x, (double *)y, s,
                                 (double *)dot_part, da);
  return launch_stream<float>(A, d_blocks, nblocks, (const float *)x, (float *)y, s, (float *)dot_part, da);
}
hpcx_status spmv_launch_fixup_tables(hpcx_csr A, const int32_t *long_row, const int32_t *long_off,
                                     int64_t nlong, void *y, cudaStream_t s, const void *x,
                                     void *dot_part) {
  if (A->dtype == HPCX_F64)
    return launch_fixup<double>(A, long_row, long_off, nlong, (double *)y, s, (const double *)x,
                                (double *)dot_part);
  return launch_fixup<float>(A, long_row, long_off, nlong, (float *)y, s, (const float *)x,
                             (float *)dot_part);
}
}  // namespace hpcx

namespace {
hpcx_status pipe_analyse(hpcx_csr A, hpcx::PipePlan &P, const std::vector<RowBlock> &blocks, int64_t nlong, int64_t nx);
hpcx_status pipe_dist_pieces(hpcx_csr A, hpcx::PipePlan &P);
hpcx_status spmv_host_pipelined(hpcx_csr A, hpcx::PipePlan &P, const RowBlock *d_blocks,
                                const std::vector<RowBlock> &blocks, int64_t nx, void *d_x, const void *x, void *y,
                                cudaStream_t s, const std::function<hpcx_status()> &dist);
}
namespace hpcx {
// Host-operand pipeline of the distributed product (dist.cu: hpcx_spmv_dist_host).  Returns
// HPCX_UNSUPPORTED when the matrix does not qualify (long rows, too few interior blocks).
hpcx_status spmv_dist_host_pipelined(hpcx_csr A, const void *x, void *y, cudaStream_t s,
                                     const std::function<hpcx_status()> &push_and_boundary) {
  if (A->pipe_dist.state == 0) {
    if (hpcx_status st = pipe_analyse(A, A->pipe_dist, A->h_blocks_int, A->nlong_dist, A->nloc)) return st;
    if (A->pipe_dist.state == 1)
      if (hpcx_status st = pipe_dist_pieces(A, A->pipe_dist)) {
        A->pipe_dist.state = 0;
        return st;
      }
  }
  if (A->pipe_dist.state != 1) return HPCX_UNSUPPORTED;
  if (!A->d_y) HPCX_CUDA(A->ctx, cudaMalloc(&A->d_y, std::max<size_t>((size_t)A->nrows * A->es, 1)));
  return spmv_host_pipelined(A, A->pipe_dist, A->d_blocks_int, A->h_blocks_int, A->nloc, A->d_xfull, x, y, s,
                             push_and_boundary);
}
}  // namespace hpcx
namespace {

// Largest column index referenced by each chunk of nonzeros [pbeg[c], pbeg[c+1]).
__global__ void __launch_bounds__(256)
chunk_colmax_kernel(const int32_t *__restrict__ col, const int64_t *__restrict__ pbeg, int nchunks,
                    int32_t *__restrict__ cmax, int32_t limit) {
  const int c = blockIdx.y;
  if (c >= nchunks) return;
  int32_t m = -1;
  // columns >= limit are halo slots of a distributed matrix: they arrive from the neighbours, not
  // with this rank's x, so they do not hold a chunk back
  for (int64_t i = pbeg[c] + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < pbeg[c + 1];
       i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t v = col[i];
    if (v < limit) m = max(m, v);
  }
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m >= 0) atomicMax(cmax + c, m);
}

// Smallest and largest column index of the matrix (create-time validation: a column outside
// [0, ncols) would become an out-of-bounds gather of x in every kernel).
__global__ void __launch_bounds__(256)
col_range_kernel(const int32_t *__restrict__ col, int64_t nnz, int32_t *__restrict__ minmax) {
  int32_t lo = INT32_MAX, hi = INT32_MIN;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t c = col[i];
    lo = min(lo, c);
    hi = max(hi, c);
  }
  for (int o = 16; o; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(minmax + 0, lo);
    atomicMax(minmax + 1, hi);
  }
}

// Piece j of x is entries [nx * j / K, nx * (j + 1) / K).  Which piece holds entry v?
__device__ __forceinline__ int piece_of(int32_t v, int64_t nx, int K) {
  int j = (int)(((int64_t)v * K) / nx);
  while (j + 1 < K && nx * (j + 1) / K <= v) ++j;
  while (j > 0 && nx * j / K > v) --j;
  return j;
}
// Pieces of the owned x (entries below nx) referenced by the nonzeros of a set of row blocks / by
// an index list: bit j of *mask = piece j.  K <= 32.
__global__ void __launch_bounds__(256)
piece_mask_blocks_kernel(const int32_t *__restrict__ col, const RowBlock *__restrict__ blocks, int64_t nblocks,
                         int32_t nx, int K, uint32_t *__restrict__ mask) {
  uint32_t m = 0;
  for (int64_t b = blockIdx.x; b < nblocks; b += gridDim.x)
    for (int64_t i = blocks[b].p0 + threadIdx.x; i < blocks[b].p1; i += blockDim.x) {
      const int32_t v = col[i];
      if (v < nx) m |= 1u << piece_of(v, nx, K);
    }
  m = __reduce_or_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicOr(mask, m);
}
__global__ void __launch_bounds__(256)
piece_mask_list_kernel(const int32_t *__restrict__ idx, int64_t n, int32_t nx, int K, uint32_t *__restrict__ mask) {
  uint32_t m = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    m |= 1u << piece_of(idx[i], nx, K);
  m = __reduce_or_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m) atomicOr(mask, m);
}

// 32 chunks: the last y chunk leaves about two chunk times after the last x piece has landed
// ((K + 2) / K of one transfer) and every chunk costs ~30 us of copy and launch set-up; measured on
// C2 (copies alone 11.0 ms): 32 chunks 12.9 ms, 64 chunks 13.5 ms, 128 chunks 15.0 ms.
constexpr int kPipeChunks = 32;

// One-time analysis for host-x / host-y products: cut the row blocks into kPipeChunks chunks and
// find the largest column each chunk reads.  A chunk can run as soon as x[0 .. cmax] has landed, so
// banded operators (the C2 / C4 stencil matrices) stream: H2D of x, the product and D2H of y overlap
// and the call takes about one PCIe transfer instead of two.  Matrices whose every chunk reads up
// to the last column degrade gracefully to copy-then-compute with the D2H still chunked.
// `blocks`: the table the chunks are cut from (all row blocks, or the interior blocks of a
// distributed matrix); nx: entries of x that travel with the call (ncols, or nloc).
hpcx_status pipe_analyse(hpcx_csr A, hpcx::PipePlan &P, const std::vector<RowBlock> &blocks, int64_t nlong, int64_t nx) {
  hpcx_ctx ctx = A->ctx;
  P.state = -1;
  const int64_t nb = (int64_t)blocks.size();
  const int K = kPipeChunks;
  if (nlong > 0 || nb < 2 * K || nx < K) return HPCX_OK;
  P.K = K;
  P.first.resize(K + 1);
  std::vector<int64_t> pbeg(K + 1);
  for (int c = 0; c <= K; ++c) {
    P.first[c] = nb * c / K;
    // a chunk's nonzeros: from its first block's to the next chunk's first block's (rows of the other
    // kind in between are covered too, which only makes the bound safer)
    pbeg[c] = c < K ? blocks[P.first[c]].p0 : blocks[nb - 1].p1;
  }
  int64_t *d_pbeg = nullptr;
  int32_t *d_cmax = nullptr;
  HPCX_CUDA(ctx, cudaMalloc((void **)&d_pbeg, (K + 1) * sizeof(int64_t)));
  cudaError_t e = cudaMalloc((void **)&d_cmax, K * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemsetAsync(d_cmax, 0xff, K * sizeof(int32_t), ctx->compute);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_pbeg, pbeg.data(), (K + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->compute);
  if (e == cudaSuccess) {
    chunk_colmax_kernel<<<dim3(64, K), 256, 0, ctx->compute>>>(A->d_col, d_pbeg, K, d_cmax, (int32_t)nx);
    ctx->launches++;
    P.cmax.resize(K);
    e = cudaMemcpyAsync(P.cmax.data(), d_cmax, K * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->compute);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->compute);
  }
  cudaFree(d_pbeg);
  cudaFree(d_cmax);
  if (e != cudaSuccess) return fail(ctx, HPCX_CUDA_ERROR, std::string("pipe_analyse: ") + cudaGetErrorString(e));
  P.ev.resize(2 * K + 2);
  for (auto &ev : P.ev) HPCX_CUDA(ctx, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  P.state = 1;
  return HPCX_OK;
}

// Distributed plan: which pieces of the owned x does the launch that pushes this rank's boundary
// entries and computes its boundary rows read?  Those pieces go up first and that launch waits
// for all of them — a boundary row or a send entry anywhere in the block is covered, not only
// those at its two ends.
hpcx_status pipe_dist_pieces(hpcx_csr A, hpcx::PipePlan &P) {
  static_assert(kPipeChunks <= 32, "the piece mask is one 32-bit word");
  hpcx_ctx ctx = A->ctx;
  uint32_t *d_mask = nullptr;
  HPCX_CUDA(ctx, cudaMalloc((void **)&d_mask, sizeof(uint32_t)));
  cudaError_t e = cudaMemsetAsync(d_mask, 0, sizeof(uint32_t), ctx->compute);
  const int64_t nbnd = (int64_t)A->h_blocks_bnd.size();
  if (e == cudaSuccess && nbnd > 0) {
    const unsigned g = (unsigned)std::min<int64_t>(nbnd, 8LL * ctx->sm_count);
    piece_mask_blocks_kernel<<<g, 256, 0, ctx->compute>>>(A->d_col, A->d_blocks_bnd, nbnd, (int32_t)A->nloc, P.K, d_mask);
    ctx->launches++;
  }
  if (e == cudaSuccess && A->nsend > 0) {
    const unsigned g = (unsigned)std::min<int64_t>((A->nsend + 255) / 256, 8LL * ctx->sm_count);
    piece_mask_list_kernel<<<g, 256, 0, ctx->compute>>>(A->d_send_idx, A->nsend, (int32_t)A->nloc, P.K, d_mask);
    ctx->launches++;
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(&P.dist_mask, d_mask, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->compute);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->compute);
  cudaFree(d_mask);
  if (e != cudaSuccess) return fail(ctx, HPCX_CUDA_ERROR, std::string("pipe_dist_pieces: ") + cudaGetErrorString(e));
  return HPCX_OK;
}

// y(host) = A x(host): x pieces on the copy-in stream, row chunks on s as their columns land, y
// chunks on the copy-out stream.  The caller synchronises s afterwards.
// Distributed matrices (dist != nullptr): the pieces of x that the push CTAs and the boundary rows
// read (P.dist_mask — for a banded operator the two ends of the owned block) go up first, then the
// others in ascending order; once the first group has landed `dist` is called — it queues the
// launch that pushes this rank's boundary entries to the neighbours and computes the boundary
// rows — and the interior chunks follow as in the single-GPU case, so the exchange overlaps the
// copies too.
hpcx_status spmv_host_pipelined(hpcx_csr A, hpcx::PipePlan &P, const RowBlock *d_blocks,
                                const std::vector<RowBlock> &blocks, int64_t nx, void *d_x, const void *x, void *y,
                                cudaStream_t s, const std::function<hpcx_status()> &dist) {
  HPCX_RANGE("spmv: host operands, chunks pipelined against PCIe");
  hpcx_ctx ctx = A->ctx;
  const int K = P.K;
  const size_t es = (size_t)A->es;
  cudaEvent_t *ex = P.ev.data(), *ey = ex + K, entry = ex[2 * K], done = ex[2 * K + 1];
  auto piece = [&](int j) { return nx * j / K; };  // first entry of x piece j
  HPCX_CUDA(ctx, cudaEventRecord(entry, s));
  HPCX_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, entry, 0));
  HPCX_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, entry, 0));
  // copy order; ex[p] is recorded after the p-th copy, and the copies share one stream, so it
  // also covers every piece copied before it
  int order[kPipeChunks], upto[kPipeChunks];  // upto[j]: position by which pieces 0..j have all been copied
  int nfirst = 0;
  if (dist)
    for (int j = 0; j < K; ++j)
      if (P.dist_mask >> j & 1u) order[nfirst++] = j;
  {
    int n = nfirst;
    for (int j = 0; j < K; ++j)
      if (!dist || !(P.dist_mask >> j & 1u)) order[n++] = j;
    int pos[kPipeChunks];
    for (int p = 0; p < K; ++p) pos[order[p]] = p;
    for (int j = 0, m = 0; j < K; ++j) upto[j] = m = std::max(m, pos[j]);
  }
  for (int p = 0; p < K; ++p) {
    const int64_t a = piece(order[p]), b = piece(order[p] + 1);
    HPCX_CUDA(ctx, cudaMemcpyAsync((char *)d_x + a * es, (const char *)x + a * es, (size_t)(b - a) * es,
                                   cudaMemcpyHostToDevice, ctx->copy_in));
    HPCX_CUDA(ctx, cudaEventRecord(ex[p], ctx->copy_in));
  }
  int waited = -1;  // position in the copy order that s has waited for
  if (dist) {
    if (nfirst > 0) {
      HPCX_CUDA(ctx, cudaStreamWaitEvent(s, ex[nfirst - 1], 0));
      waited = nfirst - 1;
    }
    if (hpcx_status st = dist()) return st;
  }
  for (int c = 0; c < K; ++c) {
    int need = 0;  // last piece this chunk reads
    while (need + 1 < K && piece(need + 1) <= P.cmax[c]) ++need;
    if (upto[need] > waited) {
      waited = upto[need];
      HPCX_CUDA(ctx, cudaStreamWaitEvent(s, ex[waited], 0));
    }
    const int64_t b0 = P.first[c], b1 = P.first[c + 1];
    hpcx_status st = hpcx::spmv_launch_blocks(A, d_blocks + b0, b1 - b0, d_x, A->d_y, s, nullptr, nullptr);
    if (st != HPCX_OK) return st;
    HPCX_CUDA(ctx, cudaEventRecord(ey[c], s));
    HPCX_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ey[c], 0));
    // rows of this chunk: up to the next chunk's first row (rows of the other kind in between — the
    // boundary rows of a distributed matrix, computed by dist() earlier on s — travel with it)
    const int64_t r0 = c == 0 ? 0 : blocks[b0].row0;
    const int64_t r1 = c + 1 < K ? blocks[b1].row0 : A->nrows;
    HPCX_CUDA(ctx, cudaMemcpyAsync((char *)y + r0 * es, (const char *)A->d_y + r0 * es, (size_t)(r1 - r0) * es,
                                   cudaMemcpyDeviceToHost, ctx->copy_out));
  }
  HPCX_CUDA(ctx, cudaEventRecord(done, ctx->copy_out));
  HPCX_CUDA(ctx, cudaStreamWaitEvent(s, done, 0));
  if (waited < K - 1) HPCX_CUDA(ctx, cudaStreamWaitEvent(s, ex[K - 1], 0));  // x fully landed before s goes on
  return HPCX_OK;
}

}  // namespace

extern "C" {

hpcx_status hpcx_csr_create(hpcx_ctx ctx, const hpcx_csr_desc *d, const hpcx_csr_opts *opts,
                            hpcx_csr *out) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_csr_create");
  *out = nullptr;
  HPCX_REQUIRE(ctx, d != nullptr, "null CSR descriptor");
  HPCX_REQUIRE(ctx, d->nrows >= 0 && d->ncols >= 0 && d->nnz >= 0, "negative CSR size");
  HPCX_REQUIRE(ctx, d->idx_bytes == 4 || d->idx_bytes == 8, "idx_bytes must be 4 or 8");
  HPCX_REQUIRE(ctx, d->dtype == HPCX_F32 || d->dtype == HPCX_F64, "bad dtype");
  HPCX_REQUIRE(ctx, d->row_ptr && (d->nnz == 0 || (d->col_idx && d->values)), "null CSR array");
  if (d->nnz > INT32_MAX || d->nrows >= INT32_MAX)
    return fail(ctx, HPCX_UNSUPPORTED, "nnz and nrows must be below 2^31 in this build");

  hpcx_csr A = new (std::nothrow) hpcx_csr_s();
  if (!A) return fail(ctx, HPCX_OOM, "host allocation failed");
  A->ctx = ctx;
  A->nrows = d->nrows;
  A->ncols = d->ncols;
  A->nnz = d->nnz;
  A->dtype = d->dtype;
  A->es = (int)dtype_size(d->dtype);
  A->idx_bytes = d->idx_bytes;
  A->kernel = opts ? opts->kernel : HPCX_SPMV_AUTO;
  A->tile = opts && opts->tile_nnz ? opts->tile_nnz : 2048;
  A->lanes_per_row = opts ? opts->lanes_per_row : 0;
  A->short_row_max = opts && opts->short_row_max > 0 ? opts->short_row_max : 48;
  A->sell_c = opts && opts->sell_c ? opts->sell_c : 32;
  A->sell_sigma = opts && opts->sell_sigma ? opts->sell_sigma : 1024;
  if (A->tile != 1024 && A->tile != 2048 && A->tile != 4096) {
    delete A;
    return fail(ctx, HPCX_INVALID_ARG, "tile_nnz must be 1024, 2048 or 4096");
  }
  if (A->kernel < HPCX_SPMV_AUTO || A->kernel > HPCX_SPMV_HUB || A->sell_c < 1 ||
      A->sell_c > kThreads || kThreads % A->sell_c != 0) {
    delete A;
    return fail(ctx, HPCX_INVALID_ARG, "bad SpMV kernel / SELL slice height");
  }

#define HPCX_TRY(expr)                 \
  do {                                 \
    hpcx_status st__ = (expr);         \
    if (st__ != HPCX_OK) {             \
      free_csr(A);                     \
      return st__;                     \
    }                                  \
  } while (0)
#define HPCX_TRY_CUDA(expr)                                                                   \
  do {                                                                                        \
    cudaError_t e__ = (expr);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      free_csr(A);                                                                            \
      return fail(ctx, e__ == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR,         \
                  std::string(#expr) + ": " + cudaGetErrorString(e__));                       \
    }                                                                                         \
  } while (0)

  // Device CSR. Padded so that 16-byte aligned bulk copies never run past the allocation.
  A->nnz_alloc = round_up(d->nnz, 4) + 8;
  HPCX_TRY_CUDA(cudaMalloc((void **)&A->d_row_ptr, (size_t)(d->nrows + 1) * 4));
  HPCX_TRY_CUDA(cudaMalloc((void **)&A->d_col, (size_t)A->nnz_alloc * 4));
  HPCX_TRY_CUDA(cudaMalloc(&A->d_val, (size_t)A->nnz_alloc * A->es));
  HPCX_TRY_CUDA(cudaMemsetAsync(A->d_col + d->nnz, 0, (size_t)(A->nnz_alloc - d->nnz) * 4, ctx->compute));
  HPCX_TRY_CUDA(cudaMemsetAsync((char *)A->d_val + (size_t)d->nnz * A->es, 0,
                                (size_t)(A->nnz_alloc - d->nnz) * A->es, ctx->compute));
  // row_ptr: validated, narrowed to int32 and binned by one kernel over the device copy
  // (csr_plan.cu); nothing loops over the rows on the host
  int64_t maxlen = 0, nnz_long_rows = 0;
  HPCX_TRY(csr_scan_rowptr(A, d->row_ptr, d->idx_bytes, d->mem, &maxlen, &nnz_long_rows));
  A->max_row_len = maxlen;
  if (d->nnz) {
    HPCX_TRY_CUDA(cudaMemcpyAsync(A->d_col, d->col_idx, (size_t)d->nnz * 4, cudaMemcpyDefault, ctx->compute));
    HPCX_TRY_CUDA(cudaMemcpyAsync(A->d_val, d->values, (size_t)d->nnz * A->es, cudaMemcpyDefault, ctx->compute));
    HPCX_TRY_CUDA(cudaStreamSynchronize(ctx->compute));
  }

  if (d->nnz) {
    // every column index must address x: one pass over the device copy
    int32_t *d_mm = nullptr, mm[2] = {INT32_MAX, INT32_MIN};
    HPCX_TRY_CUDA(cudaMalloc((void **)&d_mm, 8));
    cudaError_t e = cudaMemcpyAsync(d_mm, mm, 8, cudaMemcpyHostToDevice, ctx->compute);
    if (e == cudaSuccess) {
      const int grid = (int)std::min<int64_t>((d->nnz + 255) / 256, (int64_t)ctx->sm_count * 16);
      col_range_kernel<<<grid, 256, 0, ctx->compute>>>(A->d_col, d->nnz, d_mm);
      ctx->launches++;
      e = cudaMemcpyAsync(mm, d_mm, 8, cudaMemcpyDeviceToHost, ctx->compute);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->compute);
    }
    cudaFree(d_mm);
    HPCX_TRY_CUDA(e);
    if (mm[0] < 0 || (int64_t)mm[1] >= d->ncols) {
      free_csr(A);
      return fail(ctx, HPCX_OUT_OF_RANGE,
                  "col_idx outside [0, ncols): min " + std::to_string(mm[0]) + ", max " + std::to_string(mm[1]) +
                      ", ncols " + std::to_string(d->ncols));
    }
  }

  // STREAM tables are always built (the distributed path and the CG step use them).
  {
    int64_t n0 = 0, nchunks = 0;
    HPCX_TRY(csr_plan_rowblocks(A, nullptr, maxlen, &A->d_blocks, &A->h_blocks, &n0, &A->d_long_row, &A->d_long_off,
                                &A->nlong, &nchunks));
    HPCX_TRY_CUDA(cudaMalloc(&A->d_partials, std::max<size_t>((size_t)nchunks, 1) * A->es));
  }

  if (A->kernel == HPCX_SPMV_SELL) HPCX_TRY(sell_build(A, maxlen));  // on the device: sell_build.cu
  // the memsets and copies above ran on the legacy stream; the context's streams are
  // non-blocking, so make everything visible before anything is queued on them
  HPCX_TRY_CUDA(cudaDeviceSynchronize());
  if (A->kernel == HPCX_SPMV_AUTO || A->kernel == HPCX_SPMV_HUB) {
    // AUTO: power-law structure (a quarter of the nonzeros in rows longer than 256) AND popular
    // columns (hub coverage >= 10 %) select the HUB kernel; everything else streams.
    const bool want = A->kernel == HPCX_SPMV_HUB || nnz_long_rows * 4 > d->nnz;
    if (want && d->nnz > 0) {
      HPCX_TRY(hub_build(A));
      if (A->kernel == HPCX_SPMV_AUTO) A->kernel = A->hub.coverage >= 0.10 ? HPCX_SPMV_HUB : HPCX_SPMV_STREAM;
      if (A->kernel != HPCX_SPMV_HUB) hub_free(A);
    } else if (A->kernel == HPCX_SPMV_AUTO) {
      A->kernel = HPCX_SPMV_STREAM;
    }
  }
  *out = A;
  return HPCX_OK;
}

hpcx_status hpcx_csr_destroy(hpcx_csr A) {
  if (!A) return HPCX_INVALID_ARG;
  free_csr(A);
  return HPCX_OK;
}

hpcx_status hpcx_spmv(hpcx_csr A, const void *x, int32_t x_mem, void *y, int32_t y_mem,
                      void *stream) {
  if (!A || !x || !y) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_spmv");
  hpcx_ctx ctx = A->ctx;
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  const void *dx = x;
  void *dy = y;
  if (x_mem == HPCX_MEM_HOST && y_mem == HPCX_MEM_HOST && !A->dot_part) {
    if (A->pipe.state == 0) {
      if (A->kernel != HPCX_SPMV_STREAM) A->pipe.state = -1;
      else if (hpcx_status st = pipe_analyse(A, A->pipe, A->h_blocks, A->nlong, A->ncols)) return st;
    }
    if (A->pipe.state == 1) {
      if (!A->d_x) HPCX_CUDA(ctx, cudaMalloc(&A->d_x, std::max<size_t>((size_t)A->ncols * A->es, 1)));
      if (!A->d_y) HPCX_CUDA(ctx, cudaMalloc(&A->d_y, std::max<size_t>((size_t)A->nrows * A->es, 1)));
      hpcx_status st = spmv_host_pipelined(A, A->pipe, A->d_blocks, A->h_blocks, A->ncols, A->d_x, x, y, s, nullptr);
      if (st != HPCX_OK) return st;
      HPCX_CUDA(ctx, cudaStreamSynchronize(s));
      return HPCX_OK;
    }
  }
  if (x_mem == HPCX_MEM_HOST) {
    if (!A->d_x) HPCX_CUDA(ctx, cudaMalloc(&A->d_x, std::max<size_t>((size_t)A->ncols * A->es, 1)));
    HPCX_CUDA(ctx, cudaMemcpyAsync(A->d_x, x, (size_t)A->ncols * A->es, cudaMemcpyHostToDevice, s));
    dx = A->d_x;
  }
  if (y_mem == HPCX_MEM_HOST) {
    if (!A->d_y) HPCX_CUDA(ctx, cudaMalloc(&A->d_y, std::max<size_t>((size_t)A->nrows * A->es, 1)));
    dy = A->d_y;
  }
  hpcx_status st = A->dtype == HPCX_F64
                       ? spmv_device<double>(A, (const double *)dx, (double *)dy, s)
                       : spmv_device<float>(A, (const float *)dx, (float *)dy, s);
  if (st != HPCX_OK) return st;
  if (y_mem == HPCX_MEM_HOST)
    HPCX_CUDA(ctx, cudaMemcpyAsync(y, dy, (size_t)A->nrows * A->es, cudaMemcpyDeviceToHost, s));
  if (!stream || y_mem == HPCX_MEM_HOST) HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}

hpcx_status hpcx_csr_download(hpcx_csr A, void *row_ptr, int32_t *col_idx, void *values) {
  if (!A) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = A->ctx;
  if (row_ptr) {
    if (A->idx_bytes == 4) {
      HPCX_CUDA(ctx, cudaMemcpy(row_ptr, A->d_row_ptr, (size_t)(A->nrows + 1) * 4,
                                cudaMemcpyDeviceToHost));
    } else {
      std::vector<int32_t> t(A->nrows + 1);
      HPCX_CUDA(ctx, cudaMemcpy(t.data(), A->d_row_ptr, t.size() * 4, cudaMemcpyDeviceToHost));
      for (int64_t i = 0; i <= A->nrows; ++i) ((int64_t *)row_ptr)[i] = t[i];
    }
  }
  if (col_idx && A->nnz)
    HPCX_CUDA(ctx, cudaMemcpy(col_idx, A->d_col, (size_t)A->nnz * 4, cudaMemcpyDeviceToHost));
  if (values && A->nnz)
    HPCX_CUDA(ctx, cudaMemcpy(values, A->d_val, (size_t)A->nnz * A->es, cudaMemcpyDeviceToHost));
  return HPCX_OK;
}

hpcx_status hpcx_csr_rowblocks(hpcx_csr A, int64_t *nblocks, int64_t *out, int64_t capacity) {
  if (!A || !nblocks) return HPCX_INVALID_ARG;
  *nblocks = (int64_t)A->h_blocks.size();
  if (out) {
    HPCX_REQUIRE(A->ctx, capacity >= *nblocks, "row-block buffer too small");
    for (size_t i = 0; i < A->h_blocks.size(); ++i) {
      const RowBlock &b = A->h_blocks[i];
      int64_t *o = out + 6 * i;
      o[0] = b.row0;
      o[1] = b.nrows;
      o[2] = b.part;
      o[3] = b.lanes;
      o[4] = b.p0;
      o[5] = b.p1;
    }
  }
  return HPCX_OK;
}

hpcx_status hpcx_csr_row_bins(hpcx_csr A, int64_t *bins34) {
  if (!A || !bins34) return HPCX_INVALID_ARG;
  memcpy(bins34, A->bins, sizeof A->bins);
  return HPCX_OK;
}

hpcx_status hpcx_csr_sell_info(hpcx_csr A, int64_t *nslices, int64_t *padded_nnz, int32_t *c,
                               int32_t *sigma) {
  if (!A) return HPCX_INVALID_ARG;
  if (A->kernel != HPCX_SPMV_SELL)
    return fail(A->ctx, HPCX_UNSUPPORTED, "matrix was not created with HPCX_SPMV_SELL");
  if (nslices) *nslices = A->sell_nslices;
  if (padded_nnz) *padded_nnz = A->sell_padded;
  if (c) *c = A->sell_c;
  if (sigma) *sigma = A->sell_sigma;
  return HPCX_OK;
}

hpcx_status hpcx_csr_sell_download(hpcx_csr A, int32_t *perm, int64_t *slice_ptr,
                                   int32_t *slice_len, int32_t *col, void *val) {
  if (!A) return HPCX_INVALID_ARG;
  hpcx_ctx ctx = A->ctx;
  if (A->kernel != HPCX_SPMV_SELL)
    return fail(ctx, HPCX_UNSUPPORTED, "matrix was not created with HPCX_SPMV_SELL");
  // Structure comes back from the DEVICE copies (what the kernel reads).
  if (perm && A->nrows)
    HPCX_CUDA(ctx, cudaMemcpy(perm, A->d_perm, (size_t)A->nrows * 4, cudaMemcpyDeviceToHost));
  if (slice_ptr)
    HPCX_CUDA(ctx, cudaMemcpy(slice_ptr, A->d_slice_ptr, (size_t)(A->sell_nslices + 1) * 8,
                              cudaMemcpyDeviceToHost));
  if (slice_len && A->sell_nslices)
    HPCX_CUDA(ctx, cudaMemcpy(slice_len, A->d_slice_len, (size_t)A->sell_nslices * 4,
                              cudaMemcpyDeviceToHost));
  if (col && A->sell_padded)
    HPCX_CUDA(ctx, cudaMemcpy(col, A->d_sell_col, (size_t)A->sell_padded * 4, cudaMemcpyDeviceToHost));
  if (val && A->sell_padded)
    HPCX_CUDA(ctx, cudaMemcpy(val, A->d_sell_val, (size_t)A->sell_padded * A->es,
                              cudaMemcpyDeviceToHost));
  return HPCX_OK;
}

}  // extern "C"
