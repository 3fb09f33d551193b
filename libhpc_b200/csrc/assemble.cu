// assemble.cu — device-side CSR builders: COO -> CSR assembly with
// duplicate merging, and the counter-based generators of the synthetic configs (C5 R-MAT
// edges, 2u-1 fields).  The reference has no matrix container; the CPU
// statement these kernels are tested against is oracle/hpcx_oracle.c: orc_rmat_csr / orc_u.
//
// Assembly contract (bit-exact against the CPU builder):
//   entries sorted by (row, col, k) with k the position in the input list; equal (row, col)
//   merged by adding their values IN ASCENDING k, one rounding per add; col_idx ascending
//   within a row; row_ptr[r] = number of merged entries in rows < r.
// The sort is a stable LSD radix sort (cub::DeviceRadixSort) of the 64-bit key
// (row << 32 | col) carrying the value, so "ascending k inside a run" is the sort's stability.
#include <cub/cub.cuh>

#include "common.cuh"

namespace hpcx {
namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// u(seed, i) of SURVEY.md §8d with s = mix(seed) precomputed on the host.
__device__ __forceinline__ double u01(uint64_t s, uint64_t i) {
  return (double)(mix64(s ^ i) >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void rmat_edges_kernel(int scale, int64_t nedges, uint64_t s_edge, uint64_t s_val,
                                  int32_t *__restrict__ rows, int32_t *__restrict__ cols,
                                  float *__restrict__ vals) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nedges; k += stride) {
    uint32_t r = 0, c = 0;
    const uint64_t base = (uint64_t)k * (uint64_t)scale;
    for (int l = 0; l < scale; ++l) {
      const double u = u01(s_edge, base + (uint64_t)l);
      const uint32_t q = u < 0.57 ? 0u : u < 0.76 ? 1u : u < 0.95 ? 2u : 3u;
      r = (r << 1) | (q >> 1);
      c = (c << 1) | (q & 1u);
    }
    rows[k] = (int32_t)r;
    cols[k] = (int32_t)c;
    vals[k] = (float)__dsub_rn(__dmul_rn(2.0, u01(s_val, (uint64_t)k)), 1.0);
  }
}

template <typename T>
__global__ void field_kernel(uint64_t s, uint64_t first, int64_t n, T *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = (T)__dsub_rn(__dmul_rn(2.0, u01(s, first + (uint64_t)i)), 1.0);
}

__global__ void coo_keys_kernel(int64_t n, const int32_t *__restrict__ rows,
                                const int32_t *__restrict__ cols, uint64_t *__restrict__ keys) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    keys[i] = ((uint64_t)(uint32_t)rows[i] << 32) | (uint64_t)(uint32_t)cols[i];
}

// 1 where a run of equal keys starts.
struct HeadFlag {
  const uint64_t *keys;
  __device__ __forceinline__ int32_t operator()(int64_t i) const {
    return (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
  }
};

// pos[i] = inclusive count of run heads up to i.  The head of each run writes the merged
// entry pos[i]-1: column, and the run's values added in list order.
template <typename T>
__global__ void merge_runs_kernel(int64_t n, const uint64_t *__restrict__ keys,
                                  const T *__restrict__ vals, const int32_t *__restrict__ pos,
                                  int32_t *__restrict__ col, T *__restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t k = keys[i];
    if (i > 0 && keys[i - 1] == k) continue;
    T s = vals[i];
    for (int64_t j = i + 1; j < n && keys[j] == k; ++j) s = s + vals[j];  // --fmad=false: adds only
    const int64_t o = (int64_t)pos[i] - 1;
    col[o] = (int32_t)(uint32_t)(k & 0xffffffffull);
    out[o] = s;
  }
}

// row_ptr[r] = merged entries in rows < r = (heads before the first sorted key >= r << 32).
template <typename I>
__global__ void row_ptr_kernel(int64_t nrows, int64_t n, const uint64_t *__restrict__ keys,
                               const int32_t *__restrict__ pos, int64_t nnz,
                               I *__restrict__ row_ptr) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= nrows; r += stride) {
    const uint64_t want = (uint64_t)r << 32;
    int64_t lo = 0, hi = n;  // first index with keys[idx] >= want
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < want) lo = mid + 1; else hi = mid;
    }
    row_ptr[r] = (I)(lo == n ? nnz : (int64_t)pos[lo] - 1);
  }
}

inline int grid_for(hpcx_ctx ctx, int64_t n, int block) {
  const int64_t want = (n + block - 1) / block;
  const int64_t cap = (int64_t)ctx->sm_count * 16;
  return (int)(want < 1 ? 1 : want < cap ? want : cap);
}

struct DevBuf {  // frees on scope exit
  void *p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
};

template <typename T>
hpcx_status assemble(hpcx_ctx ctx, int64_t nrows, int64_t n, const int32_t *rows,
                     const int32_t *cols, const T *vals, int32_t idx_bytes,
                     hpcx_csr_arrays *out) {
  cudaStream_t st = ctx->compute;
  DevBuf k0, k1, v1, pos, tmp;
  const size_t nn = (size_t)(n > 0 ? n : 1);
  HPCX_CUDA(ctx, cudaMalloc(&k0.p, nn * 8));
  HPCX_CUDA(ctx, cudaMalloc(&k1.p, nn * 8));
  HPCX_CUDA(ctx, cudaMalloc(&v1.p, nn * sizeof(T)));
  HPCX_CUDA(ctx, cudaMalloc(&pos.p, nn * 4));
  int64_t nnz = 0;
  const uint64_t *keys = nullptr;
  const T *sv = nullptr;
  if (n > 0) {
    coo_keys_kernel<<<grid_for(ctx, n, 256), 256, 0, st>>>(n, rows, cols, (uint64_t *)k0.p);
    if (hpcx_status s = after_launch(ctx, "coo_keys_kernel")) return s;
    // Values ride along as the sort's payload; vals is const, so the double buffer's second
    // half doubles as the scratch copy the sort may write to.
    DevBuf v0;
    HPCX_CUDA(ctx, cudaMalloc(&v0.p, nn * sizeof(T)));
    HPCX_CUDA(ctx, cudaMemcpyAsync(v0.p, vals, nn * sizeof(T), cudaMemcpyDeviceToDevice, st));
    cub::DoubleBuffer<uint64_t> dk((uint64_t *)k0.p, (uint64_t *)k1.p);
    cub::DoubleBuffer<T> dv((T *)v0.p, (T *)v1.p);
    int row_bits = 1;
    while (row_bits < 32 && ((int64_t)1 << row_bits) < nrows) ++row_bits;
    const int end_bit = 32 + row_bits;
    size_t tb = 0;
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, n, 0, end_bit, st));
    HPCX_CUDA(ctx, cudaMalloc(&tmp.p, tb ? tb : 1));
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp.p, tb, dk, dv, n, 0, end_bit, st));
    ctx->launches += (end_bit + 7) / 8 * 3;  // histogram + onesweep passes (approximate count)
    keys = dk.Current();
    // keep the sorted values alive past v0's scope
    if (dv.Current() != (T *)v1.p)
      HPCX_CUDA(ctx, cudaMemcpyAsync(v1.p, dv.Current(), nn * sizeof(T), cudaMemcpyDeviceToDevice, st));
    sv = (const T *)v1.p;
    // inclusive scan of run heads
    cub::CountingInputIterator<int64_t> cnt(0);
    cub::TransformInputIterator<int32_t, HeadFlag, cub::CountingInputIterator<int64_t>> heads(
        cnt, HeadFlag{keys});
    size_t sb = 0;
    HPCX_CUDA(ctx, cub::DeviceScan::InclusiveSum(nullptr, sb, heads, (int32_t *)pos.p, n, st));
    DevBuf stmp;
    HPCX_CUDA(ctx, cudaMalloc(&stmp.p, sb ? sb : 1));
    HPCX_CUDA(ctx, cub::DeviceScan::InclusiveSum(stmp.p, sb, heads, (int32_t *)pos.p, n, st));
    ctx->launches += 2;
    int32_t last = 0;
    HPCX_CUDA(ctx, cudaMemcpyAsync(&last, (int32_t *)pos.p + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    HPCX_CUDA(ctx, cudaStreamSynchronize(st));  // also keeps v0/stmp alive until the work is done
    nnz = last;
  }
  DevBuf rp, col, val;  // handed to the caller on success
  HPCX_CUDA(ctx, cudaMalloc(&rp.p, (size_t)(nrows + 1) * idx_bytes));
  HPCX_CUDA(ctx, cudaMalloc(&col.p, (size_t)(nnz > 0 ? nnz : 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc(&val.p, (size_t)(nnz > 0 ? nnz : 1) * sizeof(T)));
  if (n > 0) {
    merge_runs_kernel<T><<<grid_for(ctx, n, 256), 256, 0, st>>>(n, keys, sv, (const int32_t *)pos.p,
                                                              (int32_t *)col.p, (T *)val.p);
    if (hpcx_status s = after_launch(ctx, "merge_runs_kernel")) return s;
  }
  if (idx_bytes == 4)
    row_ptr_kernel<int32_t><<<grid_for(ctx, nrows + 1, 256), 256, 0, st>>>(
        nrows, n, keys, (const int32_t *)pos.p, nnz, (int32_t *)rp.p);
  else
    row_ptr_kernel<int64_t><<<grid_for(ctx, nrows + 1, 256), 256, 0, st>>>(
        nrows, n, keys, (const int32_t *)pos.p, nnz, (int64_t *)rp.p);
  if (hpcx_status s = after_launch(ctx, "row_ptr_kernel")) return s;
  HPCX_CUDA(ctx, cudaStreamSynchronize(st));
  out->nnz = nnz;
  out->row_ptr = rp.p;
  out->col_idx = (int32_t *)col.p;
  out->values = val.p;
  rp.p = col.p = val.p = nullptr;
  return HPCX_OK;
}


// ------------------------------------------------------------------ sparse-grid bridge (§8f-4)
// Cells of a sparse 2-D grid (what sparse::RootGrid::foreach enumerates, reference
// lib/sparse/include/RootGrid.hpp:20-22, 62-75) -> 5-point operator over the active cells.
// Rows are the distinct cells in (x, y) lexicographic order (x-major like DenseBlock's
// m_block[x*B + y], DenseBlock.hpp:24-28); a neighbour that is not active is a Dirichlet zero.
__device__ __forceinline__ uint64_t cell_key(int32_t x, int32_t y) {
  return ((uint64_t)((uint32_t)x ^ 0x80000000u) << 32) | (uint64_t)((uint32_t)y ^ 0x80000000u);
}
__global__ void cell_keys_kernel(int64_t n, const int32_t *__restrict__ xs, const int32_t *__restrict__ ys,
                                 uint64_t *__restrict__ keys, int32_t *__restrict__ idx) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    keys[i] = cell_key(xs[i], ys[i]);
    idx[i] = (int32_t)i;
  }
}
// pos = inclusive head count over the sorted keys: compact the distinct keys, map cells to rows.
__global__ void cell_rows_kernel(int64_t n, const uint64_t *__restrict__ keys, const int32_t *__restrict__ idx,
                                 const int32_t *__restrict__ pos, uint64_t *__restrict__ ukeys,
                                 int32_t *__restrict__ cell_row, int32_t *__restrict__ row_x,
                                 int32_t *__restrict__ row_y) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int32_t r = pos[i] - 1;
    cell_row[idx[i]] = r;
    if (i == 0 || keys[i] != keys[i - 1]) {
      const uint64_t k = keys[i];
      ukeys[r] = k;
      row_x[r] = (int32_t)((uint32_t)(k >> 32) ^ 0x80000000u);
      row_y[r] = (int32_t)((uint32_t)k ^ 0x80000000u);
    }
  }
}
__device__ __forceinline__ int32_t find_key(const uint64_t *__restrict__ ukeys, int64_t m, uint64_t want) {
  int64_t lo = 0, hi = m;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ukeys[mid] < want) lo = mid + 1; else hi = mid;
  }
  return (lo < m && ukeys[lo] == want) ? (int32_t)lo : -1;
}
// nbr[4r..4r+3] = rows of (x-1,y), (x,y-1), (x,y+1), (x+1,y) or -1; deg[r] = 1 + active neighbours
__global__ void cell_neighbours_kernel(int64_t m, const uint64_t *__restrict__ ukeys,
                                       const int32_t *__restrict__ row_x, const int32_t *__restrict__ row_y,
                                       int32_t *__restrict__ nbr, int32_t *__restrict__ deg) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += stride) {
    const int32_t x = row_x[r], y = row_y[r];
    int32_t d = 1;
    const int32_t n0 = x > INT32_MIN ? find_key(ukeys, m, cell_key(x - 1, y)) : -1;
    const int32_t n1 = y > INT32_MIN ? find_key(ukeys, m, cell_key(x, y - 1)) : -1;
    const int32_t n2 = y < INT32_MAX ? find_key(ukeys, m, cell_key(x, y + 1)) : -1;
    const int32_t n3 = x < INT32_MAX ? find_key(ukeys, m, cell_key(x + 1, y)) : -1;
    nbr[4 * r + 0] = n0;
    nbr[4 * r + 1] = n1;
    nbr[4 * r + 2] = n2;
    nbr[4 * r + 3] = n3;
    d += (n0 >= 0) + (n1 >= 0) + (n2 >= 0) + (n3 >= 0);
    deg[r] = d;
  }
}
template <typename T, typename I>
__global__ void cell_fill_kernel(int64_t m, const int32_t *__restrict__ nbr, const int32_t *__restrict__ start,
                                 int64_t nnz, I *__restrict__ row_ptr, int32_t *__restrict__ col,
                                 T *__restrict__ val) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r <= m; r += stride) {
    if (r == m) {
      row_ptr[m] = (I)nnz;
      continue;
    }
    int64_t p = start[r];
    row_ptr[r] = (I)p;
    // ascending column = ascending (x, y): (x-1,y) < (x,y-1) < (x,y) < (x,y+1) < (x+1,y)
    if (nbr[4 * r + 0] >= 0) { col[p] = nbr[4 * r + 0]; val[p++] = T(-1); }
    if (nbr[4 * r + 1] >= 0) { col[p] = nbr[4 * r + 1]; val[p++] = T(-1); }
    col[p] = (int32_t)r;
    val[p++] = T(4);
    if (nbr[4 * r + 2] >= 0) { col[p] = nbr[4 * r + 2]; val[p++] = T(-1); }
    if (nbr[4 * r + 3] >= 0) { col[p] = nbr[4 * r + 3]; val[p++] = T(-1); }
  }
}

template <typename T>
hpcx_status cells_to_csr(hpcx_ctx ctx, int64_t n, const int32_t *xs, const int32_t *ys, int32_t idx_bytes,
                         hpcx_csr_arrays *out, hpcx_cell_map *map) {
  cudaStream_t st = ctx->compute;
  const size_t nn = (size_t)(n > 0 ? n : 1);
  DevBuf k0, k1, i0, i1, pos, tmp, ukeys, nbr, deg, start;
  DevBuf cell_row, row_x, row_y, rp, col, val;  // handed to the caller on success
  HPCX_CUDA(ctx, cudaMalloc(&cell_row.p, nn * 4));
  int64_t m = 0, nnz = 0;
  if (n > 0) {
    HPCX_CUDA(ctx, cudaMalloc(&k0.p, nn * 8));
    HPCX_CUDA(ctx, cudaMalloc(&k1.p, nn * 8));
    HPCX_CUDA(ctx, cudaMalloc(&i0.p, nn * 4));
    HPCX_CUDA(ctx, cudaMalloc(&i1.p, nn * 4));
    HPCX_CUDA(ctx, cudaMalloc(&pos.p, nn * 4));
    cell_keys_kernel<<<grid_for(ctx, n, 256), 256, 0, st>>>(n, xs, ys, (uint64_t *)k0.p, (int32_t *)i0.p);
    if (hpcx_status s = after_launch(ctx, "cell_keys_kernel")) return s;
    cub::DoubleBuffer<uint64_t> dk((uint64_t *)k0.p, (uint64_t *)k1.p);
    cub::DoubleBuffer<int32_t> di((int32_t *)i0.p, (int32_t *)i1.p);
    size_t tb = 0;
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, di, n, 0, 64, st));
    HPCX_CUDA(ctx, cudaMalloc(&tmp.p, tb ? tb : 1));
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp.p, tb, dk, di, n, 0, 64, st));
    ctx->launches += 8 * 3;
    const uint64_t *keys = dk.Current();
    cub::CountingInputIterator<int64_t> cnt(0);
    cub::TransformInputIterator<int32_t, HeadFlag, cub::CountingInputIterator<int64_t>> heads(cnt, HeadFlag{keys});
    size_t sb = 0;
    HPCX_CUDA(ctx, cub::DeviceScan::InclusiveSum(nullptr, sb, heads, (int32_t *)pos.p, n, st));
    DevBuf stmp;
    HPCX_CUDA(ctx, cudaMalloc(&stmp.p, sb ? sb : 1));
    HPCX_CUDA(ctx, cub::DeviceScan::InclusiveSum(stmp.p, sb, heads, (int32_t *)pos.p, n, st));
    ctx->launches += 2;
    int32_t last = 0;
    HPCX_CUDA(ctx, cudaMemcpyAsync(&last, (int32_t *)pos.p + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    HPCX_CUDA(ctx, cudaStreamSynchronize(st));
    m = last;
    HPCX_CUDA(ctx, cudaMalloc(&ukeys.p, (size_t)m * 8));
    HPCX_CUDA(ctx, cudaMalloc(&row_x.p, (size_t)m * 4));
    HPCX_CUDA(ctx, cudaMalloc(&row_y.p, (size_t)m * 4));
    HPCX_CUDA(ctx, cudaMalloc(&nbr.p, (size_t)m * 16));
    HPCX_CUDA(ctx, cudaMalloc(&deg.p, (size_t)m * 4));
    HPCX_CUDA(ctx, cudaMalloc(&start.p, (size_t)m * 4));
    cell_rows_kernel<<<grid_for(ctx, n, 256), 256, 0, st>>>(n, keys, di.Current(), (const int32_t *)pos.p,
                                                          (uint64_t *)ukeys.p, (int32_t *)cell_row.p,
                                                          (int32_t *)row_x.p, (int32_t *)row_y.p);
    if (hpcx_status s = after_launch(ctx, "cell_rows_kernel")) return s;
    cell_neighbours_kernel<<<grid_for(ctx, m, 256), 256, 0, st>>>(m, (const uint64_t *)ukeys.p,
                                                                (const int32_t *)row_x.p, (const int32_t *)row_y.p,
                                                                (int32_t *)nbr.p, (int32_t *)deg.p);
    if (hpcx_status s = after_launch(ctx, "cell_neighbours_kernel")) return s;
    size_t eb = 0;
    HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, eb, (int32_t *)deg.p, (int32_t *)start.p, m, st));
    DevBuf etmp;
    HPCX_CUDA(ctx, cudaMalloc(&etmp.p, eb ? eb : 1));
    HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(etmp.p, eb, (int32_t *)deg.p, (int32_t *)start.p, m, st));
    ctx->launches += 2;
    int32_t s_last = 0, d_last = 0;
    HPCX_CUDA(ctx, cudaMemcpyAsync(&s_last, (int32_t *)start.p + (m - 1), 4, cudaMemcpyDeviceToHost, st));
    HPCX_CUDA(ctx, cudaMemcpyAsync(&d_last, (int32_t *)deg.p + (m - 1), 4, cudaMemcpyDeviceToHost, st));
    HPCX_CUDA(ctx, cudaStreamSynchronize(st));
    nnz = (int64_t)s_last + d_last;
  }
  HPCX_CUDA(ctx, cudaMalloc(&rp.p, (size_t)(m + 1) * idx_bytes));
  HPCX_CUDA(ctx, cudaMalloc(&col.p, (size_t)(nnz > 0 ? nnz : 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc(&val.p, (size_t)(nnz > 0 ? nnz : 1) * sizeof(T)));
  if (idx_bytes == 4)
    cell_fill_kernel<T, int32_t><<<grid_for(ctx, m + 1, 256), 256, 0, st>>>(
        m, (const int32_t *)nbr.p, (const int32_t *)start.p, nnz, (int32_t *)rp.p, (int32_t *)col.p, (T *)val.p);
  else
    cell_fill_kernel<T, int64_t><<<grid_for(ctx, m + 1, 256), 256, 0, st>>>(
        m, (const int32_t *)nbr.p, (const int32_t *)start.p, nnz, (int64_t *)rp.p, (int32_t *)col.p, (T *)val.p);
  if (hpcx_status s = after_launch(ctx, "cell_fill_kernel")) return s;
  HPCX_CUDA(ctx, cudaStreamSynchronize(st));
  out->nnz = nnz;
  out->row_ptr = rp.p;
  out->col_idx = (int32_t *)col.p;
  out->values = val.p;
  rp.p = col.p = val.p = nullptr;
  if (map) {
    map->nrows = m;
    map->cell_row = (int32_t *)cell_row.p;
    map->row_x = (int32_t *)row_x.p;
    map->row_y = (int32_t *)row_y.p;
    cell_row.p = row_x.p = row_y.p = nullptr;
  }
  return HPCX_OK;
}

}  // namespace
}  // namespace hpcx

using namespace hpcx;

extern "C" {

hpcx_status hpcx_csr_from_coo(hpcx_ctx ctx, int64_t nrows, int64_t nedges, const int32_t *rows,
                              const int32_t *cols, const void *vals, int32_t dtype, int32_t mem,
                              int32_t idx_bytes, hpcx_csr_arrays *out) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(ctx, nrows >= 0 && nrows <= 0x7fffffffLL, "nrows must fit int32 row indices");
  HPCX_REQUIRE(ctx, nedges >= 0 && nedges < 0x7fffffffLL, "nedges must be below 2^31");
  HPCX_REQUIRE(ctx, idx_bytes == 4 || idx_bytes == 8, "idx_bytes must be 4 or 8");
  HPCX_REQUIRE(ctx, dtype == HPCX_F32 || dtype == HPCX_F64, "dtype must be HPCX_F32 or HPCX_F64");
  HPCX_REQUIRE(ctx, nedges == 0 || (rows && cols && vals), "null edge arrays");
  HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t ts = dtype_size(dtype);
  DevBuf dr, dc, dv;
  if (mem == HPCX_MEM_HOST && nedges > 0) {
    HPCX_CUDA(ctx, cudaMalloc(&dr.p, (size_t)nedges * 4));
    HPCX_CUDA(ctx, cudaMalloc(&dc.p, (size_t)nedges * 4));
    HPCX_CUDA(ctx, cudaMalloc(&dv.p, (size_t)nedges * ts));
    HPCX_CUDA(ctx, cudaMemcpy(dr.p, rows, (size_t)nedges * 4, cudaMemcpyHostToDevice));
    HPCX_CUDA(ctx, cudaMemcpy(dc.p, cols, (size_t)nedges * 4, cudaMemcpyHostToDevice));
    HPCX_CUDA(ctx, cudaMemcpy(dv.p, vals, (size_t)nedges * ts, cudaMemcpyHostToDevice));
    rows = (const int32_t *)dr.p;
    cols = (const int32_t *)dc.p;
    vals = dv.p;
  }
  return dtype == HPCX_F64
             ? assemble<double>(ctx, nrows, nedges, rows, cols, (const double *)vals, idx_bytes, out)
             : assemble<float>(ctx, nrows, nedges, rows, cols, (const float *)vals, idx_bytes, out);
}

hpcx_status hpcx_csr_from_cells(hpcx_ctx ctx, int64_t ncells, const int32_t *xs, const int32_t *ys,
                                int32_t mem, int32_t dtype, int32_t idx_bytes, hpcx_csr_arrays *out,
                                hpcx_cell_map *map) {
  if (!ctx || !out) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(ctx, ncells >= 0 && ncells < 0x19999999LL, "ncells must be below 2^31 / 5");
  HPCX_REQUIRE(ctx, idx_bytes == 4 || idx_bytes == 8, "idx_bytes must be 4 or 8");
  HPCX_REQUIRE(ctx, dtype == HPCX_F32 || dtype == HPCX_F64, "dtype must be HPCX_F32 or HPCX_F64");
  HPCX_REQUIRE(ctx, ncells == 0 || (xs && ys), "null coordinate arrays");
  HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  DevBuf dx, dy;
  if (mem == HPCX_MEM_HOST && ncells > 0) {
    HPCX_CUDA(ctx, cudaMalloc(&dx.p, (size_t)ncells * 4));
    HPCX_CUDA(ctx, cudaMalloc(&dy.p, (size_t)ncells * 4));
    HPCX_CUDA(ctx, cudaMemcpy(dx.p, xs, (size_t)ncells * 4, cudaMemcpyHostToDevice));
    HPCX_CUDA(ctx, cudaMemcpy(dy.p, ys, (size_t)ncells * 4, cudaMemcpyHostToDevice));
    xs = (const int32_t *)dx.p;
    ys = (const int32_t *)dy.p;
  }
  return dtype == HPCX_F64 ? cells_to_csr<double>(ctx, ncells, xs, ys, idx_bytes, out, map)
                           : cells_to_csr<float>(ctx, ncells, xs, ys, idx_bytes, out, map);
}

hpcx_status hpcx_synth_rmat_coo(hpcx_ctx ctx, int32_t scale, int64_t nedges, uint64_t seed,
                                int32_t *rows_dev, int32_t *cols_dev, float *vals_dev) {
  if (!ctx) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(ctx, scale >= 1 && scale <= 31, "scale must be in [1, 31]");
  HPCX_REQUIRE(ctx, nedges >= 0 && rows_dev && cols_dev && vals_dev, "null output arrays");
  HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  if (nedges == 0) return HPCX_OK;
  rmat_edges_kernel<<<grid_for(ctx, nedges, 256), 256, 0, ctx->compute>>>(
      scale, nedges, mix64(seed), mix64(seed + 1), rows_dev, cols_dev, vals_dev);
  if (hpcx_status s = after_launch(ctx, "rmat_edges_kernel")) return s;
  HPCX_CUDA(ctx, cudaStreamSynchronize(ctx->compute));
  return HPCX_OK;
}

hpcx_status hpcx_synth_field(hpcx_ctx ctx, uint64_t seed, int64_t first, int64_t n, int32_t dtype,
                             void *out_dev) {
  if (!ctx) return HPCX_INVALID_ARG;
  HPCX_REQUIRE(ctx, n >= 0 && first >= 0 && (n == 0 || out_dev), "bad range or null output");
  HPCX_REQUIRE(ctx, dtype == HPCX_F32 || dtype == HPCX_F64, "dtype must be HPCX_F32 or HPCX_F64");
  HPCX_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n == 0) return HPCX_OK;
  if (dtype == HPCX_F64)
    field_kernel<double><<<grid_for(ctx, n, 256), 256, 0, ctx->compute>>>(mix64(seed), (uint64_t)first,
                                                                        n, (double *)out_dev);
  else
    field_kernel<float><<<grid_for(ctx, n, 256), 256, 0, ctx->compute>>>(mix64(seed), (uint64_t)first,
                                                                       n, (float *)out_dev);
  if (hpcx_status s = after_launch(ctx, "field_kernel")) return s;
  HPCX_CUDA(ctx, cudaStreamSynchronize(ctx->compute));
  return HPCX_OK;
}

}  // extern "C"
