// sell_build.cu — SELL-C-sigma repack of a device CSR matrix, on the device.
//
// New relative to the reference (it has no sparse matrix format); the definition is the oracle's
// (oracle/hpcx_oracle.c: orc_sell_build) and the arrays are bit-identical to it:
//   perm       rows in windows of sigma, each window stably sorted by descending row length
//              (sigma <= 1: identity);
//   slice_len  longest row of every slice of C consecutive permuted rows;
//   slice_ptr  exclusive scan of slice_len * C (int64);
//   col / val  column-major slices, a row padded to its slice's length with value 0 and column
//              min(row, ncols - 1); lanes of the last slice beyond the last row hold 0 / 0.
// Round 1 did this on the host (std::stable_sort over row_ptr: 5-7 s for the 17-67 M rows of the
// benchmark matrices, plus a download and an upload of the whole matrix).  Here: one stable LSD
// radix sort of (window, maxlen - length) keys (cub), one thread per slice for the lengths, one
// scan, one thread per (slice, lane) for the fill — coalesced on the write side.
#include "spmv.cuh"

#include <cub/cub.cuh>

#include <algorithm>

using namespace hpcx;

namespace {

struct Scratch {
  void *p = nullptr;
  ~Scratch() { cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

inline int grid_for(hpcx_ctx ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16));
}

inline int bits_for(uint64_t v) {  // bits needed to hold values 0 .. v
  int b = 0;
  while (v) ++b, v >>= 1;
  return b;
}

// key = (window << len_bits) | (maxlen - length): ascending keys = windows in order, longest row first
__global__ void __launch_bounds__(256)
sell_key_kernel(const int32_t *__restrict__ rp, int64_t n, int64_t sigma, int32_t maxlen, int len_bits,
                uint64_t *__restrict__ keys, int32_t *__restrict__ rows) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t len = rp[i + 1] - rp[i];
    keys[i] = ((uint64_t)(i / sigma) << len_bits) | (uint64_t)(uint32_t)(maxlen - len);
    rows[i] = (int32_t)i;
  }
}

__global__ void __launch_bounds__(256)
sell_iota_kernel(int32_t *__restrict__ perm, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    perm[i] = (int32_t)i;
}

// one thread per slice; width[ns] = 0 so that the scan over ns + 1 entries ends with the total
__global__ void __launch_bounds__(256)
sell_slice_len_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ perm, int64_t n, int C, int64_t ns,
                      int32_t *__restrict__ slice_len, int64_t *__restrict__ width) {
  for (int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; s <= ns; s += (int64_t)gridDim.x * blockDim.x) {
    if (s == ns) {
      width[s] = 0;
      continue;
    }
    int32_t m = 0;
    const int64_t e = min((s + 1) * (int64_t)C, n);
    for (int64_t j = s * (int64_t)C; j < e; ++j) {
      const int32_t r = perm[j];
      m = max(m, rp[r + 1] - rp[r]);
    }
    slice_len[s] = m;
    width[s] = (int64_t)m * C;
  }
}

// one thread per (slice, lane) = per permuted row position j
template <typename V>
__global__ void __launch_bounds__(256)
sell_fill_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ col, const V *__restrict__ val,
                 const int32_t *__restrict__ perm, const int64_t *__restrict__ slice_ptr,
                 const int32_t *__restrict__ slice_len, int64_t n, int C, int64_t ns, int64_t ncols,
                 int32_t *__restrict__ scol, V *__restrict__ sval) {
  const int64_t total = ns * C;
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < total; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = j / C;
    const int lane = (int)(j - s * C);
    const int64_t base = slice_ptr[s] + lane;
    const int32_t len = slice_len[s];
    if (j < n) {
      const int32_t r = perm[j];
      const int32_t a = rp[r], rl = rp[r + 1] - a;
      const int32_t padc = (int32_t)min((int64_t)r, ncols - 1);
      for (int32_t k = 0; k < len; ++k) {
        const int64_t i = base + (int64_t)k * C;
        if (k < rl) {
          scol[i] = col[a + k];
          sval[i] = val[a + k];
        } else {
          scol[i] = padc;
          sval[i] = V(0);
        }
      }
    } else {
      for (int32_t k = 0; k < len; ++k) {
        const int64_t i = base + (int64_t)k * C;
        scol[i] = 0;
        sval[i] = V(0);
      }
    }
  }
}

}  // namespace

namespace hpcx {

// Builds A->d_perm, d_slice_len, d_slice_ptr, d_sell_col, d_sell_val from A's device CSR arrays
// (row_ptr int32).  maxlen = longest row (from the create-time scan of row_ptr).
hpcx_status sell_build(hpcx_csr A, int64_t maxlen) {
  hpcx_ctx ctx = A->ctx;
  cudaStream_t s = ctx->compute;
  const int64_t n = A->nrows;
  const int C = A->sell_c;
  const int64_t S = A->sell_sigma, ns = (n + C - 1) / C;
  A->sell_nslices = ns;
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_perm, (size_t)std::max<int64_t>(n, 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_slice_len, (size_t)std::max<int64_t>(ns, 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_slice_ptr, (size_t)(ns + 1) * 8));
  if (n == 0) {
    HPCX_CUDA(ctx, cudaMemsetAsync(A->d_slice_ptr, 0, 8, s));
    A->sell_padded = 0;
    HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_sell_col, 4));
    HPCX_CUDA(ctx, cudaMalloc(&A->d_sell_val, 8));
    return HPCX_OK;
  }
  if (S > 1) {
    Scratch keys, keys2, rows, tmp;
    HPCX_CUDA(ctx, keys.alloc((size_t)n * 8));
    HPCX_CUDA(ctx, keys2.alloc((size_t)n * 8));
    HPCX_CUDA(ctx, rows.alloc((size_t)n * 4));
    const int len_bits = std::max(1, bits_for((uint64_t)maxlen));
    const int end_bit = std::min(64, len_bits + bits_for((uint64_t)((n - 1) / S)));
    sell_key_kernel<<<grid_for(ctx, n), 256, 0, s>>>(A->d_row_ptr, n, S, (int32_t)maxlen, len_bits, (uint64_t *)keys.p,
                                                     (int32_t *)rows.p);
    size_t sb = 0;
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sb, (const uint64_t *)keys.p, (uint64_t *)keys2.p,
                                                   (const int32_t *)rows.p, A->d_perm, (int)n, 0, end_bit, s));
    HPCX_CUDA(ctx, tmp.alloc(sb));
    HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp.p, sb, (const uint64_t *)keys.p, (uint64_t *)keys2.p,
                                                   (const int32_t *)rows.p, A->d_perm, (int)n, 0, end_bit, s));
    HPCX_CUDA(ctx, cudaStreamSynchronize(s));  // the scratch of this scope is freed on exit
    ctx->launches += 2;
  } else {
    sell_iota_kernel<<<grid_for(ctx, n), 256, 0, s>>>(A->d_perm, n);
    ctx->launches += 1;
  }
  {
    Scratch width, tmp;
    HPCX_CUDA(ctx, width.alloc((size_t)(ns + 1) * 8));
    sell_slice_len_kernel<<<grid_for(ctx, ns + 1), 256, 0, s>>>(A->d_row_ptr, A->d_perm, n, C, ns, A->d_slice_len,
                                                                (int64_t *)width.p);
    size_t eb = 0;
    HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, eb, (const int64_t *)width.p, A->d_slice_ptr, (int)(ns + 1), s));
    HPCX_CUDA(ctx, tmp.alloc(eb));
    HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp.p, eb, (const int64_t *)width.p, A->d_slice_ptr, (int)(ns + 1), s));
    HPCX_CUDA(ctx, cudaMemcpyAsync(&A->sell_padded, A->d_slice_ptr + ns, 8, cudaMemcpyDeviceToHost, s));
    HPCX_CUDA(ctx, cudaStreamSynchronize(s));
    ctx->launches += 2;
  }
  HPCX_CUDA(ctx, cudaMalloc((void **)&A->d_sell_col, (size_t)std::max<int64_t>(A->sell_padded, 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc(&A->d_sell_val, (size_t)std::max<int64_t>(A->sell_padded, 1) * A->es));
  if (A->es == 8)
    sell_fill_kernel<double><<<grid_for(ctx, ns * C), 256, 0, s>>>(
        A->d_row_ptr, A->d_col, (const double *)A->d_val, A->d_perm, A->d_slice_ptr, A->d_slice_len, n, C, ns, A->ncols,
        A->d_sell_col, (double *)A->d_sell_val);
  else
    sell_fill_kernel<float><<<grid_for(ctx, ns * C), 256, 0, s>>>(
        A->d_row_ptr, A->d_col, (const float *)A->d_val, A->d_perm, A->d_slice_ptr, A->d_slice_len, n, C, ns, A->ncols,
        A->d_sell_col, (float *)A->d_sell_val);
  ctx->launches += 1;
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(ctx, HPCX_CUDA_ERROR, std::string("sell_build: ") + cudaGetErrorString(e));
  return HPCX_OK;
}

}  // namespace hpcx
