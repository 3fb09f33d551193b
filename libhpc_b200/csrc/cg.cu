// cg.cu — conjugate-gradient iteration around the CSR SpMV (the step
// either side of y = A*x in an iterative caller).  No reference counterpart (the reference has
// no solver); the CPU statement is oracle/hpcx_oracle.c: orc_cg_*.
//
// One iteration = three passes over HBM instead of the six a BLAS-1 formulation makes:
//   K1  q = A*p with <p, q> fused into the STREAM kernel's epilogue (per-CTA partials,
//       spmv.cu DOT instantiation) — no extra read of p or q;
//   K2  x += alpha*p ; r -= alpha*q ; partials of <r, r>          (4 reads, 2 writes per row)
//   K3  p = r + beta*p                                            (2 reads, 1 write per row)
// Scalars (alpha, beta, <r,r>) never leave the device: a one-CTA reduction adds the partials in
// a fixed order, the sum is all-reduced over the ranks (NCCL, one scalar) and a one-thread
// kernel derives alpha / beta and the convergence flag.  Once converged the remaining launches
// of the call are no-ops for x.  Every product and sum is rounded separately (--fmad=false).
//
// Algorithmic bytes per iteration (fp64): SpMV (12 B/nnz + 20 B/row) + 72 B/row.
#include "spmv.cuh"

#include <algorithm>

using namespace hpcx;

namespace hpcx {
hpcx_status allreduce_sum(hpcx_ctx ctx, void *buf, size_t count, int dtype, cudaStream_t s);  // dist.cu
}

struct hpcx_cg_s {
  hpcx_csr A = nullptr;
  int64_t n = 0;          // local rows
  int64_t npart = 0;      // <p, q> partials emitted by one product
  int nred = 0;           // CTAs (= partials) of the vector kernels
  void *r = nullptr, *p = nullptr, *q = nullptr;  // p aliases A's [owned | halo] buffer when distributed
  bool own_p = false;
  void *part = nullptr;   // max(npart, 2 * nred) scalars
  void *sc = nullptr;     // device scalars, see enum
  int32_t *flags = nullptr;  // [0] iterations done, [1] converged
  void *d_b = nullptr, *d_x = nullptr;  // staging for host operands
};

namespace {

enum { kRR = 0, kPQ, kAlpha, kBeta, kRRnew, kBB, kThresh, kTol2, kScalars = 8 };
constexpr int kRed = 1024;

// out[k] = sum of part[k*stride .. k*stride + n) in a fixed order, k < nout (one CTA).
template <typename T>
__global__ void __launch_bounds__(kRed) cg_reduce_kernel(const T *__restrict__ part, int64_t n,
                                                         int64_t stride, int nout, T *__restrict__ out) {
  __shared__ T sm[kRed];
  for (int k = 0; k < nout; ++k) {
    T s = T(0);
    for (int64_t i = threadIdx.x; i < n; i += kRed) s = s + part[k * stride + i];
    sm[threadIdx.x] = s;
    __syncthreads();
    for (int w = kRed / 2; w > 0; w >>= 1) {
      if ((int)threadIdx.x < w) sm[threadIdx.x] = sm[threadIdx.x] + sm[threadIdx.x + w];
      __syncthreads();
    }
    if (threadIdx.x == 0) out[k] = sm[0];
    __syncthreads();
  }
}

// phase 0: after the initial residual; 1: after <p, q>; 2: after the new <r, r>.
template <typename T>
__global__ void cg_scalar_kernel(T *__restrict__ sc, int32_t *__restrict__ flags, int phase) {
  if (phase == 0) {
    sc[kRR] = sc[kRRnew];
    sc[kThresh] = sc[kTol2] * sc[kBB];
    flags[0] = 0;
    flags[1] = sc[kRRnew] <= sc[kThresh] ? 1 : 0;
    return;
  }
  if (flags[1]) return;
  if (phase == 1) {
    // breakdown guard: <p, A p> must be positive for an SPD operator; zero, negative or NaN
    // stops the iteration (flag -1) instead of spreading NaNs through x
    if (!(sc[kPQ] > T(0))) {
      flags[1] = -1;
      return;
    }
    sc[kAlpha] = sc[kRR] / sc[kPQ];
  } else {
    sc[kBeta] = sc[kRRnew] / sc[kRR];
    sc[kRR] = sc[kRRnew];
    flags[0] += 1;
    if (sc[kRRnew] <= sc[kThresh]) flags[1] = 1;
  }
}

template <typename T>
__device__ __forceinline__ void block_partial(T v, T *__restrict__ out) {
  __shared__ T ws[8];
  for (int m = 16; m > 0; m >>= 1) v = v + __shfl_xor_sync(0xffffffffu, v, m);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    T t = ws[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) t = t + ws[w];
    *out = t;
  }
  __syncthreads();
}

// r = b - q ; p = r ; partials of <r, r> in part[0..grid) and of <b, b> in part[grid..2 grid)
template <typename T>
__global__ void __launch_bounds__(256) cg_init_kernel(const T *__restrict__ b, const T *__restrict__ q,
                                                      T *__restrict__ r, T *__restrict__ p, int64_t n,
                                                      T *__restrict__ part) {
  T rr = T(0), bb = T(0);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const T bi = b[i];
    const T ri = bi - q[i];
    r[i] = ri;
    p[i] = ri;
    rr = rr + ri * ri;
    bb = bb + bi * bi;
  }
  block_partial(rr, part + blockIdx.x);
  block_partial(bb, part + gridDim.x + blockIdx.x);
}

// x += alpha p ; r -= alpha q ; partials of <r, r>
template <typename T>
__global__ void __launch_bounds__(256) cg_update_xr_kernel(const T *__restrict__ sc,
                                                           const int32_t *__restrict__ flags,
                                                           const T *__restrict__ p, const T *__restrict__ q,
                                                           T *__restrict__ x, T *__restrict__ r, int64_t n,
                                                           T *__restrict__ part) {
  if (flags[1]) return;
  const T alpha = sc[kAlpha];
  T rr = T(0);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    x[i] = x[i] + alpha * p[i];
    const T ri = r[i] - alpha * q[i];
    r[i] = ri;
    rr = rr + ri * ri;
  }
  block_partial(rr, part + blockIdx.x);
}

// p = r + beta p
template <typename T>
__global__ void __launch_bounds__(256) cg_update_p_kernel(const T *__restrict__ sc,
                                                          const int32_t *__restrict__ flags,
                                                          const T *__restrict__ r, T *__restrict__ p,
                                                          int64_t n) {
  if (flags[1]) return;
  const T beta = sc[kBeta];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = r[i] + beta * p[i];
}

// q = A*p with the <p, q> partials in S->part (distributed or single-rank product).
hpcx_status product(hpcx_cg S, cudaStream_t s) {
  hpcx_csr A = S->A;
  A->dot_part = S->part;
  const int32_t kernel = A->kernel;
  A->kernel = HPCX_SPMV_STREAM;  // the fused dot lives in the STREAM kernel; its tables always exist
  hpcx_status rc = A->d_xfull ? hpcx_spmv_dist(A, S->p, S->q, s)
                              : hpcx_spmv(A, S->p, HPCX_MEM_DEVICE, S->q, HPCX_MEM_DEVICE, s);
  A->kernel = kernel;
  A->dot_part = nullptr;
  return rc;
}

template <typename T>
hpcx_status reduce_to(hpcx_cg S, int64_t n, int64_t stride, int nout, int slot, cudaStream_t s) {
  hpcx_ctx ctx = S->A->ctx;
  T *sc = (T *)S->sc;
  cg_reduce_kernel<T><<<1, kRed, 0, s>>>((const T *)S->part, n, stride, nout, sc + slot);
  if (hpcx_status rc = after_launch(ctx, "cg_reduce_kernel")) return rc;
  return allreduce_sum(ctx, sc + slot, (size_t)nout, S->A->dtype, s);
}

template <typename T>
hpcx_status solve_t(hpcx_cg S, const T *b, T *x, int32_t max_iters, double rel_tol, int32_t check_every,
                    cudaStream_t s) {
  hpcx_ctx ctx = S->A->ctx;
  T *sc = (T *)S->sc;
  T *r = (T *)S->r, *p = (T *)S->p, *q = (T *)S->q;
  const int64_t n = S->n;
  const T tol2 = (T)(rel_tol * rel_tol);
  HPCX_CUDA(ctx, cudaMemcpyAsync(sc + kTol2, &tol2, sizeof(T), cudaMemcpyHostToDevice, s));
  // r = b - A x0 ; p = r
  HPCX_CUDA(ctx, cudaMemcpyAsync(p, x, (size_t)n * sizeof(T), cudaMemcpyDeviceToDevice, s));
  if (hpcx_status rc = product(S, s)) return rc;
  cg_init_kernel<T><<<S->nred, 256, 0, s>>>(b, q, r, p, n, (T *)S->part);
  if (hpcx_status rc = after_launch(ctx, "cg_init_kernel")) return rc;
  // kRRnew and kBB are adjacent: one reduction launch, one all-reduce of two scalars
  static_assert(kBB == kRRnew + 1, "scalar layout");
  if (hpcx_status rc = reduce_to<T>(S, S->nred, S->nred, 2, kRRnew, s)) return rc;
  cg_scalar_kernel<T><<<1, 1, 0, s>>>(sc, S->flags, 0);
  if (hpcx_status rc = after_launch(ctx, "cg_scalar_kernel")) return rc;

  for (int it = 0; it < max_iters; ++it) {
    if (hpcx_status rc = product(S, s)) return rc;
    if (hpcx_status rc = reduce_to<T>(S, S->npart, 0, 1, kPQ, s)) return rc;
    cg_scalar_kernel<T><<<1, 1, 0, s>>>(sc, S->flags, 1);
    if (hpcx_status rc = after_launch(ctx, "cg_scalar_kernel")) return rc;
    cg_update_xr_kernel<T><<<S->nred, 256, 0, s>>>(sc, S->flags, p, q, x, r, n, (T *)S->part);
    if (hpcx_status rc = after_launch(ctx, "cg_update_xr_kernel")) return rc;
    if (hpcx_status rc = reduce_to<T>(S, S->nred, 0, 1, kRRnew, s)) return rc;
    cg_scalar_kernel<T><<<1, 1, 0, s>>>(sc, S->flags, 2);
    if (hpcx_status rc = after_launch(ctx, "cg_scalar_kernel")) return rc;
    cg_update_p_kernel<T><<<S->nred, 256, 0, s>>>(sc, S->flags, r, p, n);
    if (hpcx_status rc = after_launch(ctx, "cg_update_p_kernel")) return rc;
    if (check_every > 0 && (it + 1) % check_every == 0 && it + 1 < max_iters) {
      int32_t done = 0;  // every rank reads the same all-reduced scalars, so all stop together
      HPCX_CUDA(ctx, cudaMemcpyAsync(&done, S->flags + 1, 4, cudaMemcpyDeviceToHost, s));
      HPCX_CUDA(ctx, cudaStreamSynchronize(s));
      if (done) break;
    }
  }
  return HPCX_OK;
}

void free_cg(hpcx_cg S) {
  cudaFree(S->r);
  if (S->own_p) cudaFree(S->p);
  cudaFree(S->q);
  cudaFree(S->part);
  cudaFree(S->sc);
  cudaFree(S->flags);
  cudaFree(S->d_b);
  cudaFree(S->d_x);
  delete S;
}

}  // namespace

extern "C" {

hpcx_status hpcx_cg_create(hpcx_csr A, hpcx_cg *out) {
  if (!A || !out) return HPCX_INVALID_ARG;
  *out = nullptr;
  hpcx_ctx ctx = A->ctx;
  HPCX_REQUIRE(ctx, A->kernel == HPCX_SPMV_STREAM || A->kernel == HPCX_SPMV_HUB,
               "the CG step needs the STREAM kernel (fused <p, q>)");
  const bool dist = A->d_xfull != nullptr;
  const int64_t n = dist ? A->nloc : A->nrows;
  HPCX_REQUIRE(ctx, dist ? A->nrows == A->nloc : A->nrows == A->ncols, "CG needs a square matrix");
  hpcx_cg S = new (std::nothrow) hpcx_cg_s();
  if (!S) return fail(ctx, HPCX_OOM, "host allocation failed");
  S->A = A;
  S->n = n;
  S->npart = dist ? (int64_t)(A->h_blocks_int.size() + A->h_blocks_bnd.size()) + A->nlong_dist
                  : (int64_t)A->h_blocks.size() + A->nlong;
  S->nred = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8));
  const size_t es = (size_t)A->es, nb = (size_t)std::max<int64_t>(n, 1) * es;
  const size_t parts = (size_t)std::max<int64_t>(std::max<int64_t>(S->npart, 2 * (int64_t)S->nred), 1);
  cudaError_t e = cudaMalloc(&S->r, nb);
  if (e == cudaSuccess) e = cudaMalloc(&S->q, nb);
  if (e == cudaSuccess && !dist) {
    e = cudaMalloc(&S->p, nb);
    S->own_p = true;
  }
  if (dist) S->p = A->d_xfull;
  if (e == cudaSuccess) e = cudaMalloc(&S->part, parts * es);
  if (e == cudaSuccess) e = cudaMalloc(&S->sc, kScalars * es);
  if (e == cudaSuccess) e = cudaMalloc((void **)&S->flags, 8);
  // on the context's (non-blocking) compute stream and waited for: a legacy-stream cudaMemset would
  // not be ordered before the first solve's copies on that stream
  if (e == cudaSuccess) e = cudaMemsetAsync(S->sc, 0, kScalars * es, ctx->compute);
  if (e == cudaSuccess) e = cudaMemsetAsync(S->flags, 0, 8, ctx->compute);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->compute);
  if (e != cudaSuccess) {
    free_cg(S);
    return fail(ctx, e == cudaErrorMemoryAllocation ? HPCX_OOM : HPCX_CUDA_ERROR,
                std::string("hpcx_cg_create: ") + cudaGetErrorString(e));
  }
  *out = S;
  return HPCX_OK;
}

hpcx_status hpcx_cg_destroy(hpcx_cg S) {
  if (!S) return HPCX_INVALID_ARG;
  free_cg(S);
  return HPCX_OK;
}

hpcx_status hpcx_cg_solve(hpcx_cg S, const void *b, int32_t b_mem, void *x, int32_t x_mem,
                          int32_t max_iters, double rel_tol, int32_t check_every,
                          hpcx_cg_result *result, void *stream) {
  if (!S || !b || !x) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_cg_solve");
  hpcx_ctx ctx = S->A->ctx;
  HPCX_REQUIRE(ctx, max_iters >= 0 && rel_tol >= 0.0 && check_every >= 0, "bad iteration limits");
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;
  const size_t nb = (size_t)S->n * S->A->es;
  const void *db = b;
  void *dx = x;
  if (b_mem == HPCX_MEM_HOST) {
    if (!S->d_b) HPCX_CUDA(ctx, cudaMalloc(&S->d_b, std::max<size_t>(nb, 1)));
    HPCX_CUDA(ctx, cudaMemcpyAsync(S->d_b, b, nb, cudaMemcpyHostToDevice, s));
    db = S->d_b;
  }
  if (x_mem == HPCX_MEM_HOST) {
    if (!S->d_x) HPCX_CUDA(ctx, cudaMalloc(&S->d_x, std::max<size_t>(nb, 1)));
    HPCX_CUDA(ctx, cudaMemcpyAsync(S->d_x, x, nb, cudaMemcpyHostToDevice, s));
    dx = S->d_x;
  }
  hpcx_status rc = S->A->dtype == HPCX_F64
                       ? solve_t<double>(S, (const double *)db, (double *)dx, max_iters, rel_tol, check_every, s)
                       : solve_t<float>(S, (const float *)db, (float *)dx, max_iters, rel_tol, check_every, s);
  if (rc != HPCX_OK) return rc;
  if (x_mem == HPCX_MEM_HOST) HPCX_CUDA(ctx, cudaMemcpyAsync(x, dx, nb, cudaMemcpyDeviceToHost, s));
  if (result || !stream || x_mem == HPCX_MEM_HOST) {
    if (hpcx_status st = sync_and_check(ctx, s)) return st;
    if (result) {
      int32_t fl[2];
      HPCX_CUDA(ctx, cudaMemcpy(fl, S->flags, 8, cudaMemcpyDeviceToHost));
      double rr, bb;
      if (S->A->dtype == HPCX_F64) {
        double sc[kScalars];
        HPCX_CUDA(ctx, cudaMemcpy(sc, S->sc, sizeof sc, cudaMemcpyDeviceToHost));
        rr = sc[kRR];
        bb = sc[kBB];
      } else {
        float sc[kScalars];
        HPCX_CUDA(ctx, cudaMemcpy(sc, S->sc, sizeof sc, cudaMemcpyDeviceToHost));
        rr = sc[kRR];
        bb = sc[kBB];
      }
      result->iterations = fl[0];
      result->converged = fl[1];
      result->rr = rr;
      result->bb = bb;
    }
  }
  return HPCX_OK;
}

}  // extern "C"
