/*
 * hpcx.h — C ABI of libhpcx.so: the B200 (sm_100a) implementation of the libHPC
 * data-parallel sparse hot path (CSR SpMV y = A*x and ghost-cell stencil sweeps).
 *
 * Drop-in boundary.  The reference (Liupeter01/libHPC) has no FFI for this path: its
 * public surface is header-only C++ containers that callers index from their own
 * OpenMP loops.  Every entry point below therefore cites the reference interface or
 * loop it replaces (paths relative to the reference root):
 *
 *   grid containers   lib/hpc/include/HPCHighDimensionFlatArray.hpp:54-224
 *   aligned storage   lib/hpc/include/AlignedAlloc.hpp:29-158
 *   managed vectors   lib/gpu/util/include/cudaAllocator.hpp:39-83
 *   error convention  lib/gpu/util/include/cudaHelper.cuh:10-27,96-101
 *   x/y box sweeps    tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368,425-441,
 *                     444-457,575-601
 *   static partition  tests/test_hpc_benchmark/test_hpc_benchmark.cpp:1356,1369-1377
 *   stream overlap    lib/gpu/transfer_overlap_testsuite/src/cuda_tut_transfer_overlap.cu:92-103
 *
 * Conventions: plain pointers and sizes only; no exception crosses this boundary; every
 * call returns hpcx_status (0 = ok) and leaves a message retrievable with
 * hpcx_last_error_string().  `stream` arguments are cudaStream_t passed as void*
 * (NULL = the context's own compute stream, and the call synchronises before returning).
 * Handles are thread-compatible: one caller at a time per handle.
 * There is no CPU fallback: every compute entry point fails with HPCX_CUDA_ERROR when no
 * sm_100 device is usable.
 */
#ifndef HPCX_H_
#define HPCX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define HPCX_API __attribute__((visibility("default")))
#else
#define HPCX_API
#endif

typedef int32_t hpcx_status;
enum {
  HPCX_OK = 0,
  HPCX_INVALID_ARG = 1,
  HPCX_OOM = 2,          /* std::bad_alloc in the reference (AlignedAlloc.hpp:78-80) */
  HPCX_CUDA_ERROR = 3,   /* std::system_error, CUDA category (cudaHelper.cuh:24-27) */
  HPCX_NCCL_ERROR = 4,
  HPCX_UNSUPPORTED = 5,
  HPCX_OUT_OF_RANGE = 6  /* std::out_of_range of at() (HPCHighDimensionFlatArray.hpp:201-203) */
};

enum { HPCX_F32 = 0, HPCX_F64 = 1 };
enum { HPCX_MEM_HOST = 0, HPCX_MEM_DEVICE = 1, HPCX_MEM_MANAGED = 2 };

typedef struct hpcx_ctx_s *hpcx_ctx;
typedef struct hpcx_csr_s *hpcx_csr;
typedef struct hpcx_grid_s *hpcx_grid;

/* ------------------------------------------------------------------ context */

HPCX_API const char *hpcx_version(void);
/* CUDA devices visible to the process (what cudaGetDeviceCount reports; the reference asks the
 * runtime directly, lib/gpu/util/include/cudaHelper.cuh:96-101 being its only wrapper layer).
 * Creates no context.  Returns a negative number when the runtime reports an error. */
HPCX_API int32_t hpcx_device_count(void);
/* One context per process and device (one process per GPU). */
HPCX_API hpcx_status hpcx_init(int32_t device, hpcx_ctx *out);
HPCX_API hpcx_status hpcx_finalize(hpcx_ctx ctx);
HPCX_API const char *hpcx_last_error_string(hpcx_ctx ctx);
HPCX_API hpcx_status hpcx_synchronize(hpcx_ctx ctx);
/* Number of kernels of this library launched through ctx since hpcx_init. */
HPCX_API int64_t hpcx_launch_count(hpcx_ctx ctx);
/* Pinned host buffers (the reference hands kernels CudaAllocator-backed vectors,
 * cudaAllocator.hpp:39-83; pinned memory is what the copy pipeline needs). */
HPCX_API hpcx_status hpcx_host_alloc(hpcx_ctx ctx, size_t bytes, void **out);
HPCX_API hpcx_status hpcx_host_free(hpcx_ctx ctx, void *p);
HPCX_API hpcx_status hpcx_device_alloc(hpcx_ctx ctx, size_t bytes, void **out);
HPCX_API hpcx_status hpcx_device_free(hpcx_ctx ctx, void *p);
/* Context-free allocation by memory kind — what the reference's allocator strategies do
 * (CudaMemGPUOnly / CudaMemManaged, cudaAllocator.hpp:9-26): HPCX_MEM_DEVICE -> cudaMalloc,
 * HPCX_MEM_MANAGED -> cudaMallocManaged (valid on host and device; pass it to any entry point with
 * mem = HPCX_MEM_MANAGED), HPCX_MEM_HOST -> pinned host memory.  Backs include/hpcx/CudaAllocator.hpp.
 * Returns HPCX_OOM / HPCX_CUDA_ERROR; unlike the reference's deallocate (a no-op, :71-73) free frees. */
HPCX_API hpcx_status hpcx_mem_alloc(int32_t mem, size_t bytes, void **out);
HPCX_API hpcx_status hpcx_mem_free(int32_t mem, void *p);
/* Any direction (cudaMemcpyDefault).  Ordered after all work already issued to the device and
 * complete on return. */
HPCX_API hpcx_status hpcx_memcpy(hpcx_ctx ctx, void *dst, const void *src, size_t bytes);

/* ------------------------------------------------- ghost-padded grid (a1, a2)
 * hpcx_grid_desc describes hpc::HPCHighDimensionFlatArray<ndim,T,lo,hi>
 * (HPCHighDimensionFlatArray.hpp:54-57): dim[] are the logical sizes, slowest first
 * (row-major, last dimension fastest); lo/hi ghost cells on each side of every
 * dimension.  The HOST layout is exactly the reference's flat buffer:
 * stride[i] = prod_{j>i}(dim[j]+lo+hi), offset = sum stride[i]*(idx[i]+lo).
 * The DEVICE layout keeps the same logical indexing but pads the fastest dimension so
 * that interior x = 0 starts on a 32-byte boundary and the pitch is a multiple of
 * 32 bytes (TMA needs 16-byte strides; vector stores need sector-aligned rows). */
typedef struct {
  int32_t ndim;    /* 1..3 */
  int32_t dtype;   /* HPCX_F32 | HPCX_F64 */
  int32_t lo, hi;  /* Low_Bound, High_Bound */
  int64_t dim[3];  /* slowest .. fastest */
} hpcx_grid_desc;

/* compute_stride_and_total (HPCHighDimensionFlatArray.hpp:161-171). Host only. */
HPCX_API hpcx_status hpcx_grid_stride_total(const hpcx_grid_desc *d, int64_t *stride,
                                            int64_t *total);
/* unsafe_linearize (…:180-187) when checked == 0, safe_linearize (…:197-208) otherwise:
 * returns HPCX_OUT_OF_RANGE where at() throws std::out_of_range. Host only. */
HPCX_API hpcx_status hpcx_grid_linearize(const hpcx_grid_desc *d, const int64_t *idx,
                                         int32_t checked, int64_t *offset);

/* Value-initialised device grid (resize(dim, T{}), …:135-144: ghosts are zero). */
HPCX_API hpcx_status hpcx_grid_create(hpcx_ctx ctx, const hpcx_grid_desc *d, hpcx_grid *out);
HPCX_API hpcx_status hpcx_grid_destroy(hpcx_grid g);
/* Copy the whole padded array (ghosts included) host(reference layout) <-> device. */
HPCX_API hpcx_status hpcx_grid_upload(hpcx_grid g, const void *host_flat, void *stream);
HPCX_API hpcx_status hpcx_grid_download(hpcx_grid g, void *host_flat, void *stream);
/* Device layout introspection: x offset of interior x=0 inside a row, row pitch and
 * total element count, and the raw device pointer (data() of the device mirror). */
HPCX_API hpcx_status hpcx_grid_device_layout(hpcx_grid g, int64_t *xoff, int64_t *pitch,
                                             int64_t *total, void **dev_ptr);

/* --------------------------------------------------- stencil sweeps (a3, a4, a10) */
enum {
  /* b(y,x) = sum_{k=-r..r} a(y,x+k), k ascending from 0.0f: BM_x_blur,
   * BM_x_blur_tiling_simd_prefetch (test_hpc_benchmark.cpp:354-368, 425-441). */
  HPCX_STENCIL_BOX_X = 1,
  /* b(y,x) = sum_{k=-r..r} a(y+k,x): BM_y_blur, BM_YXx_blur_tiling_prefetch_streamed_IPL
   * (test_hpc_benchmark.cpp:444-457, 575-601). */
  HPCX_STENCIL_BOX_Y = 2,
  /* out = (in[z-1]+in[y-1]+in[x-1]+in[x+1]+in[y+1]+in[z+1]) * (1/6), adds in that
   * (ascending address) order; 3-D, ghost >= 1; ping-pong for steps > 1. */
  HPCX_STENCIL_JACOBI7 = 3
};
enum {
  HPCX_VARIANT_AUTO = 0,
  HPCX_VARIANT_TMA = 1,
  HPCX_VARIANT_DIRECT = 2,
  /* JACOBI7, fp32, steps >= 2: temporal blocking — each pass over HBM advances the field by two
   * sweeps (4 instead of 8 bytes of traffic per cell and sweep), results bit-identical to
   * single sweeps.  After the call `out` holds iterate `steps`; `in` holds an OLDER iterate
   * (steps-2, or steps-1 when steps is odd), not necessarily the previous one. */
  HPCX_VARIANT_TEMPORAL = 3
};
typedef struct {
  int32_t kind;
  int32_t radius;   /* BOX_*: r (the reference uses nblur = 8); JACOBI7: ignored */
  int32_t variant;  /* HPCX_VARIANT_* */
  int32_t zsplit;   /* JACOBI7: number of z segments per (x,y) tile; BOX_Y: output rows per
                       thread (16 or 32); 0 = auto */
} hpcx_stencil_desc;

/* Reads the (ghost-padded) grid `in`, writes the interior of `out`.
 * BOX_*: in has lo,hi >= radius along the swept axis, out may have any ghost width,
 *        steps must be 1.
 * JACOBI7: in/out same shape, lo,hi >= 1, both hold the same (Dirichlet) ghost values;
 *        after `steps` sweeps the newest iterate is in `out` and the previous one in
 *        `in` (the two handles swap device buffers when steps is even); see
 *        HPCX_VARIANT_TEMPORAL for the two-sweeps-per-pass variant. */
HPCX_API hpcx_status hpcx_stencil_sweep(hpcx_grid in, hpcx_grid out,
                                        const hpcx_stencil_desc *st, int32_t steps,
                                        void *stream);

/* 2-D convolution sweep: out(y,x) += sum_{l<kh} sum_{k<kw} in(y+l, x+k) *
 * weights[l*kw + k], l outer, k inner, every multiply and add rounded separately, accumulating
 * onto out's current value — BM_conv / BM_conv_block (test_hpc_benchmark.cpp:786-818: nkern =
 * 16, `kb` is <2,float,16>, `ka` has no ghosts).  in needs High_Bound >= kh-1 / kw-1; weights is
 * a HOST array of kh*kw values of the grids' dtype.  16 x 16 takes the shared-memory tiled
 * kernel (bit-identical to the reference loop); other extents (<= 1024 taps) or
 * HPCX_VARIANT_DIRECT the direct one. */
HPCX_API hpcx_status hpcx_conv2d(hpcx_grid in, hpcx_grid out, const void *weights, int32_t kh,
                                 int32_t kw, int32_t variant, void *stream);
/* No counterpart in the reference (it never touches a GPU on this path): the direct convolution
 * sweep of hpcx_conv2d for callers that are not allowed to allocate device memory while they run.
 * Typical cases: an application that has given all remaining HBM to its own arena after set-up,
 * so that any further cudaMalloc fails; a time-step loop that cannot tolerate the device-wide
 * synchronisation a cudaMalloc/cudaFree implies; a driver configuration without the stream-ordered
 * pool behind cudaMallocAsync (which hpcx_conv2d's direct path uses for its weights).
 * Same arguments, arithmetic and results as hpcx_conv2d(..., HPCX_VARIANT_DIRECT, stream); the
 * grids are created beforehand as usual.  The weights are staged in an array inside libhpcx.so's
 * device image instead of a per-call buffer, and the call makes no cudaMalloc, cudaFree or pool
 * request of any kind.
 *   limit    kh*kw*sizeof(dtype) <= 8192 bytes (1024 fp64 taps / 2048 fp32 taps); larger kernels
 *            are rejected with HPCX_INVALID_ARG and a message naming both sizes.
 *   cost     every GPU that loads the library holds those 8 KB until the process exits.
 *   set-up   the first call on each device looks up the array's device address with that device
 *            current; issue that first call outside any stream capture.
 *   sharing  all contexts and streams of a device use the same array: calls that overlap on the
 *            GPU race on the weights, so serialise them (one stream, or events between streams). */
HPCX_API hpcx_status hpcx_conv2d_static_scratch(hpcx_grid in, hpcx_grid out, const void *weights,
                                                int32_t kh, int32_t kw, void *stream);
/* out(x, y) = in(y, x) over the interiors of two 2-D grids of shapes (ny, nx) and (nx, ny), any
 * ghost widths — BM_transpose / BM_transpose_tiling (test_hpc_benchmark.cpp:669-700). */
HPCX_API hpcx_status hpcx_grid_transpose(hpcx_grid in, hpcx_grid out, void *stream);

/* Ghost-cell pack / unpack (a11).  axis: 0 = slowest.  side: 0 = low, 1 = high.
 * pack copies the `width` INTERIOR planes adjacent to that side into a dense buffer
 * (planes in ascending index order; each plane dense over the full padded extent of the
 * faster axes and the interior extent of the slower ones is also padded, i.e. a plane is
 * exactly what the neighbour needs as its ghost plane); unpack writes a buffer of the
 * same shape into the `width` GHOST planes on that side.  buf may be host or device. */
HPCX_API hpcx_status hpcx_grid_face_elems(hpcx_grid g, int32_t axis, int32_t width,
                                          int64_t *elems);
HPCX_API hpcx_status hpcx_grid_pack_face(hpcx_grid g, int32_t axis, int32_t side,
                                         int32_t width, void *buf, int32_t buf_mem,
                                         void *stream);
HPCX_API hpcx_status hpcx_grid_unpack_face(hpcx_grid g, int32_t axis, int32_t side,
                                           int32_t width, const void *buf, int32_t buf_mem,
                                           void *stream);

/* Halo index map of one face in the REFERENCE layout (flat offsets into data() of
 * hpc::HPCHighDimensionFlatArray, strides as compute_stride_and_total, …:161-171): the cells the
 * neighbour across (axis, side) needs — the `width` interior slices next to that side (send) —
 * and the ghost cells it fills here (recv).  Both are *nruns runs of *run_len consecutive elements;
 * send_start / recv_start (either may be NULL; *nruns entries) receive the first offset of every
 * run, slower coordinates in row-major order (ghost positions of the slower axes included, as in
 * hpcx_grid_pack_face).  For a slab split (axis 0) there is a single run: the face is contiguous
 * and is exchanged in place.  Host only, pure integer. */
HPCX_API hpcx_status hpcx_grid_export_halo_maps(const hpcx_grid_desc *d, int32_t axis, int32_t side,
                                                int32_t width, int64_t *nruns, int64_t *run_len,
                                                int64_t *send_start, int64_t *recv_start);

/* ------------------------------------------------------------- CSR matrix (a7-a9) */
typedef struct {
  int64_t nrows, ncols, nnz;
  const void *row_ptr;   /* nrows+1 entries */
  int32_t idx_bytes;     /* 4 (int32) or 8 (int64) row_ptr entries */
  int32_t dtype;         /* HPCX_F32 | HPCX_F64 */
  const int32_t *col_idx;/* nnz entries, ascending within a row */
  const void *values;    /* nnz entries */
  int32_t mem;           /* where row_ptr/col_idx/values live: HPCX_MEM_* */
  int32_t reserved;
} hpcx_csr_desc;

enum {
  HPCX_SPMV_AUTO = 0,
  HPCX_SPMV_STREAM = 1,  /* row-blocked, TMA-bulk staged tiles, adaptive row reduction */
  HPCX_SPMV_VECTOR = 2,  /* classic sub-warp vector-per-row */
  HPCX_SPMV_SELL = 3,    /* SELL-C-sigma repack */
  /* power-law matrices: nonzero-balanced persistent kernel, the x entries of the most referenced
   * ("hub") columns in shared memory, 128-bit loads of col/val, segmented warp-shuffle row sums.
   * HPCX_SPMV_AUTO picks it when more than a quarter of the nonzeros sit in rows longer than 256
   * and the hub columns receive at least 10 % of the gathers; STREAM otherwise. */
  HPCX_SPMV_HUB = 4
};
typedef struct {
  int32_t kernel;      /* HPCX_SPMV_* */
  int32_t tile_nnz;    /* STREAM: nonzeros staged per CTA (0 = default) */
  int32_t sell_c;      /* SELL slice height (0 = 32) */
  int32_t sell_sigma;  /* SELL sort window in rows (0 = 1024; 1 = no sorting) */
  int32_t lanes_per_row; /* VECTOR: 0 = from mean row length */
  int32_t short_row_max; /* STREAM: row blocks whose longest row is at most this take the
                            one-thread-per-row path (0 = 48) */
  int32_t reserved[2];
} hpcx_csr_opts;

/* Uploads (or adopts, when already on the device the arrays are copied device-to-device)
 * the matrix, builds row blocks / bins / optional SELL repack on the way. */
HPCX_API hpcx_status hpcx_csr_create(hpcx_ctx ctx, const hpcx_csr_desc *d,
                                     const hpcx_csr_opts *opts, hpcx_csr *out);
HPCX_API hpcx_status hpcx_csr_destroy(hpcx_csr A);
/* y = A*x.  x: ncols entries, y: nrows entries of A's dtype; each may be host, device or
 * managed memory (HPCX_MEM_*).  Host operands are staged through the copy pipeline; with BOTH
 * on the host the rows are cut into chunks that run as soon as the x entries they read have
 * landed, so that for banded operators H2D of x, the product and D2H of y overlap (pinned
 * buffers, hpcx_host_alloc, are what makes the copies asynchronous). */
HPCX_API hpcx_status hpcx_spmv(hpcx_csr A, const void *x, int32_t x_mem, void *y,
                               int32_t y_mem, void *stream);
/* Round trip of the device-resident CSR arrays (bit-exact structure check). Any pointer
 * may be NULL. row_ptr is written with the idx_bytes the matrix was created with. */
HPCX_API hpcx_status hpcx_csr_download(hpcx_csr A, void *row_ptr, int32_t *col_idx,
                                       void *values);
/* Row-block table of the STREAM kernel: per block {row0, nrows, part, lanes, p0, p1}
 * as 6 int64 (part >= 0 marks one chunk of a row longer than a tile). */
HPCX_API hpcx_status hpcx_csr_rowblocks(hpcx_csr A, int64_t *nblocks, int64_t *out,
                                        int64_t capacity);
/* Row-length histogram by ceil(log2(len)) bin: bins[0] = empty rows, bins[k] = rows
 * with 2^(k-1) < len <= 2^k (k = 1: len 1..2 -> see DESIGN.md), 34 entries. */
HPCX_API hpcx_status hpcx_csr_row_bins(hpcx_csr A, int64_t *bins34);
/* SELL-C-sigma arrays (only for matrices created with HPCX_SPMV_SELL): sizes first
 * (any output pointer may be NULL), then data. */
HPCX_API hpcx_status hpcx_csr_sell_info(hpcx_csr A, int64_t *nslices, int64_t *padded_nnz,
                                        int32_t *c, int32_t *sigma);
HPCX_API hpcx_status hpcx_csr_sell_download(hpcx_csr A, int32_t *perm, int64_t *slice_ptr,
                                            int32_t *slice_len, int32_t *col, void *val);

/* Device-side assembly of a CSR matrix from an edge (COO) list: entries sorted by
 * (row, col, position in the list), equal (row, col) merged by adding their values in list
 * order (one rounding per add), exactly the CPU builder's rule.  rows/cols/vals live in
 * `mem` (HPCX_MEM_HOST or device/managed).  The three result arrays are DEVICE memory owned
 * by the caller (hpcx_device_free); hand them to hpcx_csr_create with mem = HPCX_MEM_DEVICE.
 * nedges < 2^31; row_ptr is written with idx_bytes (4 or 8) per entry. */
typedef struct {
  int64_t nnz;
  void *row_ptr;
  int32_t *col_idx;
  void *values;
} hpcx_csr_arrays;
HPCX_API hpcx_status hpcx_csr_from_coo(hpcx_ctx ctx, int64_t nrows, int64_t nedges,
                                       const int32_t *rows, const int32_t *cols,
                                       const void *vals, int32_t dtype, int32_t mem,
                                       int32_t idx_bytes, hpcx_csr_arrays *out);
/* Sparse-grid bridge: the 5-point operator over the ACTIVE cells of a sparse
 * 2-D grid — the (x, y) coordinates sparse::RootGrid::foreach enumerates (reference
 * lib/sparse/include/RootGrid.hpp:20-22, 62-75), in any order, repeats allowed.  Rows are the
 * distinct cells in (x, y) lexicographic order (x-major, like DenseBlock's m_block[x*B + y],
 * DenseBlock.hpp:24-28); row r holds 4 on the diagonal and -1 for each of (x-1,y), (x,y-1),
 * (x,y+1), (x+1,y) that is active (an inactive neighbour is a Dirichlet zero), columns
 * ascending.  `map` (may be NULL) receives device arrays: cell_row[ncells] = row of each input
 * cell, row_x / row_y[nrows] = coordinates of each row.  All result arrays are DEVICE memory
 * owned by the caller (hpcx_device_free). */
typedef struct {
  int64_t nrows;
  int32_t *cell_row;
  int32_t *row_x;
  int32_t *row_y;
} hpcx_cell_map;
HPCX_API hpcx_status hpcx_csr_from_cells(hpcx_ctx ctx, int64_t ncells, const int32_t *xs,
                                         const int32_t *ys, int32_t mem, int32_t dtype,
                                         int32_t idx_bytes, hpcx_csr_arrays *out,
                                         hpcx_cell_map *map);
/* Counter-based generators of the synthetic configs, on the device (bit-identical to the
 * host generators).  R-MAT edge k: `scale` quadrant choices with
 * (a,b,c,d) = (.57,.19,.19,.05) from u(seed, k*scale + level), value (float)(2u(seed+1,k)-1).
 * Field: out[i] = 2u(seed, first+i) - 1 in dtype. */
HPCX_API hpcx_status hpcx_synth_rmat_coo(hpcx_ctx ctx, int32_t scale, int64_t nedges,
                                         uint64_t seed, int32_t *rows_dev, int32_t *cols_dev,
                                         float *vals_dev);
HPCX_API hpcx_status hpcx_synth_field(hpcx_ctx ctx, uint64_t seed, int64_t first, int64_t n,
                                      int32_t dtype, void *out_dev);

/* ------------------------------------------------- partitioner + halo maps (a12) */
/* Contiguous row blocks: bounds[p] .. bounds[p+1] for p in [0, nparts).
 * mode 0: balanced by rows with the reference's static chunking idiom
 *         chunk = ceil(n/nparts), [p*chunk, min((p+1)*chunk, n))
 *         (test_hpc_benchmark.cpp:1356,1371-1372);
 * mode 1: balanced by nonzeros (first row whose prefix nnz reaches p*nnz/nparts).
 * row_ptr is a host array. */
HPCX_API hpcx_status hpcx_partition_rows(int64_t nrows, const void *row_ptr,
                                         int32_t idx_bytes, int32_t nparts, int32_t mode,
                                         int64_t *bounds);
/* z-slab split of n planes: same static chunking as mode 0. */
HPCX_API hpcx_status hpcx_partition_slabs(int64_t n, int32_t nparts, int64_t *bounds);

/* Local view of rank `rank`'s row block of a host CSR matrix (square, ncols == nrows,
 * x partitioned like the rows).  Columns are renumbered [owned | halo], halo sorted by
 * (owner, global column).  Two-call protocol: call with all output pointers NULL to get
 * the sizes, allocate, call again.
 *   local_row_ptr[nloc+1] (int64), local_col[nnz_loc] (int32), value positions unchanged
 *   halo_gcol[nhalo] (int64 global column of each halo slot)
 *   recv_counts[nparts], send_counts[nparts] (int64), send_idx[nsend] (int32 local owned
 *   index to gather, grouped by destination rank ascending, ascending within a group)
 *   row_kind[nloc] (uint8: 0 interior = all columns owned, 1 boundary) */
typedef struct {
  int64_t nloc, nnz_loc, nhalo, nsend;
} hpcx_halo_sizes;
HPCX_API hpcx_status hpcx_csr_localize(int64_t nrows, const void *row_ptr, int32_t idx_bytes,
                                       const int32_t *col_idx, const int64_t *bounds,
                                       int32_t nparts, int32_t rank, hpcx_halo_sizes *sizes,
                                       int64_t *local_row_ptr, int32_t *local_col,
                                       int64_t *halo_gcol, int64_t *recv_counts,
                                       int64_t *send_counts, int32_t *send_idx,
                                       uint8_t *row_kind);

/* The same tables from this rank's OWN rows only (no global matrix anywhere, no table of nrows
 * entries): row_ptr has nloc + 1 entries starting at 0, col_global holds GLOBAL column indices.
 * Step 1 (local) yields local_col, halo_gcol, recv_counts, row_kind; sizes->nsend is -1 until
 * step 2.  The ranks then trade integer lists — every rank sends, to each owner q, the recv_counts[q]
 * entries of its halo_gcol that q owns (contiguous, ascending) — and step 2 turns what the others
 * asked of this rank (req_counts[q] columns from rank q, concatenated for q ascending) into
 * send_idx; send_counts = req_counts.  Bit-identical to hpcx_csr_localize on the global matrix.
 * libhpc_b200/dist.py: localize_rows_distributed does the exchange over torch.distributed. */
HPCX_API hpcx_status hpcx_csr_localize_rows(int64_t nrows_global, int64_t nloc, const void *row_ptr,
                                            int32_t idx_bytes, const int32_t *col_global,
                                            const int64_t *bounds, int32_t nparts, int32_t rank,
                                            hpcx_halo_sizes *sizes, int32_t *local_col,
                                            int64_t *halo_gcol, int64_t *recv_counts, uint8_t *row_kind);
HPCX_API hpcx_status hpcx_halo_send_lists(const int64_t *bounds, int32_t nparts, int32_t rank,
                                          const int64_t *req_counts, const int64_t *requests,
                                          int32_t *send_idx);

/* ------------------------------------------------ multi-GPU transport (a13)
 * One process per GPU.  The caller bootstraps (e.g. through torch.distributed or MPI):
 * rank 0 calls hpcx_comm_unique_id, broadcasts the 128 bytes, every rank calls
 * hpcx_comm_init.  Peer (NVLink) windows for the fused halo path are exchanged with
 * hpcx_comm_ipc_export / hpcx_comm_ipc_open. */
HPCX_API hpcx_status hpcx_comm_unique_id(hpcx_ctx ctx, void *id128);
HPCX_API hpcx_status hpcx_comm_init(hpcx_ctx ctx, int32_t nranks, int32_t rank,
                                    const void *id128);
HPCX_API hpcx_status hpcx_comm_finalize(hpcx_ctx ctx);
HPCX_API hpcx_status hpcx_comm_barrier(hpcx_ctx ctx);
HPCX_API hpcx_status hpcx_comm_ipc_export(hpcx_ctx ctx, const void *dev_ptr, void *handle64);
HPCX_API hpcx_status hpcx_comm_ipc_open(hpcx_ctx ctx, const void *handle64, void **peer_ptr);
HPCX_API hpcx_status hpcx_comm_ipc_close(hpcx_ctx ctx, void *peer_ptr);
/* The fused (peer-store) transports wait on flag words a neighbour writes.  By default such a
 * wait has no bound, like an NCCL receive.  hpcx_comm_set_timeout (or HPCX_HALO_TIMEOUT_MS in the
 * environment at hpcx_init) gives it one: a kernel that waits longer records the timeout in a
 * host-visible word and TRAPS — it never goes on to read a stale ghost plane or x-halo — so the
 * CUDA context is lost and every later call fails.  hpcx_comm_check reports (and clears) that word
 * without touching the device: HPCX_NCCL_ERROR after a timeout, HPCX_OK otherwise.  Callers that
 * pass their own stream use it after synchronising; the library's own synchronising exits
 * (stream == NULL calls, hpcx_synchronize, hpcx_cg_solve, hpcx_stencil_dist_destroy) call it. */
HPCX_API hpcx_status hpcx_comm_set_timeout(hpcx_ctx ctx, int64_t milliseconds);
HPCX_API hpcx_status hpcx_comm_check(hpcx_ctx ctx);

/* Distributed stencil: every rank owns a slab along axis 0 (z planes of a 3-D grid, rows of a 2-D
 * grid) as a grid pair created with hpcx_grid_create; neighbours are rank-1 / rank+1 (no periodic
 * wrap: the outermost ghost slices stay Dirichlet).
 * BOX_X / BOX_Y (the reference's 17-point sweeps over y-slabs, steps == 1, transport 0): BOX_X needs
 * no exchange; BOX_Y receives the `radius` ghost rows of `in` from the neighbours on the comm
 * stream while the rows that need none of them are swept, then sweeps the two boundary strips.
 * JACOBI7 transport:
 *   0  NCCL send/recv of the boundary planes on the comm stream, overlapped with the interior
 *      sweep on the compute stream (three launches per sweep; a whole call is captured into a
 *      CUDA graph once and replayed, HPCX_NCCL_GRAPHS=0 turns that off);
 *   1  fused: one kernel per sweep whose boundary CTAs store their plane straight into the
 *      neighbour's ghost plane over NVLink (CUDA IPC peer windows) and publish a sweep-number
 *      flag that the neighbour's next sweep acquires before reading the ghost (fp32 TMA sweep
 *      only; plan creation is collective over the communicator).
 * HPCX_VARIANT_TEMPORAL (either transport; plan creation is collective): two sweeps per pass,
 * TWO ghost planes traded per side and pass — every slab needs lo,hi >= 2 and >= 2 planes
 * (HPCX_INVALID_ARG on every rank otherwise).  The neighbour's boundary plane is advanced
 * redundantly on this rank; results stay bit-identical to single sweeps.  Transport 1 pushes the
 * planes with peer stores + flags (no NCCL on the data path): from inside the pass kernel on
 * slabs of at most 128 planes, with a small kernel behind each pass on thicker ones. */
typedef struct hpcx_halo_plan_s *hpcx_halo_plan;
HPCX_API hpcx_status hpcx_stencil_dist_plan(hpcx_ctx ctx, hpcx_grid in, hpcx_grid out,
                                            const hpcx_stencil_desc *st, int32_t transport,
                                            hpcx_halo_plan *out_plan);
/* BOX_X / BOX_Y over a 2-D BLOCK split: the ranks form a py x px grid (py * px == communicator
 * size, rank = ry * px + rx), every rank owns one block of the global 2-D grid as a grid pair whose
 * ghost width covers `radius` on all four sides; blocks of one block row share their y extent,
 * blocks of one block column their x extent.  BOX_Y trades `radius` ghost rows with ranks
 * rank -/+ px (contiguous rows, in place); BOX_X trades `radius` ghost columns with ranks
 * rank -/+ 1 of the same block row: the strided faces are packed and unpacked on the device
 * (hpcx_grid_pack_face's kernel) around one grouped NCCL send/recv.  Either exchange runs on the
 * comm stream beside the part of the block that needs no ghosts; the two boundary strips follow.
 * px == 1 is the y-slab split of hpcx_stencil_dist_plan.  Results are bit-identical to the
 * single-GPU sweep of the global grid, i.e. to the reference loops BM_x_blur / BM_y_blur
 * (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368, 444-457) run on the whole grid; the
 * split follows the reference's static chunking idiom (same file, 1356, 1371-1372) along both
 * axes.  Collective over the communicator. */
HPCX_API hpcx_status hpcx_stencil_dist_plan_blocks(hpcx_ctx ctx, hpcx_grid in, hpcx_grid out,
                                                   const hpcx_stencil_desc *st, int32_t py, int32_t px,
                                                   hpcx_halo_plan *out_plan);
HPCX_API hpcx_status hpcx_stencil_dist_sweep(hpcx_halo_plan plan, int32_t steps, void *stream);
HPCX_API hpcx_status hpcx_stencil_dist_destroy(hpcx_halo_plan plan);

/* Distributed SpMV: A was created from the LOCAL matrix produced by hpcx_csr_localize
 * (ncols = nloc + nhalo).  hpcx_csr_set_halo attaches the exchange lists; hpcx_spmv_dist
 * then runs {gather boundary x | interior rows} -> exchange -> boundary rows.
 * x_local has nloc entries (device), y_local nloc entries (device). */
HPCX_API hpcx_status hpcx_csr_set_halo(hpcx_csr A, int64_t nloc, int64_t nhalo,
                                       const int64_t *recv_counts, const int64_t *send_counts,
                                       const int32_t *send_idx, const uint8_t *row_kind);
/* Selects how hpcx_spmv_dist moves the boundary x entries (collective over the communicator,
 * call after hpcx_csr_set_halo):
 *   0  gather kernel -> grouped NCCL send/recv on the comm stream -> boundary rows (default);
 *   1  fused: the gather kernel stores every boundary entry straight into its slot of the
 *      destination rank's [owned | halo] buffer over NVLink (CUDA IPC peer windows) and
 *      releases a product-number flag there; the boundary rows start once the flags of all
 *      source ranks have been acquired.  One stream, no NCCL call on the data path. */
HPCX_API hpcx_status hpcx_csr_halo_transport(hpcx_csr A, int32_t transport);
/* Device buffer [owned | halo] the distributed product reads x from; callers that write
 * their owned x entries straight into it and pass it as x_local avoid a device copy. */
HPCX_API hpcx_status hpcx_csr_x_buffer(hpcx_csr A, void **x_full);
HPCX_API hpcx_status hpcx_spmv_dist(hpcx_csr A, const void *x_local, void *y_local,
                                    void *stream);

/* Row-block product with a gathered x (the fallback it prescribes for C5; no
 * reference counterpart — the reference's only partition idiom is the static chunking of
 * test_hpc_benchmark.cpp:1356,1369-1377).  A is this rank's contiguous row block created with
 * GLOBAL column ids (nrows = bounds[rank+1] - bounds[rank], ncols = bounds[nranks]; no
 * hpcx_csr_set_halo), so hpcx_csr_create picks the kernel from the block's own structure — HUB for
 * power-law blocks.  bounds (nranks + 1 entries, bounds[0] = 0, bounds[nranks] = ncols) says which
 * slice of x each rank holds; it need not follow the row blocks.  x_local: this rank's slice
 * (device); x_full: device buffer of ncols entries that receives every rank's slice (x_local may
 * already be x_full + bounds[rank]); y_local: the nrows results of this rank's rows.  Equal slices
 * travel by ncclAllGather, unequal ones by one grouped ncclSend/ncclRecv (slower: the owner of the
 * longest slice sends it nranks - 1 times); either way on `stream`, followed by the product on
 * the same stream.  For matrices whose rows reference most of x this moves fewer bytes than the
 * halo lists of hpcx_csr_set_halo and keeps the single-GPU kernel choice. */
HPCX_API hpcx_status hpcx_spmv_allgather(hpcx_csr A, const int64_t *bounds, const void *x_local,
                                         void *x_full, void *y_local, void *stream);

/* ------------------------------------------------ single process, several GPUs
 * Every caller the reference has is a single-process program driving its containers from its own
 * (OpenMP) loops (test_hpc_benchmark.cpp:356); the only multi-stream idiom it contains is driven
 * by one thread too (cuda_tut_transfer_overlap.cu:92-103).  hpcx_init_multi gives such a program N
 * GPUs without MPI or torchrun: one context per device inside the process, peer access enabled
 * between all of them, everything driven from ONE host thread.  A host CSR matrix / host grid is
 * partitioned by the library (hpcx_partition_rows + hpcx_csr_localize; even split of the slowest
 * grid axis), and products / sweeps run through the same kernels and the same fused peer-store
 * transport as the one-process-per-GPU path — the neighbours' buffers are plain peer pointers, and
 * neither NCCL nor CUDA IPC is involved.  dev_ids == NULL selects devices 0 .. ndev-1.
 * Not covered here: the CG solver across devices (its scalar all-reduce is NCCL's).
 * Flag waits of these contexts are bounded by default (120 s, or HPCX_HALO_TIMEOUT_MS /
 * hpcx_comm_set_timeout on hpcx_multi_ctx): with one thread queueing every device's kernels, a
 * launch that fails in the middle of a round would otherwise leave the kernels already queued —
 * and the thread's next synchronise — waiting for ever.
 * The fused transports assume that the CTAs of a grid are dispatched in block-index order (the
 * pushing CTAs come first in every such grid); CUDA does not guarantee that order, it is the
 * observed behaviour of the hardware scheduler, and the bounded wait is what turns a violation
 * into an error instead of a hang. */
typedef struct hpcx_multi_s *hpcx_multi;
HPCX_API hpcx_status hpcx_init_multi(int32_t ndev, const int32_t *dev_ids, hpcx_multi *out);
HPCX_API hpcx_status hpcx_multi_finalize(hpcx_multi m);
HPCX_API int32_t hpcx_multi_size(hpcx_multi m);
HPCX_API hpcx_ctx hpcx_multi_ctx(hpcx_multi m, int32_t rank);
HPCX_API const char *hpcx_multi_last_error_string(hpcx_multi m);
/* y = A*x across all devices.  The descriptor names a HOST (or managed) square matrix; rows are cut
 * into contiguous blocks (partition_mode as hpcx_partition_rows: 0 by rows, 1 by nonzeros), each
 * device gets its block in [owned | halo] numbering with the halo tables, and the exchange lists
 * are wired for the fused transport.  hpcx_multi_spmv takes HOST x and y of the global size. */
typedef struct hpcx_multi_csr_s *hpcx_multi_csr;
HPCX_API hpcx_status hpcx_multi_csr_create(hpcx_multi m, const hpcx_csr_desc *d, const hpcx_csr_opts *opts,
                                           int32_t partition_mode, hpcx_multi_csr *out);
HPCX_API hpcx_status hpcx_multi_csr_destroy(hpcx_multi_csr A);
HPCX_API hpcx_status hpcx_multi_csr_bounds(hpcx_multi_csr A, int64_t *bounds /* hpcx_multi_size + 1 */);
HPCX_API hpcx_status hpcx_multi_spmv(hpcx_multi_csr A, const void *x_host, void *y_host);
/* A ghost-padded grid (2-D or 3-D) split along its slowest axis into one slab per device.  Upload
 * hands every slab its ghost slices too (they are its neighbours' boundary slices in the host
 * array); download writes the owned slices back.  hpcx_multi_stencil_sweep: BOX_X / BOX_Y (one
 * sweep, slabs independent) and JACOBI7 over `steps` sweeps with the ghost planes traded by peer
 * stores inside the sweep kernel (HPCX_VARIANT_TEMPORAL included); results are bit-identical to
 * hpcx_stencil_sweep on one device. */
typedef struct hpcx_multi_grid_s *hpcx_multi_grid;
HPCX_API hpcx_status hpcx_multi_grid_create(hpcx_multi m, const hpcx_grid_desc *d, hpcx_multi_grid *out);
HPCX_API hpcx_status hpcx_multi_grid_destroy(hpcx_multi_grid g);
HPCX_API hpcx_status hpcx_multi_grid_upload(hpcx_multi_grid g, const void *host_flat);
HPCX_API hpcx_status hpcx_multi_grid_download(hpcx_multi_grid g, void *host_flat);
HPCX_API hpcx_status hpcx_multi_stencil_sweep(hpcx_multi_grid in, hpcx_multi_grid out,
                                              const hpcx_stencil_desc *st, int32_t steps);

/* The same product with HOST operands (nloc entries each; pinned memory makes the copies
 * asynchronous).  With the fused transport the copies, the exchange and the product overlap: the
 * last piece of x goes up first, then the others; once both ends of the owned block have landed one
 * launch pushes the boundary entries to the neighbours and computes the boundary rows, and the
 * interior rows follow in chunks whose rows of y travel back while later chunks compute.
 * Synchronous.  Collective in the sense of hpcx_spmv_dist: every rank must make the call. */
HPCX_API hpcx_status hpcx_spmv_dist_host(hpcx_csr A, const void *x_host, void *y_host, void *stream);

/* ------------------------------------------------ CG iteration around SpMV
 * Conjugate gradients for a symmetric positive definite A created with the STREAM kernel —
 * either a single-rank matrix (nrows == ncols) or a distributed one (hpcx_csr_set_halo done:
 * b and x are then the rank's nloc owned entries and the dot products are all-reduced over
 * the communicator; every rank must make the same calls).  One iteration is three passes over
 * HBM: q = A*p with <p, q> fused into the SpMV kernel, {x += alpha p, r -= alpha q, <r, r>},
 * {p = r + beta p}; the scalars stay on the device.  Starts from the x passed in.  Stops after
 * max_iters iterations or once ||r||^2 <= rel_tol^2 * ||b||^2 (later launches of the call then
 * leave x untouched); check_every > 0 makes the host poll the convergence flag every that
 * many iterations so that a converged solve returns early.  `result` (may be NULL) receives
 * the iteration count, the flag and the recurrence values of ||r||^2 and ||b||^2. */
typedef struct hpcx_cg_s *hpcx_cg;
typedef struct {
  int32_t iterations;
  int32_t converged;  /* 1 converged, 0 not yet, -1 breakdown (<p, A p> not positive: A not SPD, or NaN) */
  double rr;  /* ||r||^2 from the recurrence */
  double bb;  /* ||b||^2 */
} hpcx_cg_result;
HPCX_API hpcx_status hpcx_cg_create(hpcx_csr A, hpcx_cg *out);
HPCX_API hpcx_status hpcx_cg_destroy(hpcx_cg S);
HPCX_API hpcx_status hpcx_cg_solve(hpcx_cg S, const void *b, int32_t b_mem, void *x, int32_t x_mem,
                                   int32_t max_iters, double rel_tol, int32_t check_every,
                                   hpcx_cg_result *result, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* HPCX_H_ */
