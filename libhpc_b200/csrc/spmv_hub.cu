// spmv_hub.cu — CSR SpMV for power-law matrices (BASELINE configs[4], R-MAT): the HUB kernel.
//
// No reference implementation exists; the arithmetic contract is the oracle's
// (products rounded separately; the summation order differs from ascending p, so results agree
// with the oracle within the 1e-5 / 1e-12 bars of north_star, and are deterministic: no atomics
// on the value path).
//
// Why another kernel.  On the unpermuted R-MAT every x gather touches its own 128-byte line
// (0.93 distinct lines per gather, measured on the C5 matrix), and an SM's load/store unit retires
// about one line ("wavefront") per cycle: 263 M gathers / 148 SMs / 1.9 GHz = 0.95 ms for the
// gathers alone, whatever the DRAM does (scripts/ubench/gather_bench.cu measures exactly that:
// 0.95 ms with nothing but the gathers; the texture path is slower, 1.2 ms).  The only lever is
// to take gathers off the global-load path.  The column popularity of a power-law matrix is as
// skewed as its row lengths: the 16,384 most referenced columns of C5 receive 29 % of all
// gathers.  So:
//   * create time (device): histogram of col_idx, radix sort by descending count, the top K
//     columns become "hub" columns; a private copy of col_idx encodes them as 0x80000000 | slot;
//   * run time: ONE persistent CTA of 1024 threads per SM copies the K hub entries of x into
//     shared memory once (64 KB; the rest of the 228 KB stays L1 for the non-hub gathers — a
//     larger slice was measured slower for that reason) and serves hub gathers from there
//     (a random 32-lane shared-memory read costs ~3.5 bank cycles instead of up to 32 lines).
//
// Row-length binning (create time, device): rows longer than 256 entries ("long": 65 % of C5's
// nonzeros in 1.1 % of its rows) and rows of 1..256 entries ("short") are compacted into two
// private CSR parts, columns hub-encoded; empty rows (9.4 M of C5's 16.8 M) appear in neither and
// y is cleared by a memset first.
//
// Long part — split by nonzeros, not by rows (rows reach 238,875 entries).  The nonzeros are cut
// into tiles of 256 = two half tiles of 128; each warp owns a contiguous run of tiles; in a half
// tile a lane owns 4 consecutive nonzeros, so col and val arrive by ONE fully coalesced 128-bit
// load each ("128-bit vectorised, coalesced loads of values and column indices", north_star: a
// warp reads 512 contiguous bytes = 4 lines per instruction).  Row boundaries come from a bitmap
// (one bit per nonzero = "first of its row") and a per-tile count of the row starts before it.  A
// long row cannot start twice inside a lane's 4 entries, so a lane holds at most one (head, tail)
// pair; the warp finishes rows with a 5-step segmented shuffle scan, half tiles without a row
// start take a plain shuffle reduction, and the running sum of the row that is still open is
// carried in a register from tile to tile.  Only rows that cross a warp's range go through a
// small fix-up kernel (one partial per warp).
// Short part — one thread per row, ascending p (the oracle's order), same hub slice.
#include "spmv.cuh"

#include <cub/cub.cuh>

using namespace hpcx;

namespace {

constexpr int kHubThreads = 1024;
constexpr int kHubWarps = kHubThreads / 32;
constexpr int kTile = 256;  // nonzeros per warp step: 8 per lane
// 63 KB: with the 1 KB the runtime reserves per CTA this fits the 64 KB shared-memory carve-out and
// leaves 164 KB of L1 to the non-hub gathers (64 KB would be rounded up to the 100 KB carve-out)
constexpr int kHubBytes = 63 * 1024;

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// streaming (evict-first) 128-bit loads: the matrix is read once and must not push x out of L1
__device__ __forceinline__ int4 ld_stream(const int4 *p) { return __ldcs(p); }
__device__ __forceinline__ float4 ld_stream(const float4 *p) { return __ldcs(p); }
__device__ __forceinline__ double2 ld_stream(const double2 *p) { return __ldcs(p); }

// 4 consecutive values starting at element `at` (16-byte aligned for float, 32 for double)
__device__ __forceinline__ void load4(const float *__restrict__ val, int64_t at, float (&v)[4]) {
  const float4 a = ld_stream(reinterpret_cast<const float4 *>(val + at));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void load4(const double *__restrict__ val, int64_t at, double (&v)[4]) {
  const double2 a = ld_stream(reinterpret_cast<const double2 *>(val + at));
  const double2 b = ld_stream(reinterpret_cast<const double2 *>(val + at) + 1);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

struct HubArgs {
  // long part: tiles of 256 nonzeros, segments = long rows
  const int32_t *colL;        // hub-encoded columns, padded to whole tiles
  const void *valL;
  const uint8_t *rowstart;    // bit i: nonzero i of the long part is the first of its row
  const int32_t *tile_seg0;   // ntiles + 1: long rows that start before each tile
  const int32_t *lrow;        // original row index of each long row
  int64_t nnzL, ntiles;
  // short part: SELL-32 over the short rows sorted by descending length
  const int32_t *colS;        // column-major slices of 32 rows, hub-encoded
  const void *valS;
  const int64_t *sbase;       // first slot of each slice
  const uint16_t *slen;       // length of each (sorted) short row
  const int32_t *srow;        // original row index of each (sorted) short row
  int64_t nshort;
  const int32_t *hub_col;
  int K;
};

template <typename T>
__device__ __forceinline__ T warp_sum(T s) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) s = add_rn(s, __shfl_xor_sync(0xffffffffu, s, m));
  return s;
}

// One half tile (128 nonzeros, 4 per lane, at most one row start per lane) that contains a row
// start (`any` = ballot of the lanes that have one).  `carry` is the sum of the row open when the
// half tile begins (the lanes' private partials already folded in); `seg0` the row starts before it.
template <typename T>
__device__ __forceinline__ void hub_half_tile(const T (&p)[4], unsigned f4, unsigned any, int lane, int seg0,
                                              int range_seg0, const int32_t *__restrict__ lrow,
                                              T *__restrict__ y, T *__restrict__ head_slot, T &carry) {
  // the lane's run: head = entries before its row start, tail = from the row start on
  const int pos = f4 ? __ffs(f4) - 1 : 4;
  T head = T(0), tail = T(0);
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (u < pos) head = add_rn(head, p[u]);
    else tail = add_rn(tail, p[u]);
  }
  // segmented inclusive scan over the lanes: a lane with a row start restarts the sum with its tail
  T I = f4 ? tail : head;
  const unsigned below = any & ((2u << lane) - 1u);   // flagged lanes <= this one
  const int start = below ? 31 - __clz(below) : 0;     // first lane of this lane's scan segment
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const T o = __shfl_up_sync(0xffffffffu, I, d);
    if (lane - d >= start) I = add_rn(I, o);
  }
  const T prev = __shfl_up_sync(0xffffffffu, I, 1);  // from the previous flagged lane's tail up to lane - 1
  if (f4) {
    const unsigned before = any & ((1u << lane) - 1u);
    T closed = lane > 0 ? add_rn(prev, head) : head;
    if (before == 0) closed = add_rn(carry, closed);  // the row was open when the half tile began
    const int seg = seg0 + __popc(before) - 1;        // the row this lane's start closes
    if (seg < range_seg0)
      *head_slot = closed;  // it began in an earlier warp's range (or before nonzero 0): fix-up kernel
    else
      y[lrow[seg]] = closed;
  }
  carry = __shfl_sync(0xffffffffu, I, 31);  // tail of the last flagged lane + everything after it
}

template <typename T>
__global__ void __launch_bounds__(kHubThreads, 1)
spmv_hub_kernel(const __grid_constant__ HubArgs H, const T *__restrict__ x, T *__restrict__ y,
                T *__restrict__ head_out, T *__restrict__ tail_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T *sx = reinterpret_cast<T *>(smem_raw);
  for (int i = threadIdx.x; i < H.K; i += kHubThreads) sx[i] = __ldg(x + H.hub_col[i]);
  __syncthreads();
  auto gx = [&](int c) -> T { return c < 0 ? sx[c & 0x7fffffff] : __ldg(x + c); };

  // ---------------------------------------------------------------- long rows
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * kHubWarps + (threadIdx.x >> 5);
  const int64_t nw = (int64_t)gridDim.x * kHubWarps;
  const int64_t t0 = H.ntiles * w / nw, t1 = H.ntiles * (w + 1) / nw;
  const int range_seg0 = H.tile_seg0[t0];
  const T *__restrict__ valL = static_cast<const T *>(H.valL);
  // The row that is open: `carry` (warp-uniform) + the lanes' private partials `mine`.  Tiles
  // without a row start — most of them, known from tile_seg0 alone — only add into `mine`: no
  // bitmap load, no shuffle, no dependency between tiles; the partials are folded into `carry`
  // when a row start finally closes the row.
  T carry = T(0), mine = T(0);
  int seg0 = range_seg0;
  // the columns of a tile are fetched one tile ahead, so that its gathers can start at once
  // (3 % on C5; the kernel is bound by memory latency at 32 warps per SM)
  int4 c0 = make_int4(0, 0, 0, 0), c1 = c0;
  if (t0 < t1) {
    c0 = ld_stream(reinterpret_cast<const int4 *>(H.colL + t0 * kTile + lane * 4));
    c1 = ld_stream(reinterpret_cast<const int4 *>(H.colL + t0 * kTile + 128 + lane * 4));
  }
  for (int64_t t = t0; t < t1; ++t) {
    const int64_t e0 = t * kTile + lane * 4, e1 = e0 + 128;
    const int seg1 = H.tile_seg0[t + 1];  // warp-uniform, 32 tiles per line
    T v0[4], v1[4], p0[4], p1[4];
    load4(valL, e0, v0);
    load4(valL, e1, v1);
    const T x0[4] = {gx(c0.x), gx(c0.y), gx(c0.z), gx(c0.w)};
    const T x1[4] = {gx(c1.x), gx(c1.y), gx(c1.z), gx(c1.w)};
    if (t + 1 < t1) {  // the next tile's columns travel while this tile's gathers do
      c0 = ld_stream(reinterpret_cast<const int4 *>(H.colL + e0 + kTile));
      c1 = ld_stream(reinterpret_cast<const int4 *>(H.colL + e1 + kTile));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      p0[u] = mul_rn(v0[u], x0[u]);
      p1[u] = mul_rn(v1[u], x1[u]);
    }
    if (t == H.ntiles - 1) {  // the last tile may run past the part's end: the padding must not contribute
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (e0 + u >= H.nnzL) p0[u] = T(0);
        if (e1 + u >= H.nnzL) p1[u] = T(0);
      }
    }
    const T s0 = add_rn(add_rn(p0[0], p0[1]), add_rn(p0[2], p0[3]));
    const T s1 = add_rn(add_rn(p1[0], p1[1]), add_rn(p1[2], p1[3]));
    if (seg1 == seg0) {
      mine = add_rn(mine, add_rn(s0, s1));
      continue;
    }
    // ---- a row starts in this tile.  Bitmap bytes: half tile h, lanes 2j and 2j+1 share byte t*32 + h*16 + j
    const unsigned b0 = H.rowstart[t * 32 + (lane >> 1)], b1 = H.rowstart[t * 32 + 16 + (lane >> 1)];
    const unsigned f0 = (b0 >> ((lane & 1) * 4)) & 15u, f1 = (b1 >> ((lane & 1) * 4)) & 15u;
    const unsigned any0 = __ballot_sync(0xffffffffu, f0 != 0), any1 = __ballot_sync(0xffffffffu, f1 != 0);
    if (any0 == 0) {
      mine = add_rn(mine, s0);
    } else {
      carry = add_rn(carry, warp_sum(mine));
      mine = T(0);
      hub_half_tile<T>(p0, f0, any0, lane, seg0, range_seg0, H.lrow, y, head_out + w, carry);
    }
    if (any1 == 0) {
      mine = add_rn(mine, s1);
    } else {
      carry = add_rn(carry, warp_sum(mine));
      mine = T(0);
      hub_half_tile<T>(p1, f1, any1, lane, seg0 + __popc(any0), range_seg0, H.lrow, y, head_out + w, carry);
    }
    seg0 = seg1;
  }
  carry = add_rn(carry, warp_sum(mine));
  if (lane == 0) {
    if (H.tile_seg0[t1] == range_seg0)
      head_out[w] = carry;  // no row start in the whole range (or an empty range: carry = 0)
    else
      tail_out[w] = carry;
  }

  // ---------------------------------------------------------------- short rows
  // SELL-32 over the short rows sorted by descending length: slice s holds 32 rows column-major,
  // so every load of col / val is one coalesced line per warp; a lane walks its own row in
  // ascending p (the oracle's order, so these rows are bit-identical to it).
  const T *__restrict__ valS = static_cast<const T *>(H.valS);
  const int64_t nthreads = (int64_t)gridDim.x * kHubThreads;
  const int64_t nslots = (H.nshort + 31) / 32 * 32;
  // (Four slices per warp step — four independent rows per lane — was tried for the one- and
  // two-entry rows and lost: 0.355 vs 0.234 ms on the 33..256 rows, no gain on the 1..32 rows,
  // which sit at the one-line-per-cycle bound of their gathers plus their scattered stores of y.)
  for (int64_t k = (int64_t)blockIdx.x * kHubThreads + threadIdx.x; k < nslots; k += nthreads) {
    const int len = k < H.nshort ? H.slen[k] : 0;
    const int64_t base = H.sbase[k >> 5] + lane;
    T s = T(0);
    int j = 0;
    for (; j + 4 <= len; j += 4) {  // four gathers in flight per lane
      int c[4];
      T v[4], xv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) c[u] = __ldcs(H.colS + base + (j + u) * 32);
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldcs(valS + base + (j + u) * 32);
#pragma unroll
      for (int u = 0; u < 4; ++u) xv[u] = gx(c[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u) s = add_rn(s, mul_rn(v[u], xv[u]));
    }
    for (; j < len; ++j) s = add_rn(s, mul_rn(__ldcs(valS + base + j * 32), gx(__ldcs(H.colS + base + j * 32))));
    if (k < H.nshort) y[H.srow[k]] = s;
  }
}

// Rows that cross warp ranges: the row open at the end of range w gets its tail plus the heads of
// the following ranges up to and including the first one that contains a row start.
template <typename T>
__global__ void spmv_hub_fixup_kernel(const int32_t *__restrict__ tile_seg0, const int32_t *__restrict__ lrow,
                                      const T *__restrict__ head_out, const T *__restrict__ tail_out,
                                      T *__restrict__ y, int64_t ntiles, int64_t nw) {
  const int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (w >= nw) return;
  auto seg_at = [&](int64_t r) { return tile_seg0[ntiles * r / nw]; };
  const int s0 = seg_at(w), s1 = seg_at(w + 1);
  if (s1 == s0) return;  // no row start here: this range only contributes a head
  T total = tail_out[w];
  int64_t v = w + 1;
  while (v < nw) {
    total = add_rn(total, head_out[v]);
    if (seg_at(v + 1) != seg_at(v)) break;
    ++v;
  }
  y[lrow[s1 - 1]] = total;
}

// ------------------------------------------------------------------ create-time kernels
constexpr int kShortMax = 256;  // rows of 1..256 entries are "short" (SELL-32, one thread per row)

__global__ void __launch_bounds__(256) hub_hist_kernel(const int32_t *__restrict__ col, int64_t nnz,
                                                       uint32_t *__restrict__ freq) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(freq + col[i], 1u);
}
__global__ void __launch_bounds__(256) hub_iota_kernel(int32_t *__restrict__ v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    v[i] = (int32_t)i;
}
__global__ void __launch_bounds__(256) hub_slot_kernel(const int32_t *__restrict__ hub_col, int K,
                                                       int32_t *__restrict__ slot) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < K) slot[hub_col[k]] = k;
}
__global__ void __launch_bounds__(256) hub_cover_kernel(const uint32_t *__restrict__ freq_sorted, int K,
                                                        unsigned long long *__restrict__ out) {
  unsigned long long s = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K; i += gridDim.x * blockDim.x) s += freq_sorted[i];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}
// row classes: long (> kShortMax entries), short (1 .. kShortMax), empty
__global__ void __launch_bounds__(256) hub_class_kernel(const int32_t *__restrict__ rp, int64_t nrows,
                                                        uint8_t *__restrict__ is_long, uint8_t *__restrict__ is_short) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    const int len = rp[r + 1] - rp[r];
    is_long[r] = len > kShortMax;
    is_short[r] = len > 0 && len <= kShortMax;
  }
}
// lens[k] = length of row rows[k]; lens[n] = 0 (so an exclusive scan over n + 1 entries ends with the total)
__global__ void __launch_bounds__(256) hub_lens_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ rows,
                                                       int64_t n, int32_t *__restrict__ lens) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k <= n; k += (int64_t)gridDim.x * blockDim.x)
    lens[k] = k < n ? rp[rows[k] + 1] - rp[rows[k]] : 0;
}
__device__ __forceinline__ int32_t hub_encode(int32_t c, const int32_t *__restrict__ slot) {
  const int32_t s = slot[c];
  return s >= 0 ? (int32_t)(0x80000000u | (uint32_t)s) : c;
}
// compaction of the selected rows' entries into a private part, columns hub-encoded.
// LANES = 32: one warp per row (long rows); LANES = 1: one thread per row (short rows).
template <typename V, int LANES>
__global__ void __launch_bounds__(256)
hub_copy_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ col, const V *__restrict__ val,
                const int32_t *__restrict__ rows, const int32_t *__restrict__ prp, int64_t n,
                const int32_t *__restrict__ slot, int32_t *__restrict__ pcol, V *__restrict__ pval,
                uint32_t *__restrict__ bitmap) {
  const int64_t gt = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int l = (int)(gt % LANES);
  for (int64_t k = gt / LANES; k < n; k += (int64_t)gridDim.x * blockDim.x / LANES) {
    const int src = rp[rows[k]], len = rp[rows[k] + 1] - src, dst = prp[k];
    for (int i = l; i < len; i += LANES) {
      pcol[dst + i] = hub_encode(col[src + i], slot);
      pval[dst + i] = val[src + i];
    }
    if (bitmap && l == 0) atomicOr(bitmap + (dst >> 5), 1u << (dst & 31));  // row start
  }
}
// sort keys of the short rows: kShortMax - length, so that an ascending stable sort yields descending
// length with ascending row index among equals
__global__ void __launch_bounds__(256) hub_shortkey_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ rows,
                                                           int64_t n, int32_t *__restrict__ keys) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
    keys[k] = kShortMax - (rp[rows[k] + 1] - rp[rows[k]]);
}
// slen[k] = length of sorted short row k; width32[s] = 32 * (length of the first = longest row of slice s)
__global__ void __launch_bounds__(256) hub_slices_kernel(const int32_t *__restrict__ keys_sorted, int64_t n, int64_t nslices,
                                                         uint16_t *__restrict__ slen, int64_t *__restrict__ width32) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    const int len = kShortMax - keys_sorted[k];
    slen[k] = (uint16_t)len;
    if ((k & 31) == 0) width32[k >> 5] = 32 * (int64_t)len;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) width32[nslices] = 0;
}
// column-major fill of the SELL-32 slices: sorted short row k -> lane k % 32 of slice k / 32
template <typename V>
__global__ void __launch_bounds__(256)
hub_sell_fill_kernel(const int32_t *__restrict__ rp, const int32_t *__restrict__ col, const V *__restrict__ val,
                     const int32_t *__restrict__ rows, const uint16_t *__restrict__ slen,
                     const int64_t *__restrict__ sbase, int64_t n, const int32_t *__restrict__ slot,
                     int32_t *__restrict__ pcol, V *__restrict__ pval) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    const int src = rp[rows[k]], len = slen[k];
    const int64_t dst = sbase[k >> 5] + (k & 31);
    for (int j = 0; j < len; ++j) {
      pcol[dst + j * 32] = hub_encode(col[src + j], slot);
      pval[dst + j * 32] = val[src + j];
    }
  }
}
// tile_seg0[t] = number of long rows whose first nonzero lies before tile t
__global__ void __launch_bounds__(256) hub_tileseg_kernel(const int32_t *__restrict__ lrp, int32_t nl, int64_t ntiles,
                                                          int32_t *__restrict__ tile_seg0) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t > ntiles) return;
  const int64_t first = t * kTile;
  int32_t lo = 0, hi = nl;  // first long row k with lrp[k] >= first
  while (lo < hi) {
    const int32_t mid = lo + (hi - lo) / 2;
    if ((int64_t)lrp[mid] >= first) hi = mid; else lo = mid + 1;
  }
  tile_seg0[t] = lo;
}

struct Scratch {
  void *p = nullptr;
  ~Scratch() { cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

inline int grid_for(hpcx_ctx ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16));
}

// rows with flag set -> list (ascending) and its CSR offsets over the original row lengths
hpcx_status select_rows(hpcx_ctx ctx, const int32_t *rp, const uint8_t *flags, int64_t nrows, int32_t **rows,
                        int32_t **prp, int64_t *count, int64_t *total, cudaStream_t s) {
  Scratch nsel, tmp, tmp2;
  HPCX_CUDA(ctx, cudaMalloc((void **)rows, (size_t)std::max<int64_t>(nrows, 1) * 4));
  HPCX_CUDA(ctx, nsel.alloc(4));
  cub::CountingInputIterator<int32_t> it(0);
  size_t sb = 0;
  HPCX_CUDA(ctx, cub::DeviceSelect::Flagged(nullptr, sb, it, flags, *rows, (int32_t *)nsel.p, (int)nrows, s));
  HPCX_CUDA(ctx, tmp.alloc(sb));
  HPCX_CUDA(ctx, cub::DeviceSelect::Flagged(tmp.p, sb, it, flags, *rows, (int32_t *)nsel.p, (int)nrows, s));
  int32_t n = 0;
  HPCX_CUDA(ctx, cudaMemcpyAsync(&n, nsel.p, 4, cudaMemcpyDeviceToHost, s));
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  *count = n;
  if (!prp) return HPCX_OK;
  HPCX_CUDA(ctx, cudaMalloc((void **)prp, (size_t)(n + 1) * 4));
  hub_lens_kernel<<<grid_for(ctx, n + 1), 256, 0, s>>>(rp, *rows, n, *prp);
  size_t eb = 0;
  HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, eb, *prp, *prp, n + 1, s));
  HPCX_CUDA(ctx, tmp2.alloc(eb));
  HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp2.p, eb, *prp, *prp, n + 1, s));
  int32_t tot = 0;
  HPCX_CUDA(ctx, cudaMemcpyAsync(&tot, *prp + n, 4, cudaMemcpyDeviceToHost, s));
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  *total = tot;
  ctx->launches += 1;
  return HPCX_OK;
}

template <typename V>
hpcx_status copy_parts(hpcx_csr A, const int32_t *slot, cudaStream_t s) {
  hpcx_ctx ctx = A->ctx;
  HubTables &H = A->hub;
  if (H.nlong)
    hub_copy_kernel<V, 32><<<grid_for(ctx, H.nlong * 32), 256, 0, s>>>(A->d_row_ptr, A->d_col, (const V *)A->d_val, H.lrow,
                                                                       H.lrp, H.nlong, slot, H.colL, (V *)H.valL,
                                                                       (uint32_t *)H.rowstart);
  if (H.nshort)
    hub_sell_fill_kernel<V><<<grid_for(ctx, H.nshort), 256, 0, s>>>(A->d_row_ptr, A->d_col, (const V *)A->d_val, H.srow,
                                                                    H.slen, H.sbase, H.nshort, slot, H.colS, (V *)H.valS);
  ctx->launches += 2;
  return HPCX_OK;
}

}  // namespace

namespace hpcx {

void hub_free(hpcx_csr A) {
  HubTables &H = A->hub;
  cudaFree(H.colL);
  cudaFree(H.valL);
  cudaFree(H.rowstart);
  cudaFree(H.tile_seg0);
  cudaFree(H.lrow);
  cudaFree(H.lrp);
  cudaFree(H.colS);
  cudaFree(H.valS);
  cudaFree(H.sbase);
  cudaFree(H.slen);
  cudaFree(H.srow);
  cudaFree(H.hub_col);
  cudaFree(H.head_out);
  cudaFree(H.tail_out);
  H = HubTables{};
}

// Builds the hub tables from the device-resident CSR arrays.  On return A->hub.coverage is the
// fraction of the nonzeros whose column is a hub column (what decides HPCX_SPMV_AUTO).
hpcx_status hub_build(hpcx_csr A) {
  hpcx_ctx ctx = A->ctx;
  HubTables &H = A->hub;
  const int64_t nnz = A->nnz, nrows = A->nrows, ncols = A->ncols;
  if (nnz == 0 || ncols == 0) return HPCX_OK;
  HPCX_RANGE("hub_build");
  cudaStream_t s = ctx->compute;
  H.K = (int)std::min<int64_t>(kHubBytes / A->es, ncols);

  // ---- column popularity -> hub columns
  Scratch freq, idx, freq2, idx2, tmp, slot, cover;
  HPCX_CUDA(ctx, freq.alloc((size_t)ncols * 4));
  HPCX_CUDA(ctx, idx.alloc((size_t)ncols * 4));
  HPCX_CUDA(ctx, freq2.alloc((size_t)ncols * 4));
  HPCX_CUDA(ctx, idx2.alloc((size_t)ncols * 4));
  HPCX_CUDA(ctx, slot.alloc((size_t)ncols * 4));
  HPCX_CUDA(ctx, cover.alloc(8));
  HPCX_CUDA(ctx, cudaMemsetAsync(freq.p, 0, (size_t)ncols * 4, s));
  HPCX_CUDA(ctx, cudaMemsetAsync(slot.p, 0xff, (size_t)ncols * 4, s));
  HPCX_CUDA(ctx, cudaMemsetAsync(cover.p, 0, 8, s));
  hub_hist_kernel<<<grid_for(ctx, nnz), 256, 0, s>>>(A->d_col, nnz, (uint32_t *)freq.p);
  hub_iota_kernel<<<grid_for(ctx, ncols), 256, 0, s>>>((int32_t *)idx.p, ncols);
  size_t tb = 0;
  HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(nullptr, tb, (const uint32_t *)freq.p, (uint32_t *)freq2.p,
                                                           (const int32_t *)idx.p, (int32_t *)idx2.p, (int)ncols, 0, 32, s));
  HPCX_CUDA(ctx, tmp.alloc(tb));
  // stable: equal counts keep ascending column order, so the hub set is deterministic
  HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairsDescending(tmp.p, tb, (const uint32_t *)freq.p, (uint32_t *)freq2.p,
                                                           (const int32_t *)idx.p, (int32_t *)idx2.p, (int)ncols, 0, 32, s));
  HPCX_CUDA(ctx, cudaMalloc((void **)&H.hub_col, (size_t)H.K * 4));
  HPCX_CUDA(ctx, cudaMemcpyAsync(H.hub_col, idx2.p, (size_t)H.K * 4, cudaMemcpyDeviceToDevice, s));
  hub_cover_kernel<<<8, 256, 0, s>>>((const uint32_t *)freq2.p, H.K, (unsigned long long *)cover.p);
  hub_slot_kernel<<<(H.K + 255) / 256, 256, 0, s>>>(H.hub_col, H.K, (int32_t *)slot.p);
  ctx->launches += 4;
  unsigned long long covered = 0;
  HPCX_CUDA(ctx, cudaMemcpyAsync(&covered, cover.p, 8, cudaMemcpyDeviceToHost, s));

  // ---- row-length binning: long rows / short rows / empty rows
  Scratch is_long, is_short;
  HPCX_CUDA(ctx, is_long.alloc((size_t)nrows));
  HPCX_CUDA(ctx, is_short.alloc((size_t)nrows));
  hub_class_kernel<<<grid_for(ctx, nrows), 256, 0, s>>>(A->d_row_ptr, nrows, (uint8_t *)is_long.p, (uint8_t *)is_short.p);
  ctx->launches += 1;
  if (hpcx_status st = select_rows(ctx, A->d_row_ptr, (const uint8_t *)is_long.p, nrows, &H.lrow, &H.lrp, &H.nlong, &H.nnzL, s))
    return st;
  {
    // short rows: ascending list, then a stable sort by descending length -> SELL-32 slices
    int32_t *rows_asc = nullptr;
    int64_t dummy = 0;
    if (hpcx_status st = select_rows(ctx, A->d_row_ptr, (const uint8_t *)is_short.p, nrows, &rows_asc, nullptr, &H.nshort, &dummy, s))
      return st;
    Scratch asc, keys, keys2, stmp, w32;
    asc.p = rows_asc;  // freed with the scope
    const int64_t ns = H.nshort, nslices = (ns + 31) / 32;
    HPCX_CUDA(ctx, keys.alloc((size_t)ns * 4));
    HPCX_CUDA(ctx, keys2.alloc((size_t)ns * 4));
    HPCX_CUDA(ctx, cudaMalloc((void **)&H.srow, (size_t)std::max<int64_t>(ns, 1) * 4));
    HPCX_CUDA(ctx, cudaMalloc((void **)&H.slen, (size_t)std::max<int64_t>(ns, 1) * 2));
    HPCX_CUDA(ctx, cudaMalloc((void **)&H.sbase, (size_t)(nslices + 1) * 8));
    HPCX_CUDA(ctx, w32.alloc((size_t)(nslices + 1) * 8));
    HPCX_CUDA(ctx, cudaMemsetAsync(H.sbase, 0, (size_t)(nslices + 1) * 8, s));
    if (ns) {
      hub_shortkey_kernel<<<grid_for(ctx, ns), 256, 0, s>>>(A->d_row_ptr, rows_asc, ns, (int32_t *)keys.p);
      size_t sb = 0;
      HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, sb, (const int32_t *)keys.p, (int32_t *)keys2.p,
                                                     (const int32_t *)rows_asc, H.srow, (int)ns, 0, 9, s));
      HPCX_CUDA(ctx, stmp.alloc(sb));
      HPCX_CUDA(ctx, cub::DeviceRadixSort::SortPairs(stmp.p, sb, (const int32_t *)keys.p, (int32_t *)keys2.p,
                                                     (const int32_t *)rows_asc, H.srow, (int)ns, 0, 9, s));
      hub_slices_kernel<<<grid_for(ctx, ns), 256, 0, s>>>((const int32_t *)keys2.p, ns, nslices, H.slen, (int64_t *)w32.p);
      Scratch etmp;
      size_t eb = 0;
      HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, eb, (const int64_t *)w32.p, H.sbase, (int)(nslices + 1), s));
      HPCX_CUDA(ctx, etmp.alloc(eb));
      HPCX_CUDA(ctx, cub::DeviceScan::ExclusiveSum(etmp.p, eb, (const int64_t *)w32.p, H.sbase, (int)(nslices + 1), s));
      HPCX_CUDA(ctx, cudaMemcpyAsync(&H.nnzS, H.sbase + nslices, 8, cudaMemcpyDeviceToHost, s));
      HPCX_CUDA(ctx, cudaStreamSynchronize(s));  // scratch of this scope is freed on exit
      ctx->launches += 2;
    }
  }
  H.coverage = (double)covered / (double)nnz;
  H.ntiles = (H.nnzL + kTile - 1) / kTile;
  const int64_t nallocL = std::max<int64_t>(H.ntiles, 1) * kTile;
  // ---- the two private parts (hub-encoded columns, values), padded to whole tiles, and the
  // row-start bitmap of the long part
  HPCX_CUDA(ctx, cudaMalloc((void **)&H.colL, (size_t)nallocL * 4));
  HPCX_CUDA(ctx, cudaMalloc(&H.valL, (size_t)nallocL * A->es));
  HPCX_CUDA(ctx, cudaMalloc((void **)&H.colS, (size_t)std::max<int64_t>(H.nnzS, 1) * 4));
  HPCX_CUDA(ctx, cudaMalloc(&H.valS, (size_t)std::max<int64_t>(H.nnzS, 1) * A->es));
  HPCX_CUDA(ctx, cudaMalloc((void **)&H.rowstart, (size_t)nallocL / 8));
  HPCX_CUDA(ctx, cudaMemsetAsync(H.colL + H.nnzL, 0, (size_t)(nallocL - H.nnzL) * 4, s));
  HPCX_CUDA(ctx, cudaMemsetAsync((char *)H.valL + (size_t)H.nnzL * A->es, 0, (size_t)(nallocL - H.nnzL) * A->es, s));
  HPCX_CUDA(ctx, cudaMemsetAsync(H.rowstart, 0, (size_t)nallocL / 8, s));
  if (hpcx_status st = A->es == 8 ? copy_parts<uint64_t>(A, (const int32_t *)slot.p, s)
                                  : copy_parts<uint32_t>(A, (const int32_t *)slot.p, s))
    return st;
  HPCX_CUDA(ctx, cudaMalloc((void **)&H.tile_seg0, (size_t)(H.ntiles + 1) * 4));
  hub_tileseg_kernel<<<(unsigned)((H.ntiles + 1 + 255) / 256), 256, 0, s>>>(H.lrp, (int32_t)H.nlong, H.ntiles, H.tile_seg0);
  ctx->launches += 1;
  H.nwarps = (int64_t)ctx->sm_count * kHubWarps;
  HPCX_CUDA(ctx, cudaMalloc(&H.head_out, (size_t)H.nwarps * A->es));
  HPCX_CUDA(ctx, cudaMalloc(&H.tail_out, (size_t)H.nwarps * A->es));
  HPCX_CUDA(ctx, cudaMemsetAsync(H.head_out, 0, (size_t)H.nwarps * A->es, s));
  HPCX_CUDA(ctx, cudaMemsetAsync(H.tail_out, 0, (size_t)H.nwarps * A->es, s));
  HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(ctx, HPCX_CUDA_ERROR, std::string("hub_build: ") + cudaGetErrorString(e));
  H.ready = true;
  return HPCX_OK;
}

template <typename T>
static hpcx_status hub_launch_t(hpcx_csr A, const T *x, T *y, cudaStream_t s) {
  hpcx_ctx ctx = A->ctx;
  const HubTables &H = A->hub;
  const size_t smem = (size_t)H.K * sizeof(T);
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)spmv_hub_kernel<T>, kHubBytes)) return st;
  // empty rows are never visited by a nonzero: clear y first when there are any
  if (H.nlong + H.nshort < A->nrows) HPCX_CUDA(ctx, cudaMemsetAsync(y, 0, (size_t)A->nrows * sizeof(T), s));
  HubArgs a;
  a.colL = H.colL;
  a.valL = H.valL;
  a.rowstart = H.rowstart;
  a.tile_seg0 = H.tile_seg0;
  a.lrow = H.lrow;
  a.nnzL = H.nnzL;
  a.ntiles = H.ntiles;
  a.colS = H.colS;
  a.valS = H.valS;
  a.sbase = H.sbase;
  a.slen = H.slen;
  a.srow = H.srow;
  a.nshort = H.nshort;
  a.hub_col = H.hub_col;
  a.K = H.K;
  spmv_hub_kernel<T><<<ctx->sm_count, kHubThreads, smem, s>>>(a, x, y, (T *)H.head_out, (T *)H.tail_out);
  if (hpcx_status st = after_launch(ctx, "spmv_hub_kernel")) return st;
  if (H.nlong == 0) return HPCX_OK;
  spmv_hub_fixup_kernel<T><<<(unsigned)((H.nwarps + 255) / 256), 256, 0, s>>>(H.tile_seg0, H.lrow, (const T *)H.head_out,
                                                                             (const T *)H.tail_out, y, H.ntiles, H.nwarps);
  return after_launch(ctx, "spmv_hub_fixup_kernel");
}

hpcx_status hub_launch(hpcx_csr A, const void *x, void *y, cudaStream_t s) {
  if (A->nnz == 0) {
    HPCX_CUDA(A->ctx, cudaMemsetAsync(y, 0, (size_t)A->nrows * A->es, s));
    return HPCX_OK;
  }
  if (A->dtype == HPCX_F64) return hub_launch_t<double>(A, (const double *)x, (double *)y, s);
  return hub_launch_t<float>(A, (const float *)x, (float *)y, s);
}

}  // namespace hpcx
