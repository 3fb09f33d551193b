// stencil.cu — ghost-padded stencil sweeps for sm_100a.
//
//   BOX_X / BOX_Y  replace the loop nests of BM_x_blur / BM_x_blur_tiling_simd_prefetch and
//                  BM_y_blur / BM_YXx_blur_tiling_prefetch_streamed_IPL
//                  (reference tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368,
//                  425-441, 444-457, 575-601): res = 0.0f; for k = -r..r: res += a(.., ..+k).
//                  Adds only, in that order, so the result is bit-identical to the reference.
//   JACOBI7        new: out = (((((zm+ym)+xm)+xp)+yp)+zp) * (1/6), each
//                  operation rounded separately (the library is built with --fmad=false).
//
// Kernels:
//   *_direct      one thread per cell (quad), neighbours through L1 — any dtype / radius / shape.
//   box_x_vec / box_y_reg  register-tiled 17-point reference sweeps (radius 4 or 8).
//   jacobi7_ring  the default fp32 sweep: (TX+8)x(TY+2) plane boxes arrive in a shared-memory
//                 ring by 3-D TMA (cp.async.bulk.tensor, SASS UTMALDG) on mbarriers; 8 warps
//                 march along z with the z-1 / z centre values in registers, x neighbours by
//                 warp shuffle, y halo rows from shared memory; the per-plane CTA barrier frees
//                 the slot and thread 0 issues the next load (no producer warp).  Many short z
//                 segments; the peer-store halo transport of dist.cu lives in the same kernel.
//   jacobi7_tb2   two sweeps per pass over HBM (HPCX_VARIANT_TEMPORAL), bit-identical results.
#include "grid.cuh"

#include <algorithm>
#include <cmath>
#include <type_traits>

using namespace hpcx;

namespace {

// ------------------------------------------------------------------ direct kernels
struct GridGeom {
  int64_t n0, n1, n2;      // logical extents (z, y, x)
  int64_t glo0, glo1;      // ghost cells below (z, y)
  int64_t ext1;            // padded y extent
  int64_t pitch, xoff;     // device row pitch / offset of x = 0
};

__host__ GridGeom geom_of(const hpcx_grid_s *g) {
  return GridGeom{g->n[0], g->n[1], g->n[2], g->glo[0], g->glo[1], g->ext[1], g->pitch, g->xoff};
}

__device__ __forceinline__ int64_t gidx(const GridGeom &g, int64_t z, int64_t y, int64_t x) {
  return ((z + g.glo0) * g.ext1 + (y + g.glo1)) * g.pitch + g.xoff + x;
}

__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

// axis: 1 = along y (second fastest), 2 = along x (fastest). 2-D grids (n0 == 1).
template <typename T>
__global__ void __launch_bounds__(256)
box_direct_kernel(const T *__restrict__ in, T *__restrict__ out, GridGeom gi, GridGeom go, int axis,
                  int radius) {
  const int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t y = blockIdx.y * (int64_t)blockDim.y + threadIdx.y;
  if (x >= gi.n2 || y >= gi.n1) return;
  const int64_t step = axis == 2 ? 1 : gi.pitch;
  const T *p = in + gidx(gi, 0, y, x) - (int64_t)radius * step;
  T s = T(0);
  for (int k = -radius; k <= radius; ++k, p += step) s = add_rn(s, __ldg(p));
  out[gidx(go, 0, y, x)] = s;
}

// Register-tiled box sweeps (the 17-point reference sweeps at full HBM rate).  Every output is
// still res = 0; res += a(.. + k) for k = -R..R in that order (bit-exact with the reference
// loops); what changes is how the taps reach the registers.
//
// BOX_X: one thread per 16-byte vector of outputs.  It loads the 2R/V + 1 aligned vectors that
// cover x-R .. x+V-1+R (neighbouring threads share them through L1) and forms V ordered sums.
// (A variant that loads one vector per lane and takes the neighbours' by warp shuffle — 24
// shuffles instead of 4 extra loads — measured 119 us against 92 us on 8192^2: L1 is not the
// limit here, the shuffle pipe would be.)
template <typename T, int R>
__global__ void __launch_bounds__(256)
box_x_vec_kernel(const T *__restrict__ in, T *__restrict__ out, GridGeom gi, GridGeom go) {
  constexpr int V = 16 / (int)sizeof(T);
  static_assert(R % V == 0, "radius must be a multiple of the vector width");
  constexpr int NV = 2 * R / V + 1;
  using Vec = typename std::conditional<sizeof(T) == 4, float4, double2>::type;
  const int64_t x = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * V;
  const int64_t y = blockIdx.y;
  if (x >= gi.n2) return;
  const Vec *p = reinterpret_cast<const Vec *>(in + gidx(gi, 0, y, x) - R);
  T w[NV * V];
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const Vec v = __ldg(p + k);
    if constexpr (V == 4) {
      w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
    } else {
      w[2 * k] = v.x; w[2 * k + 1] = v.y;
    }
  }
  T o[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    T s = T(0);
#pragma unroll
    for (int k = 0; k <= 2 * R; ++k) s = add_rn(s, w[j + k]);
    o[j] = s;
  }
  Vec r;
  if constexpr (V == 4) {
    r.x = o[0]; r.y = o[1]; r.z = o[2]; r.w = o[3];
  } else {
    r.x = o[0]; r.y = o[1];
  }
  *reinterpret_cast<Vec *>(out + gidx(go, 0, y, x)) = r;
}

// BOX_Y: one thread per column and ROWS consecutive outputs.  The ROWS + 2R taps are
// independent coalesced loads (all in flight at once), each reused by up to 2R + 1 outputs from
// registers; the 2R halo rows shared with the neighbouring row blocks are L2 hits.
template <typename T, int R, int ROWS>
__global__ void __launch_bounds__(128)
box_y_reg_kernel(const T *__restrict__ in, T *__restrict__ out, GridGeom gi, GridGeom go) {
  const int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t y0 = blockIdx.y * (int64_t)ROWS;
  if (x >= gi.n2) return;
  const T *p = in + gidx(gi, 0, y0 - R, x);
  const int64_t rows = min((int64_t)ROWS, gi.n1 - y0);  // outputs of this block
  T w[ROWS + 2 * R];
#pragma unroll
  for (int k = 0; k < ROWS + 2 * R; ++k)
    w[k] = k < rows + 2 * R ? __ldg(p + k * gi.pitch) : T(0);
  T *q = out + gidx(go, 0, y0, x);
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    T s = T(0);
#pragma unroll
    for (int k = 0; k <= 2 * R; ++k) s = add_rn(s, w[j + k]);
    if (j < rows) q[j * go.pitch] = s;
  }
}

// fp32 BOX_Y, two columns per thread: the pair of ordered sums advances with one packed
// add.rn.f32x2 per tap (IEEE round-to-nearest per lane, so still bit-exact), which halves the
// issue slots of the 17 adds per cell, and a warp covers 256 contiguous bytes of every row.
__device__ __forceinline__ float2 add2_rn(float2 a, float2 b) {
  float2 d;
  asm("{\n"
      ".reg .b64 ra, rb, rd;\n"
      "mov.b64 ra, {%2, %3};\n"
      "mov.b64 rb, {%4, %5};\n"
      "add.rn.f32x2 rd, ra, rb;\n"
      "mov.b64 {%0, %1}, rd;\n"
      "}"
      : "=f"(d.x), "=f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}

// ROWS consecutive outputs of two adjacent columns per thread.  Full row blocks take the
// unpredicated path (FULL), the last partial block the guarded one.
template <int R, int ROWS, bool FULL>
__device__ __forceinline__ void box_y_reg2_body(const float *__restrict__ p, float *__restrict__ q,
                                                int64_t ip, int64_t op, int rows) {
  float2 w[ROWS + 2 * R];
#pragma unroll
  for (int k = 0; k < ROWS + 2 * R; ++k)
    w[k] = FULL || k < rows + 2 * R ? __ldg(reinterpret_cast<const float2 *>(p + k * ip))
                                    : make_float2(0.f, 0.f);
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k <= 2 * R; ++k) s = add2_rn(s, w[j + k]);
    if (FULL || j < rows) *reinterpret_cast<float2 *>(q + j * op) = s;
  }
}

template <int R, int ROWS>
__global__ void __launch_bounds__(128)
box_y_reg2_kernel(const float *__restrict__ in, float *__restrict__ out, GridGeom gi, GridGeom go) {
  const int64_t x = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * 2;
  const int64_t y0 = blockIdx.y * (int64_t)ROWS;
  if (x >= gi.n2) return;
  const float *p = in + gidx(gi, 0, y0 - R, x);
  float *q = out + gidx(go, 0, y0, x);
  const int rows = (int)min((int64_t)ROWS, gi.n1 - y0);
  if (rows == ROWS)
    box_y_reg2_body<R, ROWS, true>(p, q, gi.pitch, go.pitch, rows);
  else
    box_y_reg2_body<R, ROWS, false>(p, q, gi.pitch, go.pitch, rows);
}

// One thread per (x, y) column, marching along z over [z0, z1).
template <typename T>
__global__ void __launch_bounds__(256)
jacobi7_direct_kernel(const T *__restrict__ in, T *__restrict__ out, GridGeom g, int64_t zbeg,
                      int64_t zend, int64_t zlen, T sixth) {
  const int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t y = blockIdx.y * (int64_t)blockDim.y + threadIdx.y;
  if (x >= g.n2 || y >= g.n1) return;
  const int64_t z0 = zbeg + blockIdx.z * zlen;
  const int64_t z1 = min(z0 + zlen, zend);
  const int64_t plane = g.ext1 * g.pitch;
  const T *p = in + gidx(g, z0, y, x);
  T *q = out + gidx(g, z0, y, x);
  T zm = __ldg(p - plane), zc = __ldg(p);
  for (int64_t z = z0; z < z1; ++z, p += plane, q += plane) {
    const T zp = __ldg(p + plane);
    T s = add_rn(zm, __ldg(p - g.pitch));
    s = add_rn(s, __ldg(p - 1));
    s = add_rn(s, __ldg(p + 1));
    s = add_rn(s, __ldg(p + g.pitch));
    s = add_rn(s, zp);
    *q = mul_rn(s, sixth);
    zm = zc;
    zc = zp;
  }
}

__device__ __forceinline__ void st_cs(float4 *p, float4 v) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// ------------------------------------------------------------------ Jacobi, TMA ring
// Geometry of the ring kernel.  Tile TX = 128 (one float4 per lane) by TY = 16 rows (two adjacent
// rows per warp), S ring slots of one plane box each.
template <int S>
struct JacobiCfg {
  static constexpr int TX = 128, TY = 16, HX = 4;      // HX: x halo staged on each side
  static constexpr int BX = TX + 2 * HX, BY = TY + 2;  // box extents
  static constexpr int kPlaneBytes = BX * BY * 4;  // 9792 bytes arrive per box
  // TMA writes need a 128-byte aligned shared-memory destination: pad the slot stride.
  static constexpr int kSlotBytes = (kPlaneBytes + 127) / 128 * 128;  // 9856
  static constexpr int kSlotFloats = kSlotBytes / 4;                 // slot stride in floats
  static constexpr int kBarOff = S * kSlotBytes;
  static constexpr int kSmemBytes = kBarOff + S * 8;
};

// Work split (host planner jacobi_plan): the (x, y, z) space is cut into ntx * nty * zs CTAs of
// equal size — x tiles of TX, y tiles by even division of the rows (tile heights h and h-1 only,
// one tensor map per height), z segments by even division of the planes — with the CTA count a
// multiple of the SM count, so every SM streams the same number of cells.  All CTAs of a z
// segment march through the planes in step, which is what lets L2 serve the tile halos.
struct JacobiSplit {
  int ntx, nty, zs;   // tiles along x, y; z segments
  int hmax;           // height of the taller y tiles (<= TY)
};

// Single sweep.  There is no producer warp: the CTA's per-plane barrier — every warp has read
// plane j out of shared memory — is what frees a ring slot, and thread 0 issues the next TMA
// load right after it.  Eight warps (register allocation is per four warps), no "empty"
// mbarriers and no spinning producer competing for issue slots (a nine-warp producer/consumer
// version of this kernel measured 620 against 650 Gcell/s with the same split and was removed).
template <int S>
__global__ void __launch_bounds__(256, 3)
jacobi7_ring_kernel(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo,
                    float *__restrict__ out, GridGeom g, JacobiSplit sp, JacobiPeer pr, int zbeg, int zend,
                    float sixth) {
  using C = JacobiCfg<S>;
  extern __shared__ __align__(128) unsigned char smem[];
  float *ring = reinterpret_cast<float *>(smem);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + C::kBarOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bx = blockIdx.x % sp.ntx;
  const int by = (blockIdx.x / sp.ntx) % sp.nty;
  // With a neighbour above, the top segment runs second instead of last: both boundary planes
  // reach the neighbours within the first wave, a whole sweep before they are needed.
  const int seg = blockIdx.x / (sp.ntx * sp.nty);
  const int bz = pr.ghost_hi && sp.zs > 2 ? (seg == 0 ? 0 : seg == 1 ? sp.zs - 1 : seg - 1) : seg;
  const int x0 = bx * C::TX;
  const int y0 = (int)((int64_t)by * g.n1 / sp.nty);
  const int h = (int)((int64_t)(by + 1) * g.n1 / sp.nty) - y0;
  const int nzr = zend - zbeg;
  const int z0 = zbeg + (int)((int64_t)bz * nzr / sp.zs);
  const int z1 = zbeg + (int)((int64_t)(bz + 1) * nzr / sp.zs);
  const int np = z1 - z0 + 2;  // planes z0-1 .. z1
  if (h <= 0 || z1 <= z0) return;

  const CUtensorMap *map = h == sp.hmax ? &map_hi : &map_lo;
  const uint32_t box_bytes = (uint32_t)(C::BX * (h + 2) * 4);
  const int c0 = (int)g.xoff + x0 - C::HX;
  const int c1 = y0 + (int)g.glo1 - 1;
  auto issue = [&](int p) {  // plane p (z = z0 - 1 + p) -> slot p % S; thread 0 only
    if (pr.wait_val) {       // ghost planes filled by a neighbour: wait for its sweep number first
      const int z = z0 - 1 + p;
      if (z == -1 && pr.sig_lo) wait_flag(pr.flags + 0, pr.wait_val, pr.pw);
      if (z == (int)g.n0 && pr.sig_hi) wait_flag(pr.flags + 1, pr.wait_val, pr.pw);
    }
    mbar_expect_tx(full + p % S, box_bytes);
    tma_load_3d(ring + (p % S) * C::kSlotFloats, map, c0, c1, z0 - 1 + p + (int)g.glo0, full + p % S);
  };
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
    mbar_fence_init();
    tma_prefetch_desc(map);
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int p = 0; p < S && p < np; ++p) issue(p);

  // Warp w owns tile rows 2w and 2w+1 (box rows 2w+1, 2w+2); lane owns x0 + 4*lane .. +3.
  const int br = 2 * warp + 1;
  const int col = C::HX + 4 * lane;
  const int r0off = br * C::BX + col, r1off = r0off + C::BX;
  const int gx = x0 + 4 * lane;
  const bool row0 = 2 * warp < h, row1 = 2 * warp + 1 < h;
  const bool full4 = gx + 3 < g.n2;
  const int64_t plane_stride = g.ext1 * g.pitch;
  float *orow = out + gidx(g, z0, y0 + 2 * warp, gx);
  const int64_t inplane = (y0 + 2 * warp + g.glo1) * g.pitch + g.xoff + gx;  // offset inside a plane
  const uint32_t tiles_xy = (uint32_t)(sp.ntx * sp.nty);

  struct PL {
    float4 a, b;
  };
  PL t0, t1, t2;
  auto load_plane = [&](int j, PL &t) {
    mbar_wait(full + j % S, (j / S) & 1);
    const float *pl = ring + (j % S) * C::kSlotFloats;
    t.a = *reinterpret_cast<const float4 *>(pl + r0off);
    t.b = *reinterpret_cast<const float4 *>(pl + r1off);
  };
#define HPCX_J7(zm, ym, xm, xp, yp, zp) \
  mul_rn(add_rn(add_rn(add_rn(add_rn(add_rn(zm, ym), xm), xp), yp), zp), sixth)
  // P, Cn, N: planes j-1, j, j+1 (N is loaded here); writes plane z0 - 1 + j
  auto step = [&](int j, const PL &P, const PL &Cn, PL &N) {
    load_plane(j + 1, N);
    const float *pc = ring + (j % S) * C::kSlotFloats;
    const float4 up = *reinterpret_cast<const float4 *>(pc + r0off - C::BX);
    const float4 dn = *reinterpret_cast<const float4 *>(pc + r1off + C::BX);
    float l0 = __shfl_up_sync(0xffffffffu, Cn.a.w, 1);
    float l1 = __shfl_up_sync(0xffffffffu, Cn.b.w, 1);
    float r0 = __shfl_down_sync(0xffffffffu, Cn.a.x, 1);
    float r1 = __shfl_down_sync(0xffffffffu, Cn.b.x, 1);
    if (lane == 0) {
      l0 = pc[r0off - 1];
      l1 = pc[r1off - 1];
    }
    if (lane == 31) {
      r0 = pc[r0off + 4];
      r1 = pc[r1off + 4];
    }
    float4 o0, o1;
    o0.x = HPCX_J7(P.a.x, up.x, l0, Cn.a.y, Cn.b.x, N.a.x);
    o0.y = HPCX_J7(P.a.y, up.y, Cn.a.x, Cn.a.z, Cn.b.y, N.a.y);
    o0.z = HPCX_J7(P.a.z, up.z, Cn.a.y, Cn.a.w, Cn.b.z, N.a.z);
    o0.w = HPCX_J7(P.a.w, up.w, Cn.a.z, r0, Cn.b.w, N.a.w);
    o1.x = HPCX_J7(P.b.x, Cn.a.x, l1, Cn.b.y, dn.x, N.b.x);
    o1.y = HPCX_J7(P.b.y, Cn.a.y, Cn.b.x, Cn.b.z, dn.y, N.b.y);
    o1.z = HPCX_J7(P.b.z, Cn.a.z, Cn.b.y, Cn.b.w, dn.z, N.b.z);
    o1.w = HPCX_J7(P.b.w, Cn.a.w, Cn.b.z, r1, dn.w, N.b.w);
    // Boundary plane of the slab: the same registers also go straight into the neighbour's ghost
    // plane over NVLink; after the barrier one thread counts the tile in and the last tile of
    // the plane publishes the sweep number (CTA-uniform branches).
    const int z = z0 - 1 + j;
    float *peer0 = pr.ghost_lo && z == 0 ? pr.ghost_lo + inplane : nullptr;
    float *peer1 = pr.ghost_hi && z == (int)g.n0 - 1 ? pr.ghost_hi + inplane : nullptr;
    if (peer0 || peer1) {
      const float v0[4] = {o0.x, o0.y, o0.z, o0.w}, v1[4] = {o1.x, o1.y, o1.z, o1.w};
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        float *pp = q ? peer1 : peer0;
        if (!pp) continue;
        if (full4) {
          if (row0) *reinterpret_cast<float4 *>(pp) = o0;
          if (row1) *reinterpret_cast<float4 *>(pp + g.pitch) = o1;
        } else {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            if (gx + c < g.n2) {
              if (row0) pp[c] = v0[c];
              if (row1) pp[g.pitch + c] = v1[c];
            }
        }
      }
      __threadfence_system();
    }
    // every warp has read planes <= j out of shared memory once it passes this barrier
    __syncthreads();
    if (threadIdx.x == 0) {
      if (peer0 && (atomicAdd(pr.counters + 0, 1u) + 1) % tiles_xy == 0) st_release_sys(pr.sig_lo, pr.sig_val);
      if (peer1 && (atomicAdd(pr.counters + 1, 1u) + 1) % tiles_xy == 0) st_release_sys(pr.sig_hi, pr.sig_val);
      if (j + S - 1 < np) issue(j + S - 1);
    }
    if (full4) {
      if (row0) st_cs(reinterpret_cast<float4 *>(orow), o0);
      if (row1) st_cs(reinterpret_cast<float4 *>(orow + g.pitch), o1);
    } else {
      if (row0) {
        if (gx < g.n2) orow[0] = o0.x;
        if (gx + 1 < g.n2) orow[1] = o0.y;
        if (gx + 2 < g.n2) orow[2] = o0.z;
      }
      if (row1) {
        if (gx < g.n2) orow[g.pitch] = o1.x;
        if (gx + 1 < g.n2) orow[g.pitch + 1] = o1.y;
        if (gx + 2 < g.n2) orow[g.pitch + 2] = o1.z;
      }
    }
    orow += plane_stride;
  };
#undef HPCX_J7

  load_plane(0, t0);
  load_plane(1, t1);
  const int last = np - 2;
  for (int j = 1; j <= last; j += 3) {  // three steps per trip: the planes rotate by renaming
    step(j, t0, t1, t2);
    if (j + 1 > last) break;
    step(j + 1, t1, t2, t0);
    if (j + 2 > last) break;
    step(j + 2, t2, t0, t1);
  }
}

// ------------------------------------------------------------------ host side
hpcx_status check_same_shape(hpcx_grid a, hpcx_grid b) {
  hpcx_ctx ctx = a->ctx;
  HPCX_REQUIRE(ctx, a->ctx == b->ctx, "grids belong to different contexts");
  HPCX_REQUIRE(ctx, a->desc.ndim == b->desc.ndim && a->desc.dtype == b->desc.dtype,
               "grids differ in rank or dtype");
  for (int i = 0; i < a->desc.ndim; ++i)
    HPCX_REQUIRE(ctx, a->desc.dim[i] == b->desc.dim[i], "grids differ in logical extent");
  return HPCX_OK;
}

// The box launchers take geometries, not handles: a range of rows [y0, y1) of a grid is the same
// grid with its row origin shifted (the distributed sweep runs interior rows and boundary strips
// as separate launches, dist.cu).
struct BoxView {
  GridGeom gi, go;
  const void *in;
  void *out;
  hpcx_ctx ctx;
};
BoxView box_view(hpcx_grid in, hpcx_grid out, int64_t y0, int64_t y1) {
  BoxView v{geom_of(in), geom_of(out), in->dev, out->dev, in->ctx};
  v.gi.glo1 += y0;
  v.go.glo1 += y0;
  v.gi.n1 = v.go.n1 = y1 - y0;
  return v;
}
// ... and a range of columns [x0, x1) the same grid with its x origin shifted (block split:
// interior columns beside the exchange, boundary strips behind it).
BoxView box_view(hpcx_grid in, hpcx_grid out, int64_t y0, int64_t y1, int64_t x0, int64_t x1) {
  BoxView v = box_view(in, out, y0, y1);
  v.gi.xoff += x0;
  v.go.xoff += x0;
  v.gi.n2 = v.go.n2 = x1 - x0;
  return v;
}

template <typename T>
hpcx_status launch_box_direct(const BoxView &v, int axis, int radius, cudaStream_t s) {
  dim3 block(64, 4);
  dim3 grid((unsigned)((v.gi.n2 + 63) / 64), (unsigned)((v.gi.n1 + 3) / 4));
  box_direct_kernel<T><<<grid, block, 0, s>>>((const T *)v.in, (T *)v.out, v.gi, v.go, axis, radius);
  return after_launch(v.ctx, "box_direct_kernel");
}

// Register-tiled variants exist for the radii a 16-byte vector divides (the reference's
// nblur = 8 among them); everything else takes the direct kernel.
template <typename T, int R, int ROWS>
hpcx_status launch_box_fast_r(const BoxView &v, int axis, cudaStream_t s) {
  constexpr int V = 16 / (int)sizeof(T);
  if (axis == 2) {
    dim3 grid((unsigned)((v.gi.n2 / V + 255) / 256), (unsigned)v.gi.n1);
    box_x_vec_kernel<T, R><<<grid, 256, 0, s>>>((const T *)v.in, (T *)v.out, v.gi, v.go);
    return after_launch(v.ctx, "box_x_vec_kernel");
  }
  if constexpr (sizeof(T) == 4) {
    if (v.gi.n2 % 2 == 0) {  // pitches and x offsets are multiples of 8 floats: 8-byte aligned pairs
      dim3 grid2((unsigned)((v.gi.n2 / 2 + 127) / 128), (unsigned)((v.gi.n1 + ROWS - 1) / ROWS));
      box_y_reg2_kernel<R, ROWS><<<grid2, 128, 0, s>>>((const float *)v.in, (float *)v.out, v.gi, v.go);
      return after_launch(v.ctx, "box_y_reg2_kernel");
    }
  }
  dim3 grid((unsigned)((v.gi.n2 + 127) / 128), (unsigned)((v.gi.n1 + ROWS - 1) / ROWS));
  box_y_reg_kernel<T, R, ROWS><<<grid, 128, 0, s>>>((const T *)v.in, (T *)v.out, v.gi, v.go);
  return after_launch(v.ctx, "box_y_reg_kernel");
}

template <typename T>
hpcx_status launch_box(const BoxView &v, int axis, int radius, int variant, int rows, cudaStream_t s) {
  if (v.gi.n1 <= 0) return HPCX_OK;
  constexpr int V = 16 / (int)sizeof(T);
  const bool fast_ok = variant != HPCX_VARIANT_DIRECT && v.gi.n1 <= 65535 * 32 &&
                       (axis != 2 || (v.gi.n2 % V == 0 && v.gi.n1 <= 65535));
  if (fast_ok) {
    switch (radius) {
      case 8:
        return rows == 16 ? launch_box_fast_r<T, 8, 16>(v, axis, s) : launch_box_fast_r<T, 8, 32>(v, axis, s);
      case 4:
        return rows == 16 ? launch_box_fast_r<T, 4, 16>(v, axis, s) : launch_box_fast_r<T, 4, 32>(v, axis, s);
      default: break;
    }
  }
  if (v.gi.n1 > 65535 * 4) return fail(v.ctx, HPCX_UNSUPPORTED, "grid too tall for the box sweep");
  return launch_box_direct<T>(v, axis, radius, s);
}

int auto_zsplit(hpcx_ctx ctx, int64_t tiles_xy, int64_t nz, int ctas_per_sm) {
  // Fill the resident CTA slots of the chip once; keep z segments at least 16 planes long.
  const int64_t slots = (int64_t)ctx->sm_count * ctas_per_sm;
  int64_t zs = std::max<int64_t>(1, slots / std::max<int64_t>(tiles_xy, 1));
  zs = std::min<int64_t>(zs, std::max<int64_t>(1, nz / 16));
  return (int)zs;
}

template <typename T>
hpcx_status launch_jacobi_direct(hpcx_grid_s *g, const void *src, void *dst, int zsplit,
                                 int64_t zbeg, int64_t zend, cudaStream_t s) {
  dim3 block(64, 4);
  const int64_t nz = zend - zbeg;
  const int64_t tiles = ((g->n[2] + 63) / 64) * ((g->n[1] + 3) / 4);
  if (zsplit <= 0) zsplit = auto_zsplit(g->ctx, tiles, nz, 8);
  const int64_t zlen = (nz + zsplit - 1) / zsplit;
  dim3 grid((unsigned)((g->n[2] + 63) / 64), (unsigned)((g->n[1] + 3) / 4),
            (unsigned)((nz + zlen - 1) / zlen));
  jacobi7_direct_kernel<T><<<grid, block, 0, s>>>((const T *)src, (T *)dst, geom_of(g), zbeg, zend,
                                                  zlen, T(1) / T(6));
  return after_launch(g->ctx, "jacobi7_direct_kernel");
}

constexpr int kJacobiStages = 6;

// Equal-work split: minimise (CTAs per SM, rounded up, at least 3) x (cells one CTA loads).  `zsplit` > 0
// pins the number of z segments (tests exercise short segments that way).
JacobiSplit jacobi_plan(hpcx_ctx ctx, int64_t nx, int64_t ny, int64_t nz, int zsplit) {
  using C = JacobiCfg<kJacobiStages>;
  const int64_t ntx = (nx + C::TX - 1) / C::TX;
  const int64_t nty_min = (ny + C::TY - 1) / C::TY;
  JacobiSplit best{(int)ntx, (int)nty_min, 1, (int)((ny + nty_min - 1) / nty_min)};
  double best_cost = 1e300;
  const int64_t nty_max = std::min<int64_t>(ny, std::max<int64_t>(nty_min * 2, nty_min + 8));
  const int64_t zs_lo = zsplit > 0 ? std::min<int64_t>(zsplit, nz) : 1;
  const int64_t zs_hi = zsplit > 0 ? zs_lo : std::max<int64_t>(zs_lo, nz / 8);
  for (int64_t nty = nty_min; nty <= nty_max; ++nty) {
    const int64_t h = (ny + nty - 1) / nty;
    for (int64_t zs = zs_lo; zs <= zs_hi; ++zs) {
      const int64_t zl = (nz + zs - 1) / zs;
      const int64_t ncta = ntx * nty * zs;
      const int64_t per_sm = (ncta + ctx->sm_count - 1) / ctx->sm_count;
      // + 3 planes: ring fill before the first output plane of a segment
      // An SM reaches its streaming rate only with all 3 resident CTA slots busy (59 KB of
      // ring each), so fewer than 3 CTAs per SM cost as much as 3.
      const double cost = (double)std::max<int64_t>(per_sm, 3) * (double)(h + 2) * (double)(zl + 2 + 3);
      if (cost < best_cost * (1.0 - 1e-9)) {
        best_cost = cost;
        best = JacobiSplit{(int)ntx, (int)nty, (int)zs, (int)h};
      }
    }
  }
  return best;
}

// Tensor maps are cached on the grid handle, keyed by (buffer, box height).
const CUtensorMap *jacobi_map(hpcx_grid_s *g, void *src, int boxh, hpcx_status *st) {
  using C = JacobiCfg<kJacobiStages>;
  for (auto &e : g->maps)
    if (e.base == src && e.boxh == boxh) return &e.map;
  if (g->maps.size() >= 8) g->maps.clear();
  g->maps.reserve(8);  // entries never move while the cache holds fewer than 8
  hpcx_grid_s::MapEntry e;
  e.base = src;
  e.boxh = boxh;
  const uint64_t dims[3] = {(uint64_t)g->pitch, (uint64_t)g->ext[1], (uint64_t)g->ext[0]};
  const uint64_t strides[2] = {(uint64_t)g->pitch * 4, (uint64_t)g->pitch * g->ext[1] * 4};
  const uint32_t box[3] = {C::BX, (uint32_t)boxh, 1};
  *st = make_tensor_map(g->ctx, &e.map, HPCX_F32, 3, src, dims, strides, box);
  if (*st != HPCX_OK) return nullptr;
  g->maps.push_back(e);
  return &g->maps.back().map;
}

hpcx_status launch_jacobi_tma(hpcx_grid_s *g, void *src, void *dst, int zsplit, int64_t zbeg,
                              int64_t zend, const JacobiPeer *peer, cudaStream_t s) {
  using C = JacobiCfg<kJacobiStages>;
  hpcx_ctx ctx = g->ctx;
  const JacobiPeer pr = peer ? *peer : JacobiPeer{};
  // On grids that do not fit L2 the kernel runs fastest with MANY short z segments (about
  // 8 planes: 512^3 648 Gcell/s with one wave of 444 CTAs, 808 with 64 segments = 18 waves,
  // profiles/r22_jacobi_zsplit.txt): short-lived CTAs drift out of lock-step, so the DRAM sees a
  // steady read/write mix, and a wave still covers whole planes, so the tile halos stay L2 hits.
  int zs_pin = zsplit;
  if (zsplit <= 0 && 2 * g->ext[0] * g->ext[1] * g->pitch * 4 > ctx->l2_bytes + ctx->l2_bytes / 2)
    zs_pin = (int)std::max<int64_t>(1, (zend - zbeg + 7) / 8);
  const JacobiSplit sp = jacobi_plan(ctx, g->n[2], g->n[1], zend - zbeg, zs_pin);
  hpcx_status st = HPCX_OK;
  const CUtensorMap *hi = jacobi_map(g, src, sp.hmax + 2, &st);
  if (!hi) return st;
  const CUtensorMap *lo = sp.hmax > 1 ? jacobi_map(g, src, sp.hmax + 1, &st) : hi;
  if (!lo) return st;
  // `hi` may have moved if the cache reallocated: fetch it again (cheap, by key).
  hi = jacobi_map(g, src, sp.hmax + 2, &st);
  st = ensure_dyn_smem(ctx, (const void *)jacobi7_ring_kernel<kJacobiStages>, C::kSmemBytes);
  if (st != HPCX_OK) return st;
  const unsigned ncta = (unsigned)(sp.ntx * sp.nty * sp.zs);
  jacobi7_ring_kernel<kJacobiStages><<<ncta, 256, C::kSmemBytes, s>>>(
      *hi, *lo, (float *)dst, geom_of(g), sp, pr, (int)zbeg, (int)zend, 1.0f / 6.0f);
  return after_launch(ctx, "jacobi7_ring_kernel");
}


// ------------------------------------------------------------------ Jacobi, two sweeps per pass
// Temporal blocking (HPCX_VARIANT_TEMPORAL): one pass over HBM advances the field by TWO sweeps,
// so the traffic per sweep drops from 8 to 4 bytes per cell.  Same ring, same add order — the
// result is bit-identical to two single sweeps.
//
// A CTA owns TX = 128 columns, h <= 14 rows and a z segment.  The producer streams boxes of
// (128 + 8) x (h + 4) cells of iterate t (two halo cells per side in y, plane by plane, from two
// planes below the segment to two above).  The 8 consumer warps march along z; at step j they
//   A  compute iterate t+1 of plane j on the tile grown by one cell in x and y (their own two
//      rows x four columns from registers / shuffles / staged halo rows, plus the two edge
//      columns x0-1 and x0+128 by lanes 0 and 31); cells outside the domain keep their (fixed,
//      Dirichlet) ghost value; the plane goes into a double-buffered shared-memory array;
//   B  compute iterate t+2 of plane j-1 from the t+1 planes j-2, j-1, j (own cells in registers,
//      x neighbours by shuffle, y neighbours and edge columns from the shared t+1 plane of step
//      j-1) and stream it to HBM.
// One named barrier per step orders the t+1 planes between the warps.
template <int S_>
struct JacobiTBCfg {
  static constexpr int TX = 128, HX = 4, BX = TX + 2 * HX;  // 136 staged columns
  static constexpr int HO = 14;                               // output rows per tile (max)
  static constexpr int ER = HO + 2;                           // rows of the grown (t+1) tile
  static constexpr int BY = HO + 4;                           // staged rows of iterate t
  static constexpr int S = S_;  // ring slots: 8 -> 98.6 KB, 2 CTAs/SM; 5 -> 69.2 KB, 3 CTAs/SM
  static constexpr int kSlotBytes = (BX * BY * 4 + 127) / 128 * 128;
  static constexpr int kSlotFloats = kSlotBytes / 4;
  static constexpr int kAPlaneFloats = (ER + 2) * BX;  // one guard row above and below
  static constexpr int kAOff = S * kSlotBytes;
  static constexpr int kBarOff = kAOff + 2 * kAPlaneFloats * 4;
  static constexpr int kSmemBytes = kBarOff + S * 8;
  static constexpr int kConsumerWarps = 8;
  static constexpr int kThreads = kConsumerWarps * 32;  // no producer warp: thread 0 feeds the ring
};

struct JacobiTBSplit {
  int ntx, nty, zs, hmax;
  int zbeg, zend;  // planes written by this launch, cut into zs segments
  int zbeg2, zend2;  // optional second range, one more segment (bz == zs): the two boundary strips
                     // of a slab go out in one launch
  int zlo, zhi;    // planes whose iterate t+1 is COMPUTED: [0, n0) on one GPU; a slab with a z
                   // neighbour grows by one plane on that side (the neighbour's boundary plane,
                   // recomputed here from the two exchanged ghost planes)
};

#define HPCX_J7(zm, ym, xm, xp, yp, zp) \
  mul_rn(add_rn(add_rn(add_rn(add_rn(add_rn(zm, ym), xm), xp), yp), zp), sixth)

// One CTA's work: tile (bx, by) x z segment bz of one two-sweep pass.  WAITS (slab with z
// neighbours, fused transport): a ghost plane of the slab is loaded only after the neighbour's flag
// says it has landed (pr.wait_val, as in the one-sweep kernel).
template <int S_, bool WAITS>
__device__ __forceinline__ void
jacobi7_tb2_item(const CUtensorMap *map_hi, const CUtensorMap *map_lo, float *__restrict__ out, const GridGeom &g,
                 const JacobiTBSplit &sp, const JacobiPeer &pr, float sixth, int bx, int by, int bz,
                 unsigned char *smem) {
  using C = JacobiTBCfg<S_>;
  float *ring = reinterpret_cast<float *>(smem);
  float *abuf = reinterpret_cast<float *>(smem + C::kAOff);
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + C::kBarOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = bx * C::TX;
  const int y0 = (int)((int64_t)by * g.n1 / sp.nty);
  const int h = (int)((int64_t)(by + 1) * g.n1 / sp.nty) - y0;
  const int z0 = bz < sp.zs ? sp.zbeg + (int)((int64_t)bz * (sp.zend - sp.zbeg) / sp.zs) : sp.zbeg2;
  const int z1 = bz < sp.zs ? sp.zbeg + (int)((int64_t)(bz + 1) * (sp.zend - sp.zbeg) / sp.zs) : sp.zend2;
  if (h <= 0 || z1 <= z0) return;
  const int np = z1 - z0 + 4;  // planes z0-2 .. z1+1 of iterate t

  // The per-step barrier of the consumers (below) already proves that plane j has been read by
  // every warp, so the ring needs no "empty" barriers and no producer warp: thread 0 issues the
  // TMA load of plane j + S - 1 right after the barrier of step j.
  const CUtensorMap *map = h == sp.hmax ? map_hi : map_lo;
  const uint32_t box_bytes = (uint32_t)(C::BX * (h + 4) * 4);
  const int c0 = (int)g.xoff + x0 - C::HX;
  const int c1 = y0 + (int)g.glo1 - 2;  // may be -1: rows outside the array are zero-filled
  auto issue = [&](int p) {            // plane p -> slot p % S; thread 0 only
    if (WAITS && pr.wait_val) {
      const int z = z0 - 2 + p;
      if (z < 0 && pr.sig_lo) wait_flag(pr.flags + 0, pr.wait_val, pr.pw);
      if (z >= (int)g.n0 && pr.sig_hi) wait_flag(pr.flags + 1, pr.wait_val, pr.pw);
    }
    mbar_expect_tx(full + (p % C::S), box_bytes);
    tma_load_3d(ring + (p % C::S) * C::kSlotFloats, map, c0, c1, z0 - 2 + p + (int)g.glo0,
                full + (p % C::S));
  };
  if (threadIdx.x == 0) {
    for (int s = 0; s < C::S; ++s) mbar_init(full + s, 1);
    mbar_fence_init();
    tma_prefetch_desc(map);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int p = 0; p < C::S && p < np; ++p) issue(p);
  }

  // ---------------------------------------------------------------- consumers
  // Warp w owns rows e = 2w, 2w+1 of the grown tile (y = y0 - 1 + e), i.e. box rows 2w+1, 2w+2;
  // lane owns box columns 4 + 4*lane .. (x = x0 + 4*lane ..).  The two edge columns of the grown
  // tile (box columns 3 and 132, x = x0 - 1 and x0 + 128) are spread over four lanes: lanes 0 / 1
  // take the left edge of rows 2w / 2w+1, lanes 30 / 31 the right edge.
  const int br = 2 * warp + 1;
  const int col = C::HX + 4 * lane;
  const int r0off = br * C::BX + col, r1off = r0off + C::BX;  // own cells inside a staged plane
  const int ec = lane < 16 ? C::HX - 1 : C::HX + C::TX;
  const int eoff = (br + (lane & 1)) * C::BX + ec;            // this lane's edge cell
  const bool edge = lane < 2 || lane >= 30;
  const int a0off = (2 * warp) * C::BX + col;                 // own cells inside a t+1 plane
  const int aeoff = (2 * warp + (lane & 1)) * C::BX + ec;

  const int gx = x0 + 4 * lane;
  const int ye0 = y0 - 1 + 2 * warp, ye1 = ye0 + 1;  // y of the two rows
  const int xe = x0 - C::HX + ec;                     // x of the edge column
  const bool yin0 = ye0 >= 0 && ye0 < g.n1, yin1 = ye1 >= 0 && ye1 < g.n1;
  const bool ein = ((lane & 1) ? yin1 : yin0) && xe >= 0 && xe < g.n2;
  const bool xragged = x0 + C::TX > g.n2;  // only the last x tile of a domain with n2 % 128 != 0
  const bool orow0 = 2 * warp >= 1 && 2 * warp <= h, orow1 = 2 * warp + 1 <= h;  // rows e in [1, h]
  const int64_t plane_stride = g.ext1 * g.pitch;

  struct TP {  // iterate t of one plane: own cells of the two rows, and the edge cell
    float4 a, b;
    float e;
  };
  struct AP {  // iterate t+1 of one plane, own cells
    float4 a, b;
  };
  TP t0, t1, t2;
  AP a0, a1, a2;
  a0.a = a0.b = a1.a = a1.b = a2.a = a2.b = make_float4(0.f, 0.f, 0.f, 0.f);
  t2.a = t2.b = a0.a;
  t2.e = 0.f;
  float *orow = out + gidx(g, z0, ye0, gx);  // plane z0, this thread's first row

  auto load_plane = [&](int j, TP &t) {
    mbar_wait(full + (j % C::S), (j / C::S) & 1);
    const float *pl = ring + (j % C::S) * C::kSlotFloats;
    t.a = *reinterpret_cast<const float4 *>(pl + r0off);
    t.b = *reinterpret_cast<const float4 *>(pl + r1off);
    t.e = pl[eoff];
  };
  // One step: P, Cn, N are iterate t of planes j-1, j, j+1 (N is loaded here); aP, aC are
  // iterate t+1 of planes j-2, j-1; aN (plane j) is made here.
  auto step = [&](int j, const TP &P, const TP &Cn, TP &N, const AP &aP, const AP &aC, AP &aN) {
    const int z = z0 - 2 + j;
    load_plane(j + 1, N);
    const float *pc = ring + (j % C::S) * C::kSlotFloats;
    const float4 up = *reinterpret_cast<const float4 *>(pc + r0off - C::BX);
    const float4 dn = *reinterpret_cast<const float4 *>(pc + r1off + C::BX);
    float l0 = __shfl_up_sync(0xffffffffu, Cn.a.w, 1);
    float l1 = __shfl_up_sync(0xffffffffu, Cn.b.w, 1);
    float r0 = __shfl_down_sync(0xffffffffu, Cn.a.x, 1);
    float r1 = __shfl_down_sync(0xffffffffu, Cn.b.x, 1);
    if (lane == 0) {
      l0 = pc[r0off - 1];
      l1 = pc[r1off - 1];
    }
    if (lane == 31) {
      r0 = pc[r0off + 4];
      r1 = pc[r1off + 4];
    }
    const float eup = pc[eoff - C::BX], edn = pc[eoff + C::BX], elf = pc[eoff - 1], ert = pc[eoff + 1];

    // ---- A: iterate t+1 of plane z on the grown tile; cells outside the domain keep their value
    aN.a.x = HPCX_J7(P.a.x, up.x, l0, Cn.a.y, Cn.b.x, N.a.x);
    aN.a.y = HPCX_J7(P.a.y, up.y, Cn.a.x, Cn.a.z, Cn.b.y, N.a.y);
    aN.a.z = HPCX_J7(P.a.z, up.z, Cn.a.y, Cn.a.w, Cn.b.z, N.a.z);
    aN.a.w = HPCX_J7(P.a.w, up.w, Cn.a.z, r0, Cn.b.w, N.a.w);
    aN.b.x = HPCX_J7(P.b.x, Cn.a.x, l1, Cn.b.y, dn.x, N.b.x);
    aN.b.y = HPCX_J7(P.b.y, Cn.a.y, Cn.b.x, Cn.b.z, dn.y, N.b.y);
    aN.b.z = HPCX_J7(P.b.z, Cn.a.z, Cn.b.y, Cn.b.w, dn.z, N.b.z);
    aN.b.w = HPCX_J7(P.b.w, Cn.a.w, Cn.b.z, r1, dn.w, N.b.w);
    float aE = HPCX_J7(P.e, eup, elf, ert, edn, N.e);
    if (z < sp.zlo || z >= sp.zhi) {  // ghost plane of the domain (uniform over the CTA)
      aN.a = Cn.a;
      aN.b = Cn.b;
      aE = Cn.e;
    } else {
      if (!yin0) aN.a = Cn.a;  // ghost rows (uniform over the warp)
      if (!yin1) aN.b = Cn.b;
      if (!ein) aE = Cn.e;
      if (xragged) {           // ghost / unused columns right of the domain (uniform over the CTA)
        if (gx >= g.n2) { aN.a.x = Cn.a.x; aN.b.x = Cn.b.x; }
        if (gx + 1 >= g.n2) { aN.a.y = Cn.a.y; aN.b.y = Cn.b.y; }
        if (gx + 2 >= g.n2) { aN.a.z = Cn.a.z; aN.b.z = Cn.b.z; }
        if (gx + 3 >= g.n2) { aN.a.w = Cn.a.w; aN.b.w = Cn.b.w; }
      }
    }

    // ---- B: iterate t+2 of plane z-1 from the t+1 planes z-2 (aP), z-1 (aC), z (aN)
    if (j >= 3) {
      const float *ac = abuf + ((j - 1) & 1) * C::kAPlaneFloats + C::BX;
      const float4 bup = *reinterpret_cast<const float4 *>(ac + a0off - C::BX);
      const float4 bdn = *reinterpret_cast<const float4 *>(ac + a0off + 2 * C::BX);
      float bl0 = __shfl_up_sync(0xffffffffu, aC.a.w, 1);
      float bl1 = __shfl_up_sync(0xffffffffu, aC.b.w, 1);
      float br0 = __shfl_down_sync(0xffffffffu, aC.a.x, 1);
      float br1 = __shfl_down_sync(0xffffffffu, aC.b.x, 1);
      if (lane == 0) {
        bl0 = ac[a0off - 1];
        bl1 = ac[a0off + C::BX - 1];
      }
      if (lane == 31) {
        br0 = ac[a0off + 4];
        br1 = ac[a0off + C::BX + 4];
      }
      float4 o0, o1;
      o0.x = HPCX_J7(aP.a.x, bup.x, bl0, aC.a.y, aC.b.x, aN.a.x);
      o0.y = HPCX_J7(aP.a.y, bup.y, aC.a.x, aC.a.z, aC.b.y, aN.a.y);
      o0.z = HPCX_J7(aP.a.z, bup.z, aC.a.y, aC.a.w, aC.b.z, aN.a.z);
      o0.w = HPCX_J7(aP.a.w, bup.w, aC.a.z, br0, aC.b.w, aN.a.w);
      o1.x = HPCX_J7(aP.b.x, aC.a.x, bl1, aC.b.y, bdn.x, aN.b.x);
      o1.y = HPCX_J7(aP.b.y, aC.a.y, aC.b.x, aC.b.z, bdn.y, aN.b.y);
      o1.z = HPCX_J7(aP.b.z, aC.a.z, aC.b.y, aC.b.w, bdn.z, aN.b.z);
      o1.w = HPCX_J7(aP.b.w, aC.a.w, aC.b.z, br1, bdn.w, aN.b.w);
      if (!xragged) {
        if (orow0) st_cs(reinterpret_cast<float4 *>(orow), o0);
        if (orow1) st_cs(reinterpret_cast<float4 *>(orow + g.pitch), o1);
      } else {
        // ragged right edge of the domain (n2 % 128 != 0; rows stay 16-byte aligned)
        if (gx + 3 < g.n2) {
          if (orow0) st_cs(reinterpret_cast<float4 *>(orow), o0);
          if (orow1) st_cs(reinterpret_cast<float4 *>(orow + g.pitch), o1);
        } else {
          if (orow0) {
            if (gx < g.n2) orow[0] = o0.x;
            if (gx + 1 < g.n2) orow[1] = o0.y;
            if (gx + 2 < g.n2) orow[2] = o0.z;
          }
          if (orow1) {
            if (gx < g.n2) orow[g.pitch] = o1.x;
            if (gx + 1 < g.n2) orow[g.pitch + 1] = o1.y;
            if (gx + 2 < g.n2) orow[g.pitch + 2] = o1.z;
          }
        }
      }
      orow += plane_stride;
    }

    // ---- publish the t+1 plane of this step
    {
      float *an = abuf + (j & 1) * C::kAPlaneFloats + C::BX;
      *reinterpret_cast<float4 *>(an + a0off) = aN.a;
      *reinterpret_cast<float4 *>(an + a0off + C::BX) = aN.b;
      if (edge) an[aeoff] = aE;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      // every warp is done with planes <= j: refill the slot of plane j - 1 with plane j + S - 1
      if (j + C::S - 1 < np) issue(j + C::S - 1);
    }
  };

  load_plane(0, t0);
  load_plane(1, t1);
  // three steps per trip so that the plane registers rotate by renaming, not by moves
  const int last = np - 2;
  for (int j = 1; j <= last; j += 3) {
    step(j, t0, t1, t2, a2, a0, a1);
    if (j + 1 > last) break;
    step(j + 1, t1, t2, t0, a0, a1, a2);
    if (j + 2 > last) break;
    step(j + 2, t2, t0, t1, a1, a2, a0);
  }
}
#undef HPCX_J7

template <int S_>
__global__ void __launch_bounds__(256, S_ <= 5 ? 3 : 2)
jacobi7_tb2_kernel(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo,
                   float *__restrict__ out, GridGeom g, JacobiTBSplit sp, float sixth) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int bx = blockIdx.x % sp.ntx;
  const int by = (blockIdx.x / sp.ntx) % sp.nty;
  const int bz = blockIdx.x / (sp.ntx * sp.nty);
  jacobi7_tb2_item<S_, false>(&map_hi, &map_lo, out, g, sp, JacobiPeer{}, sixth, bx, by, bz, smem);
}

// The pass kernel of a THIN slab with z neighbours (dist.cu, peer-store transport; strong
// scaling): ghost planes are awaited inside the kernel and the CTAs that own planes 0, 1 / n-2, n-1
// copy their tile of those planes (rows y0 .. y0+h, 128 columns) from the iterate they have just
// written into the neighbour's two ghost planes with peer stores when they are done; the last
// tile of a side (arrival counter) publishes the pass number in the neighbour's flag word.  One
// launch per pass and nothing between the passes — at 64 planes per GPU the separate push kernel
// (halo_push_planes_kernel, which ends with a wait for both neighbours) cost 6 % at 8 GPUs.  On
// thick slabs a CTA's segment is hundreds of planes long, the push would leave at the very end of
// the pass, and the push kernel is the better choice (2668 against 2292 Gcell/s at 2 x 512^3).
// Segments next to a neighbour run first.  pr.ghost_lo / ghost_hi: plane base of the FIRST of the
// two ghost planes in the neighbour's destination buffer, or null.
template <int S_>
__global__ void __launch_bounds__(256, S_ <= 5 ? 3 : 2)
jacobi7_tb2_peer_kernel(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo,
                        float *__restrict__ out, GridGeom g, JacobiTBSplit sp, JacobiPeer pr, float sixth) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int bx = blockIdx.x % sp.ntx;
  const int by = (blockIdx.x / sp.ntx) % sp.nty;
  const int seg = blockIdx.x / (sp.ntx * sp.nty);
  const bool lo = pr.ghost_lo != nullptr, hi = pr.ghost_hi != nullptr;
  const int bz = !hi ? seg : !lo ? sp.zs - 1 - seg : (seg & 1) ? sp.zs - 1 - (seg >> 1) : (seg >> 1);
  jacobi7_tb2_item<S_, true>(&map_hi, &map_lo, out, g, sp, pr, sixth, bx, by, bz, smem);
  __syncthreads();  // the CTA's stores of the new iterate are visible to all of its threads (L2)
  const int64_t plane = g.ext1 * g.pitch;
  const uint32_t tiles = (uint32_t)(sp.ntx * sp.nty);
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    float *peer = side ? pr.ghost_hi : pr.ghost_lo;
    if (!peer || bz != (side ? sp.zs - 1 : 0)) continue;  // uniform over the CTA
    const int y0 = (int)((int64_t)by * g.n1 / sp.nty);
    const int h = (int)((int64_t)(by + 1) * g.n1 / sp.nty) - y0;
    const int zf = side ? (int)g.n0 - 2 : 0;
    for (int i = (int)threadIdx.x; i < 2 * h * 32; i += (int)blockDim.x) {
      const int pl = i / (h * 32), r = (i >> 5) % h, x = bx * JacobiTBCfg<S_>::TX + 4 * (i & 31);
      if (x >= g.n2) continue;
      const int64_t inplane = (int64_t)(y0 + r + g.glo1) * g.pitch + g.xoff + x;
      const float4 v = __ldcg(reinterpret_cast<const float4 *>(out + (zf + pl + g.glo0) * plane + inplane));
      *reinterpret_cast<float4 *>(peer + pl * plane + inplane) = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0 && (atomicAdd(pr.counters + side, 1u) + 1) % tiles == 0)
      st_release_sys(side ? pr.sig_hi : pr.sig_lo, pr.sig_val);
  }
}


// Two sweeps src -> dst in one pass (HPCX_VARIANT_TEMPORAL).  Tiles of 128 x <= 14 cells, z
// segments chosen so that the CTA count is about two per SM (the kernel's residency).
hpcx_status launch_jacobi_tb2(hpcx_grid_s *g, void *src, void *dst, int zsplit, int64_t zbeg, int64_t zend,
                              bool nb_lo, bool nb_hi, cudaStream_t s, int64_t zbeg2 = 0, int64_t zend2 = 0,
                              const JacobiPeer *peer = nullptr) {
  using C = JacobiTBCfg<8>;
  const int per_sm = 2;  // 98.6 KB of shared memory per CTA (five slots and 3 CTAs/SM measured 2 % slower)
  hpcx_ctx ctx = g->ctx;
  if (zend <= zbeg) {
    if (zend2 <= zbeg2) return HPCX_OK;
    zbeg = zbeg2, zend = zend2, zbeg2 = zend2 = 0;
  }
  JacobiTBSplit sp;
  sp.zbeg = (int)zbeg;
  sp.zend = (int)zend;
  sp.zbeg2 = (int)zbeg2;
  sp.zend2 = (int)zend2;

  sp.zlo = nb_lo ? -1 : 0;
  sp.zhi = (int)g->n[0] + (nb_hi ? 1 : 0);
  sp.ntx = (int)((g->n[2] + C::TX - 1) / C::TX);
  sp.nty = (int)((g->n[1] + C::HO - 1) / C::HO);
  sp.hmax = (int)((g->n[1] + sp.nty - 1) / sp.nty);
  const int64_t tiles = (int64_t)sp.ntx * sp.nty, nzr = zend - zbeg;
  int64_t zs = std::min<int64_t>(std::max(zsplit, 1), nzr);
  if (zsplit <= 0) {
    // CTAs are long-lived (2 per SM) and the halo rows of a tile are L2 hits only while its
    // neighbours run in the same wave, so whole waves matter: minimise
    //   ceil(waves) x (planes per segment + ~40 planes of ramp and tail per CTA),
    // an under-filled single wave paying for its idle slots (profiles/r22_jacobi_zsplit.txt).
    const double slots = (double)per_sm * ctx->sm_count;
    double best = 1e300;
    for (int64_t c = 1; c <= std::max<int64_t>(1, nzr / 8); ++c) {
      const double waves = (double)(tiles * c) / slots;
      const double cost = std::ceil(waves) * (waves < 1.0 ? 1.0 / waves : 1.0) * ((double)nzr / (double)c + 40.0);
      if (cost < best * (1.0 - 1e-9)) best = cost, zs = c;
    }
  }
  if (peer) zs = std::min<int64_t>(zs, std::max<int64_t>(1, nzr / 2));  // planes 0, 1 / n-2, n-1 each inside ONE segment
  sp.zs = (int)std::max<int64_t>(zs, 1);
  hpcx_status st = HPCX_OK;
  const CUtensorMap *hi = jacobi_map(g, src, sp.hmax + 4, &st);
  if (!hi) return st;
  const CUtensorMap *lo = jacobi_map(g, src, sp.hmax + 3, &st);
  if (!lo) return st;
  hi = jacobi_map(g, src, sp.hmax + 4, &st);  // may have moved if the cache reallocated
  st = ensure_dyn_smem(ctx, (const void *)jacobi7_tb2_kernel<8>, C::kSmemBytes);
  if (st != HPCX_OK) return st;
  const unsigned ncta = (unsigned)(sp.ntx * sp.nty * (sp.zs + (zend2 > zbeg2 ? 1 : 0)));
  if (peer) {
    st = ensure_dyn_smem(ctx, (const void *)jacobi7_tb2_peer_kernel<8>, C::kSmemBytes);
    if (st != HPCX_OK) return st;
    jacobi7_tb2_peer_kernel<8><<<ncta, 256, C::kSmemBytes, s>>>(*hi, *lo, (float *)dst, geom_of(g), sp, *peer,
                                                                 1.0f / 6.0f);
    return after_launch(ctx, "jacobi7_tb2_peer_kernel");
  }
  jacobi7_tb2_kernel<8><<<ncta, 256, C::kSmemBytes, s>>>(*hi, *lo, (float *)dst, geom_of(g), sp, 1.0f / 6.0f);
  return after_launch(ctx, "jacobi7_tb2_kernel");
}

}  // namespace

namespace hpcx {
// One Jacobi sweep src -> dst over the planes of grid `g` (shared with dist.cu).
hpcx_status jacobi7_launch(hpcx_grid_s *g, void *src, void *dst, int variant, int zsplit,
                           int64_t zbeg, int64_t zend, cudaStream_t s, const JacobiPeer *peer) {
  if (zend <= zbeg) return HPCX_OK;
  const bool tma_ok = g->desc.dtype == HPCX_F32 && g->ctx->encode_tiled && g->n[2] < (1 << 30);
  if (variant == HPCX_VARIANT_TMA && !tma_ok)
    return fail(g->ctx, HPCX_UNSUPPORTED, "TMA Jacobi variant needs fp32 and a tensor-map driver");
  if (peer && (variant == HPCX_VARIANT_DIRECT || !tma_ok))
    return fail(g->ctx, HPCX_UNSUPPORTED, "the peer-store transport needs the fp32 TMA sweep");
  if (variant == HPCX_VARIANT_DIRECT || !tma_ok) {
    if (g->desc.dtype == HPCX_F64)
      return launch_jacobi_direct<double>(g, src, dst, zsplit, zbeg, zend, s);
    return launch_jacobi_direct<float>(g, src, dst, zsplit, zbeg, zend, s);
  }
  return launch_jacobi_tma(g, src, dst, zsplit, zbeg, zend, peer, s);
}

// Box sweep of the rows [y0, y1) of a 2-D grid pair (shared with dist.cu).
hpcx_status box_launch_rows(hpcx_grid_s *in, hpcx_grid_s *out, const hpcx_stencil_desc *st, int64_t y0, int64_t y1,
                            cudaStream_t s) {
  if (y1 <= y0) return HPCX_OK;
  const int axis = st->kind == HPCX_STENCIL_BOX_X ? 2 : 1;
  const BoxView v = box_view(in, out, y0, y1);
  return in->desc.dtype == HPCX_F64 ? launch_box<double>(v, axis, st->radius, st->variant, st->zsplit, s)
                                    : launch_box<float>(v, axis, st->radius, st->variant, st->zsplit, s);
}

// Box sweep of the tile [y0, y1) x [x0, x1).  The vector kernels need 16-byte aligned column
// ranges; anything else takes the direct kernel (same sums, same order).
hpcx_status box_launch_tile(hpcx_grid_s *in, hpcx_grid_s *out, const hpcx_stencil_desc *st, int64_t y0, int64_t y1,
                            int64_t x0, int64_t x1, cudaStream_t s) {
  if (y1 <= y0 || x1 <= x0) return HPCX_OK;
  const int axis = st->kind == HPCX_STENCIL_BOX_X ? 2 : 1;
  const BoxView v = box_view(in, out, y0, y1, x0, x1);
  const int64_t vec = 16 / in->es;
  const int variant = (x0 % vec == 0 && (x1 - x0) % vec == 0) ? st->variant : HPCX_VARIANT_DIRECT;
  return in->desc.dtype == HPCX_F64 ? launch_box<double>(v, axis, st->radius, variant, st->zsplit, s)
                                    : launch_box<float>(v, axis, st->radius, variant, st->zsplit, s);
}

// Forces the sweep kernels onto the current device now.  CUDA loads kernels lazily, at their
// first launch, and that load synchronises with kernels already running on the device; in a
// single-process multi-GPU run (multi.cu) a kernel of device 0 may be spinning on a flag that only
// a not-yet-launched kernel of device 1 will set — if the ONE host thread is then stuck loading a
// new kernel for device 0, nothing ever launches device 1's work.  Loading everything up front
// removes the load from the launch path.
hpcx_status preload_stencil_kernels(hpcx_ctx ctx) {
  cudaFuncAttributes a;
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)jacobi7_ring_kernel<kJacobiStages>));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)jacobi7_tb2_kernel<8>));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)jacobi7_direct_kernel<float>));
  HPCX_CUDA(ctx, cudaFuncGetAttributes(&a, (const void *)jacobi7_direct_kernel<double>));
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)jacobi7_ring_kernel<kJacobiStages>, JacobiCfg<kJacobiStages>::kSmemBytes)) return st;
  if (hpcx_status st = ensure_dyn_smem(ctx, (const void *)jacobi7_tb2_peer_kernel<8>, JacobiTBCfg<8>::kSmemBytes)) return st;
  return ensure_dyn_smem(ctx, (const void *)jacobi7_tb2_kernel<8>, JacobiTBCfg<8>::kSmemBytes);
}

bool jacobi7_tb2_supported(hpcx_grid_s *g) {
  return g->desc.dtype == HPCX_F32 && g->ctx->encode_tiled && g->n[2] < (1 << 30);
}

// Two Jacobi sweeps src -> dst in one pass over planes [zbeg, zend) (shared with dist.cu).  nb_lo /
// [zbeg2, zend2): optional second strip in the same launch.  nb_lo /
// nb_lo / nb_hi: the slab has a z neighbour on that side, i.e. ghost planes -2, -1 / n0, n0+1 of `src`
// hold the neighbour's iterate t and iterate t+1 of planes -1 / n0 is computed, not copied.
hpcx_status jacobi7_tb2_launch(hpcx_grid_s *g, void *src, void *dst, int zsplit, int64_t zbeg, int64_t zend,
                               bool nb_lo, bool nb_hi, cudaStream_t s, int64_t zbeg2, int64_t zend2) {
  if (!jacobi7_tb2_supported(g))
    return fail(g->ctx, HPCX_UNSUPPORTED, "the temporally blocked Jacobi needs fp32 and a tensor-map driver");
  return launch_jacobi_tb2(g, src, dst, zsplit, zbeg, zend, nb_lo, nb_hi, s, zbeg2, zend2);
}

// The same over a whole thin slab with the halo transport inside the kernel
// (jacobi7_tb2_peer_kernel).  Needs at least two planes.
hpcx_status jacobi7_tb2_peer_launch(hpcx_grid_s *g, void *src, void *dst, int zsplit, const JacobiPeer &peer,
                                    cudaStream_t s) {
  if (!jacobi7_tb2_supported(g) || g->n[0] < 2)
    return fail(g->ctx, HPCX_UNSUPPORTED, "the fused two-sweep transport needs fp32, a tensor-map driver and two planes");
  return launch_jacobi_tb2(g, src, dst, zsplit, 0, g->n[0], peer.ghost_lo != nullptr, peer.ghost_hi != nullptr, s, 0, 0,
                           &peer);
}
}  // namespace hpcx

extern "C" hpcx_status hpcx_stencil_sweep(hpcx_grid in, hpcx_grid out, const hpcx_stencil_desc *st,
                                          int32_t steps, void *stream) {
  if (!in || !out || !st) return HPCX_INVALID_ARG;
  HPCX_RANGE("hpcx_stencil_sweep");
  hpcx_ctx ctx = in->ctx;
  hpcx_status rc = check_same_shape(in, out);
  if (rc != HPCX_OK) return rc;
  HPCX_REQUIRE(ctx, steps >= 1, "steps must be >= 1");
  HPCX_REQUIRE(ctx, in != out && in->dev != out->dev, "in and out must be different grids");
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->compute;

  if (st->kind == HPCX_STENCIL_BOX_X || st->kind == HPCX_STENCIL_BOX_Y) {
    if (in->desc.ndim != 2) return fail(ctx, HPCX_UNSUPPORTED, "box sweeps take 2-D grids");
    HPCX_REQUIRE(ctx, steps == 1, "box sweeps are single-step");
    HPCX_REQUIRE(ctx, st->radius >= 0 && st->radius <= in->desc.lo && st->radius <= in->desc.hi,
                 "radius exceeds the ghost width of the input grid");
    const int axis = st->kind == HPCX_STENCIL_BOX_X ? 2 : 1;
    rc = hpcx::box_launch_rows(in, out, st, 0, in->n[1], s);
  } else if (st->kind == HPCX_STENCIL_JACOBI7) {
    if (in->desc.ndim != 3) return fail(ctx, HPCX_UNSUPPORTED, "JACOBI7 takes 3-D grids");
    HPCX_REQUIRE(ctx, in->desc.lo >= 1 && in->desc.hi >= 1 && in->desc.lo == out->desc.lo &&
                          in->desc.hi == out->desc.hi,
                 "JACOBI7 needs identical ghost widths >= 1 on both grids");
    const bool temporal = st->variant == HPCX_VARIANT_TEMPORAL;
    if (temporal && !hpcx::jacobi7_tb2_supported(in))
      return fail(ctx, HPCX_UNSUPPORTED, "the temporally blocked Jacobi needs fp32 and a tensor-map driver");
    // TEMPORAL: steps/2 fused two-sweep passes, then one single sweep when steps is odd
    const int passes = temporal ? steps / 2 + (steps & 1) : steps;
    for (int t = 0; t < passes && rc == HPCX_OK; ++t) {
      void *src = (t & 1) ? out->dev : in->dev;
      void *dst = (t & 1) ? in->dev : out->dev;
      if (temporal && t < steps / 2)
        rc = launch_jacobi_tb2(in, src, dst, st->zsplit, 0, in->n[0], false, false, s);
      else
        rc = jacobi7_launch(in, src, dst, temporal ? HPCX_VARIANT_TMA : st->variant, temporal ? 0 : st->zsplit, 0,
                            in->n[0], s, nullptr);
    }
    if (rc == HPCX_OK && (passes & 1) == 0) std::swap(in->dev, out->dev);
  } else {
    return fail(ctx, HPCX_INVALID_ARG, "unknown stencil kind");
  }
  if (rc != HPCX_OK) return rc;
  if (!stream) HPCX_CUDA(ctx, cudaStreamSynchronize(s));
  return HPCX_OK;
}
