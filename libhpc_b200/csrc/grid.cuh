// grid.cuh — device mirror of hpc::HPCHighDimensionFlatArray
// (reference: lib/hpc/include/HPCHighDimensionFlatArray.hpp:54-224).
#pragma once
#include "common.cuh"

struct hpcx_grid_s {
  hpcx_ctx ctx = nullptr;
  hpcx_grid_desc desc{};
  int es = 4;          // element size in bytes
  // Normalised to 3-D, slowest first; missing leading dimensions have n = 1 and no ghosts.
  int64_t n[3] = {1, 1, 1};    // logical extents
  int64_t glo[3] = {0, 0, 0};  // ghost cells per side
  int64_t ghi[3] = {0, 0, 0};
  int64_t ext[3] = {1, 1, 1};  // padded extents n + glo + ghi (host layout)
  int64_t xoff = 0;            // device: element offset of interior x = 0 inside a row
  int64_t pitch = 0;           // device: row pitch in elements
  int64_t rows = 0;            // ext[0] * ext[1]
  int64_t total = 0;           // rows * pitch
  void *dev = nullptr;
  // tensor maps of the TMA sweeps, keyed by (buffer, box height); see stencil.cu
  struct MapEntry {
    void *base = nullptr;
    int boxh = 0;
    CUtensorMap map;
  };
  std::vector<MapEntry> maps;
};

namespace hpcx {
// Peer-memory halo push of the fused Jacobi sweep (dist.cu, transport 1).  All pointers are null
// on a single GPU.  Flag words hold sweep numbers (monotonic across calls).
struct JacobiPeer {
  float *ghost_lo = nullptr;     // plane base of ghost plane z = n_peer in the LOW neighbour's dst
  float *ghost_hi = nullptr;     // plane base of ghost plane z = -1 in the HIGH neighbour's dst
  uint32_t *sig_lo = nullptr;    // flag word in the low / high neighbour that receives sig_val
  uint32_t *sig_hi = nullptr;    //   once this rank's boundary plane has landed there
  const uint32_t *flags = nullptr;  // local: [0] stored by the low neighbour, [1] by the high one
  uint32_t *counters = nullptr;  // local: [0], [1] CTA arrival counters per side
  uint32_t wait_val = 0;         // ghost planes of src are valid once flags[side] >= wait_val
  uint32_t sig_val = 0;
  PeerWait pw;                   // timeout policy of the flag waits (common.cuh)
};
// Device element index of logical (z, y, x) — ghost indices are negative / >= n.
__host__ __device__ inline int64_t grid_dev_index(int64_t z, int64_t y, int64_t x, int64_t glo0,
                                                  int64_t glo1, int64_t ext1, int64_t pitch,
                                                  int64_t xoff) {
  return ((z + glo0) * ext1 + (y + glo1)) * pitch + xoff + x;
}
}  // namespace hpcx
