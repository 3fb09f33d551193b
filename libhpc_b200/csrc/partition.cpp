// partition.cpp — host partitioner and halo index maps.
//
// The reference has no multi-device code; the only partitioning idiom it contains is the
// static chunking of its multithreaded radix sort,
//   chunk = (n + nproc - 1) / nproc;  [tid*chunk, min((tid+1)*chunk, n))
// (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:1356,1371-1372), which mode 0 keeps.
// Everything here is integer work on the host and is the production partitioner: the GPU
// path consumes these tables as they are, so "bit-exact" holds by construction; the
// independent restatement in oracle/hpcx_oracle.c is what the tests compare against.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/hpcx.h"

namespace {

inline int64_t rp_at(const void *rp, int idx_bytes, int64_t i) {
  return idx_bytes == 4 ? (int64_t) static_cast<const int32_t *>(rp)[i]
                        : static_cast<const int64_t *>(rp)[i];
}

inline int owner_of(const int64_t *bounds, int nparts, int64_t col) {
  // bounds is ascending; empty parts are skipped by upper_bound.
  return (int)(std::upper_bound(bounds, bounds + nparts + 1, col) - bounds) - 1;
}

}  // namespace

extern "C" {

hpcx_status hpcx_partition_slabs(int64_t n, int32_t nparts, int64_t *bounds) {
  if (n < 0 || nparts < 1 || !bounds) return HPCX_INVALID_ARG;
  const int64_t chunk = (n + nparts - 1) / nparts;
  for (int p = 0; p <= nparts; ++p) bounds[p] = std::min<int64_t>((int64_t)p * chunk, n);
  return HPCX_OK;
}

hpcx_status hpcx_partition_rows(int64_t nrows, const void *row_ptr, int32_t idx_bytes,
                                int32_t nparts, int32_t mode, int64_t *bounds) {
  if (nrows < 0 || nparts < 1 || !bounds) return HPCX_INVALID_ARG;
  if (mode == 0) return hpcx_partition_slabs(nrows, nparts, bounds);
  if (mode != 1 || !row_ptr || (idx_bytes != 4 && idx_bytes != 8)) return HPCX_INVALID_ARG;
  const int64_t nnz = rp_at(row_ptr, idx_bytes, nrows);
  bounds[0] = 0;
  for (int p = 1; p < nparts; ++p) {
    const int64_t target = (int64_t)((__int128)nnz * p / nparts);
    // first row r with row_ptr[r] >= target
    int64_t lo = 0, hi = nrows;
    while (lo < hi) {
      const int64_t mid = lo + (hi - lo) / 2;
      if (rp_at(row_ptr, idx_bytes, mid) >= target)
        hi = mid;
      else
        lo = mid + 1;
    }
    bounds[p] = std::max(lo, bounds[p - 1]);
  }
  bounds[nparts] = nrows;
  return HPCX_OK;
}

hpcx_status hpcx_csr_localize(int64_t nrows, const void *row_ptr, int32_t idx_bytes,
                              const int32_t *col_idx, const int64_t *bounds, int32_t nparts,
                              int32_t rank, hpcx_halo_sizes *sizes, int64_t *local_row_ptr,
                              int32_t *local_col, int64_t *halo_gcol, int64_t *recv_counts,
                              int64_t *send_counts, int32_t *send_idx, uint8_t *row_kind) {
  if (nrows < 0 || !row_ptr || (idx_bytes != 4 && idx_bytes != 8) || !bounds || nparts < 1 ||
      rank < 0 || rank >= nparts || !sizes)
    return HPCX_INVALID_ARG;
  if (bounds[0] != 0 || bounds[nparts] != nrows) return HPCX_INVALID_ARG;
  for (int p = 0; p < nparts; ++p)
    if (bounds[p + 1] < bounds[p]) return HPCX_INVALID_ARG;
  const int64_t r0 = bounds[rank], r1 = bounds[rank + 1];
  const int64_t nloc = r1 - r0;
  const int64_t nnz = rp_at(row_ptr, idx_bytes, nrows);
  const int64_t p0 = rp_at(row_ptr, idx_bytes, r0), p1 = rp_at(row_ptr, idx_bytes, r1);
  if (nnz > 0 && !col_idx) return HPCX_INVALID_ARG;

  try {
    // Halo columns: every column outside [r0, r1) referenced by an owned row, ascending
    // global index (== ascending (owner, column) because owners are contiguous ranges).
    std::vector<int32_t> slot((size_t)nrows, -1);
    for (int64_t p = p0; p < p1; ++p) {
      const int64_t c = col_idx[p];
      if (c < 0 || c >= nrows) return HPCX_INVALID_ARG;
      if (c < r0 || c >= r1) slot[(size_t)c] = 0;
    }
    int64_t nhalo = 0;
    for (int64_t c = 0; c < nrows; ++c)
      if (slot[(size_t)c] == 0) slot[(size_t)c] = (int32_t)(nloc + nhalo++);
    if (nloc + nhalo > INT32_MAX) return HPCX_UNSUPPORTED;

    // Send lists: owned columns referenced by the rows of every other rank.
    std::vector<uint8_t> mark;
    std::vector<int64_t> scount((size_t)nparts, 0);
    int64_t nsend = 0;
    const bool want_send = send_idx != nullptr;
    int64_t cursor = 0;
    mark.assign((size_t)std::max<int64_t>(nloc, 1), 0);
    for (int q = 0; q < nparts; ++q) {
      if (q == rank || nloc == 0) continue;
      const int64_t a = rp_at(row_ptr, idx_bytes, bounds[q]);
      const int64_t b = rp_at(row_ptr, idx_bytes, bounds[q + 1]);
      bool any = false;
      for (int64_t p = a; p < b; ++p) {
        const int64_t c = col_idx[p];
        if (c >= r0 && c < r1) {
          mark[(size_t)(c - r0)] = 1;
          any = true;
        }
      }
      if (!any) continue;
      int64_t cnt = 0;
      for (int64_t i = 0; i < nloc; ++i)
        if (mark[(size_t)i]) {
          if (want_send) send_idx[cursor + cnt] = (int32_t)i;
          ++cnt;
          mark[(size_t)i] = 0;
        }
      scount[(size_t)q] = cnt;
      cursor += cnt;
      nsend += cnt;
    }

    sizes->nloc = nloc;
    sizes->nnz_loc = p1 - p0;
    sizes->nhalo = nhalo;
    sizes->nsend = nsend;

    if (local_row_ptr)
      for (int64_t i = 0; i <= nloc; ++i) local_row_ptr[i] = rp_at(row_ptr, idx_bytes, r0 + i) - p0;
    if (local_col)
      for (int64_t p = p0; p < p1; ++p) {
        const int64_t c = col_idx[p];
        local_col[p - p0] = (c >= r0 && c < r1) ? (int32_t)(c - r0) : slot[(size_t)c];
      }
    if (row_kind)
      for (int64_t i = 0; i < nloc; ++i) {
        uint8_t k = 0;
        for (int64_t p = rp_at(row_ptr, idx_bytes, r0 + i); p < rp_at(row_ptr, idx_bytes, r0 + i + 1);
             ++p) {
          const int64_t c = col_idx[p];
          if (c < r0 || c >= r1) {
            k = 1;
            break;
          }
        }
        row_kind[i] = k;
      }
    if (halo_gcol || recv_counts) {
      if (recv_counts) std::fill(recv_counts, recv_counts + nparts, (int64_t)0);
      int64_t h = 0;
      for (int64_t c = 0; c < nrows; ++c)
        if (slot[(size_t)c] >= 0) {
          if (halo_gcol) halo_gcol[h] = c;
          if (recv_counts) recv_counts[owner_of(bounds, nparts, c)]++;
          ++h;
        }
    }
    if (send_counts) std::copy(scount.begin(), scount.end(), send_counts);
  } catch (const std::bad_alloc &) {
    return HPCX_OOM;
  }
  return HPCX_OK;
}


// ---- the same tables without the global matrix -------------------------------------------
// hpcx_csr_localize wants the whole matrix on every rank and a table of nrows entries; a rank of a
// real distributed job holds only its own rows.  The two steps below need just those: step 1 is
// local, step 2 turns the column lists the other ranks ask for (one exchange of integer lists,
// done by the caller: torch.distributed / MPI / hpcx_comm_*) into this rank's send list.  The
// results are identical, bit for bit, to hpcx_csr_localize on the global matrix (tests/
// test_dist_cpu.py, world sizes 2 and 3).
hpcx_status hpcx_csr_localize_rows(int64_t nrows_global, int64_t nloc, const void *row_ptr, int32_t idx_bytes,
                                   const int32_t *col_global, const int64_t *bounds, int32_t nparts,
                                   int32_t rank, hpcx_halo_sizes *sizes, int32_t *local_col,
                                   int64_t *halo_gcol, int64_t *recv_counts, uint8_t *row_kind) {
  if (nrows_global < 0 || nloc < 0 || !row_ptr || (idx_bytes != 4 && idx_bytes != 8) || !bounds || nparts < 1 ||
      rank < 0 || rank >= nparts || !sizes)
    return HPCX_INVALID_ARG;
  if (bounds[0] != 0 || bounds[nparts] != nrows_global) return HPCX_INVALID_ARG;
  for (int p = 0; p < nparts; ++p)
    if (bounds[p + 1] < bounds[p]) return HPCX_INVALID_ARG;
  const int64_t r0 = bounds[rank], r1 = bounds[rank + 1];
  if (r1 - r0 != nloc) return HPCX_INVALID_ARG;
  if (rp_at(row_ptr, idx_bytes, 0) != 0) return HPCX_INVALID_ARG;
  const int64_t nnz = rp_at(row_ptr, idx_bytes, nloc);
  if (nnz > 0 && !col_global) return HPCX_INVALID_ARG;
  try {
    // halo columns: sorted distinct columns outside [r0, r1) — ascending global index, which is
    // ascending (owner, column) because owners are contiguous ranges
    std::vector<int32_t> halo;
    for (int64_t p = 0; p < nnz; ++p) {
      const int64_t c = col_global[p];
      if (c < 0 || c >= nrows_global) return HPCX_OUT_OF_RANGE;
      if (c < r0 || c >= r1) halo.push_back((int32_t)c);
    }
    std::sort(halo.begin(), halo.end());
    halo.erase(std::unique(halo.begin(), halo.end()), halo.end());
    const int64_t nhalo = (int64_t)halo.size();
    if (nloc + nhalo > INT32_MAX) return HPCX_UNSUPPORTED;
    sizes->nloc = nloc;
    sizes->nnz_loc = nnz;
    sizes->nhalo = nhalo;
    sizes->nsend = -1;  // known after the exchange (hpcx_halo_send_lists)
    if (local_col)
      for (int64_t p = 0; p < nnz; ++p) {
        const int64_t c = col_global[p];
        local_col[p] = (c >= r0 && c < r1)
                           ? (int32_t)(c - r0)
                           : (int32_t)(nloc + (std::lower_bound(halo.begin(), halo.end(), (int32_t)c) - halo.begin()));
      }
    if (halo_gcol)
      for (int64_t h = 0; h < nhalo; ++h) halo_gcol[h] = halo[(size_t)h];
    if (recv_counts) {
      std::fill(recv_counts, recv_counts + nparts, (int64_t)0);
      for (int64_t h = 0; h < nhalo; ++h) recv_counts[owner_of(bounds, nparts, halo[(size_t)h])]++;
    }
    if (row_kind)
      for (int64_t i = 0; i < nloc; ++i) {
        uint8_t k = 0;
        for (int64_t p = rp_at(row_ptr, idx_bytes, i); p < rp_at(row_ptr, idx_bytes, i + 1); ++p) {
          const int64_t c = col_global[p];
          if (c < r0 || c >= r1) {
            k = 1;
            break;
          }
        }
        row_kind[i] = k;
      }
  } catch (const std::bad_alloc &) {
    return HPCX_OOM;
  }
  return HPCX_OK;
}

hpcx_status hpcx_halo_send_lists(const int64_t *bounds, int32_t nparts, int32_t rank, const int64_t *req_counts,
                                 const int64_t *requests, int32_t *send_idx) {
  if (!bounds || nparts < 1 || rank < 0 || rank >= nparts || !req_counts) return HPCX_INVALID_ARG;
  const int64_t r0 = bounds[rank], r1 = bounds[rank + 1];
  int64_t off = 0;
  for (int q = 0; q < nparts; ++q) {
    if (req_counts[q] < 0 || (q == rank && req_counts[q] != 0)) return HPCX_INVALID_ARG;
    for (int64_t i = 0; i < req_counts[q]; ++i) {
      if (!requests || !send_idx) return HPCX_INVALID_ARG;
      const int64_t c = requests[off + i];
      if (c < r0 || c >= r1) return HPCX_OUT_OF_RANGE;            // asked for a column this rank does not own
      if (i > 0 && c <= requests[off + i - 1]) return HPCX_INVALID_ARG;  // lists are ascending and distinct
      send_idx[off + i] = (int32_t)(c - r0);
    }
    off += req_counts[q];
  }
  return HPCX_OK;
}

}  // extern "C"
