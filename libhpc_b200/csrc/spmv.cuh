// spmv.cuh — internal CSR matrix handle (new container next to the reference's lib/sparse;
// the reference has no CSR/SpMV code, SURVEY.md §0 — layout follows §8 a7).
#pragma once
#include "common.cuh"

namespace hpcx {

// One CTA of the STREAM kernel processes one row block: either a run of whole rows whose
// nonzeros fit one shared-memory tile, or one tile-sized chunk of a row longer than a tile.
struct RowBlock {
  int32_t row0;   // first row
  int32_t nrows;  // rows in the block (1 for a long-row chunk)
  int32_t part;   // -1, or slot in the partial-sum array (long-row chunk)
  int32_t lanes;  // threads cooperating on one row: 1,2,4,..,32, or 0 = whole CTA
  int64_t p0;     // first nonzero
  int64_t p1;     // one past the last nonzero
};
static_assert(sizeof(RowBlock) == 32, "RowBlock is loaded as two 16-byte words");

// Distributed product in ONE launch (transport 1, dist.cu): the grid is [push CTAs | interior row
// blocks | boundary row blocks].  Push CTAs gather this rank's boundary x entries and store them
// straight into the neighbours' halo slots over NVLink, then release the product number there;
// boundary CTAs acquire the source ranks' flags before their first halo gather; the last boundary
// CTA to finish tells the source ranks that their entries have been consumed (the ack the next
// product's push waits for).
struct SpmvDistArgs {
  int npush = 0, push_per_seg = 1, nseg = 0;  // push CTAs: push_per_seg per destination segment
  int n_int = 0, n_bnd = 0;                   // row blocks [0, n_int) read owned x only
  int nsrc = 0;
  uint32_t epoch = 0;
  const int32_t *send_idx = nullptr;   // owned entries to gather, grouped by destination
  const int32_t *seg_start = nullptr;  // nseg + 1 offsets into send_idx
  const int32_t *seg_rank = nullptr;   // destination rank of each segment
  void *const *seg_dst = nullptr;      // where each segment lands in the destination's [owned | halo] x
  uint32_t *const *seg_flag = nullptr; // the destination's data flag for this rank
  const int32_t *src_rank = nullptr;   // ranks this one receives from
  uint32_t *const *src_ack = nullptr;  // their ack word for this rank
  uint32_t *sync = nullptr;            // local block: data flags, acks, counters (dist.cu)
  PeerWait pw;
};

// Chunking of a host-operand product (spmv.cu: pipe_analyse / spmv_host_pipelined).
struct PipePlan {
  int state = 0;                 // 0 = not analysed yet, 1 = usable, -1 = not usable
  int K = 0;                     // chunks (= pieces of x)
  std::vector<int64_t> first;    // first row block of each chunk (+ end)
  std::vector<int32_t> cmax;     // largest travelling column index each chunk reads
  std::vector<cudaEvent_t> ev;   // [0, K) x pieces landed, [K, 2K) chunk computed, 2K entry, 2K+1 done
  uint32_t dist_mask = 0;        // distributed plan: bit j = piece j of the owned x is read by the push CTAs
                                 // or by a boundary row, i.e. must have landed before that launch
};

// What one rank sees of another for the fused x-halo transport (dist.cu: csr_peer_wire).
struct CsrPeerView {
  char *xfull = nullptr;        // the peer's [owned | halo] x buffer, addressable from this device
  uint32_t *sync = nullptr;     // the peer's sync block
  int64_t nloc = 0;             // the peer's owned entries
  int64_t recv_off_me = 0, recv_cnt_me = 0;  // where this rank's segment lands in the peer's halo
};
hpcx_status csr_peer_prepare(hpcx_csr A);
hpcx_status csr_peer_wire(hpcx_csr A, const std::vector<CsrPeerView> &views);

// Tables of the HUB kernel (spmv_hub.cu), built on the device at create time.
struct HubTables {
  bool ready = false;
  int K = 0;                    // hub columns (their x entries live in shared memory)
  double coverage = 0.0;        // fraction of the nonzeros whose column is a hub column
  // long part (rows of more than 256 entries): nonzero-balanced tiles of 256
  int64_t nlong = 0, nnzL = 0, ntiles = 0;
  int32_t *colL = nullptr;      // hub columns encoded as 0x80000000 | slot, padded to whole tiles
  void *valL = nullptr;
  uint8_t *rowstart = nullptr;  // bit i: nonzero i of the long part is the first of its row
  int32_t *tile_seg0 = nullptr; // ntiles + 1: long rows that start before each tile
  int32_t *lrow = nullptr;      // original row of each long row
  int32_t *lrp = nullptr;       // nlong + 1 offsets into colL / valL
  // short part (rows of 1..256 entries): SELL-32 slices over the rows sorted by descending length
  int64_t nshort = 0, nnzS = 0;  // nnzS = slots of the slices (padding included)
  int32_t *colS = nullptr;
  void *valS = nullptr;
  int64_t *sbase = nullptr;      // first slot of each slice (+ total)
  uint16_t *slen = nullptr;      // length of each sorted short row
  int32_t *srow = nullptr;       // original row of each sorted short row
  int32_t *hub_col = nullptr;   // column of each hub slot
  int64_t nwarps = 0;           // warp ranges of the persistent grid
  void *head_out = nullptr, *tail_out = nullptr;  // per warp range partials
};
hpcx_status hub_build(hpcx_csr A);
hpcx_status sell_build(hpcx_csr A, int64_t maxlen);  // sell_build.cu: SELL-C-sigma arrays on the device
hpcx_status hub_launch(hpcx_csr A, const void *x, void *y, cudaStream_t s);
void hub_free(hpcx_csr A);

}  // namespace hpcx

struct hpcx_csr_s {
  hpcx_ctx ctx = nullptr;
  int64_t nrows = 0, ncols = 0, nnz = 0;
  int32_t dtype = HPCX_F64, es = 8, idx_bytes = 4;
  int32_t kernel = HPCX_SPMV_STREAM, tile = 2048, lanes_per_row = 0, short_row_max = 48;
  // device CSR (row_ptr always int32 on the device; nnz < 2^31 is enforced at create)
  int32_t *d_row_ptr = nullptr;
  int32_t *d_col = nullptr;
  void *d_val = nullptr;
  int64_t nnz_alloc = 0;  // padded allocation, multiple of 4 plus slack for aligned bulk copies
  // STREAM kernel tables
  std::vector<hpcx::RowBlock> h_blocks;
  hpcx::RowBlock *d_blocks = nullptr;
  int64_t max_row_len = 0;
  int64_t nlong = 0;            // rows longer than a tile
  int32_t *d_long_row = nullptr;
  int32_t *d_long_off = nullptr;  // nlong + 1
  void *d_partials = nullptr;
  int64_t bins[34] = {};
  // SELL-C-sigma
  int32_t sell_c = 0, sell_sigma = 0;
  int64_t sell_nslices = 0, sell_padded = 0;
  int32_t *d_perm = nullptr, *d_slice_len = nullptr, *d_sell_col = nullptr;
  int64_t *d_slice_ptr = nullptr;
  void *d_sell_val = nullptr;
  hpcx::HubTables hub;
  // staging for host operands
  void *d_x = nullptr, *d_y = nullptr;
  // host-x / host-y products: row chunks pipelined against the PCIe copies (spmv.cu:
  // spmv_host_pipelined); pipe_dist is the same over the interior blocks of a distributed matrix
  hpcx::PipePlan pipe, pipe_dist;
  // when set (by the CG solver, cg.cu) the STREAM launches also emit per-block <x, y> partials
  void *dot_part = nullptr;
  // distributed (set by hpcx_csr_set_halo)
  int64_t nloc = 0, nhalo = 0, nsend = 0;
  std::vector<int64_t> recv_counts, send_counts;
  int32_t *d_send_idx = nullptr;
  void *d_sendbuf = nullptr;
  void *d_xfull = nullptr;  // [owned | halo]
  // one table [interior blocks | boundary blocks] (csr_plan.cu); the two pointers below view it
  hpcx::RowBlock *d_blocks_dist = nullptr;
  std::vector<hpcx::RowBlock> h_blocks_int, h_blocks_bnd;
  hpcx::RowBlock *d_blocks_int = nullptr, *d_blocks_bnd = nullptr;
  // transport 1 (peer stores, hpcx_csr_halo_transport): see dist.cu
  int32_t transport = 0;
  uint32_t epoch = 0;          // distributed products issued through this handle
  uint32_t *d_sync = nullptr;  // [0..63] data flags by source rank, [64..127] acks by destination
                               // rank, [128] CTA counter, [129] error word
  std::vector<void *> peer_open;  // cudaIpcOpenMemHandle bases to close
  int32_t nseg = 0, nsrc = 0;     // destination / source neighbours
  int32_t *d_seg_start = nullptr; // nseg + 1 offsets into send_idx
  int32_t *d_seg_rank = nullptr;  // destination rank of each segment
  void **d_seg_dst = nullptr;     // where each segment lands in the destination's [owned | halo] x
  uint32_t **d_seg_flag = nullptr;// the destination's data flag for this rank
  int32_t *d_src_rank = nullptr;  // ranks this one receives from
  uint32_t **d_src_ack = nullptr; // their ack word for this rank
  int64_t nlong_dist = 0;  // long-row tables in the numbering of the interior/boundary blocks
  int32_t *d_long_row_dist = nullptr, *d_long_off_dist = nullptr;
};
