// Distributed.hpp — one process per GPU: the C++ face of the multi-GPU entry points of hpcx.h
// (SURVEY.md §8 a11-a13, §8e).  The reference has no transport of its own; what it has is the
// static chunking idiom every parallel body uses to split an index range
// (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:1356,1369-1377: chunk = ceil(n / parts)) and the
// two-stream copy/compute overlap of lib/gpu/transfer_overlap_testsuite/src/
// cuda_tut_transfer_overlap.cu:92-103.  The classes below keep those conventions:
//
//   hpcx::Communicator      NCCL bootstrap: rank 0 makes the 128-byte id, the caller broadcasts it
//                           (MPI, a file, torch.distributed ...), every rank joins.
//   hpc::partition_slabs    the static chunking of n planes over the ranks.
//   hpc::HaloPlan           z-slab Jacobi over a DeviceGrid pair (hpcx_stencil_dist_*).
//   sparse::partition_rows  row blocks by the same chunking or by nnz prefix sums.
//   sparse::DistributedCsr  this rank's block of a host CsrMatrix: local matrix [owned | halo],
//                           exchange lists, distributed product (hpcx_csr_localize,
//                           hpcx_csr_set_halo, hpcx_spmv_dist).
//
// Errors follow Runtime.hpp: std::out_of_range for bad sizes, std::system_error otherwise.
#ifndef HPCX_DISTRIBUTED_HPP_
#define HPCX_DISTRIBUTED_HPP_

#include <array>
#include <cstdint>
#include <vector>

#include "CsrMatrix.hpp"
#include "Runtime.hpp"
#include "Stencil.hpp"

namespace hpcx {

using CommId = std::array<unsigned char, 128>;

// Joins the ranks of a job into one NCCL communicator held by the context (released by ~Context).
class Communicator {
 public:
  // rank 0 only: the id every rank needs.
  static CommId unique_id(Context &ctx) {
    CommId id{};
    HPCX_CHECK(ctx.handle(), hpcx_comm_unique_id(ctx.handle(), id.data()));
    return id;
  }
  Communicator(Context &ctx, int nranks, int rank, const CommId &id) : ctx_(ctx.handle()), nranks_(nranks), rank_(rank) {
    HPCX_CHECK(ctx_, hpcx_comm_init(ctx_, nranks, rank, id.data()));
  }
  // A job of one process: nothing to join (the context's default state; NCCL is not loaded).
  explicit Communicator(Context &ctx) : ctx_(ctx.handle()), nranks_(1), rank_(0) {}

  int size() const noexcept { return nranks_; }
  int rank() const noexcept { return rank_; }
  void barrier() const {
    if (nranks_ > 1) HPCX_CHECK(ctx_, hpcx_comm_barrier(ctx_));
  }

 private:
  hpcx_ctx ctx_;
  int nranks_, rank_;
};

}  // namespace hpcx

namespace hpc {

// bounds[r] .. bounds[r+1] = planes of rank r (static chunking, test_hpc_benchmark.cpp:1356).
inline std::vector<std::int64_t> partition_slabs(std::int64_t n, int nparts) {
  if (n < 0 || nparts < 1) throw std::out_of_range("invalid index, out of boundary");
  std::vector<std::int64_t> b(static_cast<std::size_t>(nparts) + 1);
  const hpcx_status st = hpcx_partition_slabs(n, nparts, b.data());
  if (st != HPCX_OK) hpcx::throwStatus(st, "hpcx_partition_slabs", __FILE__, __LINE__);
  return b;
}

enum class Transport : int {
  Nccl = 0,       // grouped ncclSend/ncclRecv on the comm stream, overlapped with the interior
  PeerStores = 1  // the compute kernels store into the neighbours' ghosts over NVLink, flags order it
};

// Jacobi7 over z-slabs: `in` / `out` are this rank's planes plus ghosts; neighbours are rank-1 /
// rank+1.  Jacobi7::variant = HPCX_VARIANT_TEMPORAL needs Lo, Hi >= 2 (two planes per side and pass).
template <typename T, std::size_t Lo, std::size_t Hi>
class HaloPlan {
 public:
  HaloPlan(DeviceGrid<3, T, Lo, Hi> &in, DeviceGrid<3, T, Lo, Hi> &out, Jacobi7 j = {},
           Transport transport = Transport::PeerStores)
      : ctx_(in.context()) {
    static_assert(Lo >= 1 && Hi >= 1, "Jacobi7 needs at least one ghost cell per side");
    hpcx_stencil_desc st{};
    st.kind = HPCX_STENCIL_JACOBI7;
    st.variant = j.variant;
    HPCX_CHECK(ctx_, hpcx_stencil_dist_plan(ctx_, in.handle(), out.handle(), &st, static_cast<int32_t>(transport),
                                            &plan_));
  }
  ~HaloPlan() {
    if (plan_) hpcx_stencil_dist_destroy(plan_);
  }
  HaloPlan(const HaloPlan &) = delete;
  HaloPlan &operator=(const HaloPlan &) = delete;

  // `steps` sweeps in -> out (the newest iterate ends up in `out`), ghosts exchanged on the way.
  void sweep(int steps, void *stream = nullptr) { HPCX_CHECK(ctx_, hpcx_stencil_dist_sweep(plan_, steps, stream)); }

 private:
  hpcx_ctx ctx_ = nullptr;
  hpcx_halo_plan plan_ = nullptr;
};

// The reference's box sweeps over a py x px grid of 2-D blocks (px == 1: y-slabs), one block per
// rank, rank = ry * px + rx; ghost rows / columns come from the neighbouring blocks
// (hpcx_stencil_dist_plan_blocks).  One sweep in -> out per call.
template <typename T, std::size_t Lo, std::size_t Hi, std::size_t Lo2, std::size_t Hi2>
class BoxHaloPlan {
 public:
  BoxHaloPlan(DeviceGrid<2, T, Lo, Hi> &in, DeviceGrid<2, T, Lo2, Hi2> &out, Box1D box, int py, int px)
      : ctx_(in.context()) {
    if (box.axis != 0 && box.axis != 1)
      hpcx::throwStatus(HPCX_INVALID_ARG, "Box1D axis must be 0 or 1", __FILE__, __LINE__);
    hpcx_stencil_desc st{};
    st.kind = box.axis == 1 ? HPCX_STENCIL_BOX_X : HPCX_STENCIL_BOX_Y;
    st.radius = box.radius;
    HPCX_CHECK(ctx_, hpcx_stencil_dist_plan_blocks(ctx_, in.handle(), out.handle(), &st, py, px, &plan_));
  }
  ~BoxHaloPlan() {
    if (plan_) hpcx_stencil_dist_destroy(plan_);
  }
  BoxHaloPlan(const BoxHaloPlan &) = delete;
  BoxHaloPlan &operator=(const BoxHaloPlan &) = delete;
  void sweep(void *stream = nullptr) { HPCX_CHECK(ctx_, hpcx_stencil_dist_sweep(plan_, 1, stream)); }

 private:
  hpcx_ctx ctx_ = nullptr;
  hpcx_halo_plan plan_ = nullptr;
};

}  // namespace hpc

namespace sparse {

enum class RowSplit : int { EqualRows = 0, EqualNnz = 1 };

// bounds[r] .. bounds[r+1] = rows of rank r.
template <typename T, typename Index>
std::vector<std::int64_t> partition_rows(const CsrMatrix<T, Index> &A, int nparts, RowSplit mode = RowSplit::EqualRows) {
  if (nparts < 1) throw std::out_of_range("invalid index, out of boundary");
  std::vector<std::int64_t> b(static_cast<std::size_t>(nparts) + 1);
  const hpcx_status st = hpcx_partition_rows(static_cast<std::int64_t>(A.nrows()), A.row_ptr(),
                                             static_cast<int32_t>(sizeof(Index)), nparts, static_cast<int32_t>(mode),
                                             b.data());
  if (st != HPCX_OK) hpcx::throwStatus(st, "hpcx_partition_rows", __FILE__, __LINE__);
  return b;
}

// This rank's row block of a square host matrix, resident on the device with its halo lists.
// x and y of multiply() are the rank's owned entries (rows first() .. last()), on the device.
template <typename T>
class DistributedCsr {
 public:
  template <typename Index>
  DistributedCsr(hpcx::Context &ctx, const hpcx::Communicator &comm, const CsrMatrix<T, Index> &A,
                 const std::vector<std::int64_t> &bounds, hpc::Transport transport = hpc::Transport::PeerStores,
                 const hpcx_csr_opts *opts = nullptr)
      : ctx_(ctx.handle()) {
    const int P = comm.size(), r = comm.rank();
    if (A.nrows() != A.ncols() || bounds.size() != static_cast<std::size_t>(P) + 1)
      throw std::out_of_range("invalid index, out of boundary");
    first_ = bounds[r];
    last_ = bounds[r + 1];
    const auto n = static_cast<std::int64_t>(A.nrows());
    const auto ib = static_cast<int32_t>(sizeof(Index));
    hpcx_halo_sizes sz{};
    check(hpcx_csr_localize(n, A.row_ptr(), ib, A.col_idx(), bounds.data(), P, r, &sz, nullptr, nullptr, nullptr,
                            nullptr, nullptr, nullptr, nullptr));
    std::vector<std::int64_t> rp(sz.nloc + 1), recv(P), send(P);
    std::vector<std::int32_t> col(sz.nnz_loc), send_idx(sz.nsend);
    std::vector<std::uint8_t> kind(sz.nloc);
    halo_gcol_.resize(sz.nhalo);
    check(hpcx_csr_localize(n, A.row_ptr(), ib, A.col_idx(), bounds.data(), P, r, &sz, rp.data(), col.data(),
                            halo_gcol_.data(), recv.data(), send.data(), send_idx.data(), kind.data()));
    hpcx_csr_desc d{};
    d.nrows = sz.nloc;
    d.ncols = sz.nloc + sz.nhalo;
    d.nnz = sz.nnz_loc;
    d.row_ptr = rp.data();
    d.idx_bytes = 8;
    d.dtype = hpcx::dtypeOf<T>();
    d.col_idx = col.data();
    d.values = A.values() + A.row_ptr()[first_];  // value positions are unchanged by the renumbering
    d.mem = HPCX_MEM_HOST;
    HPCX_CHECK(ctx_, hpcx_csr_create(ctx_, &d, opts, &A_));
    hpcx_status st = hpcx_csr_set_halo(A_, sz.nloc, sz.nhalo, recv.data(), send.data(), send_idx.data(), kind.data());
    if (st == HPCX_OK && transport != hpc::Transport::Nccl) st = hpcx_csr_halo_transport(A_, static_cast<int32_t>(transport));
    if (st != HPCX_OK) {
      hpcx_csr_destroy(A_);
      A_ = nullptr;
      HPCX_CHECK(ctx_, st);
    }
  }
  ~DistributedCsr() {
    if (A_) hpcx_csr_destroy(A_);
  }
  DistributedCsr(const DistributedCsr &) = delete;
  DistributedCsr &operator=(const DistributedCsr &) = delete;

  // y_local = (A x)[first() .. last()): gather / push of the boundary entries of x, interior rows,
  // exchange, boundary rows — all inside the call.  Device pointers, owned() entries each.
  void multiply(const T *x_local, T *y_local, void *stream = nullptr) const {
    HPCX_CHECK(ctx_, hpcx_spmv_dist(A_, x_local, y_local, stream));
  }

  std::int64_t first() const noexcept { return first_; }
  std::int64_t last() const noexcept { return last_; }
  std::int64_t owned() const noexcept { return last_ - first_; }
  // global column of every halo slot, in slot order (sorted by owner, then column)
  const std::vector<std::int64_t> &halo_columns() const noexcept { return halo_gcol_; }
  hpcx_csr handle() const noexcept { return A_; }
  hpcx_ctx context() const noexcept { return ctx_; }

 private:
  static void check(hpcx_status st) {
    if (st != HPCX_OK) hpcx::throwStatus(st, "hpcx_csr_localize", __FILE__, __LINE__);
  }
  hpcx_ctx ctx_ = nullptr;
  hpcx_csr A_ = nullptr;
  std::int64_t first_ = 0, last_ = 0;
  std::vector<std::int64_t> halo_gcol_;
};

}  // namespace sparse
#endif  // HPCX_DISTRIBUTED_HPP_
