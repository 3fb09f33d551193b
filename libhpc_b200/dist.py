"""One process per GPU: communicator bootstrap and the distributed sweep / product wrappers.

torch.distributed is plumbing only: it carries the 128-byte NCCL unique id from rank 0 to
the other ranks (and the barrier bench.py needs); every data-path transfer is issued by
libhpcx itself (hpcx_stencil_dist_sweep / hpcx_spmv_dist: ncclSend/ncclRecv on the
library's comm stream, overlapped with interior compute).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .api import SPMV_STREAM, Context, CsrMatrix, DeviceCsr, DeviceGrid, csr_localize, stencil_desc


def init_comm(ctx: Context, rank: int | None = None, world: int | None = None) -> None:
    """hpcx_comm_unique_id on rank 0 -> broadcast -> hpcx_comm_init everywhere."""
    import torch
    import torch.distributed as dist

    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    uid = np.zeros(128, dtype=np.uint8)
    if rank == 0:
        ctx.check(ctx.lib.hpcx_comm_unique_id(ctx.h, uid.ctypes.data))
    if world > 1:
        t = torch.from_numpy(uid)
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.broadcast(t, src=0)
        uid = t.cpu().numpy().copy()
    ctx.check(ctx.lib.hpcx_comm_init(ctx.h, world, rank, uid.ctypes.data))
    ctx.nranks, ctx.rank = world, rank


def barrier(ctx: Context) -> None:
    ctx.check(ctx.lib.hpcx_comm_barrier(ctx.h))


def block_bounds(ny: int, nx: int, py: int, px: int, rank: int):
    """(y0, y1, x0, x1) of the block rank = ry * px + rx owns in a py x px split of an ny x nx
    grid: both axes cut by the reference's static chunking idiom (hpcx_partition_slabs,
    test_hpc_benchmark.cpp:1356,1371-1372), which is what hpcx_stencil_dist_plan_blocks expects —
    blocks of one block row share their y extent, blocks of one block column their x extent."""
    from .api import partition_slabs
    if py < 1 or px < 1 or not 0 <= rank < py * px:
        raise ValueError("rank outside the py x px block grid")
    by, bx = partition_slabs(ny, py), partition_slabs(nx, px)
    ry, rx = divmod(rank, px)
    return int(by[ry]), int(by[ry + 1]), int(bx[rx]), int(bx[rx + 1])


class HaloPlan:
    """Distributed sweeps (hpcx_stencil_dist_*): JACOBI7 over z-slabs, the reference's box sweeps
    over y-slabs (radius = the reference's nblur) or, with blocks=(py, px), over a py x px grid of
    2-D blocks (rank = ry * px + rx; hpcx_stencil_dist_plan_blocks)."""

    def __init__(self, ctx: Context, gin: DeviceGrid, gout: DeviceGrid, kind: int, variant: int = 0,
                 zsplit: int = 0, transport: int = 0, radius: int = 0, blocks: tuple | None = None):
        self.ctx = ctx
        d = stencil_desc(kind, radius, variant, zsplit)
        h = C.c_void_p()
        if blocks is not None:
            ctx.check(ctx.lib.hpcx_stencil_dist_plan_blocks(ctx.h, gin.h, gout.h, C.byref(d), int(blocks[0]),
                                                            int(blocks[1]), C.byref(h)))
        else:
            ctx.check(ctx.lib.hpcx_stencil_dist_plan(ctx.h, gin.h, gout.h, C.byref(d), transport,
                                                     C.byref(h)))
        self.h = h

    def sweep(self, steps: int, stream=None):
        s = None if stream is None else int(getattr(stream, "cuda_stream", stream))
        self.ctx.check(self.ctx.lib.hpcx_stencil_dist_sweep(self.h, steps, s))

    def destroy(self):
        if self.h:
            self.ctx.lib.hpcx_stencil_dist_destroy(self.h)
            self.h = None


def distributed_csr(ctx: Context, local: dict, dtype, transport: int = 0, **opts) -> DeviceCsr:
    """Device matrix of this rank's block from the tables of csr_localize (+ halo lists).
    transport 1 selects the fused peer-store exchange (hpcx_csr_halo_transport)."""
    A = CsrMatrix(local["local_row_ptr"], local["local_col"],
                  np.ascontiguousarray(local["values"], dtype=dtype),
                  ncols=local["nloc"] + local["nhalo"])
    # every distributed product runs the STREAM kernel: AUTO would build (and keep) the HUB tables of
    # a power-law block, a second encoded copy of col / val that nothing here reads
    opts.setdefault("kernel", SPMV_STREAM)
    D = DeviceCsr(ctx, A, **opts)
    D.set_halo(local)
    if transport:
        D.halo_transport(transport)
    return D


def localize_rows_distributed(row_ptr, col_global, values, bounds, rank: int, world: int, nrows_global: int) -> dict:
    """The tables of csr_localize built from this rank's OWN rows only (hpcx_csr_localize_rows,
    one exchange of the halo column lists over torch.distributed, hpcx_halo_send_lists): no rank
    ever holds the global matrix.  row_ptr: nloc + 1 entries from 0; col_global: global columns."""
    import torch
    import torch.distributed as dist

    lib = _abi.load()
    row_ptr = np.ascontiguousarray(row_ptr)
    col_global = np.ascontiguousarray(col_global, dtype=np.int32)
    bounds = np.ascontiguousarray(bounds, dtype=np.int64)
    nloc = len(row_ptr) - 1
    sz = _abi.HaloSizes()
    head = (nrows_global, nloc, row_ptr.ctypes.data, row_ptr.dtype.itemsize, col_global.ctypes.data if len(col_global) else None,
            bounds.ctypes.data, world, rank, C.byref(sz))
    rc = lib.hpcx_csr_localize_rows(*head, None, None, None, None)
    if rc != _abi.OK:
        raise RuntimeError(f"hpcx_csr_localize_rows failed with status {rc}")
    out = dict(local_col=np.empty(sz.nnz_loc, dtype=np.int32), halo_gcol=np.empty(sz.nhalo, dtype=np.int64),
               recv_counts=np.empty(world, dtype=np.int64), row_kind=np.empty(nloc, dtype=np.uint8))
    rc = lib.hpcx_csr_localize_rows(*head, *[out[k].ctypes.data for k in ("local_col", "halo_gcol", "recv_counts", "row_kind")])
    if rc != _abi.OK:
        raise RuntimeError(f"hpcx_csr_localize_rows failed with status {rc}")
    # ---- exchange: who needs how many of whose columns, then the lists themselves
    on_gpu = world > 1 and dist.get_backend() == "nccl"
    dev = "cuda" if on_gpu else "cpu"
    req_counts = np.zeros(world, dtype=np.int64)
    requests = np.empty(0, dtype=np.int64)
    if world > 1:
        mine = torch.from_numpy(out["recv_counts"]).to(dev)
        allc = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allc, mine)
        req_counts = np.array([int(allc[q][rank]) for q in range(world)], dtype=np.int64)  # what q needs from me
        ops, recv_t, roff = [], [], 0
        for q in range(world):
            nr = int(out["recv_counts"][q])
            if nr:
                ops.append(dist.P2POp(dist.isend, torch.from_numpy(out["halo_gcol"][roff:roff + nr].copy()).to(dev), q))
            roff += nr
            if req_counts[q]:
                t = torch.empty(int(req_counts[q]), dtype=torch.int64, device=dev)
                recv_t.append(t)
                ops.append(dist.P2POp(dist.irecv, t, q))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if recv_t:
            requests = torch.cat(recv_t).cpu().numpy()
    send_idx = np.empty(int(req_counts.sum()), dtype=np.int32)
    rc = lib.hpcx_halo_send_lists(bounds.ctypes.data, world, rank, req_counts.ctypes.data,
                                  requests.ctypes.data if len(requests) else None,
                                  send_idx.ctypes.data if len(send_idx) else None)
    if rc != _abi.OK:
        raise RuntimeError(f"hpcx_halo_send_lists failed with status {rc}")
    rp64 = row_ptr.astype(np.int64)
    out.update(nloc=nloc, nnz_loc=int(sz.nnz_loc), nhalo=int(sz.nhalo), nsend=len(send_idx), send_idx=send_idx,
               send_counts=req_counts, local_row_ptr=rp64, values=values)
    return out
