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
from .api import Context, CsrMatrix, DeviceCsr, DeviceGrid, csr_localize, stencil_desc


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


class HaloPlan:
    """Distributed JACOBI7 over z-slabs (hpcx_stencil_dist_*)."""

    def __init__(self, ctx: Context, gin: DeviceGrid, gout: DeviceGrid, kind: int, variant: int = 0,
                 zsplit: int = 0, transport: int = 0):
        self.ctx = ctx
        d = stencil_desc(kind, 0, variant, zsplit)
        h = C.c_void_p()
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
    D = DeviceCsr(ctx, A, **opts)
    D.set_halo(local)
    if transport:
        D.halo_transport(transport)
    return D
