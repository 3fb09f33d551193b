"""Multi-GPU parity check, one process per GPU (launched by torchrun; see test_gpu_dist.py):
distributed JACOBI7 over z-slabs and row-partitioned SpMV with x-halo exchange, each rank
comparing its block with the oracle's GLOBAL result.  Prints 'DIST OK' on rank 0."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import libhpc_b200 as hp  # noqa: E402
import oracle  # noqa: E402
from libhpc_b200 import dist as hd  # noqa: E402
from libhpc_b200 import synth  # noqa: E402
from util import assert_spmv_close, bits_equal  # noqa: E402


def main():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = hp.Context(local)
    if world > 1:
        hd.init_comm(ctx, rank, world)

    # ---- stencil: global (nz, ny, nx) split into z-slabs
    # transport 0: NCCL send/recv on the comm stream; 1: fused peer stores + flags (TMA sweep)
    for (nz, ny, nx), steps, variant, transport in (
            ((24, 20, 140), 5, hp.VARIANT_TMA, 0), ((13, 9, 33), 4, hp.VARIANT_DIRECT, 0),
            ((64, 64, 256), 6, hp.VARIANT_AUTO, 0), ((24, 20, 140), 5, hp.VARIANT_TMA, 1),
            ((64, 64, 256), 6, hp.VARIANT_AUTO, 1), ((17, 33, 129), 3, hp.VARIANT_TMA, 1),
            ((8, 40, 64), 1, hp.VARIANT_TMA, 1), ((2 * world, 16, 32), 4, hp.VARIANT_TMA, 1),
            ((world, 16, 32), 3, hp.VARIANT_TMA, 1)):
        full = oracle.fill((nz + 2) * (ny + 2) * (nx + 2), 7, np.float32)
        want = oracle.jacobi7(full, nz, ny, nx, 1, 1, steps).reshape(nz + 2, ny + 2, nx + 2)
        b = hp.partition_slabs(nz, world)
        if int(np.diff(b).min()) < 1:
            # the static chunking idiom leaves trailing ranks without planes when nz < chunk*world;
            # the transport requires >= 1 plane per slab (HPCX_INVALID_ARG otherwise), so scale nz up
            nz = nz * world
            full = oracle.fill((nz + 2) * (ny + 2) * (nx + 2), 7, np.float32)
            want = oracle.jacobi7(full, nz, ny, nx, 1, 1, steps).reshape(nz + 2, ny + 2, nx + 2)
            b = hp.partition_slabs(nz, world)
        z0, z1 = int(b[rank]), int(b[rank + 1])
        slab = np.ascontiguousarray(full.reshape(nz + 2, ny + 2, nx + 2)[z0:z1 + 2]).ravel()
        ga = hp.DeviceGrid(ctx, (z1 - z0, ny, nx), np.float32, 1)
        gb = hp.DeviceGrid(ctx, (z1 - z0, ny, nx), np.float32, 1)
        ga.upload(slab)
        gb.upload(slab)
        plan = hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, variant, transport=transport)
        plan.sweep(steps)
        out = gb.download().reshape(z1 - z0 + 2, ny + 2, nx + 2)
        assert bits_equal(np.ascontiguousarray(out[1:-1, 1:-1, 1:-1]),
                          np.ascontiguousarray(want[z0 + 1:z1 + 1, 1:-1, 1:-1])), (rank, nz, "jacobi", transport)
        # a second call sweeps `in` (iterate steps-1) again: in -> out semantics, epoch carries on
        plan.sweep(2)
        want2 = oracle.jacobi7(full, nz, ny, nx, 1, 1, steps + 1).reshape(nz + 2, ny + 2, nx + 2)
        out = gb.download().reshape(z1 - z0 + 2, ny + 2, nx + 2)
        assert bits_equal(np.ascontiguousarray(out[1:-1, 1:-1, 1:-1]),
                          np.ascontiguousarray(want2[z0 + 1:z1 + 1, 1:-1, 1:-1])), (rank, nz, "jacobi+2", transport)
        if world > 1:
            # ghost planes of the newest iterate are the neighbours' boundary planes
            if rank > 0 and transport == 1:
                assert bits_equal(np.ascontiguousarray(out[0, 1:-1, 1:-1]),
                                  np.ascontiguousarray(want2[z0, 1:-1, 1:-1])), (rank, "ghost lo")
            if rank < world - 1 and transport == 1:
                assert bits_equal(np.ascontiguousarray(out[-1, 1:-1, 1:-1]),
                                  np.ascontiguousarray(want2[z1 + 1, 1:-1, 1:-1])), (rank, "ghost hi")
        plan.destroy()
        ga.destroy()
        gb.destroy()

    # ---- stencil, two sweeps per pass over slabs (transport 0, ghost width 2, two planes per side)
    # transport 0: two planes per side over NCCL; 1: the kernel stores them into the neighbours
    for (nz, ny, nx), steps, gw, transport in (
            ((8 * world, 20, 140), 6, 2, 0), ((3 * world - 1, 30, 260), 5, 2, 0),
            ((5 * world, 15, 128), 4, 3, 0), ((3 * world, 14, 40), 1, 2, 0),
            ((8 * world, 20, 140), 6, 2, 1), ((3 * world - 1, 30, 260), 5, 2, 1),
            ((5 * world, 15, 128), 4, 3, 1), ((2 * world, 33, 129), 7, 2, 1), ((40 * world, 64, 256), 8, 2, 1),
            # more than 128 planes per rank: the push-kernel transport (thinner slabs push from inside the pass kernel)
            ((150 * world, 16, 130), 6, 2, 1)):
        ext = (nz + 2 * gw, ny + 2 * gw, nx + 2 * gw)
        full = oracle.fill(ext[0] * ext[1] * ext[2], 11, np.float32)
        b = hp.partition_slabs(nz, world)
        assert int(np.diff(b).min()) >= 2
        z0, z1 = int(b[rank]), int(b[rank + 1])
        slab = np.ascontiguousarray(full.reshape(ext)[z0:z1 + 2 * gw]).ravel()
        ga = hp.DeviceGrid(ctx, (z1 - z0, ny, nx), np.float32, gw)
        gb = hp.DeviceGrid(ctx, (z1 - z0, ny, nx), np.float32, gw)
        ga.upload(slab)
        gb.upload(slab)
        plan = hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, hp.VARIANT_TEMPORAL, transport=transport)
        done = 0
        for k in (steps, 2, 3):  # in -> out semantics: each call sweeps `in` k times into `out`
            plan.sweep(k)
            total = done + k
            want = oracle.jacobi7(full, nz, ny, nx, gw, gw, total).reshape(ext)
            out = gb.download().reshape(z1 - z0 + 2 * gw, ny + 2 * gw, nx + 2 * gw)
            assert bits_equal(np.ascontiguousarray(out[gw:-gw, gw:-gw, gw:-gw]),
                              np.ascontiguousarray(want[z0 + gw:z1 + gw, gw:-gw, gw:-gw])), (rank, nz, "temporal", k, transport)
            done = total - (1 if k & 1 else 2)  # `in` holds the source of the call's last pass
        plan.destroy()
        ga.destroy()
        gb.destroy()
    if world > 1:
        # a slab with ghost width 1 cannot trade two planes: every rank must refuse, none may hang
        ga = hp.DeviceGrid(ctx, (4, 8, 32), np.float32, 1)
        gb = hp.DeviceGrid(ctx, (4, 8, 32), np.float32, 1)
        try:
            hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, hp.VARIANT_TEMPORAL, transport=0)
            raise AssertionError("ghost width 1 accepted for TEMPORAL over slabs")
        except hp.HpcxError:
            pass
        ga.destroy()
        gb.destroy()

    # ---- R0 across ranks: the reference's 17-point sweeps over y-slabs; BOX_Y trades 8 ghost rows
    for (ny, nx), kind, axis in (((40 * world + 7, 256), hp.STENCIL_BOX_Y, 0), ((40 * world + 7, 256), hp.STENCIL_BOX_X, 1),
                                 ((8 * world, 130), hp.STENCIL_BOX_Y, 0)):
        a_flat = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
        want = oracle.box_sweep(a_flat, ny, nx, 8, 8, 0, 0, axis, 8).reshape(ny, nx)
        if oracle.ref() is not None and nx % 64 == 0:
            assert bits_equal(want.ravel(), oracle.ref_blur("y_sse_ipl" if axis == 0 else "x_sse", a_flat, ny, nx))
        b = np.array([ny * r // world for r in range(world + 1)])
        y0, y1 = int(b[rank]), int(b[rank + 1])
        slab = np.ascontiguousarray(a_flat.reshape(ny + 16, nx + 16)[y0:y1 + 16]).copy()
        if world > 1:
            # the ghost rows shared with a neighbour must come through the exchange: poison them here
            if rank > 0:
                slab[:8] = np.nan
            if rank < world - 1:
                slab[-8:] = np.nan
        ga = hp.DeviceGrid(ctx, (y1 - y0, nx), np.float32, 8)
        gb = hp.DeviceGrid(ctx, (y1 - y0, nx), np.float32, 0)
        ga.upload(slab.ravel())
        plan = hd.HaloPlan(ctx, ga, gb, kind, transport=0, radius=8)
        plan.sweep(1)
        out = gb.download().reshape(y1 - y0, nx)
        if kind == hp.STENCIL_BOX_X and world > 1:
            pass  # BOX_X reads no ghost rows: nothing to check about them
        assert bits_equal(np.ascontiguousarray(out), np.ascontiguousarray(want[y0:y1])), (rank, "box", kind, ny)
        plan.destroy()
        ga.destroy()
        gb.destroy()

    # ---- R0 over a 2-D block split (py x px blocks, rank = ry * px + rx): BOX_X trades 8 ghost
    # columns (packed faces) inside a block row, BOX_Y 8 ghost rows between block rows
    shapes = [(py, world // py) for py in range(1, world + 1) if world % py == 0]
    for py, px in shapes:
        for (ny, nx), kind, axis in (((24 * py + 5, 64 * px), hp.STENCIL_BOX_X, 1), ((24 * py + 5, 64 * px), hp.STENCIL_BOX_Y, 0),
                                     ((9 * py, 36 * px + 2), hp.STENCIL_BOX_X, 1), ((9 * py, 36 * px + 2), hp.STENCIL_BOX_Y, 0)):
            a_flat = oracle.fill((ny + 16) * (nx + 16), 5, np.float32)
            want = oracle.box_sweep(a_flat, ny, nx, 8, 8, 0, 0, axis, 8).reshape(ny, nx)
            ry, rx = rank // px, rank % px
            by = [ny * r // py for r in range(py + 1)]
            bx = [nx * r // px for r in range(px + 1)]
            y0, y1, x0, x1 = by[ry], by[ry + 1], bx[rx], bx[rx + 1]
            block = np.ascontiguousarray(a_flat.reshape(ny + 16, nx + 16)[y0:y1 + 16, x0:x1 + 16]).copy()
            # ghosts shared with a neighbouring block must come through the exchange: poison them
            if ry > 0:
                block[:8] = np.nan
            if ry < py - 1:
                block[-8:] = np.nan
            if rx > 0:
                block[:, :8] = np.nan
            if rx < px - 1:
                block[:, -8:] = np.nan
            ga = hp.DeviceGrid(ctx, (y1 - y0, x1 - x0), np.float32, 8)
            gb = hp.DeviceGrid(ctx, (y1 - y0, x1 - x0), np.float32, 0)
            ga.upload(block.ravel())
            plan = hd.HaloPlan(ctx, ga, gb, kind, radius=8, blocks=(py, px))
            plan.sweep(1)
            out = gb.download().reshape(y1 - y0, x1 - x0)
            assert bits_equal(np.ascontiguousarray(out), np.ascontiguousarray(want[y0:y1, x0:x1])), \
                (rank, "box blocks", (py, px), kind, (ny, nx))
            plan.destroy()
            ga.destroy()
            gb.destroy()
    if world > 1:
        ga, gb = hp.DeviceGrid(ctx, (16, 64), np.float32, 8), hp.DeviceGrid(ctx, (16, 64), np.float32, 0)
        try:  # a block grid that does not cover the communicator is refused on every rank
            hd.HaloPlan(ctx, ga, gb, hp.STENCIL_BOX_X, radius=8, blocks=(world, 2))
            raise AssertionError("py * px != world accepted")
        except hp.HpcxError:
            pass
        ga.destroy()
        gb.destroy()

    # ---- SpMV: banded (C4 shape) by rows, power-law by nnz
    # op27(48): 6+ planes of 2304 rows per rank at world <= 8 — enough interior row blocks for the
    # pipelined host-operand path, with boundary planes far wider than one of its 32 x pieces
    for name, (rp, col, val), mode in (("op27", synth.op27(16, index_dtype=np.int64), 0),
                                       ("op27_48", synth.op27(48, index_dtype=np.int64), 0),
                                       ("rmat", synth.rmat(13, 16, 24, index_dtype=np.int64), 1)):
        n = len(rp) - 1
        x = synth.field(42, n, val.dtype)
        y_ref = oracle.spmv(rp, col, val, x)
        mag = oracle.spmv_abs(rp, col, val, x)
        A = hp.CsrMatrix(rp, col, val)
        bounds = hp.partition_rows(n, rp, world, mode)
        L = hp.csr_localize(A, bounds, rank)
        r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
        tol = 1e-12 if val.dtype == np.float64 else 1e-5
        for transport in (0, 1):  # NCCL send/recv, fused peer stores
            D = hd.distributed_csr(ctx, L, val.dtype, transport=transport)
            xd = torch.from_numpy(x[r0:r1].copy()).cuda()
            yd = torch.full((r1 - r0,), float("nan"), dtype=xd.dtype, device="cuda")
            torch.cuda.synchronize()
            for _ in range(3):
                D.spmv_dist(xd, yd)
            assert_spmv_close(yd.cpu().numpy(), y_ref[r0:r1], mag[r0:r1], tol)
            # a different x through the same handle: the halo slots must carry the NEW values
            x2 = synth.field(43, n, val.dtype)
            xd.copy_(torch.from_numpy(x2[r0:r1].copy()))
            torch.cuda.synchronize()
            D.spmv_dist(xd, yd)
            assert_spmv_close(yd.cpu().numpy(), oracle.spmv(rp, col, val, x2)[r0:r1],
                              oracle.spmv_abs(rp, col, val, x2)[r0:r1], tol)
            # host operands: this rank's x block in (pinned), its y block out — pipelined against PCIe
            # and the exchange with the fused transport, copy-product-copy otherwise; same bits as
            # the device-resident product
            # Every product before this one left its x in the handle's [owned | halo] buffer, so the
            # host-operand call gets an x the handle has never seen, after a product of NaNs has
            # poisoned both the owned entries and (through the exchange) the halo slots on every
            # rank: an entry read before its upload or its push has landed shows up as NaN.
            xd.fill_(float("nan"))
            torch.cuda.synchronize()
            D.spmv_dist(xd, yd)
            x3 = synth.field(44, n, val.dtype)
            xh = torch.from_numpy(x3[r0:r1].copy()).pin_memory()
            yh = torch.full((r1 - r0,), float("nan"), dtype=xh.dtype).pin_memory()
            before = ctx.launch_count()
            D.spmv_dist_host(xh.numpy(), yh.numpy())
            if name == "op27_48" and transport == 1 and world > 1:
                assert ctx.launch_count() - before >= 32, "the pipelined host-operand path was not taken"
            assert_spmv_close(yh.numpy(), oracle.spmv(rp, col, val, x3)[r0:r1],
                              oracle.spmv_abs(rp, col, val, x3)[r0:r1], tol)
            first = yh.numpy().copy()
            D.spmv_dist_host(xh.numpy(), yh.numpy())
            assert bits_equal(yh.numpy(), first), (rank, name, transport, "dist_host twice")
            # same bits as the device-resident product of the same x
            xd.copy_(xh.cuda())
            torch.cuda.synchronize()
            D.spmv_dist(xd, yd)
            assert bits_equal(yh.numpy(), yd.cpu().numpy()), (rank, name, transport, "dist_host")
            if transport == 1 and world > 1 and name == "op27":
                # rank skew: the last rank arrives 5 s late at the next product.  The flag waits of
                # the fused transport have no bound by default (like an NCCL receive), so the other
                # ranks wait and every block is still right — no stale halo is ever read.
                import time
                if rank == world - 1:
                    time.sleep(5.0)
                xd.copy_(torch.from_numpy(x[r0:r1].copy()))
                torch.cuda.synchronize()
                D.spmv_dist(xd, yd)
                assert_spmv_close(yd.cpu().numpy(), y_ref[r0:r1], mag[r0:r1], tol)
            D.destroy()

    # ---- row blocks with GLOBAL column ids and a gathered x (hpcx_spmv_allgather): the path for
    # matrices whose rows reference most of x; the block keeps the single-GPU kernel choice
    for name, (rp, col, val), mode, kernels in (
            ("op27", synth.op27(16, index_dtype=np.int64), 0, (hp.SPMV_AUTO,)),
            ("rmat", synth.rmat(13, 16, 24, index_dtype=np.int64), 1, (hp.SPMV_AUTO, hp.SPMV_HUB, hp.SPMV_STREAM))):
        n = len(rp) - 1
        bounds = hp.partition_rows(n, rp, world, mode)
        r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
        p0, p1 = int(rp[r0]), int(rp[r1])
        tol = 1e-12 if val.dtype == np.float64 else 1e-5
        for kernel in kernels:
            B = hp.CsrMatrix((rp[r0:r1 + 1] - p0).astype(np.int64), col[p0:p1].copy(), val[p0:p1].copy(), ncols=n)
            D = hp.DeviceCsr(ctx, B, kernel=kernel)
            xfull = torch.full((n,), float("nan"), dtype=torch.from_numpy(val[:1]).dtype, device="cuda")
            yd = torch.full((r1 - r0,), float("nan"), dtype=xfull.dtype, device="cuda")
            for seed in (42, 43):  # the second product must see the NEW x of every rank
                x = synth.field(seed, n, val.dtype)
                xd = torch.from_numpy(x[r0:r1].copy()).cuda()
                torch.cuda.synchronize()
                D.spmv_allgather(bounds, xd, xfull, yd)
                assert_spmv_close(yd.cpu().numpy(), oracle.spmv(rp, col, val, x)[r0:r1],
                                  oracle.spmv_abs(rp, col, val, x)[r0:r1], tol)
                assert bits_equal(xfull.cpu().numpy(), x), (rank, name, "gathered x")
            # x in equal slices (ncclAllGather) while the rows stay in their own blocks
            even = np.arange(world + 1, dtype=np.int64) * (n // world)
            assert even[-1] == n
            e0, e1 = int(even[rank]), int(even[rank + 1])
            x = synth.field(45, n, val.dtype)
            xfull.fill_(float("nan"))
            xd = torch.from_numpy(x[e0:e1].copy()).cuda()
            torch.cuda.synchronize()
            D.spmv_allgather(even, xd, xfull, yd)
            assert bits_equal(xfull.cpu().numpy(), x), (rank, name, "gathered x, equal slices")
            assert_spmv_close(yd.cpu().numpy(), oracle.spmv(rp, col, val, x)[r0:r1],
                              oracle.spmv_abs(rp, col, val, x)[r0:r1], tol)
            # owned slice already in place: x_local = x_full + bounds[rank]
            x = synth.field(44, n, val.dtype)
            xfull[r0:r1] = torch.from_numpy(x[r0:r1].copy()).cuda()
            torch.cuda.synchronize()
            D.spmv_allgather(bounds, xfull[r0:r1], xfull, yd)
            assert_spmv_close(yd.cpu().numpy(), oracle.spmv(rp, col, val, x)[r0:r1],
                              oracle.spmv_abs(rp, col, val, x)[r0:r1], tol)
            D.destroy()

    # ---- CG around the distributed product (SPD 27-point operator): iterates vs the global oracle
    rp, col, val = synth.op27(12, index_dtype=np.int64)
    n = len(rp) - 1
    b = synth.field(11, n)
    x0 = synth.field(12, n) * 0.1
    want, it, rr, bb = oracle.cg(rp, col, val, b, x0, 10)
    A = hp.CsrMatrix(rp, col, val)
    bounds = hp.partition_rows(n, rp, world, 0)
    L = hp.csr_localize(A, bounds, rank)
    r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
    for transport in (0, 1):
        D = hd.distributed_csr(ctx, L, val.dtype, transport=transport)
        S = hp.CgSolver(D)
        bd = torch.from_numpy(b[r0:r1].copy()).cuda()
        xd = torch.from_numpy(x0[r0:r1].copy()).cuda()
        torch.cuda.synchronize()
        res = S.solve(bd, xd, 10)
        assert res["iterations"] == 10, res
        assert np.abs(xd.cpu().numpy() - want[r0:r1]).max() <= 1e-10 * np.abs(want).max(), (rank, transport)
        assert abs(res["rr"] - rr) <= 1e-7 * rr and abs(res["bb"] - bb) <= 1e-10 * bb
        S.destroy()
        D.destroy()

    if world > 1:
        hd.barrier(ctx)
        dist.barrier()
    if rank == 0:
        print("DIST OK world=%d launches=%d" % (world, ctx.launch_count()), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
