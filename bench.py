#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload): configs[1] of BASELINE.json — CSR SpMV y = A*x, fp64, on the
2-D 5-point Laplacian of an 8192 x 8192 grid (67.1 M rows, 335.5 M nnz) per GPU.  At N > 1
the grid grows to (8192*N) x 8192 and is row-partitioned into N slabs (weak scaling); each
product exchanges one grid row of x (8192 doubles) with each neighbour through
hpcx_spmv_dist (fused transport: boundary x entries pushed by peer stores over NVLink; the
NCCL send/recv transport is timed beside it as an extra).

A step is one product.  `value` = effective GB/s = algorithmic bytes (12 B/nnz + 20 B/row,
SURVEY.md §8d) of all ranks / max-over-ranks device time, operands resident in HBM.
`e2e` is the same metric through the reference-facing host call (hpcx_spmv with HOST x and y at
N = 1, hpcx_spmv_dist_host per rank at N > 1: H2D of x and D2H of y pipelined against the product
inside the call and inside the timed region, matrix resident as in any iterative use), with the
time of the same copies alone (`pcie_copies_only_ms`) beside it.
The matrix (4.3 GB) is far larger than L2 (126 MB), so no explicit flush is needed.

--impl reference times the CPU restatement of the path (oracle/, OpenMP, every core of the
affinity mask whatever OMP_NUM_THREADS says) on the same GLOBAL matrix (N slabs at --gpus N): the
reference itself has no SpMV, so kind = "port".

The stencil half of the metric (Jacobi 512^3, R0 box sweeps) is in `extra` and, compactly, in the
LAST key of the line (`stencil`), each with its own cpu_baseline (the reference's own loops for R0,
kind "reference"; the Jacobi restatement, kind "port") and e2e (host array in, host array out).
At N > 1 the line carries `dist_parity`: before anything is timed, small distributed Jacobi
problems run through both transports and the temporal variant and every rank compares its slab
bit-for-bit with the single-GPU sweep of the same global field; the timed distributed SpMV is
checked against closed forms (A*1 and A*linear field) that a stale halo cannot satisfy.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRID = 8192  # configs[1]
METRIC = "spmv_effective_GBps"
UNIT = "GB/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0: float, t1: float):
        if not self.proc:
            return None
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for _, r in self.rows]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        if not sm:
            return None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i] == "Active" for r in rows)]
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "samples": len(sm)}


def build_local_problem(world: int, rank: int):
    """This rank's rows of the (GRID*world) x GRID Laplacian (global column indices), then the
    library's own partitioner: hpcx_partition_rows for the bounds and the distributed localiser
    (hpcx_csr_localize_rows + one exchange of column lists + hpcx_halo_send_lists) for the local
    numbering and the halo tables.  No rank ever holds the global matrix."""
    import libhpc_b200 as hp
    from libhpc_b200 import dist as hd
    from libhpc_b200 import synth
    t = time.time()
    nrows_global = GRID * GRID * world
    bounds = hp.partition_rows(nrows_global, None, world, 0)  # static chunking: one slab of grid rows per rank
    assert int(bounds[rank]) == rank * GRID * GRID
    rp, col, val = synth.laplace2d_rows(GRID * world, GRID, rank * GRID, (rank + 1) * GRID)
    t1 = time.time()
    L = hd.localize_rows_distributed(rp, col, val, bounds, rank, world, nrows_global)
    L["local_row_ptr"] = L["local_row_ptr"].astype(np.int32)
    L["partition_s"] = time.time() - t1
    log(f"[rank {rank}] built slab: {L['nloc']} rows, {L['nnz_loc']} nnz in {t1 - t:.1f}s; partitioner "
        f"{L['partition_s']:.2f}s ({L['nhalo']} halo columns, {L['nsend']} sends)")
    return L


def algorithmic_bytes(nrows: int, nnz: int, es: int = 8) -> int:
    return nnz * (es + 4) + nrows * (4 + 2 * es) + 4


def cpu_spmv_baseline(L: dict, x: np.ndarray, budget_s: float = 12.0, min_reps: int = 3, max_reps: int = 20):
    """The CPU restatement (oracle/hpcx_oracle.c: orc_spmv_i32_f64, OpenMP static rows, product
    and sum rounded separately) timed on this box's host cores on the SAME matrix.
    This is the one place bench.py executes oracle/: as the measured CPU baseline."""
    import oracle
    host_threads()
    rp, col, val = L["local_row_ptr"], L["local_col"], np.ascontiguousarray(L["values"])
    xfull = np.zeros(L["nloc"] + L["nhalo"])
    xfull[:L["nloc"]] = x
    y = np.empty(L["nloc"])
    oracle.spmv(rp, col, val, xfull, out=y)  # warm-up (first touch of y, page faults)
    times = []
    t_start = time.time()
    while len(times) < min_reps or (time.time() - t_start < budget_s and len(times) < max_reps):
        t = time.perf_counter()
        oracle.spmv(rp, col, val, xfull, out=y)
        times.append(time.perf_counter() - t)
    return times


def host_threads() -> int:
    """OpenMP team of the CPU legs: every core of the affinity mask.  Launchers export
    OMP_NUM_THREADS=1 (torch.distributed.run does), so the count is set explicitly."""
    import oracle
    return oracle.set_threads()


def mem_available_bytes() -> int:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 1 << 62


def run_reference(args, rank: int, world: int):
    """--impl reference: the CPU path on the host cores, rank 0 only, on the SAME global problem as
    our arm at this N (an (8192*N) x 8192 grid = N slabs), all host cores."""
    if rank != 0:
        return
    import oracle
    cores = host_threads()
    rows_slab, nnz_slab = GRID * GRID, 5 * GRID * GRID - 2 * GRID  # upper bound per slab
    need = world * (nnz_slab * 12 + rows_slab * 24)
    slabs = world
    avail = mem_available_bytes()
    if need * 1.2 > avail:  # bounded sample: as many slabs as the host memory holds
        slabs = max(1, int(world * avail / (need * 1.2)))
    t = time.time()
    rp, col, val = oracle.laplace2d(GRID * slabs, GRID)  # int64 row_ptr: nnz exceeds 2^31 from 7 slabs on
    n, nnz = len(rp) - 1, len(col)
    x = oracle.fill(n, 42)
    y = np.empty(n)
    log(f"[reference] built {slabs}-slab matrix ({n} rows, {nnz} nnz) in {time.time() - t:.1f}s, {cores} threads")
    for _ in range(max(args.warmup, 1)):
        oracle.spmv(rp, col, val, x, out=y)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.spmv(rp, col, val, x, out=y)
    dt = time.perf_counter() - t0
    nbytes = algorithmic_bytes(n, nnz)
    gbs = nbytes * args.steps / dt / 1e9
    sample = (f"the full {world}-slab global matrix of our arm's config ({n} rows, {nnz} nnz), {args.steps} products"
              if slabs == world else
              f"{slabs} of {world} slabs ({n} rows, {nnz} nnz: what host memory holds), {args.steps} products")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(world), "same_config": slabs == world,
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "note": "the reference has no SpMV (SURVEY.md §0); this is the OpenMP CPU "
                                 "restatement in oracle/hpcx_oracle.c"},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "checksum": float(np.abs(y).sum()),
    }), flush=True)


def workload_config(world: int) -> dict:
    return {"workload": "BASELINE configs[1]: CSR SpMV fp64, 2-D 5-point Laplacian 8192x8192 per GPU "
                        "(67,108,864 rows, 335,511,552 nnz at N=1)",
            "grid": [GRID * world, GRID], "rows_per_gpu": GRID * GRID,
            "partition": "1 slab" if world == 1 else f"{world} row slabs, x-halo 8192 doubles per neighbour, "
                         "pushed by peer stores over NVLink (hpcx_csr_halo_transport 1)",
            "l2": "inputs (4.3 GB matrix per GPU) larger than L2; no explicit flush",
            "algorithmic_bytes": "12 B/nnz + 20 B/row (SURVEY.md §8d)"}


def gpu_warm(torch, stream, step, min_ms: float = 30.0, max_steps: int = 2000):
    """Runs `step` until about min_ms of GPU time have passed: a timed region that follows a CPU leg
    (cpu_baseline, host-side set-up) would otherwise start at idle clocks — a 1.7 ms region right
    after a one-second CPU baseline measured the box-y sweep at half its rate."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    step()
    e1.record(stream)
    e1.synchronize()
    one = max(e0.elapsed_time(e1), 1e-3)
    for _ in range(min(max_steps, int(min_ms / one))):
        step()


def time_steps(torch, stream, step, steps: int, barrier):
    """K steps between two events on `stream`, bracketed by barrier + synchronize."""
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    w1 = time.time()
    barrier()
    return e0.elapsed_time(e1) * 1e-3, w0, w1


def dist_parity_check(ctx, torch, dist, world, rank, hp, hd, synth):
    """N > 1, before anything is timed: a (16*N) x 64 x 256 Jacobi problem through the NCCL
    transport, the fused peer-store transport and the two-sweeps-per-pass variant over both; every
    rank compares its slab BIT FOR BIT with the single-GPU hpcx_stencil_sweep of the same global
    field on its own GPU (product vs product: the single-GPU kernel is what the oracle-checked
    GPU tests cover).  Any mismatch on any rank ends the run non-zero."""
    nz, ny, nx = 16 * world, 64, 256
    cases = ((1, hp.VARIANT_TMA, 0, 5, "one_sweep_nccl"), (1, hp.VARIANT_TMA, 1, 6, "one_sweep_peer"),
             (2, hp.VARIANT_TEMPORAL, 0, 6, "temporal_nccl"), (2, hp.VARIANT_TEMPORAL, 1, 5, "temporal_peer"))
    verdict = {}
    for gw, variant, transport, steps, name in cases:
        ext = (nz + 2 * gw, ny + 2 * gw, nx + 2 * gw)
        full = synth.field(7, ext[0] * ext[1] * ext[2], np.float32)
        ga, gb = hp.DeviceGrid(ctx, (nz, ny, nx), np.float32, gw), hp.DeviceGrid(ctx, (nz, ny, nx), np.float32, gw)
        ga.upload(full)
        gb.upload(full)
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=steps, variant=hp.VARIANT_TMA)
        want = gb.download().reshape(ext)
        ga.destroy()
        gb.destroy()
        b = hp.partition_slabs(nz, world)
        z0, z1 = int(b[rank]), int(b[rank + 1])
        slab = np.ascontiguousarray(full.reshape(ext)[z0:z1 + 2 * gw]).ravel()
        sa, sb = (hp.DeviceGrid(ctx, (z1 - z0, ny, nx), np.float32, gw) for _ in range(2))
        sa.upload(slab)
        sb.upload(slab)
        plan = hd.HaloPlan(ctx, sa, sb, hp.STENCIL_JACOBI7, variant, transport=transport)
        plan.sweep(steps)
        out = sb.download().reshape(z1 - z0 + 2 * gw, ext[1], ext[2])
        same = np.array_equal(np.ascontiguousarray(out[gw:-gw, gw:-gw, gw:-gw]).view(np.uint32),
                              np.ascontiguousarray(want[z0 + gw:z1 + gw, gw:-gw, gw:-gw]).view(np.uint32))
        plan.destroy()
        sa.destroy()
        sb.destroy()
        t = torch.tensor([1.0 if same else 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        verdict[name] = "ok" if float(t.item()) == 1.0 else "MISMATCH"
    # R0 box sweeps over y-slabs (ghost rows) and x-slabs (ghost columns, packed faces): every rank's
    # block against the single-GPU sweep of the global grid
    for kind, blocks, name in ((hp.STENCIL_BOX_Y, (world, 1), "box_y_yslabs"), (hp.STENCIL_BOX_X, (1, world), "box_x_xslabs")):
        py, px = blocks
        ny, nx = 40 * py, 64 * px
        full = synth.field(5, (ny + 16) * (nx + 16), np.float32)
        ga, gb = hp.DeviceGrid(ctx, (ny, nx), np.float32, 8), hp.DeviceGrid(ctx, (ny, nx), np.float32, 0)
        ga.upload(full)
        hp.stencil_sweep(ga, gb, kind, radius=8)
        want = gb.download().reshape(ny, nx)
        ga.destroy()
        gb.destroy()
        ry, rx = rank // px, rank % px
        y0, y1, x0, x1 = ny * ry // py, ny * (ry + 1) // py, nx * rx // px, nx * (rx + 1) // px
        block = np.ascontiguousarray(full.reshape(ny + 16, nx + 16)[y0:y1 + 16, x0:x1 + 16]).copy()
        if ry > 0:
            block[:8] = np.nan  # must come through the exchange
        if ry < py - 1:
            block[-8:] = np.nan
        if rx > 0:
            block[:, :8] = np.nan
        if rx < px - 1:
            block[:, -8:] = np.nan
        sa, sb = hp.DeviceGrid(ctx, (y1 - y0, x1 - x0), np.float32, 8), hp.DeviceGrid(ctx, (y1 - y0, x1 - x0), np.float32, 0)
        sa.upload(block.ravel())
        plan = hd.HaloPlan(ctx, sa, sb, kind, radius=8, blocks=blocks)
        plan.sweep(1)
        out = sb.download().reshape(y1 - y0, x1 - x0)
        same = np.array_equal(np.ascontiguousarray(out).view(np.uint32), np.ascontiguousarray(want[y0:y1, x0:x1]).view(np.uint32))
        plan.destroy()
        sa.destroy()
        sb.destroy()
        t = torch.tensor([1.0 if same else 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        verdict[name] = "ok" if float(t.item()) == 1.0 else "MISMATCH"
    if any(v != "ok" for v in verdict.values()):
        raise SystemExit(f"[rank {rank}] bench.py: a distributed sweep differs from the single-GPU sweep: {verdict}")
    return verdict


def laplacian_of_linear(torch, rank: int, world: int):
    """x = f and the exact A*f of the (8192*world) x 8192 5-point Laplacian on this rank's slab for
    f(Y, X) = 3Y + 5X + 1 (small integers: every product and partial sum is exact in fp64, so the
    result is order-independent).  Zero in the interior, the missing neighbours' values on the
    domain boundary — a stale or misplaced halo entry cannot reproduce it."""
    Y = torch.arange(rank * GRID, (rank + 1) * GRID, dtype=torch.float64, device="cuda")[:, None]
    X = torch.arange(GRID, dtype=torch.float64, device="cuda")[None, :]
    f = 3.0 * Y + 5.0 * X + 1.0
    want = torch.zeros_like(f)
    want[:, 0] += f[:, 0] - 5.0   # missing left neighbour f(Y, -1) would have cancelled: add it back
    want[:, -1] += f[:, -1] + 5.0
    if rank == 0:
        want[0] += f[0] - 3.0
    if rank == world - 1:
        want[-1] += f[-1] + 3.0
    return f.reshape(-1), want.reshape(-1)


def run_ours(args, rank: int, world: int, local: int):
    import torch
    import torch.distributed as dist

    import libhpc_b200 as hp
    from libhpc_b200 import dist as hd
    from libhpc_b200 import synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libhpcx has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = hp.Context(local)
    if world > 1:
        hd.init_comm(ctx, rank, world)

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    dist_parity = None
    if world > 1:
        dist_parity = dist_parity_check(ctx, torch, dist, world, rank, hp, hd, synth)
        log(f"[rank {rank}] dist_parity {dist_parity}")

    t = time.time()
    L = build_local_problem(world, rank)
    build_s = time.time() - t
    nloc, nnz = L["nloc"], L["nnz_loc"]
    A = hp.CsrMatrix(L["local_row_ptr"], L["local_col"], L["values"], ncols=nloc + L["nhalo"])
    t = time.time()
    D = hp.DeviceCsr(ctx, A)
    create_s = time.time() - t
    if world > 1:
        D.set_halo(L)
        D.halo_transport(1)  # fused gather + peer stores + flags (NCCL variant timed below as extra)
    halo_s = time.time() - t - create_s
    log(f"[rank {rank}] hpcx_csr_create {create_s:.2f}s, halo tables {halo_s:.2f}s")
    setup_ms = {"hpcx_csr_create": max_over_ranks(create_s) * 1e3, "halo_tables": max_over_ranks(halo_s) * 1e3,
                "partitioner_localize_rows": max_over_ranks(L["partition_s"]) * 1e3,
                "host_matrix_build_python": max_over_ranks(build_s - L["partition_s"]) * 1e3}

    stream = torch.cuda.Stream()
    x_pin = torch.empty(nloc, dtype=torch.float64).pin_memory()
    y_pin = torch.empty(nloc, dtype=torch.float64).pin_memory()
    x_pin.numpy()[:] = synth.field(42, nloc, first=rank * nloc)
    yd = torch.empty(nloc, dtype=torch.float64, device="cuda")
    if world == 1:
        xd = x_pin.cuda()
        x_arg = xd
        step = lambda: D.spmv(x_arg, yd, stream)  # noqa: E731
    else:
        xd = torch.empty(nloc, dtype=torch.float64, device="cuda")
        xd.copy_(x_pin)
        xfull = D.x_buffer()  # [owned | halo]: write owned x in place, no per-step device copy
        ctx.memcpy((xfull, hp.MEM_DEVICE), xd, nloc * 8)
        x_arg = (xfull, hp.MEM_DEVICE)
        step = lambda: D.spmv_dist(x_arg, yd, stream)  # noqa: E731
    torch.cuda.synchronize()

    # ---- device-resident timing
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
        if world == 1:
            gpu_warm(torch, stream, step)  # on top of the W warm-up steps: the matrix build left the GPU idle
        sampler = ClockSampler(local) if rank == 0 else None
        n0 = ctx.launch_count()
        dt, w0, w1 = time_steps(torch, stream, step, args.steps, barrier)
        launches = ctx.launch_count() - n0
        clocks = sampler.stop(w0, w1) if sampler else None
    dt = max_over_ranks(dt)
    nbytes_rank = algorithmic_bytes(nloc, nnz)
    value = nbytes_rank * world * args.steps / dt / 1e9

    # ---- property check of the timed operator (no oracle here): A*1 = missing-neighbour count
    ones = torch.ones(nloc + L["nhalo"], dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()  # `ones` was filled on torch's default stream, the product runs on `stream`
    if world == 1:
        D.spmv(ones, yd, stream)
    else:
        ctx.memcpy((xfull, hp.MEM_DEVICE), ones, nloc * 8)
        D.spmv_dist(x_arg, yd, stream)
    stream.synchronize()
    yv = yd.view(GRID, GRID)
    expect = torch.zeros(GRID, GRID, dtype=torch.float64, device="cuda")
    expect[:, 0] += 1
    expect[:, -1] += 1
    if rank == 0:
        expect[0] += 1
    if rank == world - 1:
        expect[-1] += 1
    if not torch.equal(yv, expect):
        raise SystemExit(f"[rank {rank}] bench.py: A*1 does not match the closed form")
    del ones, expect
    # ... and A*f for a linear field f: the halo entries must be the neighbours' CURRENT values
    f, want = laplacian_of_linear(torch, rank, world)
    torch.cuda.synchronize()
    if world == 1:
        D.spmv(f, yd, stream)
    else:
        ctx.memcpy((xfull, hp.MEM_DEVICE), f, nloc * 8)
        barrier()
        D.spmv_dist(x_arg, yd, stream)
    stream.synchronize()
    if not torch.equal(yd, want):
        raise SystemExit(f"[rank {rank}] bench.py: A*(linear field) does not match the closed form")
    del f, want
    spmv_check = "A*1 and A*(3Y+5X+1) equal their closed forms on every rank (exact in fp64)"
    if world > 1:
        barrier()
        ctx.memcpy((xfull, hp.MEM_DEVICE), xd, nloc * 8)

    # ---- end to end: host x in, host y out, every step
    def e2e_step():
        if world == 1:
            D.spmv(x_pin.numpy(), y_pin.numpy(), stream)  # hpcx_spmv(HOST, HOST): H2D + kernel + D2H + sync
        else:
            # hpcx_spmv_dist_host: pinned x in, pinned y out; copies, exchange and product overlap
            D.spmv_dist_host(x_pin.numpy(), y_pin.numpy(), stream)

    with torch.cuda.stream(stream):
        for _ in range(min(args.warmup, 3)):
            e2e_step()
        k_e2e = max(3, min(args.steps, 10))
        dt_e2e, _, _ = time_steps(torch, stream, e2e_step, k_e2e, barrier)
    dt_e2e = max_over_ranks(dt_e2e)
    e2e_value = nbytes_rank * world * k_e2e / dt_e2e / 1e9
    checksum = float(np.abs(y_pin.numpy()).sum())
    if world > 1:
        # the distributed e2e result must be the device-resident one — checked on a product issued
        # after NaNs have gone through the handle (owned entries and, by the exchange, every halo
        # slot), so that an entry read before its upload or push had landed could not go unnoticed
        y_dev = torch.empty_like(yd)
        nan = torch.full_like(xd, float("nan"))
        ctx.memcpy((xfull, hp.MEM_DEVICE), nan, nloc * 8)
        barrier()
        D.spmv_dist(x_arg, y_dev, stream)
        stream.synchronize()
        barrier()
        del nan
        D.spmv_dist_host(x_pin.numpy(), y_pin.numpy(), stream)
        stream.synchronize()
        barrier()
        ctx.memcpy((xfull, hp.MEM_DEVICE), xd, nloc * 8)
        barrier()
        D.spmv_dist(x_arg, y_dev, stream)
        stream.synchronize()
        if not torch.equal(y_dev.cpu(), y_pin):
            raise SystemExit(f"[rank {rank}] bench.py: hpcx_spmv_dist_host differs from the device-resident product")
        barrier()
    # the bound of that leg: the same bytes over PCIe in both directions at once with every rank
    # copying, and nothing else (host memory / root-complex bandwidth shared by the ranks)
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    y_scratch = torch.empty_like(y_pin).pin_memory()

    def copies_only():
        with torch.cuda.stream(s_in):
            xd.copy_(x_pin, non_blocking=True)
        with torch.cuda.stream(s_out):
            y_scratch.copy_(yd, non_blocking=True)
        s_in.synchronize()
        s_out.synchronize()
    copies_only()
    barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        copies_only()
    torch.cuda.synchronize()
    pcie_ms = max_over_ranks((time.perf_counter() - t0) / 5) * 1e3

    extra, stencil = stencil_extra(args, ctx, torch, stream, world, rank, barrier, max_over_ranks, hp, hd, synth)
    extra["cg_iteration_c2_fp64"] = cg_extra(args, ctx, torch, stream, world, barrier, max_over_ranks, hp, D,
                                             nloc, nnz, xd)
    if args.configs == "all":
        if world == 1:
            extra["c1_laplace_1024sq_fp64"] = c1_extra(args, ctx, torch, stream, barrier, hp, synth)
        extra.update(sparse_extra(args, ctx, torch, stream, world, rank, barrier, max_over_ranks, hp, hd, synth))
    if world > 1:
        # the same product with the NCCL send/recv exchange, for comparison
        D.halo_transport(0)
        # checked before it is timed: A*(3Y+5X+1) through gather_boundary_x_kernel -> grouped
        # ncclSend/ncclRecv -> boundary rows must equal the closed form exactly on every rank
        f, want = laplacian_of_linear(torch, rank, world)
        torch.cuda.synchronize()
        ctx.memcpy((xfull, hp.MEM_DEVICE), f, nloc * 8)
        barrier()
        yd.fill_(float("nan"))
        torch.cuda.synchronize()  # the fill ran on torch's stream, the product runs on `stream`
        D.spmv_dist(x_arg, yd, stream)
        stream.synchronize()
        if not torch.equal(yd, want):
            raise SystemExit(f"[rank {rank}] bench.py: NCCL-transport A*(linear field) does not match the closed form")
        del f, want
        barrier()
        ctx.memcpy((xfull, hp.MEM_DEVICE), xd, nloc * 8)
        barrier()
        if isinstance(dist_parity, dict):
            dist_parity["spmv_nccl_transport_closed_form"] = "ok"
            dist_parity["spmv_peer_transport_closed_forms"] = "ok"
            dist_parity["spmv_dist_host_after_nan_poison"] = "ok"
        with torch.cuda.stream(stream):
            for _ in range(3):
                step()
            dt_n, _, _ = time_steps(torch, stream, step, args.steps, barrier)
        dt_n = max_over_ranks(dt_n)
        extra["spmv_nccl_transport"] = {"metric": METRIC, "value": nbytes_rank * world * args.steps / dt_n / 1e9,
                                        "ms_per_step": dt_n / args.steps * 1e3}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        times = cpu_spmv_baseline(L, x_pin.numpy())
        best = min(times)
        cpu = {"value": nbytes_rank / best / 1e9, "unit": UNIT, "cores": host_threads(), "kind": "port",
               "sample": f"full configs[1] matrix, best of {len(times)} products (mean "
                         f"{nbytes_rank / np.mean(times) / 1e9:.1f} GB/s) after 1 warm-up",
               "note": "OpenMP CPU restatement oracle/hpcx_oracle.c:orc_spmv_i32_f64; the reference has no SpMV"}

    if rank == 0:
        peak, how = measured_peak()
        per_launch_s = dt / args.steps
        achieved = nbytes_rank / per_launch_s / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("spmv_stream_kernel_c2_dram_bytes_per_launch")
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(world),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": how,
                         "kernel": "spmv_stream_kernel<double,2048>",
                         "frac_of_nominal_8000": achieved / 8000.0,
                         "algorithmic_bytes_per_launch": nbytes_rank,
                         "note": "per-rank step time / per-rank bytes" + (
                             "" if world == 1 else " (step = push + interior + flag wait + boundary + ack kernels)")},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nloc * 8 * world,
                    "d2h_bytes_per_step": nloc * 8 * world, "steps": k_e2e,
                    "ms_per_step": dt_e2e / k_e2e * 1e3,
                    "pcie_copies_only_ms": pcie_ms,
                    "note": "pcie_copies_only_ms = the same H2D + D2H with every rank copying at once and no product: "
                            "the bound of this leg",
                    "call": "hpcx_spmv(A, x HOST, y HOST)" if world == 1 else
                            "hpcx_spmv_dist_host(A, x HOST, y HOST) per rank"},
            "gpu_launches": launches, "clocks": clocks, "checksum": checksum, "setup_ms": setup_ms,
            "spmv_check": spmv_check,
            "extra": extra,
            # last keys on purpose: the stencil half of BASELINE.json's metric and the list of
            # distributed checks (each one ends the run non-zero on a mismatch) survive a truncated tail
            "stencil": stencil,
            "dist_parity": dist_parity if world > 1 else "n/a (one GPU)",
        }), flush=True)
    D.destroy()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def stencil_extra(args, ctx, torch, stream, world, rank, barrier, max_over_ranks, hp, hd, synth):
    """Second half of BASELINE.json's metric: stencil Gcell-updates/s.
    configs[2]: 3-D 7-point Jacobi, 512^3 fp32 per GPU (z-slabs of a 512*N x 512 x 512 domain at
    N > 1, ghost planes exchanged every sweep), 8 B/cell algorithmic; and at N = 1 the
    reference-native 17-point box sweeps on 8192^2 fp32 (R0), each with a cpu_baseline (the
    reference's own loops for R0, the restatement for Jacobi) and an e2e leg (host array in, host
    array out through hpcx_grid_upload / hpcx_stencil_sweep / hpcx_grid_download).
    Returns (detailed entries for `extra`, compact summary for the line's last key)."""
    out, summary = {}, {"unit": "Gcell-updates/s", "n_gpus": world}
    n, T = 512, 50
    peak, _ = measured_peak()

    def entry(cells_all_ranks, dt, launches=None, **kw):
        per_rank_gbs = cells_all_ranks * 8 / world / dt / 1e9
        e = {"metric": "stencil_Gcell_updates_per_s", "value": cells_all_ranks / dt / 1e9,
             "roofline": {"bound": "hbm", "achieved": per_rank_gbs, "peak": peak, "unit": "GB/s",
                          "frac": per_rank_gbs / peak, "algorithmic_bytes_per_cell": 8}}
        if launches is not None:
            e["gpu_launches"] = launches
        e.update(kw)
        return e

    def timed(step, reps=2):
        with torch.cuda.stream(stream):
            step()
            step()  # twice: an even number of sweeps swaps the two buffers, and the NCCL transport keeps
            #         one captured graph per orientation
            if world == 1:
                gpu_warm(torch, stream, step)  # (N > 1: every rank must make the same number of calls)
            n0 = ctx.launch_count()
            dt, _, _ = time_steps(torch, stream, step, reps, barrier)
            launches = ctx.launch_count() - n0
        return max_over_ranks(dt), launches

    f = synth.field(7, (n + 2) ** 3, np.float32, first=rank * (n + 2) ** 3)
    ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, 1), hp.DeviceGrid(ctx, (n, n, n), np.float32, 1)
    ga.upload(f)
    gb.upload(f)
    # N > 1: transport 1 = fused peer stores (one kernel per sweep), transport 0 = NCCL send/recv
    plans = {"tma": hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, hp.VARIANT_TMA, transport=1),
             "tma_nccl": hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, hp.VARIANT_TMA, transport=0)} if world > 1 else {}
    note2 = ("two sweeps per pass over HBM (HPCX_VARIANT_TEMPORAL): the 8 B/cell/sweep figure is kept as the "
             "algorithmic unit, the kernel moves 4 B/cell/sweep, so frac may exceed 1")
    for variant, name in ((hp.VARIANT_TMA, "tma"), (hp.VARIANT_TMA, "tma_nccl"), (hp.VARIANT_DIRECT, "direct"),
                          (hp.VARIANT_TEMPORAL, "temporal")):
        if (world > 1) != (name in plans) or (world == 1 and name == "tma_nccl"):
            continue
        if world > 1:
            plan = plans[name]
            step = lambda: plan.sweep(T, stream)  # noqa: E731
        else:
            step = lambda: hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=T, variant=variant,  # noqa: E731
                                            stream=stream)
        dt, launches = timed(step)
        cells = n ** 3 * world * T * 2
        e = entry(cells, dt, launches, sweeps=2 * T, ms_per_sweep=dt / (2 * T) * 1e3,
                  transport={"tma": "peer stores + flags (fused)", "tma_nccl": "NCCL send/recv"}.get(name) if world > 1 else None)
        if name == "temporal":
            e["roofline"]["note"] = note2
        out[f"jacobi7_512cubed_fp32_{name}"] = e
        summary["jacobi7_512c_" + {"tma": "1sweep" if world == 1 else "1sweep_peer", "tma_nccl": "1sweep_nccl",
                                   "direct": "direct", "temporal": "2sweep"}[name]] = round(e["value"], 1)
        if name == "tma":
            summary["jacobi7_frac_of_hbm"] = round(e["roofline"]["frac"], 4)
    for plan in plans.values():
        plan.destroy()

    if world == 1 and rank == 0:
        # ---- e2e: host array in, host array out, every step (one sweep per step: the worst case for
        # the copies; an iterative caller amortises them over its T sweeps)
        host_in = ctx.pinned(((n + 2) ** 3,), np.float32)
        host_out = ctx.pinned(((n + 2) ** 3,), np.float32)
        host_in[:] = f

        def j_e2e():
            ga.upload(host_in, stream)
            hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=1, variant=hp.VARIANT_TMA, stream=stream)
            gb.download(host_out, stream)
            stream.synchronize()
        dt, _ = timed(j_e2e, 5)
        out["jacobi7_512cubed_fp32_tma"]["e2e"] = {
            "value": n ** 3 * 5 / dt / 1e9, "unit": "Gcell-updates/s", "ms_per_step": dt / 5 * 1e3,
            "h2d_bytes_per_step": host_in.nbytes, "d2h_bytes_per_step": host_out.nbytes,
            "call": "hpcx_grid_upload(pinned) + hpcx_stencil_sweep(1 sweep) + hpcx_grid_download(pinned)"}
        summary["jacobi7_e2e_1sweep_per_copy"] = round(n ** 3 * 5 / dt / 1e9, 2)
        if not args.no_cpu:
            # ---- cpu_baseline (kind "port": the reference has no 3-D stencil): orc_jacobi7_f32, all cores
            import oracle
            cores = host_threads()
            src = np.array(host_in)
            dst = src.copy()
            lib = oracle.lib()
            import ctypes as C
            call = lambda: lib.orc_jacobi7_f32(src.ctypes.data_as(C.c_void_p), dst.ctypes.data_as(C.c_void_p),  # noqa: E731
                                               C.c_int64(n), C.c_int64(n), C.c_int64(n), C.c_int64(1), C.c_int64(1))
            call()
            ts = []
            while len(ts) < 3 or (sum(ts) < 6.0 and len(ts) < 12):
                t0 = time.perf_counter()
                call()
                ts.append(time.perf_counter() - t0)
            out["jacobi7_512cubed_fp32_tma"]["cpu_baseline"] = {
                "value": n ** 3 / min(ts) / 1e9, "unit": "Gcell-updates/s", "cores": cores, "kind": "port",
                "sample": f"full 512^3 sweep, best of {len(ts)} after 1 warm-up (oracle/hpcx_oracle.c:orc_jacobi7_f32)"}
            summary["jacobi7_cpu"] = {"value": round(n ** 3 / min(ts) / 1e9, 3), "cores": cores, "kind": "port"}
            del src, dst
    del f
    ga.destroy()
    gb.destroy()
    if world > 1:
        # two sweeps per pass over slabs: ghost width 2, two planes per side and pass
        f = synth.field(7, (n + 4) ** 3, np.float32, first=rank * (n + 4) ** 3)
        ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, 2), hp.DeviceGrid(ctx, (n, n, n), np.float32, 2)
        ga.upload(f)
        gb.upload(f)
        del f
        for transport, name in ((1, "temporal"), (0, "temporal_nccl")):
            plan = hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, hp.VARIANT_TEMPORAL, transport=transport)
            dt, launches = timed(lambda: plan.sweep(T, stream))
            cells = n ** 3 * world * T * 2
            e = entry(cells, dt, launches, sweeps=2 * T, ms_per_sweep=dt / (2 * T) * 1e3,
                      transport=("peer stores + flags (fused), two planes per side and two-sweep pass" if transport
                                 else "NCCL send/recv, two planes per side and two-sweep pass"))
            e["roofline"]["note"] = note2
            out[f"jacobi7_512cubed_fp32_{name}"] = e
            summary["jacobi7_512c_2sweep_" + ("peer" if transport else "nccl")] = round(e["value"], 1)
            plan.destroy()
        ga.destroy()
        gb.destroy()
        # strong scaling: ONE 512^3 domain cut into world z-slabs of 512/world planes
        nzl = n // world
        f = synth.field(7, (nzl + 4) * (n + 4) ** 2, np.float32, first=rank * nzl * (n + 4) ** 2)
        ga = hp.DeviceGrid(ctx, (nzl, n, n), np.float32, 2)
        gb = hp.DeviceGrid(ctx, (nzl, n, n), np.float32, 2)
        ga.upload(f)
        gb.upload(f)
        del f
        for variant, name in ((hp.VARIANT_TMA, "tma"), (hp.VARIANT_TEMPORAL, "temporal")):
            plan = hd.HaloPlan(ctx, ga, gb, hp.STENCIL_JACOBI7, variant, transport=1)
            dt, _ = timed(lambda: plan.sweep(T, stream))
            cells = nzl * n * n * world * T * 2
            out[f"jacobi7_512cubed_total_fp32_{name}_strong"] = {
                "metric": "stencil_Gcell_updates_per_s", "value": cells / dt / 1e9, "scaling": "strong",
                "planes_per_gpu": nzl, "ms_per_sweep": dt / (2 * T) * 1e3,
                "transport": "peer stores + flags"}
            summary["jacobi7_512c_total_strong_" + ("1sweep" if name == "tma" else "2sweep")] = round(cells / dt / 1e9, 1)
            plan.destroy()
        ga.destroy()
        gb.destroy()
    if world > 1:
        # R0 across GPUs, weak: every rank owns an 8192 x 8192 block of the reference's blur grid.
        # y-slabs (world x 1 blocks): the y sweep trades 8 ghost rows; x-slabs (1 x world blocks): the
        # x sweep trades 8 ghost columns as packed faces.  The other sweep of each split needs no
        # exchange and runs at the one-GPU rate.
        m = 8192
        a = hp.DeviceGrid(ctx, (m, m), np.float32, 8)
        b = hp.DeviceGrid(ctx, (m, m), np.float32, 0)
        a.upload(synth.field(3, (m + 16) ** 2, np.float32, first=rank))
        for kind, name, blocks, what in ((hp.STENCIL_BOX_Y, "box_y_yslabs", (world, 1), "8 ghost rows per side (NCCL, in place)"),
                                         (hp.STENCIL_BOX_X, "box_x_xslabs", (1, world), "8 ghost columns per side (packed faces, NCCL)")):
            plan = hd.HaloPlan(ctx, a, b, kind, radius=8, blocks=blocks)
            dt, launches = timed(lambda: [plan.sweep(1, stream) for _ in range(20)])
            out[f"r0_{name}_8192sq_per_gpu_fp32"] = entry(m * m * world * 40, dt, launches, ms_per_sweep=dt / 40 * 1e3,
                                                          exchange=what, blocks=list(blocks))
            summary[f"r0_{name}"] = round(m * m * world * 40 / dt / 1e9, 1)
            plan.destroy()
        a.destroy()
        b.destroy()
    if world == 1:
        m = 8192
        a = hp.DeviceGrid(ctx, (m, m), np.float32, 8)
        b = hp.DeviceGrid(ctx, (m, m), np.float32, 0)
        fa = synth.field(3, (m + 16) ** 2, np.float32)
        a.upload(fa)
        host_a = ctx.pinned(((m + 16) ** 2,), np.float32)
        host_b = ctx.pinned((m * m,), np.float32)
        host_a[:] = fa
        for kind, name, ref_which in ((hp.STENCIL_BOX_X, "box_x", "x_sse"), (hp.STENCIL_BOX_Y, "box_y", "y_sse_ipl")):
            dt, _ = timed(lambda: hp.stencil_sweep(a, b, kind, radius=8, stream=stream), 20)
            e = entry(m * m * 20, dt, None, ms_per_sweep=dt / 20 * 1e3)

            def box_e2e():
                a.upload(host_a, stream)
                hp.stencil_sweep(a, b, kind, radius=8, stream=stream)
                b.download(host_b, stream)
                stream.synchronize()
            dte, _ = timed(box_e2e, 5)
            e["e2e"] = {"value": m * m * 5 / dte / 1e9, "unit": "Gcell-updates/s", "ms_per_step": dte / 5 * 1e3,
                        "h2d_bytes_per_step": host_a.nbytes, "d2h_bytes_per_step": host_b.nbytes,
                        "call": "hpcx_grid_upload(pinned) + hpcx_stencil_sweep + hpcx_grid_download(pinned): "
                                "what hpc::stencil_sweep(a, b, Box1D) does on host containers"}
            summary[f"r0_{name}"] = round(e["value"], 1)
            summary[f"r0_{name}_e2e"] = round(e["e2e"]["value"], 2)
            if not args.no_cpu:
                # cpu_baseline, kind "reference": the reference's OWN loop (BM_x_blur_tiling_simd_prefetch /
                # BM_YXx_blur_tiling_prefetch_streamed_IPL, test_hpc_benchmark.cpp:425-441, 575-601) compiled
                # into oracle/_ref, on its own containers, all host cores
                import oracle
                cores = host_threads()
                if oracle.ref() is not None:
                    ts, bref = oracle.ref_blur_timed(ref_which, fa, m, m, 6)
                    ts = ts[1:]
                    same = bool(np.array_equal(bref.view(np.uint32), np.asarray(host_b).view(np.uint32)))
                    e["cpu_baseline"] = {"value": m * m / min(ts) / 1e9, "unit": "Gcell-updates/s", "cores": cores,
                                         "kind": "reference", "bit_identical_to_gpu": same,
                                         "sample": f"full 8192^2 sweep, best of {len(ts)} after 1 warm-up, the "
                                                   f"reference's own {ref_which} loop (oracle/_ref/libhpcref.so)"}
                    summary[f"r0_{name}_cpu"] = {"value": round(m * m / min(ts) / 1e9, 3), "cores": cores,
                                                 "kind": "reference", "bit_identical_to_gpu": same}
                    if not same:
                        raise SystemExit(f"bench.py: GPU {name} sweep differs from the reference's own loop")
            out[f"r0_{name}_8192sq_fp32"] = e
        del fa
        # transpose (BM_transpose, test_hpc_benchmark.cpp:672-682) on the same 8192^2 shape: 8 B/cell
        c = hp.DeviceGrid(ctx, (m, m), np.float32, 0)
        dt, _ = timed(lambda: hp.grid_transpose(b, c, stream=stream), 20)
        out["transpose_8192sq_fp32"] = entry(m * m * 20, dt, None, ms_per_sweep=dt / 20 * 1e3)
        a.destroy()
        b.destroy()
        c.destroy()
        # 16 x 16 convolution (BM_conv, test_hpc_benchmark.cpp:786-806) on the reference's 1024^2 grid:
        # 512 separately rounded fp32 operations per cell (bit-exact order, no FMA) -> FP32-issue bound
        k = 1024
        kb = hp.DeviceGrid(ctx, (k, k), np.float32, 16)
        ka = hp.DeviceGrid(ctx, (k, k), np.float32, 0)
        kb.upload(synth.field(32, (k + 32) ** 2, np.float32))
        w = synth.field(33, 256, np.float32).reshape(16, 16)
        dt, _ = timed(lambda: hp.conv2d(kb, ka, w, stream=stream), 20)
        sm = torch.cuda.get_device_properties(0).multi_processor_count
        fp32_peak = sm * 128 * 1.965  # Gop/s: 128 FP32 lanes per SM at the 1965 MHz boost clock
        out["conv16x16_1024sq_fp32"] = {
            "metric": "stencil_Gcell_updates_per_s", "value": k * k * 20 / dt / 1e9, "ms_per_sweep": dt / 20 * 1e3,
            "roofline": {"bound": "fp32-issue (separately rounded mul + add, no FMA)",
                         "achieved": k * k * 512 * 20 / dt / 1e9, "peak": fp32_peak, "unit": "Gop/s",
                         "frac": k * k * 512 * 20 / dt / 1e9 / fp32_peak,
                         "hbm_GBps": k * k * 8 * 20 / dt / 1e9}}
        ka.destroy()
        kb.destroy()
    return out, summary


def cg_extra(args, ctx, torch, stream, world, barrier, max_over_ranks, hp, D, nloc, nnz, b):
    """CG iterations around the configs[1] operator (fused SpMV + <p,q>, then
    {x, r, <r,r>} and {p} passes; scalars on the device, all-reduced over the ranks).
    Algorithmic bytes per iteration and rank: SpMV (12 B/nnz + 20 B/row) + 72 B/row."""
    peak, _ = measured_peak()
    S = hp.CgSolver(D)
    x = torch.zeros(nloc, dtype=torch.float64, device="cuda")
    iters = max(args.steps, 10)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        S.solve(b, x, 3, 0.0, stream=stream, want_result=False)  # warm-up
        n0 = ctx.launch_count()
        x.zero_()
        dt, _, _ = time_steps(torch, stream,
                              lambda: S.solve(b, x, iters, 0.0, stream=stream, want_result=False), 1, barrier)
        launches = ctx.launch_count() - n0
    res = S.solve(b, x, 0)  # reads back the recurrence scalars of the timed solve's final state
    dt = max_over_ranks(dt)
    per_iter = dt / (iters + 1)  # the initial residual costs one more product + vector pass
    nbytes = algorithmic_bytes(nloc, nnz) + 72 * nloc
    S.destroy()
    return {"metric": "cg_effective_GBps", "value": nbytes * world / per_iter / 1e9, "unit": "GB/s",
            "iterations": iters, "ms_per_iteration": per_iter * 1e3, "gpu_launches": launches,
            "rr_over_bb_after_restart": res["rr"] / res["bb"] if res["bb"] else None,
            "roofline": {"bound": "hbm", "achieved": nbytes / per_iter / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": nbytes / per_iter / 1e9 / peak, "algorithmic_bytes_per_iteration": nbytes,
                         "note": "three passes per iteration: fused SpMV+dot, x/r update + <r,r>, p update"}}


def c1_extra(args, ctx, torch, stream, barrier, hp, synth):
    """configs[0] (C1, "reference CPU SIMD path (plumbing)"): 2-D 5-point Laplacian 1024 x 1024, fp64 —
    the one config the CPU runs comfortably.  84 MB of algorithmic bytes: L2-resident on the GPU, so it
    is reported as plumbing and excluded from roofline claims."""
    n = 1024
    rp, col, val = synth.laplace2d(n, n)
    A = hp.CsrMatrix(rp, col, val)
    t0 = time.perf_counter()
    D = hp.DeviceCsr(ctx, A)
    create_ms = (time.perf_counter() - t0) * 1e3
    x = synth.field(42, n * n)
    xd = torch.from_numpy(x).cuda()
    yd = torch.empty(n * n, dtype=torch.float64, device="cuda")
    xh, yh = ctx.pinned((n * n,), np.float64), ctx.pinned((n * n,), np.float64)
    xh[:] = x
    torch.cuda.synchronize()
    nbytes = algorithmic_bytes(n * n, len(col))
    with torch.cuda.stream(stream):
        for _ in range(5):
            D.spmv(xd, yd, stream)
        dt, _, _ = time_steps(torch, stream, lambda: D.spmv(xd, yd, stream), 50, barrier)
        for _ in range(3):
            D.spmv(xh, yh, stream)
        dte, _, _ = time_steps(torch, stream, lambda: D.spmv(xh, yh, stream), 20, barrier)
    e = {"metric": METRIC, "value": nbytes * 50 / dt / 1e9, "unit": UNIT, "us_per_step": dt / 50 * 1e6,
         "rows": n * n, "nnz": int(len(col)), "setup_ms": {"hpcx_csr_create": create_ms},
         "note": "L2-resident plumbing config: no roofline claim",
         "e2e": {"value": nbytes * 20 / dte / 1e9, "unit": UNIT, "us_per_step": dte / 20 * 1e6,
                 "h2d_bytes_per_step": n * n * 8, "d2h_bytes_per_step": n * n * 8,
                 "call": "hpcx_spmv(A, x HOST, y HOST)"}}
    if not args.no_cpu:
        import oracle
        cores = host_threads()
        y = np.empty(n * n)
        oracle.spmv(rp, col, val, x, out=y)
        ts = []
        for _ in range(20):
            t0 = time.perf_counter()
            oracle.spmv(rp, col, val, x, out=y)
            ts.append(time.perf_counter() - t0)
        same = bool(np.array_equal(y.view(np.uint64), np.asarray(yh).view(np.uint64)))
        e["cpu_baseline"] = {"value": nbytes / min(ts) / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
                             "us_per_step": min(ts) * 1e6, "bit_identical_to_gpu": same,
                             "sample": "full C1 matrix, best of 20 products (oracle/hpcx_oracle.c:orc_spmv_i32_f64)"}
        if not same:
            raise SystemExit("bench.py: C1 GPU product differs from the CPU restatement (expected bit-identical)")
    D.destroy()
    return e


def c5_distributed(args, ctx, torch, stream, world, rank, barrier, max_over_ranks, hp, hd):
    """configs[4] across GPUs (STRONG scaling, the global matrix fixed): nnz-balanced contiguous row
    blocks (hpcx_partition_rows mode 1).  The exchange is all-to-all in shape — every rank needs most
    of x — and runs through the same fused peer-store transport as the banded operators (one launch
    per product: push | interior | boundary + ack); no ncclAllGather fallback is needed.  Every rank
    generates and assembles the global R-MAT on its own GPU (0.1 s), keeps its row block, and the
    distributed localiser derives the halo tables; the product of a random x is checked against the
    single-GPU HUB kernel's result on the same rank."""
    scale = 24
    nr = 1 << scale
    t = time.time()
    arrs = hp.synth_rmat_csr(ctx, scale, 16, 24, 4)
    R = hp.DeviceCsr.from_device_arrays(ctx, arrs)  # single-GPU reference product (AUTO -> HUB)
    rp, col, val = arrs.download()
    arrs.destroy()
    x = torch.empty(nr, dtype=torch.float32, device="cuda")
    hp.synth_field(ctx, 42, nr, x, np.float32)
    y_one = torch.empty(nr, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    R.spmv(x, y_one, stream)
    stream.synchronize()
    R.destroy()
    bounds = hp.partition_rows(nr, rp, world, 1)
    r0, r1 = int(bounds[rank]), int(bounds[rank + 1])
    p0, p1 = int(rp[r0]), int(rp[r1])
    L = hd.localize_rows_distributed((rp[r0:r1 + 1] - p0).astype(np.int64), col[p0:p1], val[p0:p1], bounds, rank, world, nr)
    nnz_total = int(rp[-1])
    # the same rows with GLOBAL column ids, for the gathered-x variant below
    rp_blk, col_blk, val_blk = (rp[r0:r1 + 1] - p0).astype(np.int32), col[p0:p1].copy(), val[p0:p1].copy()
    del rp, col, val
    L["local_row_ptr"] = L["local_row_ptr"].astype(np.int32)
    A = hp.CsrMatrix(L["local_row_ptr"], L["local_col"], np.ascontiguousarray(L["values"]), ncols=L["nloc"] + L["nhalo"])
    D = hp.DeviceCsr(ctx, A, kernel=hp.SPMV_STREAM)
    D.set_halo(L)
    D.halo_transport(1)
    log(f"[rank {rank}] C5 block: rows [{r0}, {r1}), {L['nnz_loc']} nnz, {L['nhalo']} halo columns, "
        f"{L['nsend']} sends, set up in {time.time() - t:.1f}s")
    halo_cols = int(L["nhalo"])
    xfull = D.x_buffer()
    ctx.memcpy((xfull, hp.MEM_DEVICE), x[r0:r1].contiguous(), (r1 - r0) * 4)
    y = torch.empty(r1 - r0, dtype=torch.float32, device="cuda")
    x_arg = (xfull, hp.MEM_DEVICE)
    step = lambda: D.spmv_dist(x_arg, y, stream)  # noqa: E731
    barrier()
    step()
    stream.synchronize()
    # same matrix, same x: the distributed rows against the single-GPU product (different summation
    # orders in fp32: bound 3e-5 of max|y|, value reported)
    err = float((y - y_one[r0:r1]).abs().max() / y_one.abs().max())
    if not err <= 3e-5:
        raise SystemExit(f"[rank {rank}] bench.py: distributed C5 product differs from the single-GPU one ({err:.2e})")
    with torch.cuda.stream(stream):
        for _ in range(3):
            step()
        dt, _, _ = time_steps(torch, stream, step, args.steps, barrier)
    dt_push = max_over_ranks(dt)
    D.destroy()
    del D, A, L
    # ---- variant 2 (SURVEY.md 8(e)'s prescription for C5): hpcx_spmv_allgather — the row block keeps
    # GLOBAL column ids, every rank's slice of x is gathered into a full-length x (64 MiB in all, one
    # grouped ncclSend/ncclRecv), and the block runs the kernel hpcx_csr_create picks for it (HUB)
    t = time.time()
    G = hp.DeviceCsr(ctx, hp.CsrMatrix(rp_blk, col_blk, val_blk, ncols=nr))
    setup_gather = time.time() - t
    del rp_blk, col_blk, val_blk
    # x travels in EQUAL slices (ncclAllGather: balanced traffic) while the rows stay nnz-balanced;
    # with x split like the rows the rank owning the longest slice (56 % of x at N = 8) sends it N - 1
    # times and the gather alone takes 0.68 ms (profiles/r3f_bench_n8.json)
    xbounds = np.arange(world + 1, dtype=np.int64) * (nr // world)
    e0, e1 = int(xbounds[rank]), int(xbounds[rank + 1])
    xg = torch.full((nr,), float("nan"), dtype=torch.float32, device="cuda")
    xg[e0:e1] = x[e0:e1]
    x_own = xg[e0:e1]
    stepg = lambda: G.spmv_allgather(xbounds, x_own, xg, y, stream)  # noqa: E731
    torch.cuda.synchronize()
    barrier()
    y.fill_(float("nan"))
    torch.cuda.synchronize()  # the fill ran on torch's stream, the product runs on `stream`
    stepg()
    stream.synchronize()
    err_g = float((y - y_one[r0:r1]).abs().max() / y_one.abs().max())
    if not err_g <= 3e-5 or not torch.equal(xg, x):
        raise SystemExit(f"[rank {rank}] bench.py: gathered-x C5 product differs from the single-GPU one ({err_g:.2e})")
    with torch.cuda.stream(stream):
        for _ in range(3):
            stepg()
        dt, _, _ = time_steps(torch, stream, stepg, args.steps, barrier)
    dt_gather = max_over_ranks(dt)
    # the same block's kernel alone on the gathered x (no exchange): what the gather costs on top
    stepk = lambda: G.spmv(xg, y, stream)  # noqa: E731
    with torch.cuda.stream(stream):
        stepk()
        dt, _, _ = time_steps(torch, stream, stepk, args.steps, barrier)
    dt_kernel = max_over_ranks(dt)
    G.destroy()
    log(f"[rank {rank}] C5 gathered-x block set up in {setup_gather:.1f}s; {dt_gather / args.steps * 1e3:.3f} ms per product "
        f"(block kernel alone {dt / args.steps * 1e3:.3f} ms on this rank; halo push: {dt_push / args.steps * 1e3:.3f} ms)")
    dt = min(dt_push, dt_gather)
    gbytes = algorithmic_bytes(nr, nnz_total, 4)
    peak, _ = measured_peak()
    e = {"metric": METRIC, "value": gbytes * args.steps / dt / 1e9, "unit": UNIT, "scaling": "strong",
         "ms_per_step": dt / args.steps * 1e3, "rows": nr, "nnz": nnz_total,
         "partition": f"{world} nnz-balanced row blocks; rank 0 owns rows [0, {int(bounds[1])})",
         "halo_columns_rank0": halo_cols if rank == 0 else None,
         "max_rel_diff_vs_single_gpu": max(err, err_g),
         "variants": {
             "halo_push_stream_kernel": {"value": gbytes * args.steps / dt_push / 1e9, "ms_per_step": dt_push / args.steps * 1e3,
                                         "max_rel_diff_vs_single_gpu": err},
             "allgather_x_block_kernel": {"value": gbytes * args.steps / dt_gather / 1e9, "ms_per_step": dt_gather / args.steps * 1e3,
                                          "max_rel_diff_vs_single_gpu": err_g,
                                          "kernel_only_ms_max_over_ranks": dt_kernel / args.steps * 1e3,
                                          "call": "hpcx_spmv_allgather (x in equal slices by ncclAllGather, then the block's AUTO kernel on its nnz-balanced rows)"}},
         "roofline": {"bound": "hbm", "achieved": gbytes / world * args.steps / dt / 1e9, "peak": peak, "unit": "GB/s",
                      "frac": gbytes / world * args.steps / dt / 1e9 / peak,
                      "note": "global algorithmic bytes / N per rank, the faster of the two variants"}}
    return e


def sparse_extra(args, ctx, torch, stream, world, rank, barrier, max_over_ranks, hp, hd, synth):
    """The other SpMV configs of BASELINE.json as `extra` entries.
    C4 (configs[3]): 27-point operator on 256^3, fp64, row-partitioned into z-slab blocks with
    an x-halo of one plane (65,536 doubles) per neighbour — STRONG scaling (global size fixed).
    C5 (configs[4]): R-MAT scale 24, average degree 16, fp32, built on the device; N = 1 only."""
    out = {}
    peak, _ = measured_peak()
    n = 256
    zb = hp.partition_slabs(n, world)
    t = time.time()
    rp_s, col_s, val_s = synth.op27_rows(n, int(zb[rank]), int(zb[rank + 1]))
    L = hd.localize_rows_distributed(rp_s, col_s, val_s, zb * n * n, rank, world, n ** 3)
    del rp_s, col_s, val_s
    L["local_row_ptr"] = L["local_row_ptr"].astype(np.int32)
    nloc = L["nloc"]
    A = hp.CsrMatrix(L["local_row_ptr"], L["local_col"], L["values"], ncols=nloc + L["nhalo"])
    D = hp.DeviceCsr(ctx, A)
    log(f"[rank {rank}] C4 slab: {nloc} rows, {L['nnz_loc']} nnz built + uploaded in {time.time() - t:.1f}s")
    yd = torch.empty(nloc, dtype=torch.float64, device="cuda")
    xd = torch.ones(nloc + L["nhalo"], dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    if world > 1:
        D.set_halo(L)
        D.halo_transport(1)
        xfull = D.x_buffer()
        ctx.memcpy((xfull, hp.MEM_DEVICE), xd, nloc * 8)
        x_arg = (xfull, hp.MEM_DEVICE)
        step = lambda: D.spmv_dist(x_arg, yd, stream)  # noqa: E731
    else:
        step = lambda: D.spmv(xd, yd, stream)  # noqa: E731
    # property check (no oracle): row sums of the operator are 27 - (entries in the row)
    step()
    stream.synchronize()
    want = 27.0 - torch.from_numpy(np.diff(L["local_row_ptr"]).astype(np.float64)).cuda()
    if not torch.equal(yd, want):
        raise SystemExit(f"[rank {rank}] bench.py: C4 A*1 does not match the closed form")
    x = torch.empty(nloc, dtype=torch.float64, device="cuda")
    hp.synth_field(ctx, 42, nloc, x, np.float64, first=int(zb[rank]) * n * n)
    if world > 1:
        ctx.memcpy((xfull, hp.MEM_DEVICE), x, nloc * 8)
    else:
        xd[:nloc].copy_(x)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        for _ in range(3):
            step()
        n0 = ctx.launch_count()
        dt, _, _ = time_steps(torch, stream, step, args.steps, barrier)
        launches = ctx.launch_count() - n0
    dt = max_over_ranks(dt)
    nnz_total = (3 * n - 2) ** 3
    gbytes = algorithmic_bytes(n ** 3, nnz_total)
    rbytes = algorithmic_bytes(nloc, L["nnz_loc"])
    out["c4_op27_256cubed_fp64"] = {
        "metric": METRIC, "value": gbytes * args.steps / dt / 1e9, "unit": UNIT, "scaling": "strong",
        "ms_per_step": dt / args.steps * 1e3, "rows": n ** 3, "nnz": nnz_total, "gpu_launches": launches,
        "partition": f"{world} z-slab row blocks, x-halo {n * n} doubles per neighbour" +
                     (" (peer stores + flags)" if world > 1 else ""),
        "roofline": {"bound": "hbm", "achieved": rbytes * args.steps / dt / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": rbytes * args.steps / dt / 1e9 / peak, "algorithmic_bytes_per_launch": rbytes,
                     "note": "rank 0's bytes / max-over-ranks step time"}}
    D.destroy()
    del xd, yd, x, A, L
    if world > 1:
        out["c5_rmat_scale24_fp32"] = c5_distributed(args, ctx, torch, stream, world, rank, barrier, max_over_ranks,
                                                      hp, hd)
        return out

    scale = 24
    t = time.time()
    arrs = hp.synth_rmat_csr(ctx, scale, 16, 24, 4)
    R = hp.DeviceCsr.from_device_arrays(ctx, arrs)
    log(f"C5 R-MAT scale {scale}: {arrs.nnz} nnz after merging, generated + assembled + row-blocked "
        f"on the device in {time.time() - t:.1f}s")
    nr = 1 << scale
    rp, _, val = arrs.download()
    arrs.destroy()
    x = torch.ones(nr, dtype=torch.float32, device="cuda")
    y = torch.empty(nr, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    R.spmv(x, y, stream)
    stream.synchronize()
    # property check (no oracle): A*1 = row sums, against float64 prefix sums of the values
    cs = np.concatenate(([0.0], np.cumsum(val, dtype=np.float64)))
    ca = np.concatenate(([0.0], np.cumsum(np.abs(val), dtype=np.float64)))
    rs, ra = cs[rp[1:]] - cs[rp[:-1]], ca[rp[1:]] - ca[rp[:-1]]
    err = np.abs(y.cpu().numpy().astype(np.float64) - rs)
    if not np.all(err <= 1e-5 * ra + 1e-7 * np.abs(ca[-1]) / len(val)):
        raise SystemExit("bench.py: C5 A*1 does not match the row sums within 1e-5")
    lens = np.diff(rp)
    del cs, ca, rs, ra, val
    hp.synth_field(ctx, 42, nr, x, np.float32)
    torch.cuda.synchronize()
    step = lambda: R.spmv(x, y, stream)  # noqa: E731
    with torch.cuda.stream(stream):
        for _ in range(3):
            step()
        n0 = ctx.launch_count()
        dt, _, _ = time_steps(torch, stream, step, args.steps, barrier)
        launches = ctx.launch_count() - n0
    rbytes = algorithmic_bytes(nr, R.nnz, 4)
    bins = R.row_bins()
    out["c5_rmat_scale24_fp32"] = {
        "metric": METRIC, "value": rbytes * args.steps / dt / 1e9, "unit": UNIT,
        "ms_per_step": dt / args.steps * 1e3, "rows": nr, "nnz": int(R.nnz), "gpu_launches": launches,
        "row_length": {"max": int(lens.max()), "mean": float(lens.mean()), "empty_rows": int((lens == 0).sum()),
                       "ceil_log2_bins": [int(b) for b in bins[:int(np.flatnonzero(bins)[-1]) + 1]]},
        "roofline": {"bound": "hbm", "achieved": rbytes * args.steps / dt / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": rbytes * args.steps / dt / 1e9 / peak, "algorithmic_bytes_per_launch": rbytes,
                     "note": "8 B/nnz + 12 B/row: x counted as compulsory only (4 B/row); the gather's "
                             "real DRAM traffic is in profiles/"}}
    R.destroy()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--configs", default="all", choices=["all", "headline"],
                    help="all: also time configs[3] (27-point 256^3) and configs[4] (R-MAT) as extras")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port", "29533",
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local)


if __name__ == "__main__":
    main()
