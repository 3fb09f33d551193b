#!/usr/bin/env python
"""A/B sweep of the STREAM kernel's create-time knobs (tile_nnz, short_row_max) on the three SpMV
configs at full size.  Run on the GPU box: python scripts/spmv_sweep.py > gpurun_out/<tag>/sweep.txt"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp  # noqa: E402
from libhpc_b200 import synth  # noqa: E402

ctx = hp.Context(0)
stream = torch.cuda.Stream()
which = sys.argv[1:] or ["c2", "c4", "c5"]


def timeit(D, x, y, reps=20):
    with torch.cuda.stream(stream):
        for _ in range(3):
            D.spmv(x, y, stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            D.spmv(x, y, stream)
        e1.record(stream)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def sweep(name, make, dtype, combos):
    nrows = None
    for tile, srm in combos:
        D = make(tile, srm)
        nrows = D.nrows
        tdt = torch.float64 if dtype == np.float64 else torch.float32
        x = torch.empty(D.ncols, dtype=tdt, device="cuda")
        hp.synth_field(ctx, 42, D.ncols, x, dtype)
        y = torch.empty(nrows, dtype=tdt, device="cuda")
        ms = timeit(D, x, y)
        gbs = D.algorithmic_bytes() / ms / 1e6
        bl = D.rowblocks()
        lanes, cnt = np.unique(bl[:, 3], return_counts=True)
        print(f"{name} tile={tile} short_row_max={srm}: {ms:.4f} ms  {gbs:.0f} GB/s  blocks={len(bl)} "
              f"lanes={dict(zip(lanes.tolist(), cnt.tolist()))} checksum={float(y.double().abs().sum()):.6e}",
              flush=True)
        D.destroy()


if "c2" in which:
    t = time.time()
    rp, col, val = synth.laplace2d_big(8192, 8192)
    A = hp.CsrMatrix(rp, col, val)
    print(f"c2 built in {time.time() - t:.1f}s", flush=True)
    sweep("c2", lambda tile, srm: hp.DeviceCsr(ctx, A, tile_nnz=tile, short_row_max=srm), np.float64,
          [(1024, 48), (2048, 48), (4096, 48)])
    del A, rp, col, val
if "c4" in which:
    t = time.time()
    L = synth.op27_slab_local(256, 0, 256)
    A = hp.CsrMatrix(L["local_row_ptr"].astype(np.int32), L["local_col"], L["values"])
    print(f"c4 built in {time.time() - t:.1f}s", flush=True)
    sweep("c4", lambda tile, srm: hp.DeviceCsr(ctx, A, tile_nnz=tile, short_row_max=srm), np.float64,
          [(2048, 8), (1024, 48), (2048, 48), (4096, 48)])
    del A, L
if "c5" in which:
    arrs = hp.synth_rmat_csr(ctx, 24, 16, 24, 4)
    sweep("c5", lambda tile, srm: hp.DeviceCsr.from_device_arrays(ctx, arrs, tile_nnz=tile, short_row_max=srm),
          np.float32, [(2048, 8), (1024, 48), (2048, 48), (4096, 48), (2048, 16), (2048, 128), (4096, 128)])
    arrs.destroy()
ctx.close()
