"""Runs a few temporally blocked Jacobi passes on 512^3 (for ncu captures and quick timing)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libhpc_b200 as hp
from libhpc_b200 import synth
ctx = hp.Context(0)
n = 512
f = synth.field(7, (n + 2) ** 3, np.float32)
ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, 1), hp.DeviceGrid(ctx, (n, n, n), np.float32, 1)
ga.upload(f); gb.upload(f)
st = torch.cuda.Stream()
for variant, name in ((hp.VARIANT_TEMPORAL, "temporal"), (hp.VARIANT_TMA, "tma")):
    for zs in ([0, 1, 2, 3, 4] if variant == hp.VARIANT_TEMPORAL else [0]):
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=4, variant=variant, zsplit=zs, stream=st)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=40, variant=variant, zsplit=zs, stream=st)
        e1.record(st)
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 40
        print(f"{name} zsplit={zs}: {us:.1f} us/sweep  {n**3 / us / 1e3:.1f} Gcell/s", flush=True)
