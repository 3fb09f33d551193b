"""Times the single-sweep Jacobi variants on 512^3 (A/B: HPCX_JACOBI_PRODUCER=1 selects the
producer-warp kernel) over a few z splits."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libhpc_b200 as hp
from libhpc_b200 import synth
ctx = hp.Context(0)
n = int(os.environ.get("N", 512))
f = synth.field(7, (n + 2) ** 3, np.float32)
ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, 1), hp.DeviceGrid(ctx, (n, n, n), np.float32, 1)
ga.upload(f); gb.upload(f)
st = torch.cuda.Stream()
variant = hp.VARIANT_TEMPORAL if os.environ.get("TB") == "1" else hp.VARIANT_TMA
tag = "temporal" if os.environ.get("TB") == "1" else "producer" if os.environ.get("HPCX_JACOBI_PRODUCER") == "1" else "ring"
for zs in [int(v) for v in os.environ.get('ZS', '0,9,12,15,18,24,30,36').split(',')]:
    hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=4, variant=variant, zsplit=zs, stream=st)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=40, variant=variant, zsplit=zs, stream=st)
        e1.record(st)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / 40)
    print(f"{tag} n={n} zsplit={zs}: {best:.1f} us/sweep  {n**3 / best / 1e3:.1f} Gcell/s", flush=True)
