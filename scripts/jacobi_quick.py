import os, sys
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import libhpc_b200 as hp
from libhpc_b200 import synth
ctx = hp.Context(0)
n = 512
f = synth.field(7, (n + 2) ** 3, np.float32)
ga, gb = hp.DeviceGrid(ctx, (n, n, n), np.float32, 1), hp.DeviceGrid(ctx, (n, n, n), np.float32, 1)
ga.upload(f); gb.upload(f)
stream = torch.cuda.Stream()
for rep in range(6):
    with torch.cuda.stream(stream):
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=50, variant=hp.VARIANT_TMA, stream=stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        hp.stencil_sweep(ga, gb, hp.STENCIL_JACOBI7, steps=100, variant=hp.VARIANT_TMA, stream=stream)
        e1.record(stream)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 10
    print(f"rep {rep}: {us:.1f} us/sweep {n**3/us/1e3:.1f} Gcell/s", flush=True)
os.system("nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_throttle_reasons.active --format=csv")
