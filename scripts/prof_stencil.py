#!/usr/bin/env python
"""Short stencil-only driver for ncu captures (Jacobi 512^3 fp32, R0 box sweeps 8192^2).
Prints CUDA-event timings when run plain; the same command line is what ncu profiles."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402

import libhpc_b200 as hp  # noqa: E402
from libhpc_b200 import synth  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ctx = hp.Context(0)
stream = torch.cuda.Stream()


def timed(fn, n):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(n):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


if which in ("all", "jacobi"):
    n = 512
    f = synth.field(7, (n + 2) ** 3, np.float32)
    a, b = hp.DeviceGrid(ctx, (n, n, n), np.float32, 1), hp.DeviceGrid(ctx, (n, n, n), np.float32, 1)
    a.upload(f)
    b.upload(f)
    for variant, name in ((hp.VARIANT_TMA, "tma"), (hp.VARIANT_DIRECT, "direct")):
        ms = timed(lambda: hp.stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=2, variant=variant, stream=stream), reps) / 2
        print(f"jacobi7 512^3 {name}: {ms * 1e3:.1f} us/sweep, {n ** 3 / ms / 1e6:.1f} Gcell/s, "
              f"{n ** 3 * 8 / ms / 1e6:.0f} GB/s")
    a.destroy()
    b.destroy()
if which in ("all", "box"):
    m = 8192
    a = hp.DeviceGrid(ctx, (m, m), np.float32, 8)
    b = hp.DeviceGrid(ctx, (m, m), np.float32, 0)
    a.upload(synth.field(3, (m + 16) ** 2, np.float32))
    for kind, name in ((hp.STENCIL_BOX_X, "box_x"), (hp.STENCIL_BOX_Y, "box_y")):
        for variant, vn, zs in ((hp.VARIANT_AUTO, "fast", 0), (hp.VARIANT_AUTO, "fast16", 16),
                                (hp.VARIANT_DIRECT, "direct", 0)):
            ms = timed(lambda: hp.stencil_sweep(a, b, kind, radius=8, variant=variant, zsplit=zs, stream=stream), reps)
            print(f"{name} 8192^2 {vn}: {ms * 1e3:.1f} us/sweep, {m * m / ms / 1e6:.1f} Gcell/s, "
                  f"{m * m * 8 / ms / 1e6:.0f} GB/s")
ctx.close()
