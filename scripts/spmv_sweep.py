"""Every SpMV kernel of the library on the three matrix shapes of BASELINE.json at full size
(C2 Laplacian 8192^2 fp64, C4 27-point 256^3 fp64, C5 R-MAT scale 24 fp32): create time, time per
product, effective GB/s on the algorithmic bytes of SURVEY.md 8(d).  SURVEY.md 8 a9 asks for the
row-binned and SELL-C-sigma variants; this is where their numbers come from (profiles/*_spmv_kernels.txt)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp  # noqa: E402
from libhpc_b200 import synth  # noqa: E402

ctx = hp.Context(0)
PEAK = 6543.7
if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
    import json
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", PEAK))


def bench(name, make, n, ncols, dtype, nbytes, variants):
    x = torch.from_numpy(synth.field(42, ncols, dtype)).cuda()
    y = torch.empty(n, dtype=x.dtype, device="cuda")
    ref = None
    for label, kw in variants:
        t = time.time()
        try:
            D = make(**kw)
        except Exception as e:  # a variant the library refuses for this shape
            print(f"{name:4s} {label:28s} refused: {str(e)[:80]}", flush=True)
            continue
        tc = time.time() - t
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            for _ in range(3):
                D.spmv(x, y, stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                D.spmv(x, y, stream)
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        out = y.cpu().numpy().astype(np.float64)
        if ref is None:
            ref = out
        err = np.abs(out - ref).max() / max(np.abs(ref).max(), 1e-300)
        print(f"{name:4s} {label:28s} create {tc:6.2f} s  {ms:7.3f} ms  {nbytes / ms / 1e6:7.0f} GB/s  "
              f"{nbytes / ms / 1e6 / PEAK:5.3f} of HBM  max diff vs first {err:.1e}", flush=True)
        D.destroy()


GENERIC = [("STREAM", dict(kernel=hp.SPMV_STREAM)), ("VECTOR", dict(kernel=hp.SPMV_VECTOR)),
           ("SELL-32-1024", dict(kernel=hp.SPMV_SELL)), ("SELL-32-1 (unsorted)", dict(kernel=hp.SPMV_SELL, sell_sigma=1))]

# C2
rp, col, val = synth.laplace2d(8192, 8192)
A = hp.CsrMatrix(rp, col, val)
bench("C2", lambda **kw: hp.DeviceCsr(ctx, A, **kw), A.nrows, A.nrows, np.float64, 12 * A.nnz + 20 * A.nrows + 4, GENERIC)
del A, rp, col, val
# C4
rp, col, val = synth.op27(256)
A = hp.CsrMatrix(rp, col, val)
bench("C4", lambda **kw: hp.DeviceCsr(ctx, A, **kw), A.nrows, A.nrows, np.float64, 12 * A.nnz + 20 * A.nrows + 4, GENERIC)
del A, rp, col, val
# C5 (assembled on the device; SELL is built from the downloaded arrays)
arrs = hp.synth_rmat_csr(ctx, 24, 16, 24, 4)
n = 1 << 24
rp, col, val = arrs.download()
nnz = len(col)
A = hp.CsrMatrix(rp, col, val, ncols=n)
variants = GENERIC[:2] + [("HUB", dict(kernel=hp.SPMV_HUB)), ("AUTO", dict(kernel=hp.SPMV_AUTO))] + \
    [(f"SELL-32-{s}", dict(kernel=hp.SPMV_SELL, sell_sigma=s)) for s in (1024, 65536, 1 << 24)]
bench("C5", lambda **kw: hp.DeviceCsr(ctx, A, **kw), n, n, np.float32, 8 * nnz + 12 * n + 4, variants)
