"""Where the HUB kernel's time goes on C5: the long-row part and the short-row part timed alone."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp  # noqa: E402

ctx = hp.Context(0)
arrs = hp.synth_rmat_csr(ctx, 24, 16, 24, 4)
rp, col, val = arrs.download()
arrs.destroy()
n = 1 << 24
lens = np.diff(rp).astype(np.int64)
x = torch.rand(n, device="cuda")
y = torch.empty(n, device="cuda")


def timed(name, keep):
    l2 = np.where(keep, lens, 0)
    rp2 = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(l2, out=rp2[1:])
    mask = np.repeat(keep, lens)
    c2, v2 = col[mask], val[mask]
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp2.astype(np.int32), c2, v2), kernel=hp.SPMV_HUB)
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(s):
        for _ in range(3):
            D.spmv(x, y, s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(10):
            D.spmv(x, y, s)
        e1.record(s)
    torch.cuda.synchronize()
    print(f"{name:28s} rows {int(keep.sum()):9d} nnz {len(c2):10d}  {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
    D.destroy()


timed("all rows", lens >= 0)
timed("rows 33..256 only", (lens > 32) & (lens <= 256))
timed("short rows only (1..32)", (lens > 0) & (lens <= 32))
timed("rows > 256 only", lens > 256)
ctx.close()
