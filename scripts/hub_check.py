"""Quick GPU check + timing of the HUB SpMV kernel (not a test: tests/test_gpu_spmv.py covers it)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import libhpc_b200 as hp  # noqa: E402
import oracle  # noqa: E402
from libhpc_b200 import synth  # noqa: E402
from util import assert_spmv_close, random_csr  # noqa: E402

ctx = hp.Context(0)
ok = True


def check(name, rp, col, val, x, ncols=None, tol=None):
    global ok
    tol = tol or (1e-12 if val.dtype == np.float64 else 1e-5)
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp, col, val, ncols=ncols), kernel=hp.SPMV_HUB)
    y = np.full(len(rp) - 1, np.nan, dtype=val.dtype)
    D.spmv(x, y)
    D.destroy()
    try:
        assert_spmv_close(y, oracle.spmv(rp, col, val, x), oracle.spmv_abs(rp, col, val, x), tol)
        print("ok  ", name, flush=True)
    except AssertionError as e:
        ok = False
        print("FAIL", name, e, flush=True)


for scale in (8, 10, 13, 16):
    rp, col, val = synth.rmat(scale, 16, 24)
    check(f"rmat{scale} f32", rp, col, val, synth.field(42, 1 << scale, np.float32))
    check(f"rmat{scale} f64", rp, col, val.astype(np.float64), synth.field(42, 1 << scale, np.float64))
lens = ([0, 0, 1, 2, 0, 5000, 3, 0, 33, 64, 65, 127, 128, 129, 0, 2047, 2048, 2049, 1, 0] * 3 + [9000, 0, 0, 0, 1, 1, 1, 1, 17])
for dt in (np.float32, np.float64):
    rp, col, val = random_csr(len(lens), 9500, lens, 5, dt)
    check(f"ragged {np.dtype(dt).name}", rp, col, val, oracle.fill(9500, 6, dt), ncols=9500)
rp, col, val = synth.laplace2d(300, 200)
check("laplace 300x200", rp, col, val, synth.field(1, 60000))
rp, col, val = synth.op27(24)
check("op27 24", rp, col, val, synth.field(1, 24 ** 3))
n = 10000
check("one dense row", np.array([0, n], dtype=np.int64), np.arange(n, dtype=np.int32), oracle.fill(n, 3, np.float64),
      oracle.fill(n, 4, np.float64), ncols=n)
check("single entry", np.array([0, 1]), np.array([2], np.int32), np.array([3.0]), np.array([1.0, 2.0, 4.0]), ncols=3)
check("trailing empty rows", np.array([0, 3, 3, 3, 3]), np.array([0, 1, 2], np.int32), np.array([1.0, 2.0, 3.0]),
      np.array([1.0, 1.0, 1.0]), ncols=3)

# ---- C5 at full size: time STREAM vs HUB, check HUB against the oracle
if os.environ.get("FULL", "1") == "1":
    arrs = hp.synth_rmat_csr(ctx, 24, 16, 24, 4)
    rp, col, val = arrs.download()
    n = 1 << 24
    x = oracle.fill(n, 42, np.float32)
    xd = torch.from_numpy(x).cuda()
    yd = torch.empty(n, dtype=torch.float32, device="cuda")
    nbytes = 8 * len(col) + 12 * n + 4
    y_ref = oracle.spmv_wide(rp, col, val, x)
    mag = oracle.spmv_abs(rp, col, val, x)
    for kernel, name in ((hp.SPMV_STREAM, "STREAM"), (hp.SPMV_HUB, "HUB"), (hp.SPMV_AUTO, "AUTO")):
        t = time.time()
        D = hp.DeviceCsr.from_device_arrays(ctx, arrs, kernel=kernel)
        tc = time.time() - t
        stream = torch.cuda.Stream()
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            for _ in range(3):
                D.spmv(xd, yd, stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                D.spmv(xd, yd, stream)
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        try:
            assert_spmv_close(yd.cpu().numpy(), y_ref, mag, 1e-5)
            verdict = "parity ok"
        except AssertionError as e:
            verdict = f"PARITY FAIL {e}"
            ok = False
        print(f"C5 {name:6s} create {tc:.2f}s  {ms:.3f} ms  {nbytes / ms / 1e6:.0f} GB/s  {verdict}", flush=True)
        D.destroy()
    arrs.destroy()
print("HUB CHECK", "OK" if ok else "FAILED", flush=True)
ctx.close()
