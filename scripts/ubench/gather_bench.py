"""x-gather strategies on the C5 matrix (R-MAT scale 24): LDG vs texture path vs hub slice in
shared memory, CSR order vs column-sorted tiles.  Microbenchmark, not product code."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp  # noqa: E402

lib = C.CDLL(os.path.join(os.path.dirname(__file__), "libgather.so"))
lib.gb_run.restype = C.c_float
lib.gb_run.argtypes = [C.c_int] + [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int]

scale = int(os.environ.get("SCALE", 24))
ctx = hp.Context(0)
arrs = hp.synth_rmat_csr(ctx, scale, 16, 24, 4)
rp, col_h, val_h = arrs.download()
arrs.destroy()
n = 1 << scale
nnz = len(col_h) // 4 * 4
col = torch.from_numpy(col_h[:nnz]).cuda()
val = torch.from_numpy(val_h[:nnz]).cuda()
x = torch.rand(n, device="cuda")
out = torch.empty(148 * 8 * 1024, device="cuda")
freq = torch.bincount(col, minlength=n)
order = torch.argsort(freq, descending=True, stable=True)
sms = torch.cuda.get_device_properties(0).multi_processor_count
algo_bytes = 8 * nnz + 12 * n


def run(which, c, ch, hub, K, blocks=sms):
    ms = lib.gb_run(which, c.data_ptr(), ch.data_ptr() if ch is not None else None, val.data_ptr(), x.data_ptr(), n,
                    hub.data_ptr() if hub is not None else None, K, nnz, out.data_ptr(), blocks)
    torch.cuda.synchronize()
    return ms


def hub_encode(c, K):
    hub = order[:K].to(torch.int32).contiguous()
    slot = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    slot[hub.long()] = torch.arange(K, dtype=torch.int32, device="cuda")
    s = slot[c.long()]
    ch = torch.where(s >= 0, s | torch.tensor(-2147483648, dtype=torch.int32, device="cuda"), c)
    return ch.contiguous(), hub


names = {0: "stream only", 1: "ldg vec4", 2: "tex vec4", 3: "ldg lane U4", 4: "ldg lane U8", 5: "tex lane U4", 6: "tex lane U8"}
orders = {"csr": col}
if os.environ.get("QUICK") == "1":
    ch, hub = hub_encode(col, 16384)
    for w, nm in ((7, "hub ldg U8 1024t grid-stride"), (12, "hub contiguous warp ranges, scalar loads"),
                  (13, "hub contiguous warp ranges, 128-bit loads")):
        print(f"K=16384 {nm:44s} {run(w, col, ch, hub, 16384):7.3f} ms", flush=True)
    ctx.close()
    sys.exit(0)
T = 32768
m = nnz // T * T
cs = col.clone()
cs[:m] = col[:m].view(-1, T).sort(dim=1).values.view(-1)
orders["sorted32k"] = cs
for oname, c in orders.items():
    for w, nm in names.items():
        ms = run(w, c, None, None, 0)
        print(f"{oname:10s} {nm:14s} {ms:7.3f} ms  {algo_bytes / ms / 1e6:7.0f} GB/s-eff", flush=True)
    for K in (16384, 32768, 49152):
        ch, hub = hub_encode(c, K)
        frac = float((ch < 0).float().mean())
        for w, nm in ((7, "hub ldg U8 1024t"), (8, "hub tex U8 1024t"), (9, "hub ldg U16 1024t"), (10, "hub ldg U8 512t")):
            ms = run(w, c, ch, hub, K)
            print(f"{oname:10s} K={K:6d} ({frac:.2f} hub) {nm:18s} {ms:7.3f} ms  {algo_bytes / ms / 1e6:7.0f} GB/s-eff", flush=True)
    for K, ctas in ((8192, 4), (12288, 4), (16384, 3)):
        ch, hub = hub_encode(c, K)
        frac = float((ch < 0).float().mean())
        ms = run(11, c, ch, hub, K, sms * ctas)
        print(f"{oname:10s} K={K:6d} ({frac:.2f} hub) hub ldg U8 512t x{ctas} CTAs/SM {ms:7.3f} ms  {algo_bytes / ms / 1e6:7.0f} GB/s-eff", flush=True)
ctx.close()
