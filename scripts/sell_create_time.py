import sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import libhpc_b200 as hp
from libhpc_b200 import synth
ctx = hp.Context(0)
rp, col, val = synth.op27(256)
A = hp.CsrMatrix(rp, col, val)
for sigma in (1024, 1):
    t = time.time(); D = hp.DeviceCsr(ctx, A, kernel=hp.SPMV_SELL, sell_sigma=sigma); print(f"C4 SELL-32-{sigma} create {time.time()-t:.2f} s", flush=True); D.destroy()
t = time.time(); D = hp.DeviceCsr(ctx, A, kernel=hp.SPMV_STREAM); print(f"C4 STREAM create {time.time()-t:.2f} s", flush=True); D.destroy()
