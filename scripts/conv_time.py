"""Times the 16 x 16 convolution on the reference's 1024^2 grid and checks it against the oracle."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import libhpc_b200 as hp
import oracle
from libhpc_b200 import synth
ctx = hp.Context(0)
k = 1024
kb = hp.DeviceGrid(ctx, (k, k), np.float32, 16); ka = hp.DeviceGrid(ctx, (k, k), np.float32, 0)
fb = synth.field(32, (k + 32) ** 2, np.float32)
kb.upload(fb)
w = synth.field(33, 256, np.float32).reshape(16, 16)
hp.conv2d(kb, ka, w)
ok = np.array_equal(ka.download(), oracle.conv2d(fb, np.zeros(k * k, np.float32), w, k, k, 16, 16))
st = torch.cuda.Stream()
for _ in range(3): hp.conv2d(kb, ka, w, stream=st)
torch.cuda.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(20): hp.conv2d(kb, ka, w, stream=st)
    e1.record(st); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) * 1e3 / 20)
print(f"{os.environ.get('TAG','')} bit_exact={ok}: {best:.1f} us  {k*k/best/1e3:.1f} Gcell/s", flush=True)
