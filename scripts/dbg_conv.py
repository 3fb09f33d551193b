import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import libhpc_b200 as hp
import oracle
from test_gpu_dense2d import _conv
ctx = hp.Context(0)
for (ny, nx), ks, ilo, ihi in (((64, 64), (16, 16), 16, 16), ((70, 45), (16, 16), 16, 16), ((64, 45), (16, 16), 16, 16),
                               ((70, 64), (16, 16), 16, 16), ((70, 45), (16, 16), 0, 15)):
    kb = oracle.fill((ny + ilo + ihi) * (nx + ilo + ihi), 21, np.float32)
    ka = oracle.fill(ny * nx, 22, np.float32)
    w = oracle.fill(ks[0] * ks[1], 23, np.float32).reshape(ks)
    want = oracle.conv2d(kb, ka, w, ny, nx, ilo, ihi).reshape(ny, nx)
    for v in (hp.VARIANT_AUTO, hp.VARIANT_DIRECT):
        got = _conv(ctx, kb, ka, w, ny, nx, ilo, ihi, variant=v).reshape(ny, nx)
        bad = got != want
        print((ny, nx), ilo, ihi, "variant", v, "bad", int(bad.sum()), "of", bad.size,
              "rows", np.flatnonzero(bad.any(axis=1))[:5], "cols", np.flatnonzero(bad.any(axis=0))[:5],
              "maxabs", float(np.abs(got).max()), flush=True)
