"""Run under HPCX_JACOBI_DOWN=1 (see test_gpu_stencil.py): the top z segment of the TMA Jacobi
sweep marches downwards, as it does when it feeds a high neighbour in the peer-store transport;
results must stay bit-identical to the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import libhpc_b200 as hp  # noqa: E402
import oracle  # noqa: E402
from util import bits_equal  # noqa: E402

assert os.environ.get("HPCX_JACOBI_DOWN") == "1"
with hp.Context(0) as ctx:
    for dims, steps, zsplit in (((37, 20, 140), 3, 2), ((37, 20, 140), 2, 5), ((64, 64, 128), 4, 0),
                                ((2, 9, 33), 3, 0), ((16, 100, 256), 2, 0)):
        nz, ny, nx = dims
        f = oracle.fill((nz + 2) * (ny + 2) * (nx + 2), 7, np.float32)
        a, b = hp.DeviceGrid(ctx, dims, np.float32, 1), hp.DeviceGrid(ctx, dims, np.float32, 1)
        a.upload(f)
        b.upload(f)
        hp.stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=steps, variant=hp.VARIANT_TMA, zsplit=zsplit)
        assert bits_equal(b.download(), oracle.jacobi7(f, nz, ny, nx, 1, 1, steps)), dims
        a.destroy()
        b.destroy()
print("DOWN OK")
