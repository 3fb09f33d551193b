"""Golden vectors of the conv / transpose sweeps from the REFERENCE's own loops (run where
/root/reference is mounted, after `make -C oracle`):   python tests/golden/make_golden_dense2d.py

  dense2d_ref.npz   BM_conv and BM_conv_block (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:
                    793-806, 808-818) on a 64 x 64 float32 grid: ka = 2u(seed 31)-1 (accumulated
                    onto), kb = 2u(seed 32)-1 over the <2,float,16> padded buffer (ghosts
                    included), kc = 2u(seed 33)-1; sha256 digests of the same at 256 x 256.
                    BM_transpose and BM_transpose_tiling (:672-682, 684-698) digests on 128 x 128
                    of a = 2u(seed 34)-1.
The reference loops are compiled with -ffp-contract=off (oracle/Makefile): one rounding per
multiply and per add, the arithmetic contract of the whole build.
Inputs are regenerated from the seeds by the tests; only outputs are stored.
"""
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def conv_inputs(n):
    return (oracle.fill(n * n, 31, np.float32), oracle.fill((n + 32) ** 2, 32, np.float32),
            oracle.fill(256, 33, np.float32))


def main():
    assert oracle.ref() is not None, "build oracle/_ref first (make -C oracle)"
    n = 64
    ka, kb, kc = conv_inputs(n)
    out = {w: oracle.ref_conv(w, ka, kb, kc, n) for w in ("scalar", "block")}
    n2 = 256
    ka2, kb2, kc2 = conv_inputs(n2)
    dig = {"conv_" + w: hashlib.sha256(oracle.ref_conv(w, ka2, kb2, kc2, n2).tobytes()).hexdigest()
           for w in ("scalar", "block")}
    nt = 128
    a = oracle.fill(nt * nt, 34, np.float32)
    for w in ("scalar", "tiling"):
        dig["transpose_" + w] = hashlib.sha256(oracle.ref_transpose(w, a, nt).tobytes()).hexdigest()
    np.savez_compressed(os.path.join(HERE, "dense2d_ref.npz"), n=np.array(n), n_big=np.array(n2),
                        n_transpose=np.array(nt), digests=np.array(json.dumps(dig)),
                        conv_scalar=out["scalar"], conv_block=out["block"])
    print("wrote dense2d_ref.npz", dig)


if __name__ == "__main__":
    main()
