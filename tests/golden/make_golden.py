"""Generates the golden vectors of tests/golden/ from the REFERENCE's own code.

Run in the authoring container (where /root/reference is mounted and oracle/_ref has been
built by `make -C oracle`):   python tests/golden/make_golden.py

  blur_ref.npz    outputs of the reference loop bodies BM_x_blur, BM_x_blur_tiling_simd_prefetch,
                  BM_y_blur, BM_YXx_blur_tiling_prefetch_streamed_IPL
                  (tests/test_hpc_benchmark/test_hpc_benchmark.cpp:354-368,425-441,444-457,575-601)
                  on a 64 x 128 float32 grid, ghost 8, a = 2u(seed=3, padded index) - 1,
                  ghosts included; plus sha256 digests of the same sweeps at 512 x 1024.
  index_ref.json  compute_stride_and_total / unsafe_linearize / safe_linearize of
                  hpc::HPCHighDimensionFlatArray (lib/hpc/include/HPCHighDimensionFlatArray.hpp:
                  161-208) for a list of (dims, lo, hi, idx) cases incl. ghost and out-of-range.
Inputs are regenerated from the seed by the tests; only outputs are stored.
"""
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    assert oracle.ref() is not None, "build oracle/_ref first (make -C oracle)"
    out = {}
    ny, nx = 64, 128
    a = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    for which in oracle.REF_BLUR:
        out[which] = oracle.ref_blur(which, a, ny, nx)
    ny2, nx2 = 512, 1024
    a2 = oracle.fill((ny2 + 16) * (nx2 + 16), 3, np.float32)
    digests = {w: hashlib.sha256(oracle.ref_blur(w, a2, ny2, nx2).tobytes()).hexdigest()
               for w in oracle.REF_BLUR}
    np.savez_compressed(os.path.join(HERE, "blur_ref.npz"), shape=np.array([ny, nx]),
                        big_shape=np.array([ny2, nx2]),
                        big_digests=np.array(json.dumps(digests)), **out)

    cases = []
    shapes = [([7], 0, 0), ([7], 2, 2), ([5], 1, 3), ([4, 6], 0, 0), ([4, 6], 8, 8), ([3, 5], 1, 1),
              ([3, 5], 1, 3), ([2, 9], 3, 0), ([3, 4, 5], 0, 0), ([3, 4, 5], 1, 1), ([2, 3, 4], 2, 2),
              ([2, 3, 4], 1, 3), ([4, 2, 3], 0, 2), ([8192, 8192], 8, 8), ([512, 512, 512], 1, 1)]
    rng = np.random.default_rng(0)
    for dim, lo, hi in shapes:
        idxs = [[0] * len(dim), [-lo] * len(dim), [d + hi - 1 for d in dim], [d + hi for d in dim],
                [-lo - 1] + [0] * (len(dim) - 1)]
        for _ in range(6):
            idxs.append([int(rng.integers(-lo - 1, d + hi + 1)) for d in dim])
        for idx in idxs:
            rc_u, st, total, off_u = oracle.ref_index(dim, lo, hi, idx, 0)
            rc_s, _, _, off_s = oracle.ref_index(dim, lo, hi, idx, 1)
            assert rc_u == 0
            cases.append(dict(dim=dim, lo=lo, hi=hi, idx=idx, stride=[int(s) for s in st],
                              total=total, unsafe=off_u, safe=(off_s if rc_s == 0 else None)))
    with open(os.path.join(HERE, "index_ref.json"), "w") as f:
        json.dump(cases, f, indent=0)
    print("wrote", len(cases), "index cases and", len(out), "blur outputs")


if __name__ == "__main__":
    main()
