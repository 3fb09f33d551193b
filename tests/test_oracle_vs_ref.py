"""Pins the oracle: against the committed golden vectors produced by the reference's own
code (tests/golden/make_golden.py) and, where oracle/_ref was built, against the reference
library itself."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from util import bits_equal

G = os.path.join(os.path.dirname(__file__), "golden")
BLUR = np.load(os.path.join(G, "blur_ref.npz"))
INDEX = json.load(open(os.path.join(G, "index_ref.json")))
AXIS = {"x_scalar": 1, "x_sse": 1, "y_scalar": 0, "y_sse_ipl": 0}


@pytest.mark.parametrize("which", sorted(AXIS))
def test_box_sweep_matches_reference_golden(which):
    ny, nx = (int(v) for v in BLUR["shape"])
    a = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    got = oracle.box_sweep(a, ny, nx, 8, 8, 0, 0, AXIS[which], 8)
    assert bits_equal(got, BLUR[which])  # bit-exact: adds only, same order


@pytest.mark.parametrize("which", sorted(AXIS))
def test_box_sweep_matches_reference_digest_512x1024(which):
    ny, nx = (int(v) for v in BLUR["big_shape"])
    dig = json.loads(str(BLUR["big_digests"]))
    a = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    got = oracle.box_sweep(a, ny, nx, 8, 8, 0, 0, AXIS[which], 8)
    assert hashlib.sha256(got.tobytes()).hexdigest() == dig[which]


def test_reference_scalar_and_sse_agree_in_golden():
    assert bits_equal(BLUR["x_scalar"], BLUR["x_sse"])
    assert bits_equal(BLUR["y_scalar"], BLUR["y_sse_ipl"])


def test_index_math_matches_reference_golden():
    for c in INDEX:
        st, total = oracle.grid_stride_total(c["dim"], c["lo"], c["hi"])
        assert list(st) == c["stride"] and total == c["total"]
        assert oracle.grid_linearize(c["dim"], c["lo"], c["hi"], c["idx"], False) == c["unsafe"]
        assert oracle.grid_linearize(c["dim"], c["lo"], c["hi"], c["idx"], True) == c["safe"]


@pytest.mark.skipif(oracle.ref() is None, reason="oracle/_ref not built (reference tree absent)")
@pytest.mark.parametrize("shape", [(64, 64), (192, 320)])
def test_box_sweep_matches_live_reference(shape):
    ny, nx = shape
    a = oracle.fill((ny + 16) * (nx + 16), 11, np.float32)
    for which, axis in AXIS.items():
        assert bits_equal(oracle.box_sweep(a, ny, nx, 8, 8, 0, 0, axis, 8),
                          oracle.ref_blur(which, a, ny, nx))


@pytest.mark.skipif(oracle.ref() is None, reason="oracle/_ref not built (reference tree absent)")
def test_index_math_matches_live_reference():
    rng = np.random.default_rng(5)
    for dim, lo, hi in [([9], 2, 2), ([6, 7], 8, 8), ([5, 4], 1, 3), ([3, 4, 5], 1, 1), ([4, 2, 3], 0, 2)]:
        for _ in range(50):
            idx = [int(rng.integers(-lo - 2, d + hi + 2)) for d in dim]
            rc, st, total, off = oracle.ref_index(dim, lo, hi, idx, 1)
            assert rc in (0, 1)
            want = None if rc else off
            assert oracle.grid_linearize(dim, lo, hi, idx, True) == want
            st2, tot2 = oracle.grid_stride_total(dim, lo, hi)
            assert list(st2) == list(st) and tot2 == total
