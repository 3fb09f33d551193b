"""Pins the conv2d / transpose restatements (oracle/hpcx_oracle.c: orc_conv2d_*, orc_transpose)
against golden vectors from the reference's own BM_conv / BM_conv_block / BM_transpose loops
(tests/golden/make_golden_dense2d.py) and, where oracle/_ref was built, the live reference."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from util import bits_equal

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "dense2d_ref.npz"))
DIG = json.loads(str(G["digests"]))


def conv_inputs(n):
    return (oracle.fill(n * n, 31, np.float32), oracle.fill((n + 32) ** 2, 32, np.float32),
            oracle.fill(256, 33, np.float32).reshape(16, 16))


def test_conv_matches_reference_golden():
    n = int(G["n"])
    ka, kb, kc = conv_inputs(n)
    got = oracle.conv2d(kb, ka, kc, n, n, 16, 16)
    assert bits_equal(got, G["conv_scalar"]) and bits_equal(got, G["conv_block"])


def test_conv_matches_reference_digest_256():
    n = int(G["n_big"])
    ka, kb, kc = conv_inputs(n)
    got = oracle.conv2d(kb, ka, kc, n, n, 16, 16)
    assert hashlib.sha256(got.tobytes()).hexdigest() == DIG["conv_scalar"] == DIG["conv_block"]


def test_transpose_matches_reference_digest():
    n = int(G["n_transpose"])
    a = oracle.fill(n * n, 34, np.float32)
    got = oracle.transpose(a, n, n)
    assert hashlib.sha256(got.tobytes()).hexdigest() == DIG["transpose_scalar"] == DIG["transpose_tiling"]
    assert np.array_equal(got.reshape(n, n), a.reshape(n, n).T)


def test_conv_known_answers():
    # a single tap at (l, k) shifts the field; a ones kernel on a ones field counts the taps
    n = 12
    kb = oracle.fill((n + 6) ** 2, 5, np.float64)
    w = np.zeros((3, 4))
    w[2, 1] = 1.0
    got = oracle.conv2d(kb, np.zeros(n * n), w, n, n, 3, 3).reshape(n, n)
    assert np.array_equal(got, kb.reshape(n + 6, n + 6)[3 + 2:3 + 2 + n, 3 + 1:3 + 1 + n])
    got = oracle.conv2d(np.ones((n + 6) ** 2), np.full(n * n, 0.5), np.ones((3, 4)), n, n, 3, 3)
    assert np.array_equal(got, np.full(n * n, 12.5))
    # ghost widths of the output are honoured and its ghosts untouched
    out0 = oracle.fill((n + 3) ** 2, 6, np.float64)
    got = oracle.conv2d(kb, out0, w, n, n, 3, 3, 1, 2).reshape(n + 3, n + 3)
    ref = out0.reshape(n + 3, n + 3).copy()
    ref[1:1 + n, 1:1 + n] += kb.reshape(n + 6, n + 6)[5:5 + n, 4:4 + n]
    assert np.array_equal(got, ref)


def test_transpose_rectangular_with_ghosts():
    ny, nx = 5, 9
    a = oracle.fill((ny + 3) * (nx + 3), 8, np.float64)
    got = oracle.transpose(a, ny, nx, 1, 2, 2, 0).reshape(nx + 2, ny + 2)
    assert np.array_equal(got[2:, 2:], a.reshape(ny + 3, nx + 3)[1:1 + ny, 1:1 + nx].T)
    assert not got[:2].any() and not got[:, :2].any()


@pytest.mark.skipif(oracle.ref() is None, reason="oracle/_ref not built (reference tree absent)")
def test_conv_and_transpose_match_live_reference():
    n = 128
    ka, kb, kc = (oracle.fill(n * n, 41, np.float32), oracle.fill((n + 32) ** 2, 42, np.float32),
                  oracle.fill(256, 43, np.float32).reshape(16, 16))
    want = oracle.ref_conv("scalar", ka, kb, kc, n)
    assert bits_equal(want, oracle.ref_conv("block", ka, kb, kc, n))
    assert bits_equal(oracle.conv2d(kb, ka, kc, n, n, 16, 16), want)
    assert bits_equal(oracle.transpose(ka, n, n), oracle.ref_transpose("scalar", ka, n))
    assert bits_equal(oracle.transpose(ka, n, n), oracle.ref_transpose("tiling", ka, n))
