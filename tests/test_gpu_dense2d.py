"""conv2d and transpose sweeps through the C ABI vs the oracle and the reference golden vectors
(tests/golden/dense2d_ref.npz: outputs of the reference's BM_conv / BM_conv_block / BM_transpose
loops).  Both are BIT-EXACT: the convolution adds its 256 separately rounded products in the
reference's (l, k) order (library built with --fmad=false); the transpose only moves data."""
import hashlib
import json
import os

import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from util import bits_equal

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "dense2d_ref.npz"))
DIG = json.loads(str(G["digests"]))


def _conv(ctx, kb, ka, w, ny, nx, ilo, ihi, olo=0, ohi=0, dtype=np.float32, variant=hp.VARIANT_AUTO, reps=1):
    gi = hp.DeviceGrid(ctx, (ny, nx), dtype, ilo, ihi)
    go = hp.DeviceGrid(ctx, (ny, nx), dtype, olo, ohi)
    gi.upload(np.ascontiguousarray(kb, dtype=dtype))
    go.upload(np.ascontiguousarray(ka, dtype=dtype))
    for _ in range(reps):
        hp.conv2d(gi, go, w, variant)
    out = go.download()
    gi.destroy()
    go.destroy()
    return out


def _inputs(n):
    return (oracle.fill(n * n, 31, np.float32), oracle.fill((n + 32) ** 2, 32, np.float32),
            oracle.fill(256, 33, np.float32).reshape(16, 16))


@pytest.mark.parametrize("variant", [hp.VARIANT_AUTO, hp.VARIANT_DIRECT])
def test_conv_matches_reference_golden(ctx, variant):
    n = int(G["n"])
    ka, kb, kc = _inputs(n)
    assert bits_equal(_conv(ctx, kb, ka, kc, n, n, 16, 16, variant=variant), G["conv_scalar"])


def test_conv_matches_reference_digest_256(ctx):
    n = int(G["n_big"])
    ka, kb, kc = _inputs(n)
    got = _conv(ctx, kb, ka, kc, n, n, 16, 16)
    assert hashlib.sha256(got.tobytes()).hexdigest() == DIG["conv_scalar"]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape,kshape,ilo,ihi,olo,ohi", [
    ((1, 1), (16, 16), 0, 15, 0, 0), ((70, 45), (16, 16), 16, 16, 0, 0), ((64, 32), (16, 16), 3, 15, 2, 1),
    ((129, 200), (16, 16), 0, 17, 0, 3), ((33, 65), (3, 4), 1, 3, 0, 0), ((20, 31), (1, 1), 0, 0, 1, 1),
    ((40, 40), (5, 16), 2, 15, 0, 0), ((17, 300), (32, 32), 0, 31, 0, 0)])
def test_conv_ragged_shapes_and_kernels(ctx, dtype, shape, kshape, ilo, ihi, olo, ohi):
    ny, nx = shape
    kb = oracle.fill((ny + ilo + ihi) * (nx + ilo + ihi), 21, dtype)
    ka = oracle.fill((ny + olo + ohi) * (nx + olo + ohi), 22, dtype)
    w = oracle.fill(kshape[0] * kshape[1], 23, dtype).reshape(kshape)
    want = oracle.conv2d(kb, ka, w, ny, nx, ilo, ihi, olo, ohi)
    assert bits_equal(_conv(ctx, kb, ka, w, ny, nx, ilo, ihi, olo, ohi, dtype), want)
    assert bits_equal(_conv(ctx, kb, ka, w, ny, nx, ilo, ihi, olo, ohi, dtype, hp.VARIANT_DIRECT), want)


def test_conv_accumulates_across_calls_and_kernel_changes(ctx):
    n = 48
    ka, kb, kc = oracle.fill(n * n, 1, np.float32), oracle.fill((n + 32) ** 2, 2, np.float32), \
        oracle.fill(256, 3, np.float32).reshape(16, 16)
    want = oracle.conv2d(kb, oracle.conv2d(kb, ka, kc, n, n, 16, 16), kc, n, n, 16, 16)
    assert bits_equal(_conv(ctx, kb, ka, kc, n, n, 16, 16, reps=2), want)
    # a different kernel right after: the cached constant-memory weights must be replaced
    kc2 = oracle.fill(256, 4, np.float32).reshape(16, 16)
    assert bits_equal(_conv(ctx, kb, ka, kc2, n, n, 16, 16), oracle.conv2d(kb, ka, kc2, n, n, 16, 16))
    # fp64 with other weights in between, then the fp32 kernel again (one constant symbol serves both)
    kb64, ka64, kc64 = kb.astype(np.float64), ka.astype(np.float64), oracle.fill(256, 5, np.float64).reshape(16, 16)
    assert bits_equal(_conv(ctx, kb64, ka64, kc64, n, n, 16, 16, dtype=np.float64),
                      oracle.conv2d(kb64, ka64, kc64, n, n, 16, 16))
    assert bits_equal(_conv(ctx, kb, ka, kc2, n, n, 16, 16), oracle.conv2d(kb, ka, kc2, n, n, 16, 16))


def test_conv_full_reference_size_1024(ctx):
    n = 1024  # conv_n of the reference (test_hpc_benchmark.cpp:787)
    ka, kb, kc = _inputs(n)
    assert bits_equal(_conv(ctx, kb, ka, kc, n, n, 16, 16), oracle.conv2d(kb, ka, kc, n, n, 16, 16))


def test_conv_rejects_thin_ghosts_and_shape_mismatch(ctx):
    gi = hp.DeviceGrid(ctx, (8, 8), np.float32, 4, 4)
    go = hp.DeviceGrid(ctx, (8, 8), np.float32, 0, 0)
    g3 = hp.DeviceGrid(ctx, (8, 9), np.float32, 0, 0)
    with pytest.raises(hp.HpcxError):
        hp.conv2d(gi, go, np.ones((16, 16), np.float32))
    with pytest.raises(hp.HpcxError):
        hp.conv2d(gi, g3, np.ones((2, 2), np.float32))
    for g in (gi, go, g3):
        g.destroy()


def _transpose(ctx, a, ny, nx, ilo, ihi, olo, ohi, dtype):
    gi = hp.DeviceGrid(ctx, (ny, nx), dtype, ilo, ihi)
    go = hp.DeviceGrid(ctx, (nx, ny), dtype, olo, ohi)
    gi.upload(a)
    hp.grid_transpose(gi, go)
    out = go.download()
    gi.destroy()
    go.destroy()
    return out


def test_transpose_matches_reference_digest(ctx):
    n = int(G["n_transpose"])
    a = oracle.fill(n * n, 34, np.float32)
    got = _transpose(ctx, a, n, n, 0, 0, 0, 0, np.float32)
    assert hashlib.sha256(got.tobytes()).hexdigest() == DIG["transpose_scalar"]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape,ilo,ihi,olo,ohi", [((1, 1), 0, 0, 0, 0), ((5, 9), 1, 2, 2, 0), ((64, 64), 0, 0, 0, 0),
                                                  ((65, 130), 8, 8, 0, 0), ((200, 67), 0, 0, 3, 1),
                                                  ((256, 1024), 0, 0, 0, 0), ((1000, 31), 2, 2, 2, 2)])
def test_transpose_ragged_shapes(ctx, dtype, shape, ilo, ihi, olo, ohi):
    ny, nx = shape
    a = oracle.fill((ny + ilo + ihi) * (nx + ilo + ihi), 24, dtype)
    assert bits_equal(_transpose(ctx, a, ny, nx, ilo, ihi, olo, ohi, dtype),
                      oracle.transpose(a, ny, nx, ilo, ihi, olo, ohi))


def test_transpose_8192_is_an_involution(ctx):
    n = 8192  # nx, ny of the reference (test_hpc_benchmark.cpp:30-31)
    a = oracle.fill(n * n, 35, np.float32)
    ga, gb, gc = (hp.DeviceGrid(ctx, (n, n), np.float32, 0, 0) for _ in range(3))
    ga.upload(a)
    hp.grid_transpose(ga, gb)
    hp.grid_transpose(gb, gc)
    assert bits_equal(gc.download(), a)
    assert np.array_equal(gb.download().reshape(n, n)[:64], a.reshape(n, n)[:, :64].T)
    for g in (ga, gb, gc):
        g.destroy()
