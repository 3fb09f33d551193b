"""CUDA stencil sweeps through the C ABI vs the oracle and the reference golden vectors.
Box sweeps (adds only, fixed order) and Jacobi-7 (library built with --fmad=false, fixed
order) are BIT-EXACT; the stated fp32 tolerance 1e-5 is therefore not needed."""
import hashlib
import json
import os

import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from util import bits_equal

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
BLUR = np.load(os.path.join(G, "blur_ref.npz"))
KIND = {"x_scalar": hp.STENCIL_BOX_X, "x_sse": hp.STENCIL_BOX_X, "y_scalar": hp.STENCIL_BOX_Y,
        "y_sse_ipl": hp.STENCIL_BOX_Y}


def _box(ctx, a_flat, ny, nx, kind, radius=8, alo=8, blo=0, dtype=np.float32, variant=hp.VARIANT_AUTO):
    a = hp.HPCHighDimensionFlatArray(ny, nx, dtype=dtype, lo=alo, buffer=a_flat)
    b = hp.HPCHighDimensionFlatArray(ny, nx, dtype=dtype, lo=blo)
    hp.stencil_sweep_host(ctx, a, b, kind, radius, 1, variant)
    return b.data()


@pytest.mark.parametrize("which", sorted(KIND))
def test_box_sweeps_match_reference_golden(ctx, which):
    ny, nx = (int(v) for v in BLUR["shape"])
    a = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    assert bits_equal(_box(ctx, a, ny, nx, KIND[which]), BLUR[which])


@pytest.mark.parametrize("which", ["x_scalar", "y_scalar"])
def test_box_sweeps_match_reference_digest(ctx, which):
    ny, nx = (int(v) for v in BLUR["big_shape"])
    dig = json.loads(str(BLUR["big_digests"]))
    a = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    assert hashlib.sha256(_box(ctx, a, ny, nx, KIND[which]).tobytes()).hexdigest() == dig[which]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape,radius,alo,blo", [((1, 1), 0, 0, 0), ((5, 7), 2, 2, 0), ((33, 129), 8, 8, 3),
                                                  ((64, 250), 3, 5, 1), ((257, 31), 8, 9, 0),
                                                  ((40, 256), 8, 8, 0), ((100, 64), 4, 4, 2),
                                                  ((31, 36), 8, 9, 1), ((67, 1028), 8, 8, 0),
                                                  ((32, 12), 4, 6, 0),
                                                  # whole warps along the row: the shuffle-window BOX_X kernel
                                                  ((19, 2048), 8, 8, 0), ((7, 1024), 4, 4, 1), ((5, 3072), 8, 10, 0)])
def test_box_sweeps_ragged_shapes(ctx, dtype, shape, radius, alo, blo):
    ny, nx = shape
    a = oracle.fill((ny + 2 * alo) * (nx + 2 * alo), 13, dtype)
    for kind, axis in ((hp.STENCIL_BOX_X, 1), (hp.STENCIL_BOX_Y, 0)):
        want = oracle.box_sweep(a, ny, nx, alo, alo, blo, blo, axis, radius)
        # AUTO takes the register-tiled kernels where they apply, DIRECT the one-thread-per-cell one
        assert bits_equal(_box(ctx, a, ny, nx, kind, radius, alo, blo, dtype), want)
        assert bits_equal(_box(ctx, a, ny, nx, kind, radius, alo, blo, dtype, hp.VARIANT_DIRECT), want)


def test_r0_full_size_8192_matches_oracle(ctx):
    """R0: the reference's own shape, 8192^2 fp32 ghost 8 (the oracle takes ~1 s here)."""
    n = 8192
    a = oracle.fill((n + 16) * (n + 16), 3, np.float32)
    for kind, axis in ((hp.STENCIL_BOX_X, 1), (hp.STENCIL_BOX_Y, 0)):
        assert bits_equal(_box(ctx, a, n, n, kind), oracle.box_sweep(a, n, n, 8, 8, 0, 0, axis, 8))


def _jacobi(ctx, f, dims, lo, hi, steps, variant, zsplit=0):
    a = hp.DeviceGrid(ctx, dims, f.dtype, lo, hi)
    b = hp.DeviceGrid(ctx, dims, f.dtype, lo, hi)
    try:
        a.upload(f)
        b.upload(f)
        hp.stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=steps, variant=variant, zsplit=zsplit)
        return b.download(), a.download()
    finally:
        a.destroy()
        b.destroy()


@pytest.mark.parametrize("variant", [hp.VARIANT_TMA, hp.VARIANT_DIRECT, hp.VARIANT_AUTO])
@pytest.mark.parametrize("dims,lo,hi", [((1, 1, 1), 1, 1), ((3, 5, 7), 1, 1), ((16, 16, 128), 1, 1),
                                        ((17, 33, 129), 1, 1), ((40, 50, 300), 2, 3), ((64, 64, 64), 1, 1),
                                        ((24, 100, 256), 1, 1), ((5, 37, 64), 1, 2)])
@pytest.mark.parametrize("steps", [1, 2, 8])
def test_jacobi7_fp32_bit_exact(ctx, variant, dims, lo, hi, steps):
    nz, ny, nx = dims
    f = oracle.fill((nz + lo + hi) * (ny + lo + hi) * (nx + lo + hi), 7, np.float32)
    newest, prev = _jacobi(ctx, f, dims, lo, hi, steps, variant)
    assert bits_equal(newest, oracle.jacobi7(f, nz, ny, nx, lo, hi, steps))
    assert bits_equal(prev, oracle.jacobi7(f, nz, ny, nx, lo, hi, steps - 1))


@pytest.mark.parametrize("dims,lo,hi", [((1, 1, 1), 1, 1), ((2, 3, 5), 1, 1), ((3, 5, 7), 1, 1), ((16, 14, 128), 1, 1),
                                        ((17, 33, 129), 1, 1), ((40, 50, 300), 2, 3), ((64, 64, 64), 1, 1),
                                        ((24, 100, 256), 1, 1), ((5, 37, 64), 1, 2), ((33, 15, 260), 1, 1)])
@pytest.mark.parametrize("steps", [2, 3, 4, 7])
def test_jacobi7_temporal_blocking_bit_exact(ctx, dims, lo, hi, steps):
    """HPCX_VARIANT_TEMPORAL: two sweeps per pass over HBM, bit-identical to single sweeps; `in`
    ends up holding an older iterate (steps-2, or steps-1 when steps is odd)."""
    nz, ny, nx = dims
    f = oracle.fill((nz + lo + hi) * (ny + lo + hi) * (nx + lo + hi), 7, np.float32)
    newest, older = _jacobi(ctx, f, dims, lo, hi, steps, hp.VARIANT_TEMPORAL)
    assert bits_equal(newest, oracle.jacobi7(f, nz, ny, nx, lo, hi, steps))
    assert bits_equal(older, oracle.jacobi7(f, nz, ny, nx, lo, hi, steps - 1 if steps & 1 else steps - 2))


@pytest.mark.parametrize("zsplit", [1, 2, 5, 37])
def test_jacobi7_temporal_blocking_zsplit(ctx, zsplit):
    dims = (37, 20, 140)
    f = oracle.fill(39 * 22 * 142, 8, np.float32)
    newest, _ = _jacobi(ctx, f, dims, 1, 1, 4, hp.VARIANT_TEMPORAL, zsplit)
    assert bits_equal(newest, oracle.jacobi7(f, *dims, 1, 1, 4))


def test_jacobi7_temporal_blocking_512_cubed_and_fp64_rejected(ctx):
    n = 512
    f = oracle.fill((n + 2) ** 3, 7, np.float32)
    tb, _ = _jacobi(ctx, f, (n, n, n), 1, 1, 8, hp.VARIANT_TEMPORAL)
    tma, _ = _jacobi(ctx, f, (n, n, n), 1, 1, 8, hp.VARIANT_TMA)
    assert bits_equal(tb, tma)
    f64 = oracle.fill(11 * 12 * 72, 9, np.float64)
    with pytest.raises(hp.HpcxError) as e:
        _jacobi(ctx, f64, (9, 10, 70), 1, 1, 2, hp.VARIANT_TEMPORAL)
    assert e.value.status == hp._abi.UNSUPPORTED


@pytest.mark.parametrize("zsplit", [1, 2, 5])
def test_jacobi7_zsplit(ctx, zsplit):
    dims = (37, 20, 140)
    f = oracle.fill(39 * 22 * 142, 8, np.float32)
    for variant in (hp.VARIANT_TMA, hp.VARIANT_DIRECT):
        newest, _ = _jacobi(ctx, f, dims, 1, 1, 3, variant, zsplit)
        assert bits_equal(newest, oracle.jacobi7(f, *dims, 1, 1, 3))


def test_jacobi7_fp64(ctx):
    dims = (9, 10, 70)
    f = oracle.fill(11 * 12 * 72, 9, np.float64)
    newest, _ = _jacobi(ctx, f, dims, 1, 1, 4, hp.VARIANT_AUTO)
    assert bits_equal(newest, oracle.jacobi7(f, *dims, 1, 1, 4))
    with pytest.raises(hp.HpcxError) as e:
        _jacobi(ctx, f, dims, 1, 1, 1, hp.VARIANT_TMA)
    assert e.value.status == hp._abi.UNSUPPORTED


def test_c3_full_size_512_cubed(ctx):
    """configs[2] at full size: T=1 against the oracle (≈1 s of CPU), T=8 by digest of checksums:
    TMA and direct kernels must agree bit-for-bit, and a linear field stays a fixed point."""
    n = 512
    f = oracle.fill((n + 2) ** 3, 7, np.float32)
    tma1, _ = _jacobi(ctx, f, (n, n, n), 1, 1, 1, hp.VARIANT_TMA)
    assert bits_equal(tma1, oracle.jacobi7(f, n, n, n, 1, 1, 1))
    tma8, _ = _jacobi(ctx, f, (n, n, n), 1, 1, 8, hp.VARIANT_TMA)
    dir8, _ = _jacobi(ctx, f, (n, n, n), 1, 1, 8, hp.VARIANT_DIRECT)
    assert bits_equal(tma8, dir8)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("dims,lo,hi", [((6, 7), 2, 2), ((4, 5, 6), 1, 1), ((3, 4, 5), 1, 3), ((9,), 2, 2)])
def test_upload_download_pack_unpack(ctx, dtype, dims, lo, hi):
    ext = [d + lo + hi for d in dims]
    f = oracle.fill(int(np.prod(ext)), 21, dtype)
    g = hp.DeviceGrid(ctx, dims, dtype, lo, hi)
    try:
        assert not g.download().any()  # value-initialised
        g.upload(f)
        assert bits_equal(g.download(), f)  # ghosts included, reference layout round trip
        nd = len(dims)
        n3 = [1] * (3 - nd) + list(dims)
        gl3 = [0] * (3 - nd) + [lo] * nd
        gh3 = [0] * (3 - nd) + [hi] * nd
        for axis in range(nd):
            for side in (0, 1):
                for width in range(1, min(lo if side == 0 else hi, dims[axis]) + 1):
                    want = np.empty(g.face_elems(axis, width), dtype=dtype)
                    oracle.face_copy(f.copy(), want, n3, gl3, gh3, axis + 3 - nd, side, width, False, True)
                    got = g.pack_face(axis, side, width)
                    assert bits_equal(got, want)
                    # unpack into the ghost planes of the same side and compare whole arrays
                    g.unpack_face(axis, side, width, got)
                    ref = f.copy()
                    oracle.face_copy(ref, want, n3, gl3, gh3, axis + 3 - nd, side, width, True, False)
                    assert bits_equal(g.download(), ref)
                    g.upload(f)
    finally:
        g.destroy()


def test_stencil_error_convention(ctx):
    a = hp.DeviceGrid(ctx, (8, 8), np.float32, 2)
    b = hp.DeviceGrid(ctx, (8, 8), np.float32, 0)
    c = hp.DeviceGrid(ctx, (8, 9), np.float32, 0)
    try:
        with pytest.raises(hp.HpcxError) as e:
            hp.stencil_sweep(a, b, hp.STENCIL_BOX_X, radius=3)  # radius > ghost width
        assert e.value.status == hp._abi.INVALID_ARG
        with pytest.raises(hp.HpcxError):
            hp.stencil_sweep(a, c, hp.STENCIL_BOX_X, radius=1)  # shape mismatch
        with pytest.raises(hp.HpcxError):
            hp.stencil_sweep(a, b, hp.STENCIL_JACOBI7)  # 2-D grid
    finally:
        for g in (a, b, c):
            g.destroy()
