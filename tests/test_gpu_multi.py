"""Single-process multi-GPU driver (hpcx_init_multi): one host thread, every visible
GPU, peer stores between them.  On a one-GPU box the same code runs with one device (the split,
upload / download and sweep paths are all exercised; the peer transport needs >= 2 GPUs and is
covered when they are there).  Results are compared with the oracle on the GLOBAL problem:
Jacobi and box sweeps bit-exact, SpMV within tolerance (bit-exact for short rows)."""
import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from libhpc_b200 import synth
from util import assert_spmv_close, bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mctx():
    m = hp.MultiContext()  # every visible GPU
    yield m
    m.close()


def _device_sets():
    import torch
    n = torch.cuda.device_count()
    return [list(range(k)) for k in (1, 2, 4, 8) if k <= n]


def test_multi_context_basics(mctx):
    assert mctx.size >= 1
    with pytest.raises(hp.HpcxError):
        hp.MultiContext([0, 0])  # a device listed twice


@pytest.mark.parametrize("name", ["op27", "laplace", "rmat_by_nnz"])
def test_multi_spmv_vs_global_oracle(mctx, name):
    if name == "op27":
        rp, col, val = synth.op27(20, index_dtype=np.int64)
        mode = 0
    elif name == "laplace":
        rp, col, val = synth.laplace2d(300, 257)
        mode = 0
    else:
        rp, col, val = synth.rmat(13, 16, 24)
        val = val.astype(np.float64)
        mode = 1
    n = len(rp) - 1
    A = hp.CsrMatrix(rp, col, val)
    M = hp.MultiCsr(mctx, A, mode)
    try:
        b = M.bounds()
        assert b[0] == 0 and b[-1] == n and (np.diff(b) >= 0).all()
        assert np.array_equal(b, oracle.partition_rows(rp, n, mctx.size, mode))
        for seed in (42, 43, 44):  # the halo slots must carry each product's own x
            x = synth.field(seed, n)
            y = M.spmv(x)
            y_ref = oracle.spmv(rp, col, val, x)
            assert_spmv_close(y, y_ref, oracle.spmv_abs(rp, col, val, x), 1e-12)
            if name != "rmat_by_nnz":
                assert bits_equal(y, y_ref)  # short rows: the oracle's summation order
    finally:
        M.destroy()


@pytest.mark.parametrize("variant,gw,steps", [(hp.VARIANT_TMA, 1, 5), (hp.VARIANT_TMA, 1, 2), (hp.VARIANT_TEMPORAL, 2, 6),
                                              (hp.VARIANT_TEMPORAL, 2, 5), (hp.VARIANT_TMA, 1, 131),
                                              (hp.VARIANT_TEMPORAL, 2, 70)])  # > 64: chunked launches
def test_multi_jacobi_bit_exact(mctx, variant, gw, steps):
    nz, ny, nx = 16 * mctx.size + 3, 40, 260
    ext = (nz + 2 * gw, ny + 2 * gw, nx + 2 * gw)
    full = oracle.fill(ext[0] * ext[1] * ext[2], 7, np.float32)
    want = oracle.jacobi7(full, nz, ny, nx, gw, gw, steps)
    a, b = hp.MultiGrid(mctx, (nz, ny, nx), np.float32, gw), hp.MultiGrid(mctx, (nz, ny, nx), np.float32, gw)
    try:
        a.upload(full)
        b.upload(full)
        hp.multi_stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=steps, variant=variant)
        out = b.download(full.copy())
        assert bits_equal(out, want)
        # a second call continues from `in` (iterate steps-1 or steps-2): in -> out semantics
        hp.multi_stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=2, variant=variant)
        prev = steps - 1 if (variant != hp.VARIANT_TEMPORAL or steps % 2) else steps - 2
        assert bits_equal(b.download(full.copy()), oracle.jacobi7(full, nz, ny, nx, gw, gw, prev + 2))
    finally:
        a.destroy()
        b.destroy()


@pytest.mark.parametrize("kind,axis", [(hp.STENCIL_BOX_X, 1), (hp.STENCIL_BOX_Y, 0)])
def test_multi_box_sweeps_match_reference_loops(mctx, kind, axis):
    """R0 across devices: y-slabs of the reference's 17-point sweeps, ghost rows from the host array."""
    ny, nx = 64 * mctx.size + 24, 512
    a_flat = oracle.fill((ny + 16) * (nx + 16), 3, np.float32)
    a, b = hp.MultiGrid(mctx, (ny, nx), np.float32, 8), hp.MultiGrid(mctx, (ny, nx), np.float32, 0)
    try:
        a.upload(a_flat)
        hp.multi_stencil_sweep(a, b, kind, radius=8)
        out = b.download(np.zeros(ny * nx, dtype=np.float32))
        assert bits_equal(out, oracle.box_sweep(a_flat, ny, nx, 8, 8, 0, 0, axis, 8))
        if oracle.ref() is not None:
            assert bits_equal(out, oracle.ref_blur("x_sse" if kind == hp.STENCIL_BOX_X else "y_sse_ipl", a_flat, ny, nx))
    finally:
        a.destroy()
        b.destroy()


def test_multi_on_every_device_count():
    """The same problem on 1, 2, 4, 8 devices (as many as the box has): identical bits."""
    nz, ny, nx, steps = 48, 24, 136, 4
    full = oracle.fill((nz + 2) * (ny + 2) * (nx + 2), 9, np.float32)
    want = oracle.jacobi7(full, nz, ny, nx, 1, 1, steps)
    rp, col, val = synth.laplace2d(96, 64)
    x = synth.field(5, 96 * 64)
    y_ref = oracle.spmv(rp, col, val, x)
    for devs in _device_sets():
        with hp.MultiContext(devs) as m:
            a, b = hp.MultiGrid(m, (nz, ny, nx), np.float32, 1), hp.MultiGrid(m, (nz, ny, nx), np.float32, 1)
            a.upload(full)
            b.upload(full)
            hp.multi_stencil_sweep(a, b, hp.STENCIL_JACOBI7, steps=steps, variant=hp.VARIANT_TMA)
            assert bits_equal(b.download(full.copy()), want), devs
            a.destroy()
            b.destroy()
            M = hp.MultiCsr(m, hp.CsrMatrix(rp, col, val))
            assert bits_equal(M.spmv(x), y_ref), devs
            M.destroy()
