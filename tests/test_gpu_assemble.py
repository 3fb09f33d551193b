"""Device CSR assembly (hpcx_csr_from_coo) and the device generators against the CPU builder
(oracle/hpcx_oracle.c: orc_rmat_csr, orc_fill_*): structure and merged values bit-exact."""
import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from libhpc_b200 import synth
from util import assert_spmv_close, bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = hp.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("scale,ef,idx_bytes", [(10, 16, 4), (14, 16, 8), (16, 8, 4), (6, 64, 4)])
def test_rmat_device_matches_oracle(ctx, scale, ef, idx_bytes):
    rp, col, val = oracle.rmat(scale, ef, 24)
    D = hp.synth_rmat_csr(ctx, scale, ef, 24, idx_bytes)
    rp_d, col_d, val_d = D.download()
    assert D.nnz == len(col)
    assert rp_d.dtype.itemsize == idx_bytes
    assert np.array_equal(rp_d.astype(np.int64), rp)
    assert np.array_equal(col_d, col)
    assert bits_equal(val_d, val)
    # straight into a device matrix, no host round trip
    A = hp.DeviceCsr.from_device_arrays(ctx, D)
    x = synth.field(42, 1 << scale, np.float32)
    y = np.empty(1 << scale, np.float32)
    A.spmv(x, y)
    assert_spmv_close(y, oracle.spmv(rp, col, val, x), oracle.spmv_abs(rp, col, val, x), 1e-5)
    A.destroy()
    D.destroy()


def _cpu_assemble(nrows, rows, cols, vals):
    """Plain restatement of the assembly rule: sort by (row, col, k), add runs in k order."""
    order = np.lexsort((np.arange(len(rows)), cols, rows))
    r, c, v = rows[order], cols[order], vals[order]
    out_r, out_c, out_v = [], [], []
    for i in range(len(r)):
        if out_r and out_r[-1] == r[i] and out_c[-1] == c[i]:
            out_v[-1] = vals.dtype.type(out_v[-1] + v[i])
        else:
            out_r.append(r[i]); out_c.append(c[i]); out_v.append(v[i])
    rp = np.zeros(nrows + 1, np.int64)
    np.cumsum(np.bincount(np.array(out_r, np.int64), minlength=nrows), out=rp[1:])
    return rp, np.array(out_c, np.int32), np.array(out_v, vals.dtype)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("nrows,ne,ncolmax", [(1, 1, 1), (37, 500, 11), (1000, 20000, 50), (5, 0, 5),
                                              (300, 4000, 2 ** 31 - 1)])
def test_from_coo_host_lists(ctx, dtype, nrows, ne, ncolmax):
    rng = np.random.default_rng(nrows * 7 + ne)
    rows = rng.integers(0, nrows, ne).astype(np.int32)
    cols = rng.integers(0, ncolmax, ne).astype(np.int32)
    if ne > 10:  # make sure some rows are empty and some runs are long
        rows[rows == nrows // 2] = 0
        cols[: ne // 4] = cols[0]
    vals = (rng.standard_normal(ne) * 10.0 ** rng.integers(-6, 6, ne)).astype(dtype)
    D = hp.csr_from_coo(ctx, nrows, rows, cols, vals, idx_bytes=8)
    rp_d, col_d, val_d = D.download()
    rp, col, val = _cpu_assemble(nrows, rows, cols, vals)
    assert np.array_equal(rp_d, rp) and np.array_equal(col_d, col) and bits_equal(val_d, val)
    D.destroy()


def test_from_coo_rejects_bad_arguments(ctx):
    z = np.zeros(1, np.int32)
    with pytest.raises(hp.HpcxError):
        hp.csr_from_coo(ctx, 4, z, z, np.zeros(1), idx_bytes=3)
    with pytest.raises(hp.HpcxError):
        hp.csr_from_coo(ctx, -1, z, z, np.zeros(1))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_device_field_matches_oracle(ctx, dtype):
    import torch
    n, first = 100003, 12345
    out = torch.empty(n, dtype=torch.float32 if dtype == np.float32 else torch.float64, device="cuda")
    hp.synth_field(ctx, 7, n, out, dtype, first)
    want = oracle.fill(first + n, 7, dtype)[first:]
    assert bits_equal(out.cpu().numpy(), want)


# ---- sparse-grid bridge (hpcx_csr_from_cells)
def _cells(seed, n, span, dup):
    rng = np.random.default_rng(seed)
    xs = rng.integers(-span, span, n).astype(np.int32)
    ys = rng.integers(-span, span, n).astype(np.int32)
    if dup and n:
        xs[: n // 3] = xs[n // 3: 2 * (n // 3)][: n // 3]
        ys[: n // 3] = ys[n // 3: 2 * (n // 3)][: n // 3]
    return xs, ys


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,span,dup,idx_bytes", [(0, 1, False, 4), (1, 1, False, 8), (500, 12, True, 4),
                                                  (20000, 90, True, 8), (3000, 2 ** 30, False, 4)])
def test_cells_bridge_matches_oracle(ctx, dtype, n, span, dup, idx_bytes):
    xs, ys = _cells(n + span, n, span, dup)
    if n > 5:  # the int32 corners too
        xs[:2], ys[:2] = (-2 ** 31, 2 ** 31 - 1), (-2 ** 31, 2 ** 31 - 1)
    rp, col, val, cell_row, rx, ry = oracle.cells_csr(xs, ys)
    D, m = hp.csr_from_cells(ctx, xs, ys, dtype, idx_bytes)
    rp_d, col_d, val_d = D.download()
    assert m["nrows"] == len(rx) and D.nnz == len(col)
    assert np.array_equal(rp_d.astype(np.int64), rp) and np.array_equal(col_d, col)
    assert bits_equal(val_d, val.astype(dtype))
    assert np.array_equal(m["cell_row"], cell_row) and np.array_equal(m["row_x"], rx) and np.array_equal(m["row_y"], ry)
    if len(rx):
        A = hp.DeviceCsr.from_device_arrays(ctx, D)
        x = synth.field(42, len(rx), dtype)
        y = np.empty(len(rx), dtype)
        A.spmv(x, y)
        assert np.array_equal(y, oracle.spmv(rp, col, val.astype(dtype), x))  # rows of <= 5 entries: bit-exact
        A.destroy()
    D.destroy()


def test_cells_bridge_full_rectangle_is_the_laplacian(ctx):
    nx, ny = 300, 200
    xs, ys = np.meshgrid(np.arange(nx, dtype=np.int32), np.arange(ny, dtype=np.int32), indexing="ij")
    D, m = hp.csr_from_cells(ctx, np.ascontiguousarray(xs.ravel()), np.ascontiguousarray(ys.ravel()))
    rp, col, val = D.download()
    rp2, col2, val2 = synth.laplace2d(nx, ny)
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2) and np.array_equal(val, val2)
    D.destroy()
