"""Analytic known-answer tests for the parts of the oracle the reference cannot pin
(CSR builders, SpMV, Jacobi, SELL, partitioner)."""
import numpy as np
import pytest

import oracle
from libhpc_b200 import synth
from util import random_csr


@pytest.mark.parametrize("ny,nx", [(1, 1), (1, 9), (7, 1), (5, 8), (33, 17)])
def test_laplace2d_closed_forms(ny, nx):
    rp, col, val = oracle.laplace2d(ny, nx)
    n = ny * nx
    assert rp[0] == 0 and rp[-1] == len(col) == 5 * n - 2 * nx - 2 * ny
    for i in range(n):
        c = col[rp[i]:rp[i + 1]]
        assert (np.diff(c) > 0).all() and i in c
    # A * 1 = number of missing neighbours (0 interior, 1 edge, 2 corner)
    y = oracle.spmv(rp, col, val, np.ones(n))
    yy, xx = np.divmod(np.arange(n), nx)
    missing = (yy == 0).astype(float) + (yy == ny - 1) + (xx == 0) + (xx == nx - 1)
    assert np.array_equal(y, missing)
    # A * (linear field) = 0 in the interior
    lin = 3.0 * xx + 2.0 * yy + 1.0
    y = oracle.spmv(rp, col, val, lin)
    inner = (yy > 0) & (yy < ny - 1) & (xx > 0) & (xx < nx - 1)
    assert np.array_equal(y[inner], np.zeros(inner.sum()))


@pytest.mark.parametrize("n", [1, 2, 3, 6])
def test_op27_closed_forms(n):
    rp, col, val = oracle.op27(n)
    N = n ** 3
    assert rp[-1] == len(col) == (3 * n - 2) ** 3
    y = oracle.spmv(rp, col, val, np.ones(N))
    nbrs = np.diff(rp) - 1
    assert np.array_equal(y, 26.0 - nbrs)
    if n >= 3:
        z, r = np.divmod(np.arange(N), n * n)
        yy, xx = np.divmod(r, n)
        inner = (z > 0) & (z < n - 1) & (yy > 0) & (yy < n - 1) & (xx > 0) & (xx < n - 1)
        lin = 1.0 * xx - 2.0 * yy + 5.0 * z
        assert np.array_equal(oracle.spmv(rp, col, val, lin)[inner], np.zeros(inner.sum()))


def test_spmv_identity_and_permutation():
    n = 257
    x = oracle.fill(n, 9, np.float64)
    rp = np.arange(n + 1, dtype=np.int64)
    assert np.array_equal(oracle.spmv(rp, np.arange(n, dtype=np.int32), np.ones(n), x), x)
    perm = np.random.default_rng(1).permutation(n).astype(np.int32)
    assert np.array_equal(oracle.spmv(rp, perm, np.ones(n), x), x[perm])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_spmv_matches_sequential_python(dtype):
    rp, col, val = random_csr(40, 50, [0, 1, 3, 50, 7] * 8, 3, dtype)
    x = oracle.fill(50, 4, dtype)
    y = oracle.spmv(rp, col, val, x)
    for i in range(40):
        s = dtype(0)
        for p in range(rp[i], rp[i + 1]):
            s = dtype(s + dtype(val[p] * x[col[p]]))
        assert y[i] == s
    assert np.array_equal(oracle.spmv(rp.astype(np.int32), col, val, x), y)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_jacobi_linear_field_is_a_fixed_point(dtype):
    nz, ny, nx, lo = 5, 6, 7, 1
    z, y, x = np.meshgrid(np.arange(-lo, nz + lo), np.arange(-lo, ny + lo), np.arange(-lo, nx + lo),
                          indexing="ij")
    f = (2.0 * x + 4.0 * y - 6.0 * z + 12.0).astype(dtype).ravel()  # small integers: exact sums
    out = oracle.jacobi7(f, nz, ny, nx, lo, lo, steps=3)
    # (sum of 6 neighbours) = 6 f exactly; times fl(1/6) may round: allow 1 ulp per step
    assert np.allclose(out, f, rtol=4 * np.finfo(dtype).eps, atol=0)


def test_jacobi_single_cell_by_hand():
    f = np.zeros(27, dtype=np.float32)
    f[[4, 10, 12, 14, 16, 22]] = [1, 2, 3, 4, 5, 6]  # z-1, y-1, x-1, x+1, y+1, z+1 of the centre
    out = oracle.jacobi7(f, 1, 1, 1, 1, 1)
    want = np.float32(np.float32(1 + 2 + 3 + 4 + 5 + 6) * (np.float32(1) / np.float32(6)))
    assert out[13] == want and np.array_equal(np.delete(out, 13), np.delete(f, 13))


def test_rmat_reproducible_and_sorted():
    rp, col, val = oracle.rmat(10, 8, 24)
    rp2, col2, val2 = oracle.rmat(10, 8, 24)
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2) and np.array_equal(val, val2)
    assert rp[-1] == len(col) <= 8 << 10
    for i in range(1 << 10):
        assert (np.diff(col[rp[i]:rp[i + 1]]) > 0).all()
    # skew: the heaviest row is far above the mean
    assert np.diff(rp).max() > 8 * np.diff(rp).mean()


def test_sell_structure():
    lens = [3, 0, 9, 1, 1, 7, 2, 2, 5, 4, 0, 6, 1]
    rp, col, val = random_csr(len(lens), 20, lens, 7, np.float64)
    perm, sptr, slen, scol, sval = oracle.sell_build(rp, col, val, 20, c=4, sigma=8)
    assert sorted(perm) == list(range(len(lens)))
    assert list(perm[:8]) == [2, 5, 0, 6, 7, 3, 4, 1]  # desc length, ties by row id
    assert list(slen) == [9, 2, 6, 0] and sptr[-1] == 4 * (9 + 2 + 6 + 0)
    x = oracle.fill(20, 8, np.float64)
    y = np.zeros(len(lens))
    for s in range(len(slen)):
        for lane in range(4):
            j = s * 4 + lane
            if j < len(lens):
                acc = 0.0
                for k in range(slen[s]):
                    i = sptr[s] + k * 4 + lane
                    acc = acc + sval[i] * x[scol[i]]
                y[perm[j]] = acc
    assert np.array_equal(y, oracle.spmv(rp, col, val, x))  # zero padding adds exactly 0


def test_partition_rows_modes():
    assert list(oracle.partition_rows(None, 10, 4, 0)) == [0, 3, 6, 9, 10]
    assert list(oracle.partition_rows(None, 3, 8, 0)) == [0, 1, 2, 3, 3, 3, 3, 3, 3]
    rp = np.array([0, 10, 10, 11, 12, 40, 41], dtype=np.int64)
    b = oracle.partition_rows(rp, 6, 2, 1)
    assert list(b) == [0, 5, 6] or list(b) == [0, 4, 6]
    assert list(b) == [0, 5, 6]  # first row with row_ptr >= 20 is row 5 (row_ptr[5] = 40)


def test_localize_reassembles_global_product():
    rp, col, val = oracle.op27(6)
    n = 6 ** 3
    x = oracle.fill(n, 2, np.float64)
    y = oracle.spmv(rp, col, val, x)
    bounds = oracle.partition_rows(rp, n, 4, 1)
    for r in range(4):
        L = oracle.localize(rp, col, bounds, r)
        xl = np.concatenate([x[bounds[r]:bounds[r + 1]], x[L["halo_gcol"]]])
        p0 = rp[bounds[r]]
        yl = oracle.spmv(L["local_row_ptr"], L["local_col"], val[p0:p0 + L["nnz_loc"]], xl)
        assert np.array_equal(yl, y[bounds[r]:bounds[r + 1]])
        assert L["recv_counts"].sum() == L["nhalo"] and L["send_counts"].sum() == L["nsend"]
        assert L["recv_counts"][r] == 0 and L["send_counts"][r] == 0


# ---- CG step: analytic checks of the CPU restatement
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 2e-4)])
def test_cg_oracle_recovers_known_solution(dtype, tol):
    rp, col, val = synth.laplace2d(24, 24, dtype, np.int64)
    xs = synth.field(5, 24 * 24, dtype)
    b = oracle.spmv(rp, col, val, xs)
    x, it, rr, bb = oracle.cg(rp, col, val, b, np.zeros_like(b), 500, tol / 10)
    assert 0 < it < 500 and rr <= (tol / 10) ** 2 * bb
    assert np.abs(x - xs).max() <= 50 * tol * np.abs(xs).max()


def test_cg_oracle_identity_and_zero_iterations():
    n = 50
    rp = np.arange(n + 1, dtype=np.int64)
    col = np.arange(n, dtype=np.int32)
    val = np.full(n, 2.0)
    b = synth.field(3, n)
    x, it, rr, bb = oracle.cg(rp, col, val, b, np.zeros(n), 10, 1e-14)
    assert it == 1 and np.allclose(x, b / 2, rtol=1e-15) and rr == 0.0
    x0 = synth.field(4, n)
    x, it, _, _ = oracle.cg(rp, col, val, b, x0, 0)
    assert it == 0 and np.array_equal(x, x0)


# ---- sparse-grid bridge
def test_cells_operator_on_a_full_rectangle_is_the_laplacian():
    nx, ny = 7, 5  # rows are x-major: row = x*ny + y, i.e. laplace2d with x as the slow index
    xs, ys = np.meshgrid(np.arange(nx, dtype=np.int32), np.arange(ny, dtype=np.int32), indexing="ij")
    perm = np.random.default_rng(1).permutation(nx * ny)
    rp, col, val, cell_row, rx, ry = oracle.cells_csr(xs.ravel()[perm], ys.ravel()[perm])
    rp2, col2, val2 = synth.laplace2d(nx, ny, index_dtype=np.int64)
    assert np.array_equal(rp, rp2) and np.array_equal(col, col2) and np.array_equal(val, val2)
    assert np.array_equal(cell_row, (xs.ravel() * ny + ys.ravel())[perm])
    assert np.array_equal(rx * ny + ry, np.arange(nx * ny))


def test_cells_operator_handles_repeats_islands_and_negative_coordinates():
    xs = np.array([3, -2, 3, 100, -2, -2147483648, 2147483647], dtype=np.int32)
    ys = np.array([4, 7, 4, -100, 8, 0, 2147483647], dtype=np.int32)
    rp, col, val, cell_row, rx, ry = oracle.cells_csr(xs, ys)
    assert list(zip(rx, ry)) == [(-2147483648, 0), (-2, 7), (-2, 8), (3, 4), (100, -100), (2147483647, 2147483647)]
    assert list(cell_row) == [3, 1, 3, 4, 2, 0, 5]
    assert list(rp) == [0, 1, 3, 5, 6, 7, 8]
    assert list(col) == [0, 1, 2, 1, 2, 3, 4, 5] and list(val) == [4, 4, -1, -1, 4, 4, 4, 4]
    rp, col, val, cell_row, rx, ry = oracle.cells_csr(np.empty(0, np.int32), np.empty(0, np.int32))
    assert list(rp) == [0] and len(col) == 0
