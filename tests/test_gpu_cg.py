"""CG step around the SpMV (hpcx_cg_*) against the CPU restatement
(oracle/hpcx_oracle.c: orc_cg_*).  The per-element arithmetic is the oracle's; only the order of
the dot-product reductions differs, so iterates agree to a stated relative tolerance:
1e-10 of max|x| after 12 fp64 iterations, 2e-4 for fp32."""
import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from libhpc_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = hp.Context(0)
    yield c
    c.close()


def arrow(n, dtype):
    """SPD arrow matrix: dense first row/column (a row longer than any tile) + dominant diagonal."""
    rows = np.concatenate([np.zeros(n - 1, np.int64), np.arange(1, n), np.arange(n)])
    cols = np.concatenate([np.arange(1, n), np.zeros(n - 1, np.int64), np.arange(n)])
    # spread diagonal (many distinct eigenvalues, so CG does not finish in two steps)
    vals = np.concatenate([np.full(2 * (n - 1), -0.01), 1.0 + 3.0 * synth.uniform(9, n)])
    vals[-n] = 0.01 * n + 1.0  # diagonal of row 0 dominates its n-1 off-diagonals
    order = np.lexsort((cols, rows))
    rp = np.zeros(n + 1, np.int64)
    np.cumsum(np.bincount(rows, minlength=n), out=rp[1:])
    return rp, cols[order].astype(np.int32), vals[order].astype(dtype)


CASES = {
    "lap2d": lambda dt: synth.laplace2d(40, 56, dt, np.int64),
    "op27": lambda dt: synth.op27(12, dt, np.int64),
    "arrow": lambda dt: arrow(6000, dt),
}


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("tile", [0, 1024])
def test_cg_iterates_match_oracle(ctx, name, dtype, tile):
    rp, col, val = CASES[name](dtype)
    n = len(rp) - 1
    b = synth.field(11, n, dtype)
    x0 = synth.field(12, n, dtype) * dtype(0.1)
    iters = 2 if name == "arrow" else 12  # the arrow matrix amplifies rounding quickly: it checks the long-row dot
    want, it, rr, bb = oracle.cg(rp, col, val, b, x0, iters)
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp, col, val), tile_nnz=tile)
    S = hp.CgSolver(D)
    x = x0.copy()
    res = S.solve(b, x, iters)  # host operands
    assert res["iterations"] == it == iters and not res["converged"]
    tol = 1e-10 if dtype == np.float64 else 2e-4
    if name == "arrow":
        tol *= 10
    assert np.abs(x - want).max() <= tol * np.abs(want).max()
    assert abs(res["rr"] - rr) <= 1e3 * tol * max(rr, 1e-300) and abs(res["bb"] - bb) <= tol * bb
    # the plain product through the same handle is untouched by the solver
    y = np.empty(n, dtype)
    D.spmv(b, y)
    if name != "arrow":  # short rows: bit-exact
        assert np.array_equal(y, oracle.spmv(rp, col, val, b))
    S.destroy()
    D.destroy()


@pytest.mark.parametrize("dtype,rel_tol", [(np.float64, 1e-10), (np.float32, 1e-4)])
def test_cg_converges_and_freezes_x(ctx, dtype, rel_tol):
    import torch
    rp, col, val = synth.laplace2d(48, 48, dtype, np.int64)
    n = len(rp) - 1
    xs = synth.field(5, n, dtype)
    b = oracle.spmv(rp, col, val, xs)
    _, it_ref, _, _ = oracle.cg(rp, col, val, b, np.zeros(n, dtype), 2000, rel_tol)
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp, col, val))
    S = hp.CgSolver(D)
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    bd = torch.from_numpy(b).cuda()
    xd = torch.zeros(n, dtype=tdt, device="cuda")
    torch.cuda.synchronize()
    res = S.solve(bd, xd, it_ref + 40, rel_tol)  # device operands, runs past convergence
    assert res["converged"] and abs(res["iterations"] - it_ref) <= 3
    assert res["rr"] <= rel_tol ** 2 * res["bb"]
    x = xd.cpu().numpy()
    r = b - oracle.spmv(rp, col, val, x)  # true residual
    assert np.linalg.norm(r.astype(np.float64)) <= 20 * rel_tol * np.linalg.norm(b.astype(np.float64))
    # polling the flag returns early with the same iterate
    xd2 = torch.zeros(n, dtype=tdt, device="cuda")
    torch.cuda.synchronize()
    res2 = S.solve(bd, xd2, 100000, rel_tol, check_every=8)
    assert res2["converged"] and res2["iterations"] == res["iterations"]
    assert torch.equal(xd, xd2)
    S.destroy()
    D.destroy()


def test_cg_rejects_non_stream_and_rectangular(ctx):
    rp, col, val = synth.laplace2d(8, 8, np.float64, np.int64)
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp, col, val), kernel=hp.SPMV_VECTOR)
    with pytest.raises(hp.HpcxError):
        hp.CgSolver(D)
    D.destroy()
    D = hp.DeviceCsr(ctx, hp.CsrMatrix(rp, col, val, ncols=100))
    with pytest.raises(hp.HpcxError):
        hp.CgSolver(D)
    D.destroy()
