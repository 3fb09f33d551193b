"""Shared helpers of the parity tests (checker side)."""
import numpy as np

import oracle


def assert_spmv_close(y, y_ref, mag, tol):
    """norm-wise max|dy| / max(|y_ref|_inf, tiny) <= tol and row-wise
    |dy_i| <= tol * sum_p |a_p x_p| (mag)."""
    y = np.asarray(y, dtype=np.float64)
    y_ref = np.asarray(y_ref, dtype=np.float64)
    d = np.abs(y - y_ref)
    scale = max(np.abs(y_ref).max(initial=0.0), np.finfo(np.float64).tiny)
    assert d.max(initial=0.0) / scale <= tol, f"norm-wise error {d.max() / scale:.3e} > {tol}"
    bad = d > tol * mag + np.finfo(np.float64).tiny
    assert not bad.any(), f"{bad.sum()} rows exceed the row-wise bound, worst {d[bad].max():.3e}"


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and a.tobytes() == b.tobytes()


def random_csr(nrows, ncols, lens, seed, dtype):
    """CSR with prescribed row lengths, ascending distinct columns, values 2u-1."""
    rng = np.random.default_rng(seed)
    lens = np.minimum(np.asarray(lens, dtype=np.int64), ncols)
    rp = np.zeros(nrows + 1, dtype=np.int64)
    np.cumsum(lens, out=rp[1:])
    col = np.empty(rp[-1], dtype=np.int32)
    for i in range(nrows):
        if lens[i]:
            if lens[i] * 4 > ncols:
                c = rng.permutation(ncols)[:lens[i]]
            else:
                c = np.unique(rng.integers(0, ncols, size=lens[i] * 2))[:lens[i]]
                while len(c) < lens[i]:
                    c = np.unique(np.concatenate([c, rng.integers(0, ncols, size=lens[i])]))[:lens[i]]
            col[rp[i]:rp[i + 1]] = np.sort(c)
    val = oracle.fill(int(rp[-1]), seed + 1, dtype)
    return rp, col, val
