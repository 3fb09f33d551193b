"""Deterministic synthetic inputs of BASELINE.json's configs (host side, numpy).

These are input generators, not compute: the matrices and fields they build are what
bench.py and the tests hand to the CUDA library.  The oracle has its own C generators
(oracle/hpcx_oracle.c); tests/test_host_logic.py checks the two agree bit-for-bit.

RNG: counter-based, u(seed, i) = (mix(mix(seed) ^ i) >> 11) * 2^-53 with
mix = the splitmix64 finaliser, so any element can be generated independently.
"""
from __future__ import annotations

import numpy as np

_G = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def _mix(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = z + _G
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def bits(seed: int, first: int, n: int) -> np.ndarray:
    s = _mix(np.array([seed], dtype=np.uint64))[0]
    idx = np.arange(first, first + n, dtype=np.uint64)
    return _mix(s ^ idx)


def uniform(seed: int, n: int, first: int = 0) -> np.ndarray:
    """u(seed, first .. first+n) in [0, 1) as float64."""
    return (bits(seed, first, n) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def field(seed: int, n: int, dtype=np.float64, first: int = 0, chunk: int = 1 << 24) -> np.ndarray:
    """2u - 1 in `dtype` (x vectors, grid fills), generated in chunks to bound memory."""
    out = np.empty(n, dtype=dtype)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        out[a:b] = (2.0 * uniform(seed, b - a, first + a) - 1.0).astype(dtype)
    return out


def laplace2d(ny: int, nx: int, dtype=np.float64, index_dtype=np.int32):
    """2-D 5-point Laplacian (C1/C2): row i = y*nx + x, columns ascending
    {i-nx, i-1, i, i+1, i+nx} with out-of-grid neighbours omitted, values {-1,-1,4,-1,-1}."""
    n = ny * nx
    y = np.repeat(np.arange(ny, dtype=np.int64), nx)
    x = np.tile(np.arange(nx, dtype=np.int64), ny)
    has = np.stack([y > 0, x > 0, np.ones(n, dtype=bool), x < nx - 1, y < ny - 1], axis=1)
    del y, x
    i = np.arange(n, dtype=np.int64)[:, None]
    off = np.array([-nx, -1, 0, 1, nx], dtype=np.int64)[None, :]
    cols = (i + off)[has].astype(np.int32)
    vals = np.broadcast_to(np.array([-1, -1, 4, -1, -1], dtype=dtype)[None, :], has.shape)[has]
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(has.sum(axis=1), out=rp[1:])
    return rp.astype(index_dtype), cols, np.ascontiguousarray(vals)


def laplace2d_big(ny: int, nx: int, dtype=np.float64, index_dtype=np.int32, rows_per_chunk=1024):
    """Same matrix as laplace2d, built in row-slab chunks (bounded temporaries; 8192^2)."""
    nnz = 5 * nx * ny - 2 * nx - 2 * ny
    rp = np.empty(ny * nx + 1, dtype=index_dtype)
    col = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=dtype)
    p = 0
    for y0 in range(0, ny, rows_per_chunk):
        y1 = min(ny, y0 + rows_per_chunk)
        m = (y1 - y0) * nx
        y = np.repeat(np.arange(y0, y1, dtype=np.int64), nx)
        x = np.tile(np.arange(nx, dtype=np.int64), y1 - y0)
        has = np.stack([y > 0, x > 0, np.ones(m, dtype=bool), x < nx - 1, y < ny - 1], axis=1)
        i = (y * nx + x)[:, None]
        off = np.array([-nx, -1, 0, 1, nx], dtype=np.int64)[None, :]
        c = (i + off)[has]
        k = len(c)
        col[p:p + k] = c
        val[p:p + k] = np.broadcast_to(np.array([-1, -1, 4, -1, -1], dtype=dtype)[None, :],
                                       has.shape)[has]
        cnt = has.sum(axis=1)
        rp[y0 * nx:y1 * nx] = p + np.cumsum(cnt) - cnt
        p += k
    rp[ny * nx] = p
    assert p == nnz
    return rp, col, val


def laplace2d_rows(ny: int, nx: int, y0: int, y1: int, dtype=np.float64, index_dtype=np.int64, rows_per_chunk=1024):
    """Grid rows [y0, y1) of the ny x nx 5-point Laplacian as a CSR row slab with GLOBAL column
    indices (what one rank of a distributed job holds): (row_ptr from 0, col, val)."""
    m = (y1 - y0) * nx
    nnz = 5 * m - 2 * (y1 - y0) - (nx if y0 == 0 else 0) - (nx if y1 == ny else 0)
    rp = np.empty(m + 1, dtype=index_dtype)
    col = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=dtype)
    p = 0
    for a in range(y0, y1, rows_per_chunk):
        b = min(y1, a + rows_per_chunk)
        k = (b - a) * nx
        y = np.repeat(np.arange(a, b, dtype=np.int64), nx)
        x = np.tile(np.arange(nx, dtype=np.int64), b - a)
        has = np.stack([y > 0, x > 0, np.ones(k, dtype=bool), x < nx - 1, y < ny - 1], axis=1)
        i = (y * nx + x)[:, None]
        off = np.array([-nx, -1, 0, 1, nx], dtype=np.int64)[None, :]
        c = (i + off)[has]
        q = len(c)
        col[p:p + q] = c
        val[p:p + q] = np.broadcast_to(np.array([-1, -1, 4, -1, -1], dtype=dtype)[None, :], has.shape)[has]
        cnt = has.sum(axis=1)
        rp[(a - y0) * nx:(b - y0) * nx] = p + np.cumsum(cnt) - cnt
        p += q
    rp[m] = p
    assert p == nnz
    return rp, col, val


def op27_rows(n: int, z0: int, z1: int, dtype=np.float64, planes_per_chunk: int = 16):
    """Planes [z0, z1) of the n^3 27-point operator as a CSR row slab with GLOBAL column indices."""
    rps, cols, vals, base = [np.zeros(1, np.int64)], [], [], 0
    for a in range(z0, z1, planes_per_chunk):
        b = min(z1, a + planes_per_chunk)
        rp, c, v = op27(n, dtype, np.int64, a, b)
        cols.append(c)
        vals.append(v)
        rps.append(rp[1:] + base)
        base += int(rp[-1])
    return np.concatenate(rps), np.concatenate(cols), np.concatenate(vals)


def op27(n: int, dtype=np.float64, index_dtype=np.int32, z_lo: int = 0, z_hi: int | None = None):
    """3-D 27-point operator on n^3 (C4): 27 neighbours ascending, Dirichlet truncation,
    centre 26, others -1.  [z_lo, z_hi) selects a slab of rows (global column numbering)."""
    z_hi = n if z_hi is None else z_hi
    zz, yy, xx = np.meshgrid(np.arange(z_lo, z_hi, dtype=np.int64), np.arange(n, dtype=np.int64),
                             np.arange(n, dtype=np.int64), indexing="ij")
    zz, yy, xx = zz.ravel(), yy.ravel(), xx.ravel()
    m = len(zz)
    has = np.empty((m, 27), dtype=bool)
    off = np.empty(27, dtype=np.int64)
    k = 0
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                has[:, k] = ((zz + dz >= 0) & (zz + dz < n) & (yy + dy >= 0) & (yy + dy < n) &
                             (xx + dx >= 0) & (xx + dx < n))
                off[k] = (dz * n + dy) * n + dx
                k += 1
    i = ((zz * n + yy) * n + xx)[:, None]
    cols = (i + off[None, :])[has].astype(np.int32)
    v = np.full(27, -1, dtype=dtype)
    v[13] = 26
    vals = np.ascontiguousarray(np.broadcast_to(v[None, :], has.shape)[has])
    rp = np.zeros(m + 1, dtype=np.int64)
    np.cumsum(has.sum(axis=1), out=rp[1:])
    return rp.astype(index_dtype), cols, vals


def rmat(scale: int, edge_factor: int = 16, seed: int = 24, index_dtype=np.int32):
    """R-MAT (C5): (a,b,c,d) = (.57,.19,.19,.05), no vertex permutation, level l of edge k
    from u(seed, k*scale + l); values (float32)(2u(seed+1, k) - 1); sorted by (row, col, k),
    duplicates merged by adding in ascending k."""
    n = 1 << scale
    ne = n * edge_factor
    r = np.zeros(ne, dtype=np.uint64)
    c = np.zeros(ne, dtype=np.uint64)
    k = np.arange(ne, dtype=np.uint64)
    s = _mix(np.array([seed], dtype=np.uint64))[0]
    for l in range(scale):
        with np.errstate(over="ignore"):
            u = (_mix(s ^ (k * np.uint64(scale) + np.uint64(l))) >> np.uint64(11)).astype(
                np.float64) * (1.0 / 9007199254740992.0)
        q = (u >= 0.57).astype(np.uint64) + (u >= 0.76) + (u >= 0.95)
        r = (r << np.uint64(1)) | (q >> np.uint64(1))
        c = (c << np.uint64(1)) | (q & np.uint64(1))
    key = (r << np.uint64(32)) | c
    order = np.argsort(key, kind="stable")  # stable => ties stay in ascending k
    key = key[order]
    v = (2.0 * uniform(seed + 1, ne) - 1.0).astype(np.float32)[order]
    first = np.ones(ne, dtype=bool)
    first[1:] = key[1:] != key[:-1]
    starts = np.flatnonzero(first)
    col = (key[starts] & np.uint64(0xFFFFFFFF)).astype(np.int32)
    rows = (key[starts] >> np.uint64(32)).astype(np.int64)
    # sequential float32 accumulation inside each duplicate run, ascending k
    val = v[starts].copy()
    run = np.diff(np.append(starts, ne))
    j = 1
    todo = np.flatnonzero(run > j)
    while len(todo):
        val[todo] = val[todo] + v[starts[todo] + j]
        j += 1
        todo = todo[run[todo] > j]
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows, minlength=n), out=rp[1:])
    return rp.astype(index_dtype), col, val


def laplace2d_slab_local(ny_total: int, nx: int, y0: int, y1: int, dtype=np.float64):
    """Rank-local view of grid rows [y0, y1) of the ny_total x nx 5-point Laplacian, built
    without ever forming the global matrix.  Returns exactly the tables hpcx_csr_localize
    produces for a contiguous row partition whose block is [y0*nx, y1*nx)
    (tests/test_host_logic.py checks the two agree): columns renumbered [owned | halo], halo
    = grid row y0-1 (from the rank below) then grid row y1 (from the rank above)."""
    ny = y1 - y0
    nloc = ny * nx
    has_lo, has_hi = y0 > 0, y1 < ny_total
    # Build the slab as the interior rows of a (ny + has_lo + has_hi)-row grid, then cut.
    pad_lo, pad_hi = int(has_lo), int(has_hi)
    rp_e, col_e, val_e = laplace2d_big(ny + pad_lo + pad_hi, nx, dtype, np.int64)
    a, b = rp_e[pad_lo * nx], rp_e[(pad_lo + ny) * nx]
    local_row_ptr = rp_e[pad_lo * nx:(pad_lo + ny) * nx + 1] - a
    c = col_e[a:b].astype(np.int64) - pad_lo * nx  # owned -> [0, nloc); below -> [-nx, 0); above -> [nloc, nloc+nx)
    del rp_e, col_e
    local_col = np.where(c < 0, c + nx + nloc, np.where(c >= nloc, c + pad_lo * nx, c)).astype(np.int32)
    values = val_e[a:b]
    nhalo = (pad_lo + pad_hi) * nx
    P_lo = np.arange(nx, dtype=np.int32)
    send_idx = np.concatenate(([P_lo] if has_lo else []) + ([nloc - nx + P_lo] if has_hi else []) +
                              [np.empty(0, np.int32)]).astype(np.int32)
    row_kind = np.zeros(nloc, dtype=np.uint8)
    if has_lo:
        row_kind[:nx] = 1
    if has_hi:
        row_kind[nloc - nx:] = 1
    halo_gcol = np.concatenate(([(y0 - 1) * nx + np.arange(nx, dtype=np.int64)] if has_lo else []) +
                               ([y1 * nx + np.arange(nx, dtype=np.int64)] if has_hi else []) +
                               [np.empty(0, np.int64)])
    return dict(nloc=nloc, nnz_loc=int(b - a), nhalo=nhalo, nsend=len(send_idx),
                local_row_ptr=local_row_ptr, local_col=local_col, values=values,
                halo_gcol=halo_gcol, send_idx=send_idx, row_kind=row_kind, has_lo=has_lo,
                has_hi=has_hi)


def slab_counts(local: dict, nparts: int, rank: int, nx: int):
    """recv_counts / send_counts of a slab partition (neighbours rank-1 and rank+1)."""
    recv = np.zeros(nparts, dtype=np.int64)
    send = np.zeros(nparts, dtype=np.int64)
    if local["has_lo"]:
        recv[rank - 1] = send[rank - 1] = nx
    if local["has_hi"]:
        recv[rank + 1] = send[rank + 1] = nx
    return recv, send


def op27_slab_local(n: int, z0: int, z1: int, dtype=np.float64, planes_per_chunk: int = 16):
    """Rank-local view of planes [z0, z1) of the n^3 27-point operator (C4) without forming the
    global matrix: the tables hpcx_csr_localize yields for the row block [z0*n^2, z1*n^2).
    Halo = plane z0-1 (owned by the rank below) then plane z1 (rank above)."""
    pl = n * n
    nloc = (z1 - z0) * pl
    has_lo, has_hi = z0 > 0, z1 < n
    rps, cols, vals = [np.zeros(1, np.int64)], [], []
    base = 0
    for a in range(z0, z1, planes_per_chunk):
        b = min(z1, a + planes_per_chunk)
        rp, c, v = op27(n, dtype, np.int64, a, b)
        c = c.astype(np.int64) - z0 * pl  # owned -> [0, nloc); below -> [-pl, 0); above -> [nloc, nloc+pl)
        c = np.where(c < 0, c + pl + nloc, np.where(c >= nloc, c + int(has_lo) * pl, c))
        cols.append(c.astype(np.int32))
        vals.append(v)
        rps.append(rp[1:] + base)
        base += int(rp[-1])
    P = np.arange(pl, dtype=np.int32)
    send_idx = np.concatenate(([P] if has_lo else []) + ([nloc - pl + P] if has_hi else []) +
                              [np.empty(0, np.int32)]).astype(np.int32)
    row_kind = np.zeros(nloc, dtype=np.uint8)
    if has_lo:
        row_kind[:pl] = 1
    if has_hi:
        row_kind[nloc - pl:] = 1
    halo_gcol = np.concatenate(([(z0 - 1) * pl + np.arange(pl, dtype=np.int64)] if has_lo else []) +
                               ([z1 * pl + np.arange(pl, dtype=np.int64)] if has_hi else []) +
                               [np.empty(0, np.int64)])
    return dict(nloc=nloc, nnz_loc=base, nhalo=(int(has_lo) + int(has_hi)) * pl, nsend=len(send_idx),
                local_row_ptr=np.concatenate(rps), local_col=np.concatenate(cols),
                values=np.concatenate(vals), halo_gcol=halo_gcol, send_idx=send_idx,
                row_kind=row_kind, has_lo=has_lo, has_hi=has_hi)
