"""CPU oracle for libhpcx — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (libhpc_b200/, include/) never does.

`lib()` is oracle/liboracle.so (hpcx_oracle.c: the C restatement); `ref()` is
oracle/_ref/libhpcref.so (the reference's own header + loop bodies, built where the
reference tree is mounted; see oracle/Makefile) or None when it was never built.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile liboracle.so (and oracle/_ref when /root/reference is mounted)."""
    args = ["make", "-C", HERE] + (["-B"] if force else [])
    p = subprocess.run(args, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + p.stdout + p.stderr)


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        L = _LIB
        L.orc_u.restype = C.c_double
        L.orc_u.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_bits.restype = C.c_uint64
        L.orc_bits.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_grid_stride_total.restype = C.c_int64
        L.orc_laplace2d_nnz.restype = C.c_int64
        L.orc_laplace2d_nnz.argtypes = [C.c_int64, C.c_int64]
        L.orc_op27_nnz.restype = C.c_int64
        L.orc_op27_nnz.argtypes = [C.c_int64]
        L.orc_rmat_csr.restype = C.c_int64
        L.orc_sell_build.restype = C.c_int64
    return _LIB


def ref():
    global _REF
    if _REF is None:
        path = os.path.join(HERE, "_ref", "libhpcref.so")
        if not os.path.exists(path):
            return None
        _REF = C.CDLL(path)
    return _REF


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def set_threads(n: int | None = None) -> int:
    """OpenMP team size of the oracle and of oracle/_ref (one libgomp per process).  Default: every
    core of the affinity mask, whatever OMP_NUM_THREADS says (torchrun exports 1)."""
    n = len(os.sched_getaffinity(0)) if n is None else int(n)
    lib().orc_set_threads(C.c_int(n))
    return int(lib().orc_max_threads())


# ------------------------------------------------------------------ RNG
def fill(n: int, seed: int, dtype=np.float64, first: int = 0) -> np.ndarray:
    out = np.empty(n, dtype=dtype)
    fn = lib().orc_fill_f64 if out.dtype == np.float64 else lib().orc_fill_f32
    fn(_p(out), C.c_int64(n), C.c_uint64(seed), C.c_uint64(first))
    return out


# ------------------------------------------------------------------ grid
def grid_stride_total(dim, lo, hi):
    d = np.asarray(dim, dtype=np.int64)
    st = np.zeros(len(d), dtype=np.int64)
    total = lib().orc_grid_stride_total(C.c_int(len(d)), _p(d), C.c_int64(lo), C.c_int64(hi), _p(st))
    return st, int(total)


def grid_linearize(dim, lo, hi, idx, checked: bool):
    """Returns the flat offset, or None where the reference throws std::out_of_range."""
    d = np.asarray(dim, dtype=np.int64)
    ix = np.asarray(idx, dtype=np.int64)
    off = C.c_int64(0)
    rc = lib().orc_grid_linearize(C.c_int(len(d)), _p(d), C.c_int64(lo), C.c_int64(hi), _p(ix),
                                  C.c_int(int(checked)), C.byref(off))
    return None if rc else int(off.value)


def box_sweep(a: np.ndarray, ny, nx, alo, ahi, blo, bhi, axis: int, radius: int) -> np.ndarray:
    """a: flat <2,T,alo,ahi> buffer; returns the flat <2,T,blo,bhi> buffer (ghosts zero)."""
    b = np.zeros((ny + blo + bhi) * (nx + blo + bhi), dtype=a.dtype)
    fn = lib().orc_box_sweep_f32 if a.dtype == np.float32 else lib().orc_box_sweep_f64
    fn(_p(np.ascontiguousarray(a)), _p(b), C.c_int64(ny), C.c_int64(nx), C.c_int64(alo),
       C.c_int64(ahi), C.c_int64(blo), C.c_int64(bhi), C.c_int(axis), C.c_int(radius))
    return b


def jacobi7(src: np.ndarray, nz, ny, nx, lo, hi, steps: int = 1) -> np.ndarray:
    """`steps` ping-pong sweeps; both buffers start as copies of src (same Dirichlet ghosts)."""
    a = np.ascontiguousarray(src).copy()
    b = a.copy()
    fn = lib().orc_jacobi7_f32 if a.dtype == np.float32 else lib().orc_jacobi7_f64
    for _ in range(steps):
        fn(_p(a), _p(b), C.c_int64(nz), C.c_int64(ny), C.c_int64(nx), C.c_int64(lo), C.c_int64(hi))
        a, b = b, a
    return a


def face_copy(grid: np.ndarray, buf: np.ndarray, n3, glo3, ghi3, axis3, side, width, ghost, pack):
    n = np.asarray(n3, dtype=np.int64)
    gl = np.asarray(glo3, dtype=np.int64)
    gh = np.asarray(ghi3, dtype=np.int64)
    lib().orc_face_copy(_p(grid), _p(buf), _p(n), _p(gl), _p(gh), C.c_int(axis3), C.c_int(side),
                        C.c_int(width), C.c_int(int(ghost)), C.c_int(int(pack)),
                        C.c_int(grid.dtype.itemsize))


# ------------------------------------------------------------------ CSR
def laplace2d(ny: int, nx: int, with_values: bool = True):
    nnz = lib().orc_laplace2d_nnz(ny, nx)
    rp = np.empty(ny * nx + 1, dtype=np.int64)
    col = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=np.float64) if with_values else None
    lib().orc_laplace2d_csr(C.c_int64(ny), C.c_int64(nx), _p(rp), _p(col), _p(val))
    return rp, col, val


def op27(n: int, with_values: bool = True):
    nnz = lib().orc_op27_nnz(n)
    rp = np.empty(n ** 3 + 1, dtype=np.int64)
    col = np.empty(nnz, dtype=np.int32)
    val = np.empty(nnz, dtype=np.float64) if with_values else None
    lib().orc_op27_csr(C.c_int64(n), _p(rp), _p(col), _p(val))
    return rp, col, val


def rmat(scale: int, edge_factor: int = 16, seed: int = 24):
    n = 1 << scale
    ne = n * edge_factor
    rp = np.empty(n + 1, dtype=np.int64)
    col = np.empty(ne, dtype=np.int32)
    val = np.empty(ne, dtype=np.float32)
    m = lib().orc_rmat_csr(C.c_int(scale), C.c_int64(ne), C.c_uint64(seed), _p(rp), _p(col), _p(val))
    if m < 0:
        raise MemoryError("orc_rmat_csr")
    return rp, col[:m].copy(), val[:m].copy()


def spmv(rp, col, val, x, out=None):
    y = np.empty(len(rp) - 1, dtype=val.dtype) if out is None else out
    rp = np.ascontiguousarray(rp)
    if rp.dtype == np.int32:
        fn = lib().orc_spmv_i32_f64 if val.dtype == np.float64 else lib().orc_spmv_i32_f32
    else:
        rp = rp.astype(np.int64, copy=False)
        fn = lib().orc_spmv_f64 if val.dtype == np.float64 else lib().orc_spmv_f32
    fn(C.c_int64(len(y)), _p(rp), _p(col), _p(val), _p(np.ascontiguousarray(x, dtype=val.dtype)), _p(y))
    return y


def spmv_wide(rp, col, val, x):
    """fp32 matrix and vector, rows accumulated in double (orc_spmv_wide_f32): the yardstick for
    rows too long for an fp32 running sum."""
    assert val.dtype == np.float32
    y = np.empty(len(rp) - 1, dtype=np.float64)
    lib().orc_spmv_wide_f32(C.c_int64(len(y)), _p(np.ascontiguousarray(rp, dtype=np.int64)), _p(col), _p(val),
                            _p(np.ascontiguousarray(x, dtype=np.float32)), _p(y))
    return y


def spmv_abs(rp, col, val, x):
    mag = np.empty(len(rp) - 1, dtype=np.float64)
    fn = lib().orc_spmv_abs_f64 if val.dtype == np.float64 else lib().orc_spmv_abs_f32
    fn(C.c_int64(len(mag)), _p(np.ascontiguousarray(rp, dtype=np.int64)), _p(col), _p(val),
       _p(np.ascontiguousarray(x, dtype=val.dtype)), _p(mag))
    return mag


def row_bins(rp):
    bins = np.zeros(34, dtype=np.int64)
    rp = np.ascontiguousarray(rp, dtype=np.int64)
    lib().orc_row_bins(C.c_int64(len(rp) - 1), _p(rp), _p(bins))
    return bins


def sell_build(rp, col, val, ncols, c=32, sigma=1024):
    rp = np.ascontiguousarray(rp, dtype=np.int64)
    n = len(rp) - 1
    ns = (n + c - 1) // c
    perm = np.empty(n, dtype=np.int32)
    sptr = np.empty(ns + 1, dtype=np.int64)
    slen = np.empty(ns, dtype=np.int32)
    args = [C.c_int64(n), C.c_int64(ncols), _p(rp), _p(col), _p(val), C.c_int(val.dtype.itemsize),
            C.c_int(c), C.c_int(sigma), _p(perm), _p(sptr), _p(slen)]
    padded = lib().orc_sell_build(*args, None, None)
    scol = np.empty(padded, dtype=np.int32)
    sval = np.empty(padded, dtype=val.dtype)
    lib().orc_sell_build(*args, _p(scol), _p(sval))
    return perm, sptr, slen, scol, sval


def partition_rows(rp, nrows, nparts, mode):
    b = np.empty(nparts + 1, dtype=np.int64)
    rp64 = None if rp is None else np.ascontiguousarray(rp, dtype=np.int64)
    lib().orc_partition_rows(C.c_int64(nrows), _p(rp64), C.c_int(nparts), C.c_int(mode), _p(b))
    return b


def localize(rp, col, bounds, rank):
    rp = np.ascontiguousarray(rp, dtype=np.int64)
    bounds = np.ascontiguousarray(bounds, dtype=np.int64)
    P = len(bounds) - 1
    n = len(rp) - 1
    sizes = np.zeros(4, dtype=np.int64)
    head = [C.c_int64(n), _p(rp), _p(col), _p(bounds), C.c_int(P), C.c_int(rank), _p(sizes)]
    lib().orc_csr_localize(*head, None, None, None, None, None, None, None)
    nloc, nnz_loc, nhalo, nsend = (int(v) for v in sizes)
    out = dict(
        local_row_ptr=np.empty(nloc + 1, dtype=np.int64), local_col=np.empty(nnz_loc, dtype=np.int32),
        halo_gcol=np.empty(nhalo, dtype=np.int64), recv_counts=np.empty(P, dtype=np.int64),
        send_counts=np.empty(P, dtype=np.int64), send_idx=np.empty(nsend, dtype=np.int32),
        row_kind=np.empty(nloc, dtype=np.uint8))
    lib().orc_csr_localize(*head, *[_p(out[k]) for k in (
        "local_row_ptr", "local_col", "halo_gcol", "recv_counts", "send_counts", "send_idx", "row_kind")])
    out.update(nloc=nloc, nnz_loc=nnz_loc, nhalo=nhalo, nsend=nsend)
    return out


# ------------------------------------------------------------------ reference (oracle/_ref)
REF_BLUR = {"x_scalar": 0, "x_sse": 1, "y_scalar": 2, "y_sse_ipl": 3}


def ref_blur(which: str, a_flat: np.ndarray, ny: int, nx: int) -> np.ndarray:
    r = ref()
    if r is None:
        raise RuntimeError("oracle/_ref/libhpcref.so was not built (reference tree absent)")
    a_flat = np.ascontiguousarray(a_flat, dtype=np.float32)
    assert a_flat.size == (ny + 16) * (nx + 16)
    b = np.empty(ny * nx, dtype=np.float32)
    r.hpcref_blur(C.c_int(REF_BLUR[which]), _p(a_flat), _p(b), C.c_int64(ny), C.c_int64(nx))
    return b


def ref_blur_timed(which: str, a_flat: np.ndarray, ny: int, nx: int, reps: int):
    """The reference's own sweep loop (BM_x_blur ...) timed `reps` times on persistent containers
    (hpcref_blur_setup / _run / _result, oracle/ref_harness.cpp).  Returns (seconds list, b_flat)."""
    import time
    r = ref()
    if r is None:
        raise RuntimeError("oracle/_ref/libhpcref.so was not built (reference tree absent)")
    a_flat = np.ascontiguousarray(a_flat, dtype=np.float32)
    assert a_flat.size == (ny + 16) * (nx + 16)
    r.hpcref_blur_setup(_p(a_flat), C.c_int64(ny), C.c_int64(nx))
    times = []
    for _ in range(reps):
        t = time.perf_counter()
        r.hpcref_blur_run(C.c_int(REF_BLUR[which]))
        times.append(time.perf_counter() - t)
    b = np.empty(ny * nx, dtype=np.float32)
    r.hpcref_blur_result(_p(b))
    return times, b


def ref_index(dim, lo, hi, idx, checked):
    """(rc, stride, total, offset) from the reference container; rc -1 = not instantiated."""
    r = ref()
    d = np.asarray(dim, dtype=np.int64)
    st = np.zeros(len(d), dtype=np.int64)
    total = C.c_int64(0)
    off = C.c_int64(0)
    ix = None if idx is None else np.asarray(idx, dtype=np.int64)
    rc = r.hpcref_index(C.c_int(len(d)), C.c_int(lo), C.c_int(hi), _p(d), _p(ix),
                        C.c_int(int(checked)), _p(st), C.byref(total), C.byref(off))
    return rc, st, int(total.value), int(off.value)


def cg(rp, col, val, b, x0, max_iters: int, tol: float = 0.0):
    """orc_cg_*: returns (x, iterations, rr, bb)."""
    rp = np.ascontiguousarray(rp, dtype=np.int64)
    x = np.ascontiguousarray(x0, dtype=val.dtype).copy()
    b = np.ascontiguousarray(b, dtype=val.dtype)
    fn = lib().orc_cg_f64 if val.dtype == np.float64 else lib().orc_cg_f32
    fn.restype = C.c_int64
    rr, bb = C.c_double(), C.c_double()
    it = fn(C.c_int64(len(rp) - 1), _p(rp), _p(np.ascontiguousarray(col, dtype=np.int32)),
            _p(np.ascontiguousarray(val)), _p(b), _p(x), C.c_int64(max_iters), C.c_double(tol),
            C.byref(rr), C.byref(bb))
    return x, int(it), rr.value, bb.value


# ------------------------------------------------------------------ conv2d / transpose (8f-2, 8f-3)
def conv2d(a_in: np.ndarray, a_out: np.ndarray, w: np.ndarray, ny, nx, ilo, ihi, olo=0, ohi=0):
    """out(y,x) += sum in(y+l,x+k) w(l,k) on flat reference-layout buffers; returns the new out."""
    out = np.ascontiguousarray(a_out).copy()
    w = np.ascontiguousarray(w, dtype=out.dtype)
    fn = lib().orc_conv2d_f32 if out.dtype == np.float32 else lib().orc_conv2d_f64
    fn(_p(np.ascontiguousarray(a_in, dtype=out.dtype)), _p(out), _p(w), C.c_int64(ny), C.c_int64(nx),
       C.c_int64(ilo), C.c_int64(ihi), C.c_int64(olo), C.c_int64(ohi), C.c_int(w.shape[0]), C.c_int(w.shape[1]))
    return out


def transpose(a_in: np.ndarray, ny, nx, ilo=0, ihi=0, olo=0, ohi=0) -> np.ndarray:
    a_in = np.ascontiguousarray(a_in)
    out = np.zeros((nx + olo + ohi) * (ny + olo + ohi), dtype=a_in.dtype)
    lib().orc_transpose(_p(a_in), _p(out), C.c_int64(ny), C.c_int64(nx), C.c_int64(ilo), C.c_int64(ihi),
                        C.c_int64(olo), C.c_int64(ohi), C.c_int(a_in.dtype.itemsize))
    return out


def ref_conv(which: str, ka: np.ndarray, kb: np.ndarray, kc: np.ndarray, n: int) -> np.ndarray:
    """The reference's BM_conv ("scalar") / BM_conv_block ("block") loop on an n x n grid:
    ka flat n*n, kb flat <2,float,16> of (n, n), kc flat 16*16.  Returns the updated ka."""
    r = ref()
    if r is None:
        raise RuntimeError("oracle/_ref/libhpcref.so was not built (reference tree absent)")
    ka = np.ascontiguousarray(ka, dtype=np.float32).copy()
    r.hpcref_conv(C.c_int({"scalar": 0, "block": 1}[which]), _p(ka),
                  _p(np.ascontiguousarray(kb, dtype=np.float32)),
                  _p(np.ascontiguousarray(kc, dtype=np.float32)), C.c_int64(n))
    return ka


def ref_transpose(which: str, a: np.ndarray, n: int) -> np.ndarray:
    """BM_transpose ("scalar") / BM_transpose_tiling ("tiling", n % 64 == 0) on an n x n grid."""
    r = ref()
    if r is None:
        raise RuntimeError("oracle/_ref/libhpcref.so was not built (reference tree absent)")
    b = np.empty(n * n, dtype=np.float32)
    r.hpcref_transpose(C.c_int({"scalar": 0, "tiling": 1}[which]),
                       _p(np.ascontiguousarray(a, dtype=np.float32)), _p(b), C.c_int64(n))
    return b


def cells_csr(xs, ys):
    """orc_cells_csr: (row_ptr, col, val, cell_row, row_x, row_y) of the active-cell operator."""
    xs = np.ascontiguousarray(xs, dtype=np.int32)
    ys = np.ascontiguousarray(ys, dtype=np.int32)
    n = len(xs)
    rp = np.empty(n + 1, dtype=np.int64)
    col = np.empty(5 * max(n, 1), dtype=np.int32)
    val = np.empty(5 * max(n, 1), dtype=np.float64)
    cell_row = np.empty(max(n, 1), dtype=np.int32)
    rx = np.empty(max(n, 1), dtype=np.int32)
    ry = np.empty(max(n, 1), dtype=np.int32)
    nnz = C.c_int64()
    fn = lib().orc_cells_csr
    fn.restype = C.c_int64
    m = fn(C.c_int64(n), _p(xs), _p(ys), _p(rp), _p(col), _p(val), _p(cell_row), _p(rx), _p(ry), C.byref(nnz))
    return rp[:m + 1].copy(), col[:nnz.value].copy(), val[:nnz.value].copy(), cell_row[:n].copy(), rx[:m].copy(), ry[:m].copy()
