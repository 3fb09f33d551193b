"""Host-side logic of the product through the C ABI (no GPU): grid index math, partitioner,
halo maps, synthetic generators — bit-exact against the oracle and the reference golden."""
import json
import os

import numpy as np
import pytest

import libhpc_b200 as hp
import oracle
from libhpc_b200 import synth
from util import random_csr

G = os.path.join(os.path.dirname(__file__), "golden")
INDEX = json.load(open(os.path.join(G, "index_ref.json")))


def test_grid_index_math_matches_reference_golden():
    for c in INDEX:
        a = hp.HPCHighDimensionFlatArray.__new__(hp.HPCHighDimensionFlatArray)
        if c["total"] > 1 << 26:  # do not allocate the 8192^2 / 512^3 buffers here
            hp.HPCHighDimensionFlatArray.__init__(a, *c["dim"], lo=c["lo"], hi=c["hi"],
                                                  buffer=np.lib.stride_tricks.as_strided(
                                                      np.zeros(1, np.float32), (c["total"],), (0,)))
        else:
            hp.HPCHighDimensionFlatArray.__init__(a, *c["dim"], lo=c["lo"], hi=c["hi"])
        assert list(a.stride) == c["stride"] and a.total == c["total"]
        assert a._lin(tuple(c["idx"]), 0) == c["unsafe"]
        if c["safe"] is None:
            with pytest.raises(IndexError):  # std::out_of_range in the reference
                a.at(c["idx"])
        else:
            assert a._lin(tuple(c["idx"]), 1) == c["safe"]


def test_array_container_semantics():
    a = hp.HPCHighDimensionFlatArray(4, 6, dtype=np.float32, lo=2)
    assert a.data().shape == (8 * 10,) and not a.data().any()  # value-initialised, ghosts zero
    a.set((-2, -2), 5.0)
    a.set((3, 5), 7.0)
    assert a.data()[0] == 5.0 and a(-2, -2) == 5.0 and a.at((3, 5)) == 7.0
    assert a.interior()[3, 5] == 7.0 and a.padded_view()[0, 0] == 5.0
    with pytest.raises(IndexError):
        a.at((4 + 2, 0))
    with pytest.raises(ValueError):
        hp.HPCHighDimensionFlatArray(0, 4)
    with pytest.raises(TypeError):
        hp.HPCHighDimensionFlatArray(4, dtype=np.int32)
    b = hp.HPCHighDimensionFlatArray(3, 4, 5, dtype=np.float64, lo=1, hi=3)
    assert b.stride == (8 * 9, 9, 1) and b.total == 7 * 8 * 9


@pytest.mark.parametrize("n,p", [(10, 4), (3, 8), (0, 2), (512, 8), (513, 8), (1, 1)])
def test_partition_slabs_static_chunking(n, p):
    chunk = (n + p - 1) // p
    want = [min(i * chunk, n) for i in range(p + 1)]
    assert list(hp.partition_slabs(n, p)) == want
    assert list(hp.partition_rows(n, None, p, 0)) == want
    assert list(oracle.partition_rows(None, n, p, 0)) == want


@pytest.mark.parametrize("idx", [np.int32, np.int64])
def test_partition_rows_by_nnz_matches_oracle(idx):
    rp, col, val = oracle.rmat(11, 8, 3)
    for p in (1, 2, 3, 4, 8):
        got = hp.partition_rows(len(rp) - 1, rp.astype(idx), p, 1)
        assert np.array_equal(got, oracle.partition_rows(rp, len(rp) - 1, p, 1))
        assert got[0] == 0 and got[-1] == len(rp) - 1 and (np.diff(got) >= 0).all()


def _check_localize(rp, col, val, bounds):
    A = hp.CsrMatrix(rp, col, val)
    for r in range(len(bounds) - 1):
        got = hp.csr_localize(A, bounds, r)
        want = oracle.localize(rp, col, bounds, r)
        for k in ("nloc", "nnz_loc", "nhalo", "nsend"):
            assert got[k] == want[k], k
        for k in ("local_row_ptr", "local_col", "halo_gcol", "recv_counts", "send_counts",
                  "send_idx", "row_kind"):
            assert np.array_equal(got[k], want[k]), (k, r)  # bit-exact integer structure


def test_localize_matches_oracle_banded():
    rp, col, val = oracle.op27(6)
    _check_localize(rp, col, val, hp.partition_rows(6 ** 3, rp, 4, 0))
    rp, col, val = oracle.laplace2d(9, 7)
    _check_localize(rp.astype(np.int32), col, val, hp.partition_rows(63, rp, 3, 1))


def test_localize_matches_oracle_irregular_and_empty_parts():
    rp, col, val = oracle.rmat(9, 8, 5)
    _check_localize(rp, col, val, hp.partition_rows(512, rp, 8, 1))
    _check_localize(rp, col, val, np.array([0, 0, 100, 100, 512], dtype=np.int64))
    lens = [0, 3, 0, 0, 12, 1, 0, 40, 2, 0]
    rp, col, val = random_csr(10, 10, lens, 2, np.float32)
    _check_localize(rp, col, val, np.array([0, 4, 5, 10], dtype=np.int64))


def test_localize_send_recv_are_mutually_consistent():
    rp, col, val = oracle.rmat(9, 8, 6)
    A = hp.CsrMatrix(rp, col, val)
    P = 4
    bounds = hp.partition_rows(512, rp, P, 1)
    L = [hp.csr_localize(A, bounds, r) for r in range(P)]
    for r in range(P):
        for q in range(P):
            assert L[r]["send_counts"][q] == L[q]["recv_counts"][r]
        # what r sends to q are exactly q's halo columns owned by r, in q's halo order
        off = 0
        for q in range(P):
            n = L[r]["send_counts"][q]
            sent = L[r]["send_idx"][off:off + n] + bounds[r]
            roff = L[q]["recv_counts"][:r].sum()
            assert np.array_equal(sent, L[q]["halo_gcol"][roff:roff + n])
            off += n


def test_synth_generators_match_oracle_bit_for_bit():
    for ny, nx in ((1, 1), (6, 5), (1, 7), (64, 33)):
        a = oracle.laplace2d(ny, nx)
        for b in (synth.laplace2d(ny, nx, index_dtype=np.int64),
                  synth.laplace2d_big(ny, nx, index_dtype=np.int64, rows_per_chunk=5)):
            assert all(np.array_equal(u, v) and u.dtype == v.dtype for u, v in zip(a, b))
    a, b = oracle.op27(5), synth.op27(5, index_dtype=np.int64)
    assert all(np.array_equal(u, v) for u, v in zip(a, b))
    a, b = oracle.rmat(11, 16, 24), synth.rmat(11, 16, 24, index_dtype=np.int64)
    assert all(np.array_equal(u, v) and u.dtype == v.dtype for u, v in zip(a, b))
    assert np.array_equal(oracle.fill(1000, 42, np.float64, 5), synth.field(42, 1000, np.float64, 5))
    assert np.array_equal(oracle.fill(1000, 7, np.float32), synth.field(7, 1000, np.float32, chunk=64))


def test_op27_slab_rows_equal_global_rows():
    rp, col, val = synth.op27(6, index_dtype=np.int64)
    rs, cs, vs = synth.op27(6, index_dtype=np.int64, z_lo=2, z_hi=4)
    a, b = rp[2 * 36], rp[4 * 36]
    assert np.array_equal(cs, col[a:b]) and np.array_equal(vs, val[a:b])
    assert np.array_equal(rs, rp[2 * 36:4 * 36 + 1] - a)


def test_compute_entry_points_fail_loudly_without_gpu():
    """No CPU fallback: without a device the context cannot be created."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(hp.HpcxError) as e:
        hp.Context(0)
    assert e.value.status == hp._abi.CUDA_ERROR and "no CPU fallback" in str(e.value)


@pytest.mark.parametrize("P", [1, 2, 3, 4])
def test_slab_local_builder_equals_localize_of_global(P):
    ny_per, nx = 5, 7
    ny = ny_per * P
    rp, col, val = synth.laplace2d(ny, nx, index_dtype=np.int64)
    A = hp.CsrMatrix(rp, col, val)
    bounds = np.arange(P + 1, dtype=np.int64) * ny_per * nx
    for r in range(P):
        want = hp.csr_localize(A, bounds, r)
        got = synth.laplace2d_slab_local(ny, nx, r * ny_per, (r + 1) * ny_per)
        for k in ("nloc", "nnz_loc", "nhalo", "nsend"):
            assert got[k] == want[k], k
        for k in ("local_row_ptr", "local_col", "halo_gcol", "send_idx", "row_kind", "values"):
            assert np.array_equal(got[k], want[k]), (k, r)
        recv, send = synth.slab_counts(got, P, r, nx)
        assert np.array_equal(recv, want["recv_counts"]) and np.array_equal(send, want["send_counts"])


@pytest.mark.parametrize("P", [1, 2, 3])
def test_op27_slab_local_builder_equals_localize_of_global(P):
    n = 6
    rp, col, val = synth.op27(n, index_dtype=np.int64)
    A = hp.CsrMatrix(rp, col, val)
    zb = hp.partition_slabs(n, P)
    bounds = zb * n * n
    for r in range(P):
        want = hp.csr_localize(A, bounds, r)
        got = synth.op27_slab_local(n, int(zb[r]), int(zb[r + 1]), planes_per_chunk=1 + r)
        for k in ("nloc", "nnz_loc", "nhalo", "nsend"):
            assert got[k] == want[k], k
        for k in ("local_row_ptr", "local_col", "halo_gcol", "send_idx", "row_kind"):
            assert np.array_equal(got[k], want[k]), (k, r)
        assert np.array_equal(got["values"], val[rp[bounds[r]]:rp[bounds[r + 1]]])
        recv, send = synth.slab_counts(got, P, r, n * n)
        assert np.array_equal(recv, want["recv_counts"]) and np.array_equal(send, want["send_counts"])


def test_halo_index_maps_match_reference_index_math():
    """hpcx_grid_export_halo_maps against a brute-force enumeration with the reference's index math
    (oracle.grid_linearize, pinned to the reference header): every cell of the `width` interior
    slices next to a face (send) and of the ghost slices behind it (recv), in row-major order."""
    import itertools
    for dims, lo, hi in (((5, 7), 2, 2), ((4, 3, 6), 1, 1), ((6, 5, 4), 1, 3), ((3, 9), 8, 8), ((9,), 2, 2)):
        for axis in range(len(dims)):
            for side in (0, 1):
                for width in range(1, min((hi if side else lo), dims[axis]) + 1):
                    run_len, send, recv = hp.grid_halo_maps(dims, lo, hi, axis, side, width)
                    slower = [range(-lo, dims[i] + hi) for i in range(axis)]
                    faster = [range(-lo, dims[i] + hi) for i in range(axis + 1, len(dims))]
                    s0 = dims[axis] - width if side else 0
                    r0 = dims[axis] if side else -width
                    want_s, want_r = [], []
                    for outer in itertools.product(*slower):
                        for k in range(width):
                            for inner in itertools.product(*faster):
                                want_s.append(oracle.grid_linearize(dims, lo, hi, list(outer) + [s0 + k] + list(inner), True))
                                want_r.append(oracle.grid_linearize(dims, lo, hi, list(outer) + [r0 + k] + list(inner), True))
                    got_s = np.concatenate([np.arange(a, a + run_len) for a in send])
                    got_r = np.concatenate([np.arange(a, a + run_len) for a in recv])
                    assert got_s.tolist() == want_s and got_r.tolist() == want_r, (dims, lo, hi, axis, side, width)
    with pytest.raises(hp.HpcxError):
        hp.grid_halo_maps((5, 7), 1, 1, 0, 0, 2)  # wider than the ghost region


@pytest.mark.parametrize("py,px", [(1, 1), (2, 1), (1, 3), (2, 2), (4, 2), (3, 5)])
def test_block_bounds_tile_the_grid(py, px):
    """dist.block_bounds: the py x px blocks cover every cell of the grid exactly once, blocks of one
    block row share their rows and blocks of one block column their columns (what
    hpcx_stencil_dist_plan_blocks requires of neighbouring blocks)."""
    from libhpc_b200 import dist as hd
    ny, nx = 53, 130
    seen = np.zeros((ny, nx), dtype=np.int32)
    rows, cols = {}, {}
    for rank in range(py * px):
        y0, y1, x0, x1 = hd.block_bounds(ny, nx, py, px, rank)
        seen[y0:y1, x0:x1] += 1
        ry, rx = divmod(rank, px)
        assert rows.setdefault(ry, (y0, y1)) == (y0, y1)
        assert cols.setdefault(rx, (x0, x1)) == (x0, x1)
    assert (seen == 1).all()
    with pytest.raises(ValueError):
        hd.block_bounds(ny, nx, py, px, py * px)
