"""The sparse-grid bridge against the reference's REAL sparse::RootGrid
(lib/sparse/include/RootGrid.hpp:14-22, 62-75; HashBlock / PointerBlock / DenseBlock), compiled
through a test-only stand-in for oneTBB (oracle/tbb_shim: concurrent_vector -> std::vector,
concurrent_hash_map -> mutex + unordered_map, parallel_for -> loop).

CPU: what the true foreach enumerates for the reference's own workload (golden fixture, made by
tests/golden/make_golden_rootgrid.py, re-checked live where the reference tree is mounted) is what
the bridge assumes: every cell of every allocated 16 x 16 leaf, global coordinates
(leaf << 4) + local, values where they were written.  GPU: sparse::csr_from_grid on the true tree."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
FIX = json.load(open(os.path.join(ROOT, "tests", "golden", "rootgrid_cells.json")))
BIN = os.path.join(ROOT, "tests", "cpp", "rootgrid_check")
HAVE_REF = os.path.isdir("/root/reference/lib/sparse/include")


def _workload(n):
    """The writes of rootgrid_check.cpp: fill, in fp32 like the reference's loop."""
    t = np.arange(n)
    tf = t.astype(np.float32)
    x = np.floor(np.float32(-100.0) + np.float32(0.2) * tf).astype(np.int64)
    y = np.floor(np.float32(100.0) + np.float32(-0.6) * tf).astype(np.int64)
    v = np.float32(1.0) + np.float32(0.001) * (t % 97).astype(np.float32)
    return x, y, v


def test_true_foreach_enumerates_whole_leaves_in_global_coordinates():
    x, y, v = _workload(400)
    leaves = sorted({(int(a) >> 4, int(b) >> 4) for a, b in zip(x, y)})
    assert [tuple(l) for l in FIX["leaves"]] == leaves  # coordinate -> leaf by >> BShift (RootGrid.hpp:79-84)
    assert FIX["ncells"] == 256 * len(leaves)            # foreach visits every cell of an allocated leaf
    last = {}
    for a, b, c in zip(x, y, v):                          # later writes win
        last[(int(a), int(b))] = float(c)
    got = {(a, b): c for a, b, c in FIX["written"]}
    assert got.keys() == last.keys()
    assert all(np.float32(got[k]) == np.float32(last[k]) for k in last)


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not mounted")
def test_fixture_is_what_the_reference_enumerates_here():
    import importlib.util
    spec = importlib.util.spec_from_file_location("mk", os.path.join(ROOT, "tests", "golden", "make_golden_rootgrid.py"))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    mk.build()
    cells = mk.enumerate_cells()
    assert len(cells) == FIX["ncells"]
    assert sorted({(a >> 4, b >> 4) for a, b, _ in cells}) == [tuple(l) for l in FIX["leaves"]]
    assert [[a, b, c] for a, b, c in cells if c != 0.0] == FIX["written"]


def test_bridge_rows_are_the_enumerated_cells_x_major():
    """orc_cells_csr (the oracle the device bridge is checked against) on the true enumeration:
    rows are the cells in (x, y) lexicographic order — DenseBlock's x-major m_block[x*B + y]."""
    cells = sorted((16 * lx + i, 16 * ly + j) for lx, ly in FIX["leaves"] for i in range(16) for j in range(16))
    xs = np.array([c[0] for c in cells], dtype=np.int32)
    ys = np.array([c[1] for c in cells], dtype=np.int32)
    perm = np.random.default_rng(0).permutation(len(cells))
    rp, col, val, cell_row, rx, ry = oracle.cells_csr(xs[perm], ys[perm])
    assert np.array_equal(rx, xs) and np.array_equal(ry, ys)
    assert np.array_equal(np.stack([rx[cell_row], ry[cell_row]]), np.stack([xs[perm], ys[perm]]))
    # interior cells of a leaf have all four neighbours active: 5 entries, diagonal 4
    inner = (xs % 16 > 0) & (xs % 16 < 15) & (ys % 16 > 0) & (ys % 16 < 15)
    assert (np.diff(rp)[inner] == 5).all()


@pytest.mark.gpu
def test_csr_from_grid_on_the_true_rootgrid():
    if not os.path.exists(BIN):
        pytest.skip("tests/cpp/rootgrid_check was not built (needs the reference tree at build time)")
    p = subprocess.run([BIN, "gpu"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "ROOTGRID OK" in p.stdout, p.stdout + p.stderr
    assert "rootgrid hash_dense16" in p.stdout and "rootgrid hash_pointer8_dense4" in p.stdout
