"""The C++ face of libhpcx (include/hpcx/*.hpp) — compiled with g++ here, linked against
libhpcx.so.

CPU part: hpc::HPCHighDimensionFlatArray's strides, totals, offsets and out_of_range verdicts,
index by index, against the oracle's restatement of the reference index math and — where
oracle/_ref was built from /root/reference — against the reference container itself; the
sparse::CsrMatrix builders against the oracle's CSR builders (FNV-1a digests, bit-exact).
GPU part (-m gpu): the reference's sweep loop nests and a CPU SpMV loop, written in the test
program, against hpc::stencil_sweep / sparse::spmv."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
SRC = os.path.join(ROOT, "tests", "cpp", "host_check.cpp")
BIN = os.path.join(ROOT, "tests", "cpp", "host_check")


@pytest.fixture(scope="module")
def binary():
    deps = [SRC] + [os.path.join(ROOT, "include", "hpcx", f) for f in os.listdir(os.path.join(ROOT, "include", "hpcx"))]
    deps.append(os.path.join(ROOT, "include", "hpcx.h"))
    if not os.path.exists(BIN) or any(os.path.getmtime(d) > os.path.getmtime(BIN) for d in deps):
        lib = os.path.join(ROOT, "libhpc_b200")
        subprocess.run(["g++", "-std=c++17", "-O1", "-fopenmp", "-ffp-contract=off", "-Wall",
                        "-I" + os.path.join(ROOT, "include"), SRC, "-L" + lib, "-lhpcx",
                        "-Wl,-rpath," + lib, "-o", BIN], check=True)
    return BIN


def _run(binary, mode, timeout=300):
    p = subprocess.run([binary, mode], capture_output=True, text=True, timeout=timeout)
    return p.returncode, p.stdout, p.stderr


def _fnv1a(a: np.ndarray) -> str:
    h = 1469598103934665603
    # vectorised FNV is not possible; the arrays here are small enough for a byte loop in chunks
    for b in a.tobytes():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


def test_container_index_math_matches_reference(binary):
    rc, out, err = _run(binary, "index")
    assert rc == 0, err
    have_ref = oracle.ref() is not None
    n_checked = 0
    for line in out.splitlines():
        m = re.match(r"layout D=(\d+) lo=(\d+) hi=(\d+) dims ([\d ]+) stride ([\d ]+) total (\d+)", line)
        if m:
            lo, hi = int(m.group(2)), int(m.group(3))
            dims = [int(v) for v in m.group(4).split()]
            stride = [int(v) for v in m.group(5).split()]
            st, total = oracle.grid_stride_total(dims, lo, hi)
            assert list(st) == stride and total == int(m.group(6)), line
            if have_ref:
                r, rst, rtot, _ = oracle.ref_index(dims, lo, hi, None, False)
                if r >= 0:
                    assert list(rst) == stride and rtot == total, line
            n_checked += 1
            continue
        m = re.match(r"(at2|op3) lo=(\d+) hi=(\d+) dims ([\d ]+) idx ([-\d ]+) off (-?\d+)", line)
        if m:
            checked = m.group(1) == "at2"
            lo, hi = int(m.group(2)), int(m.group(3))
            dims = [int(v) for v in m.group(4).split()]
            idx = [int(v) for v in m.group(5).split()]
            off = int(m.group(6))
            want = oracle.grid_linearize(dims, lo, hi, idx, checked)
            assert (want if want is not None else -1) == off, line
            if have_ref:
                r, _, _, roff = oracle.ref_index(dims, lo, hi, idx, checked)
                if r == 0:
                    assert roff == off, line
                elif r > 0:  # the reference threw std::out_of_range
                    assert off == -1, line
            n_checked += 1
    assert n_checked > 500
    assert "value_initialised 1" in out


def test_csr_builders_match_oracle(binary):
    rc, out, err = _run(binary, "csr")
    assert rc == 0, err
    rows = {m.group(1): m.groups() for m in re.finditer(
        r"csr (\S+) nrows (\d+) nnz (\d+) row_ptr (\w+) col (\w+) val (\w+) bytes (\d+)", out)}
    rp, col, val = oracle.laplace2d(64, 48)
    g = rows["laplace2d_64x48_f64_i32"]
    assert int(g[1]) == 64 * 48 and int(g[2]) == len(col) == 5 * 64 * 48 - 2 * 64 - 2 * 48
    assert g[3] == _fnv1a(rp.astype(np.int32)) and g[4] == _fnv1a(col) and g[5] == _fnv1a(val)
    assert int(g[6]) == 12 * len(col) + 20 * 64 * 48 + 4
    g = rows["laplace2d_1024_f64_i32"]  # C1: nnz = 5n^2 - 4n, bytes per SURVEY.md §8d
    assert int(g[2]) == 5238784 and int(g[6]) == 83836928 + 4
    rp, col, val = oracle.op27(12)
    g = rows["op27_12_f64_i64"]
    assert int(g[2]) == (3 * 12 - 2) ** 3
    assert g[3] == _fnv1a(rp.astype(np.int64)) and g[4] == _fnv1a(col) and g[5] == _fnv1a(val)
    assert "triplets nnz 4 rp 0 2 3 4 col 0 3 0 1 val 4 2 -1 1.75" in out
    assert "triplets_oob out_of_range" in out
    # partition helpers of include/hpcx/Distributed.hpp against the oracle's partitioner
    rp, col, val = oracle.laplace2d(10, 10)
    m = re.search(r"partition_rows ([\d ]+) \| by_nnz ([\d ]+) \| slabs ([\d ]+)", out)
    assert m, out
    assert [int(v) for v in m.group(1).split()] == list(oracle.partition_rows(rp, 100, 3, 0))
    assert [int(v) for v in m.group(2).split()] == list(oracle.partition_rows(rp, 100, 3, 1))
    assert [int(v) for v in m.group(3).split()] == list(oracle.partition_rows(None, 13, 8, 0))  # static chunking


@pytest.mark.gpu
def test_cpp_entry_points_on_gpu(binary):
    rc, out, err = _run(binary, "gpu", 600)
    assert rc == 0 and "CPP GPU OK" in out, out + err
    assert "box axis 0 bit_exact 1" in out and "box axis 1 bit_exact 1" in out
    assert "jacobi7 bit_exact 1" in out
    assert "jacobi7_temporal bit_exact 1" in out
    assert "box_radius system_error" in out and "spmv_short_x out_of_range" in out
    assert "managed_vectors bit_exact 1" in out
    assert "conv2d bit_exact 1" in out and "transpose bit_exact 1" in out
    assert "from_coo product bit_exact 1" in out and "converged 1" in out
    assert "grid_bridge ok 1" in out
    assert "distributed_face ok 1" in out


@pytest.mark.gpu
def test_cpp_single_process_multi_gpu(binary):
    """hpcx::MultiContext from C++: one process, one thread, every visible GPU."""
    rc, out, err = _run(binary, "multi", 600)
    assert rc == 0 and "CPP MULTI OK" in out, out + err
    assert "multi box axis 0 bit_exact 1" in out and "multi box axis 1 bit_exact 1" in out
    assert "multi jacobi7 bit_exact 1" in out and "multi spmv bit_exact 1" in out
