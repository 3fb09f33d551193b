"""Builds tests/cpp/rootgrid_check against the REFERENCE's lib/sparse headers (oracle/tbb_shim stands
in for oneTBB) and records what the true sparse::RootGrid::foreach enumerates for the reference's
line-writing workload (test_hpc_benchmark.cpp:859-880) in tests/golden/rootgrid_cells.json.
Run in the authoring container (needs /root/reference); the fixture and the built binary travel."""
import json
import os
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
REF = "/root/reference/lib/sparse/include"
SRC = os.path.join(ROOT, "tests", "cpp", "rootgrid_check.cpp")
BIN = os.path.join(ROOT, "tests", "cpp", "rootgrid_check")


def build():
    lib = os.path.join(ROOT, "libhpc_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-fopenmp", "-ffp-contract=off", "-Wall", "-Wno-unknown-pragmas", "-Wno-delete-non-virtual-dtor",
                    "-I" + os.path.join(ROOT, "oracle", "tbb_shim"), "-I" + REF, "-I" + os.path.join(ROOT, "include"),
                    SRC, "-L" + lib, "-lhpcx", "-Wl,-rpath,$ORIGIN/../../libhpc_b200", "-o", BIN], check=True)
    return BIN


def enumerate_cells():
    out = subprocess.run([BIN, "cells"], capture_output=True, text=True, check=True).stdout
    cells = [l.split()[1:] for l in out.splitlines() if l.startswith("cell ")]
    return [(int(x), int(y), float(v)) for x, y, v in cells]


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference tree absent")
    build()
    cells = enumerate_cells()
    leaves = sorted({(x >> 4, y >> 4) for x, y, _ in cells})
    written = [[x, y, v] for x, y, v in cells if v != 0.0]
    json.dump({"workload": "px=-100, py=100, vx=0.2, vy=-0.6, 400 steps, value 1 + 0.001*(t % 97), fp32 (rootgrid_check.cpp: fill)",
               "layout": "RootGrid<float, HashBlock<DenseBlock<16, float>>>", "ncells": len(cells), "leaves": leaves,
               "written": written}, open(os.path.join(ROOT, "tests", "golden", "rootgrid_cells.json"), "w"))
    print(len(cells), "cells,", len(leaves), "leaves,", len(written), "written")
