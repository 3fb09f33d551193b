"""In-tree build of libhpcx.so (sm_100a only) with plain nvcc.

    python -m libhpc_b200.build            # incremental
    python -m libhpc_b200.build --force    # rebuild everything

The shared library lands next to this file (libhpc_b200/libhpcx.so) so that it travels
with the repository snapshot to the GPU box; objects go to libhpc_b200/csrc/build/.
There is no JIT and no torch extension: the product is a C-ABI library (include/hpcx.h).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "libhpcx.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

CU_SOURCES = ["context.cu", "grid.cu", "spmv.cu", "spmv_hub.cu", "csr_plan.cu", "sell_build.cu", "stencil.cu", "dist.cu", "multi.cu", "assemble.cu", "cg.cu", "dense2d.cu"]
CXX_SOURCES = ["partition.cpp"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--fmad=false",  # arithmetic contract: separately rounded mul/add (see oracle/)
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-Wall,-Wno-unknown-pragmas",
    "-Xptxas", "-v",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-ffp-contract=off"]


def _deps_newer(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _headers() -> list[str]:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".hpp"))]
    hs.append(os.path.join(HERE, "..", "include", "hpcx.h"))
    hs.append(os.path.abspath(__file__))
    return hs


def _run(cmd: list[str], log: str | None = None) -> None:
    p = subprocess.run(cmd, capture_output=True, text=True)
    if log:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + p.stdout + p.stderr)
    if p.returncode != 0:
        sys.stderr.write(p.stdout + p.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs = []
    objs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _deps_newer(o, [s] + hdrs):
            jobs.append(([NVCC] + NVCC_FLAGS + ["-c", s, "-o", o], o + ".log"))
    for src in CXX_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _deps_newer(o, [s] + hdrs):
            jobs.append((["g++"] + CXX_FLAGS + ["-I/usr/local/cuda/include", "-c", s, "-o", o],
                         o + ".log"))
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _run(*j), jobs))
    if jobs or force or _deps_newer(LIB, objs):
        _run([NVCC, "-shared", "-o", LIB] + objs + ["-ldl", "-lpthread"])
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
