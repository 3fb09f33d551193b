"""The C-ABI library loads and exports exactly what include/hpcx.h declares (no compute)."""
import os
import re
import subprocess

import pytest

from libhpc_b200 import _abi

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _declared():
    src = open(os.path.join(ROOT, "include", "hpcx.h")).read()
    return sorted(set(re.findall(r"HPCX_API[^;(]*?\b(hpcx_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_typed():
    lib = _abi.load()
    decl = _declared()
    assert len(decl) >= 40
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in hpcx.h but not exported"
    assert sorted(_abi.PROTOTYPES) == decl, "ctypes prototypes and hpcx.h disagree"


def test_no_undeclared_exports():
    out = subprocess.run(["nm", "-D", "--defined-only", _abi.LIB_PATH], capture_output=True,
                         text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and "hpcx_" in l)
    assert exported == _declared()


def test_header_compiles_as_c():
    p = subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c",
                        os.path.join(ROOT, "include", "hpcx.h")], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr


def test_sass_has_tma_and_no_legacy_tensor_path():
    """TMA bulk copies (UBLKCP) and tensor loads (UTMALDG) are in the sm_100a SASS."""
    p = subprocess.run(["cuobjdump", "-sass", _abi.LIB_PATH], capture_output=True, text=True)
    if p.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in p.stdout
    assert "UBLKCP" in p.stdout and "UTMALDG" in p.stdout


def test_version_string():
    assert b"sm_100a" in _abi.load().hpcx_version()


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "libhpc_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "liboracle" not in text and "hpcx_oracle.c\"" not in text, f
