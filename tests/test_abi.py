"""CPU: the C-ABI library loads and exports every symbol include/spif_b200.h declares; argument errors are returned
as codes; compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "spif_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spif_[a-z0-9_]+)\s*\(", src)))


def test_header_compiles_as_c():
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER], check=True)


def test_library_exports_every_declared_symbol():
    from sparkinfer_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/spif_b200.h but not exported"
    # and the Python mirror binds all of them
    assert set(syms) <= set(_lib.SYMBOLS), set(syms) - set(_lib.SYMBOLS)
    L = _lib.load()
    assert L.spif_abi_version() >= 1
    assert L.spif_row_size(1, 5120) == 10240 and L.spif_row_size(8, 5120) == 5440 and L.spif_row_size(2, 5120) == 2880
    assert L.spif_workspace_size(5120, 13824, 1) > 0
    assert b"workspace" in L.spif_error_string(-4)


def test_no_cpu_fallback():
    """without a CUDA device compute entry points return an error instead of computing on the host"""
    torch = pytest.importorskip("torch")
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from sparkinfer_b200 import _lib
    L = _lib.load()
    n = ctypes.c_int(0)
    assert L.spif_device_sm_count(ctypes.byref(n)) != 0
    buf = (ctypes.c_float * 64)()
    rc = L.spif_fatrelu(buf, buf, 64, 0.01, None)
    assert rc != 0


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sparkinfer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                s = open(os.path.join(dp, f)).read()
                assert "oracle" not in s.lower() or f == "__init__.py" and "oracle" not in s, (dp, f)
