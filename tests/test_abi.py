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


def test_ctypes_mirrors_match_the_header_layout(tmp_path):
    """every struct of include/spif_b200.h: field names, order, offsets and size as gcc lays them out == the ctypes mirrors the
    Python host passes through the C-ABI (a field appended on one side only -- e.g. spif_shard_comm.mc, ABI 4 -- shows up here)"""
    from sparkinfer_b200 import _lib
    from sparkinfer_b200.balancer import Move
    mirrors = {"spif_ffn_args": _lib.FfnArgs, "spif_pred_args": _lib.PredArgs, "spif_layer_io": _lib.LayerIo,
               "spif_shard_comm": _lib.ShardCommStruct, "spif_move": Move, "spif_gguf_tensor": _lib.GgufTensor}
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    structs = {}
    for name, body in re.findall(r"typedef struct (spif_[a-z_]+) \{(.*?)\}\s*\1;", src, flags=re.S):
        fields = []
        for decl in body.split(";"):
            for part in decl.split(","):  # "int32_t rank, world" declares two
                m = re.search(r"([A-Za-z_][A-Za-z0-9_]*)\s*(\[[^\]]*\])?\s*$", part.strip())
                if m and part.strip():
                    fields.append(m.group(1))
        structs[name] = fields
    assert set(structs) == set(mirrors), (sorted(structs), sorted(mirrors))
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {']
    for name, fields in structs.items():
        lines.append(f'  printf("{name} size %zu\\n", sizeof({name}));')
        for f in fields:
            lines.append(f'  printf("{name} {f} %zu\\n", offsetof({name}, {f}));')
    lines += ['  printf("abi %d\\n", SPIF_B200_ABI_VERSION);', '  return 0;', '}']
    c = tmp_path / "layout.c"
    c.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-o", str(exe), str(c)], check=True)
    got = {}
    for ln in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines():
        parts = ln.split()
        got[tuple(parts[:-1])] = int(parts[-1])
    for name, fields in structs.items():
        cls = mirrors[name]
        assert [f[0] for f in cls._fields_] == fields, (name, [f[0] for f in cls._fields_], fields)
        assert ctypes.sizeof(cls) == got[(name, "size")], name
        for f in fields:
            assert getattr(cls, f).offset == got[(name, f)], (name, f)
    assert _lib.load().spif_abi_version() == got[("abi",)]
