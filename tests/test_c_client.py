"""CPU: a plain C99 program (tests/c/host_abi.c) compiled by gcc against include/spif_b200.h and linked with libspif_b200.so --
the boundary as a C / cgo / JNI host sees it, no Python in between.  It drives the host-only entry points for real (model file,
model split, load-time budget + reorder, placement, balancer state) and checks that compute entry points refuse to run without
a device."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_client_builds_and_runs(tmp_path):
    exe = tmp_path / "host_abi"
    libdir = os.path.join(ROOT, "sparkinfer_b200")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "host_abi.c"), "-o", str(exe), "-L", libdir, "-lspif_b200", f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([str(exe), os.path.join(ROOT, "tests", "golden", "model_tiny_ref.gguf"), str(tmp_path / "ms.gguf")],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "host_abi ok" in r.stdout, r.stdout + r.stderr
