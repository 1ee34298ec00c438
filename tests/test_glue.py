"""The ggml-cuda glue (glue/ggml_cuda_spif.cu, shown in INTEGRATION.md) must compile against the reference's own headers:
every ggml_tensor / ggml_backend_cuda_context field it reads exists there with that type.  CPU test (nvcc cross-compiles);
skipped where /root/reference is absent (the GPU box)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("SPIF_REFERENCE", "/root/reference")


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "ggml", "src", "ggml-cuda", "common.cuh")), reason="reference tree absent")
def test_glue_compiles_against_reference_headers(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    g = os.path.join(REF, "ggml")
    cmd = [nvcc, "-std=c++17", "-gencode", "arch=compute_100,code=sm_100", "-c", os.path.join(ROOT, "glue", "ggml_cuda_spif.cu"),
           "-o", str(tmp_path / "glue.o"), "-I", os.path.join(g, "include"), "-I", os.path.join(g, "src"),
           "-I", os.path.join(g, "src", "ggml-cuda"), "-I", os.path.join(ROOT, "include"), "-DGGML_USE_CUDA",
           "-DGGML_CUDA_PEER_MAX_BATCH_SIZE=128", "-extended-lambda", "-w"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-4000:]
    syms = subprocess.run(["nm", "-C", str(tmp_path / "glue.o")], capture_output=True, text=True).stdout
    for name in ("spif_glue_mul_mat_sparse", "spif_glue_axpy_sparse", "spif_glue_fatrelu", "spif_glue_shifted_step",
                 "spif_glue_try_fuse_sparse_ffn", "spif_glue_lookahead"):
        assert name in syms, name
    for name in ("spif_mul_mat_sparse", "spif_axpy_sparse", "spif_sparse_ffn", "spif_sparse_ffn_lookahead"):
        assert f"U {name}" in syms, f"glue must bind the C-ABI symbol {name}"
