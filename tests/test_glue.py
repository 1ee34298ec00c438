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


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "ggml", "src", "ggml-cuda", "ggml-cuda.cu")), reason="reference tree absent")
def test_reference_tree_patch_applies_and_compiles(tmp_path):
    """glue/reference_tree.patch (the patched reference tree): applies cleanly to the reference's ggml-cuda.cu
    and CMakeLists.txt, and the patched dispatch file compiles with -DGGML_CUDA_SPIF_B200, leaving exactly the glue's entry
    points undefined; without the define the file is the reference's own."""
    d = tmp_path / "tree" / "ggml" / "src" / "ggml-cuda"
    d.mkdir(parents=True)
    for f in ("ggml-cuda.cu", "CMakeLists.txt"):
        shutil.copy(os.path.join(REF, "ggml", "src", "ggml-cuda", f), d / f)
    r = subprocess.run(["patch", "-p1", "--no-backup-if-mismatch", "-i", os.path.join(ROOT, "glue", "reference_tree.patch")],
                       cwd=tmp_path / "tree", capture_output=True, text=True)
    assert r.returncode == 0 and "FAILED" not in r.stdout and "fuzz" not in r.stdout, r.stdout + r.stderr
    cm = (d / "CMakeLists.txt").read_text()
    assert "GGML_CUDA_SPIF_B200_DIR" in cm and "glue/ggml_cuda_spif.cu" in cm and "spif_b200" in cm
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    g = os.path.join(REF, "ggml")
    cmd = [nvcc, "-std=c++17", "-gencode", "arch=compute_100,code=sm_100", "-c", str(d / "ggml-cuda.cu"), "-o",
           str(tmp_path / "ggml-cuda.o"), "-I", os.path.join(g, "include"), "-I", os.path.join(g, "src"),
           "-I", os.path.join(g, "src", "ggml-cuda"), "-DGGML_USE_CUDA", "-DGGML_CUDA_USE_GRAPHS", "-DGGML_CUDA_SPIF_B200",
           "-DGGML_CUDA_PEER_MAX_BATCH_SIZE=128", "-w"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-4000:]
    syms = subprocess.run(["nm", "-C", str(tmp_path / "ggml-cuda.o")], capture_output=True, text=True).stdout
    for name in ("spif_glue_mul_mat_sparse", "spif_glue_axpy_sparse", "spif_glue_fatrelu", "spif_glue_shifted_step",
                 "spif_glue_try_fuse_sparse_ffn"):
        assert f"U {name}(" in syms, name
    # the reference's own sparse kernels are no longer reached from the dispatch
    assert "ggml_cuda_mul_mat_sparse_q" not in syms and "ggml_cuda_op_fatrelu" not in syms


# ---- the glue, executed on a GPU (oracle/glue_harness.cu): the reference's graph, context and pool; its kernels beside ours
GLUETEST = os.path.join(ROOT, "oracle", "_ref", "libspif_gluetest.so")


@pytest.mark.gpu
@pytest.mark.parametrize("arch,wtype", [(0, 1), (1, 1), (2, 1), (0, 8)])   # F16 for the three recipes, Q8_0 for ProSparse
@pytest.mark.parametrize("n_embd,n_ff", [(1024, 4096), (5120, 13824)])
def test_glue_runs_reference_graph_on_gpu(arch, wtype, n_embd, n_ff):
    """One layer of build_sparse_ffn's graph (built with the reference's constructors, in its node order) three ways: the
    reference's CUDA kernels node by node, the same loop with the sparse ops routed through glue/ggml_cuda_spif.cu, and the
    loop with the fusion hook (one spif_sparse_ffn call).  Tolerances: against the reference kernels 2e-3 (they accumulate
    in half2 at GGML_PREC_DEFAULT, mm-sparse.cu:355-361, and add with float atomics); against the CPU oracle 1e-3.
    Q8_0: same activation quantiser as the reference's CUDA path (quantize.cu) -> 1e-4 against its kernels, the quant format's
    bound 2e-2 against the CPU ops (which quantise x with round-half-even of x * id)."""
    import ctypes

    import numpy as np
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    if not os.path.exists(GLUETEST):
        pytest.skip("oracle/_ref/libspif_gluetest.so not built (needs /root/reference at build time)")
    import synth
    from oracle_lib import Oracle
    tol_ref, tol_cpu = (2e-3, 1e-3) if wtype == 1 else (1e-4, 2e-2)
    lib = ctypes.CDLL(GLUETEST)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.spif_gluetest_layer.restype = ci
    lib.spif_gluetest_layer.argtypes = [ci, ci] + [vp] * 5 + [ci, ci] + [vp] * 5 + [ctypes.POINTER(ci), ctypes.POINTER(ci)]
    o = Oracle()
    seed = synth.SEED0 + 5 * n_embd + n_ff + arch
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (o.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, 1, seed + 1)
    p = synth.direct_mask(n_ff, 1, 0.10, seed + 2)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    du, dg, dd, dx, dp = dev(qu), dev(qg), dev(qd), dev(x[0]), dev(p[0])
    up_ref, up_glue = torch.empty(n_ff, device="cuda"), torch.empty(n_ff, device="cuda")
    y_ref, y_ops, y_fused = (torch.empty(n_embd, device="cuda") for _ in range(3))
    n_total, n_fused = ci(0), ci(0)
    ptr = lambda t: vp(t.data_ptr())
    rc = lib.spif_gluetest_layer(arch, wtype, ptr(du), ptr(dg), ptr(dd), ptr(dx), ptr(dp), n_embd, n_ff, ptr(up_ref), ptr(up_glue),
                                 ptr(y_ref), ptr(y_ops), ptr(y_fused), ctypes.byref(n_total), ctypes.byref(n_fused))
    assert rc == 0
    torch.cuda.synchronize()
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
    up_ref, up_glue, y_ref, y_ops, y_fused = (t.cpu().numpy() for t in (up_ref, up_glue, y_ref, y_ops, y_fused))
    act = ~(p[0] < 0.5)
    assert np.all(up_glue[~act] == 0.0) and np.all(up_ref[~act] == 0.0)          # dense, zero-filled dst (mm-sparse.cu:397)
    assert rel(up_glue, up_ref) < tol_ref, rel(up_glue, up_ref)
    assert rel(y_ops, y_ref) < tol_ref, rel(y_ops, y_ref)
    assert rel(y_fused, y_ref) < tol_ref, rel(y_fused, y_ref)
    assert n_total.value == {0: 5, 1: 6, 2: 3}[arch] and n_fused.value == n_total.value  # the whole layer became one call
    y_cpu = o.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff)[0][0]
    assert rel(y_ops, y_cpu) < tol_cpu and rel(y_fused, y_cpu) < tol_cpu
    assert rel(up_glue, o.mul_mat_sparse(wtype, qu, x, p, n_embd, n_ff)[0]) < tol_cpu


@pytest.mark.gpu
@pytest.mark.parametrize("n_embd,r,n_ff", [(1024, 256, 4096), (5120, 1024, 13824)])
def test_glue_lookahead_chain_on_gpu(n_embd, r, n_ff):
    """spif_glue_lookahead (the single-launch fast path) on a two-layer chain of the reference's tensors: predictor(0); FFN(0) +
    predictor(1) from x_0; FFN(1) -- against build_predictor's and build_sparse_ffn's nodes run through the reference's own CUDA
    kernels in the reference's order.  Masks may differ only where the probability is within 1e-3 of the threshold (the reference's
    predictor runs mul_mat_vec_f with half2 products); if they agree, y must agree to 2e-3."""
    import ctypes

    import numpy as np
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    if not os.path.exists(GLUETEST):
        pytest.skip("oracle/_ref/libspif_gluetest.so not built (needs /root/reference at build time)")
    import synth
    from oracle_lib import F16, Oracle
    lib = ctypes.CDLL(GLUETEST)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    lib.spif_gluetest_lookahead.restype = ci
    lib.spif_gluetest_lookahead.argtypes = [ci] + [vp] * 7 + [ci, ci, ci] + [vp] * 4
    o = Oracle()
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    keep, cols = [], [[] for _ in range(6)]
    for l in range(2):
        seed = synth.SEED0 + 9 * n_embd + n_ff + 100 * l
        Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
        wpu, wpd, bpd = synth.predictor_weights(n_embd, r, n_ff, 0.1, seed + 3)
        ts = [dev(o.quantize(F16, w)) for w in (Wu, Wg, Wd, wpu, wpd)] + [dev(bpd)]
        keep.append(ts)
        for k, t in enumerate(ts):
            cols[k].append(t.data_ptr())
    arrs = [(vp * 2)(*c) for c in cols]
    x = dev(np.concatenate([synth.activations(n_embd, 1, synth.SEED0 + 77 + l) for l in range(2)]))
    p_ref, p_glue = torch.zeros((2, n_ff), device="cuda"), torch.zeros((2, n_ff), device="cuda")
    y_ref, y_glue = torch.zeros((2, n_embd), device="cuda"), torch.zeros((2, n_embd), device="cuda")
    ptr = lambda t: vp(t.data_ptr())
    rc = lib.spif_gluetest_lookahead(F16, *[ctypes.cast(a, vp) for a in arrs], ptr(x), n_embd, n_ff, r, ptr(p_ref), ptr(p_glue),
                                     ptr(y_ref), ptr(y_glue))
    assert rc == 0, rc
    torch.cuda.synchronize()
    p_ref, p_glue, y_ref, y_glue = (t.cpu().numpy() for t in (p_ref, p_glue, y_ref, y_glue))
    assert np.all(np.isfinite(p_glue)) and np.all(np.isfinite(y_glue))
    assert np.abs(p_glue - p_ref).max() < 2e-3
    flips = (p_glue < 0.5) != (p_ref < 0.5)
    assert np.all(np.abs(p_ref[flips] - 0.5) < 1e-3), int(flips.sum())
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
    for l in range(2):
        if not flips[l].any():
            assert rel(y_glue[l], y_ref[l]) < 2e-3, (l, rel(y_glue[l], y_ref[l]))
    assert 0.03 < float((~(p_glue < 0.5)).mean()) < 0.3       # a real mask, not all-or-nothing
