"""GPU parity, round-2 additions (VERDICT r1 "parity gaps"): the exact BASELINE.json shapes through spif_sparse_ffn, the
n_tok = 1 neuron_idx / MASK_BY_ROW / unaligned-mask branch of the in-kernel mask scan, a mask longer than 64 x threads
(second half of the packed activity bits), the balancer's only device kernel (spif_dfr_update) against a numpy
restatement of build_dfr, and the two unary ops against the reference formulas.  -m gpu, through the C-ABI."""
import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, BF16, F16, Q4_0, Q8_0, TYPE_NAMES

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = {F16: 1e-3, BF16: 8e-3}


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


# ---- the headline shapes (BASELINE.json configs[1], [4], [0]) + Qwen2-7B's n_ff (m > 64 * threads: bits[1] of the scan)
@pytest.mark.parametrize("wtype", [F16, BF16])
@pytest.mark.parametrize("n_embd,n_ff", [(5120, 13824), (8192, 28672), (4096, 11008), (3584, 18944)])
def test_sparse_ffn_fused_baseline_shapes(oracle, wtype, n_embd, n_ff):
    import sparkinfer_b200 as sp
    seed = synth.SEED0 + 17 * n_embd + n_ff
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, 1, seed + 1)
    p = synth.direct_mask(n_ff, 1, 0.10, seed + 2)
    ws = sp.Workspace(n_embd, n_ff, 1)
    y_ref, _, _, hid_ref = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, qu, qg, qd, x, p, n_embd, n_ff)
    y, hid, stats = sp.sparse_ffn(ARCH_PROSPARSE, wtype, dev(qu), dev(qg), dev(qd), dev(x), dev(p), n_ff, want_hidden=True,
                                  want_stats=True, ws=ws)
    act = ~(p < 0.5)
    assert rel_err(y.cpu().numpy(), y_ref) < TOL[wtype]
    assert rel_err(hid.cpu().numpy() * act, hid_ref * act) < TOL[wtype]
    assert int(stats.cpu().numpy()[0, 0]) == act.sum()


# the quant format's own bound: |dw| <= d/2 per weight with d = max|w| / 127 (Q8_0) or / 8 (Q4_0) per block, plus
# the 8-bit activation quantisation of the integer dot; for N(0, 0.02^2) weights and N(0,1) x this is ~0.4 % (Q8_0) and ~7 %
# (Q4_0) of a single product, averaged down by sqrt(n_embd) in a dot.  The CUDA path and the oracle see the SAME quantised
# weights, so what is compared here is only activation quantiser + summation order: the bars below are those of
# tests/test_gpu_parity.py (2e-4 with the CPU quantiser, 2e-2 with the CUDA reference's).
@pytest.mark.parametrize("wtype", [Q8_0, Q4_0])
@pytest.mark.parametrize("n_tok", [1, 8])
def test_quant_fused_13b_shape(oracle, wtype, n_tok):
    """BASELINE.json configs[2]: ProSparse-13B quantised, batch 1 and 8 (32 is in tests/test_gpu_batched.py)"""
    import sparkinfer_b200 as sp
    n_embd, n_ff = 5120, 13824
    seed = synth.SEED0 + 23 * n_tok + wtype
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, n_tok, seed + 1)
    p = synth.direct_mask(n_ff, n_tok, 0.10, seed + 2)
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    y_ref = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, qu, qg, qd, x, p, n_embd, n_ff)[0]
    y64 = oracle.sparse_ffn_f64(ARCH_PROSPARSE, wtype, qu, qg, qd, x, p, n_embd, n_ff)[0].astype(np.float32)
    y = sp.sparse_ffn(ARCH_PROSPARSE, wtype, dev(qu), dev(qg), dev(qd), dev(x), dev(p), n_ff, ws=ws).cpu().numpy()
    assert rel_err(y, y_ref) < 2e-2, (TYPE_NAMES[wtype], n_tok, rel_err(y, y_ref))
    assert rel_err(y, y64) < 2e-2
    if n_tok == 1:
        y = sp.sparse_ffn(ARCH_PROSPARSE, wtype, dev(qu), dev(qg), dev(qd), dev(x), dev(p), n_ff, flags=sp.FLAG_CPU_ROUNDING,
                          ws=ws).cpu().numpy()
        assert rel_err(y, y_ref) < 2e-4, (TYPE_NAMES[wtype], rel_err(y, y_ref))


# ---- n_tok = 1 with a row -> neuron map (GPU cache / shard rows): the scalar branch of build_my_list (GEMV, AXPY, FUSED)
@pytest.mark.parametrize("by_row", [False, True])
@pytest.mark.parametrize("wtype", [F16, Q8_0])
def test_neuron_idx_single_token(oracle, wtype, by_row):
    import sparkinfer_b200 as sp
    n_embd, n_ff, m = 1024, 4096, 1531 if wtype == F16 else 1536   # odd m: also the unaligned-mask branch
    seed = synth.SEED0 + 31 + wtype + by_row
    rng = np.random.default_rng(seed)
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    nidx = np.sort(rng.choice(n_ff, m, replace=False)).astype(np.int32)
    x = synth.activations(n_embd, 1, seed + 1)
    p_full = synth.direct_mask(n_ff, 1, 0.15, seed + 2)
    outside = np.ones(n_ff, bool)
    outside[nidx] = False
    p_oracle = p_full.copy()
    p_oracle[0, outside] = 0.1                       # neurons not held here contribute nothing
    p_gpu = p_full[:, nidx].copy() if by_row else p_full
    flags = sp.FLAG_MASK_BY_ROW if by_row else 0
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    qu_s, qg_s, qd_s = (oracle.quantize(wtype, w[nidx]) for w in (Wu, Wg, Wd))
    tol = 1e-3 if wtype == F16 else 2e-2
    ws = sp.Workspace(n_embd, n_ff, 1)
    # GEMV: dst is dense {n_ff}, scattered to neuron ids (mm-sparse.cu:20,101), zero elsewhere
    ref_u = oracle.mul_mat_sparse(wtype, qu, x, p_oracle, n_embd, n_ff)
    got_u = sp.mul_mat_sparse(dev(qu_s), wtype, dev(x), dev(p_gpu), neu_info=dev(nidx), n_ff=n_ff, flags=flags, ws=ws).cpu().numpy()
    assert rel_err(got_u, ref_u) < tol
    assert np.all(got_u[0, outside] == 0.0) and np.all(got_u[p_oracle < 0.5] == 0.0)
    # AXPY
    hid = (ref_u * rng.normal(0, 1, ref_u.shape)).astype(np.float32)
    hid[:, ::5] = 0.0
    ref_y = oracle.axpy_sparse(wtype, qd, hid, p_oracle, n_embd, n_ff)
    got_y = sp.axpy_sparse(dev(qd_s), wtype, dev(hid), dev(p_gpu), neu_info=dev(nidx), n_embd=n_embd, flags=flags, ws=ws).cpu().numpy()
    assert rel_err(got_y, ref_y) < tol
    # FUSED
    y_ref, _, _, hid_ref = oracle.sparse_ffn(ARCH_RELU_GATED, wtype, qu, qg, qd, x, p_oracle, n_embd, n_ff)
    y, hidg = sp.sparse_ffn(ARCH_RELU_GATED, wtype, dev(qu_s), dev(qg_s), dev(qd_s), dev(x), dev(p_gpu), n_ff, neuron_idx=dev(nidx),
                            flags=flags, want_hidden=True, ws=ws)
    act = ~(p_oracle < 0.5)
    assert rel_err(y.cpu().numpy(), y_ref) < tol
    assert rel_err(hidg.cpu().numpy() * act, hid_ref * act) < tol


# ---- spif_dfr_update against build_dfr (src/llama-graph.cpp:910-918) restated in numpy, EMA and non-EMA
@pytest.mark.parametrize("ema", [True, False])
@pytest.mark.parametrize("n_ff,g,n_tok", [(13824, 16, 1), (18944, 32, 3), (1024, 8, 5)])
def test_dfr_update_vs_build_dfr(n_ff, g, n_tok, ema):
    import sparkinfer_b200 as sp
    rng = np.random.default_rng(synth.SEED0 + n_ff + g)
    p = synth.direct_mask(n_ff, n_tok, 0.2, synth.SEED0 + n_ff, near=0.02)
    p[0, :6] = [0.5, np.nextafter(np.float32(0.5), np.float32(1)), np.nextafter(np.float32(0.5), np.float32(0)), np.nan, 1.0, 0.0]
    lam = np.float32(0.9)
    dfr0 = rng.random(n_ff // g).astype(np.float32)
    # reference: mask = shifted_step(sparse_idx, -0.5) = (p + (-0.5)) > 0 (unary.cu:611-613: strictly above; 0.5 and NaN are 0),
    # sum_cols over tokens, reshape {g, n_g}, sum_rows -> deltas; scale_add: l*a + [(1-l)*] (b / (n_tok*g)) (binbcast.cu:28-34)
    mask = ((p + np.float32(-0.5)) > 0).astype(np.float32)
    deltas = mask.sum(axis=0, dtype=np.float32).reshape(n_ff // g, g).sum(axis=1, dtype=np.float32)
    norm = np.float32(n_tok * g)
    ref = lam * dfr0 + ((np.float32(1) - lam) * (deltas / norm) if ema else deltas / norm)
    d = dev(dfr0.copy())
    sp.dfr_update(dev(p), d, g, float(lam), ema=ema)
    got = d.cpu().numpy()
    assert np.allclose(got, ref, rtol=2e-7, atol=1e-8), np.abs(got - ref).max()   # an FMA may replace mul + add


def test_unary_ops_vs_reference_formulas():
    """ggml_fatrelu: x > t ? x : 0 (unary.cu:566-568); ggml_shifted_step: (x + t) > 0 (unary.cu:611-613); bit-exact"""
    import sparkinfer_b200 as sp
    rng = np.random.default_rng(7)
    x = rng.normal(0, 0.05, 100003).astype(np.float32)
    x[:8] = [0.01, np.nextafter(np.float32(0.01), np.float32(1)), -0.0, 0.0, np.nan, np.inf, -np.inf, 0.5]
    for t in (0.01, 0.0, -0.5):
        tf = np.float32(t)
        ref = np.where(x > tf, x, np.float32(0))
        got = sp.fatrelu(dev(x), t).cpu().numpy()
        assert np.array_equal(got, ref, equal_nan=True)
        with np.errstate(invalid="ignore"):
            ref2 = ((x + tf) > 0).astype(np.float32)
        got2 = sp.shifted_step(dev(x), t).cpu().numpy()
        assert np.array_equal(got2, ref2)


# ---- more active neurons per CTA than shared-memory slots: the quantised warp-per-neuron kernel (spif_ffn_qwarp.cuh) works in
# rounds (gate/up prefetch of the next round, slot and barrier-phase reuse) -- 60 % activity gives ~33 neurons per CTA on 8 slots
@pytest.mark.parametrize("arch", [ARCH_PROSPARSE, ARCH_RELU_UP])
@pytest.mark.parametrize("wtype", [Q8_0, Q4_0])
def test_quant_many_rounds(oracle, wtype, arch):
    import sparkinfer_b200 as sp
    n_embd, n_ff, n_tok = 512, 8192, 2
    seed = synth.SEED0 + 77 + wtype + arch
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, n_tok, seed + 1)
    p = synth.direct_mask(n_ff, n_tok, 0.60, seed + 2)
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    gate = None if arch == ARCH_RELU_UP else dev(qg)
    y_ref, _, _, hid_ref = oracle.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff)
    y, hid, stats = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff,
                                  flags=sp.FLAG_CPU_ROUNDING | sp.FLAG_PER_TOKEN, want_hidden=True, want_stats=True, ws=ws)
    act = ~(p < 0.5)
    assert rel_err(y.cpu().numpy(), y_ref) < 2e-4
    assert rel_err(hid.cpu().numpy() * act, hid_ref * act) < 2e-4
    assert np.array_equal(stats.cpu().numpy()[:, 0], act.sum(axis=1))
    assert np.array_equal(stats.cpu().numpy()[:, 1], np.count_nonzero(hid.cpu().numpy(), axis=1))
    # the two single ops through the same kernel
    ref_u = oracle.mul_mat_sparse(wtype, qu, x, p, n_embd, n_ff)
    got_u = sp.mul_mat_sparse(dev(qu), wtype, dev(x), dev(p), flags=sp.FLAG_CPU_ROUNDING, ws=ws).cpu().numpy()
    assert rel_err(got_u, ref_u) < 3e-5
    hidv = ref_u.copy()
    hidv[:, ::3] = 0.0
    ref_y = oracle.axpy_sparse(wtype, qd, hidv, p, n_embd, n_ff)
    got_y = sp.axpy_sparse(dev(qd), wtype, dev(hidv), dev(p), n_embd=n_embd, ws=ws).cpu().numpy()
    assert rel_err(got_y, ref_y) < 3e-5
