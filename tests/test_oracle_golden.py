"""CPU: pin the oracle (oracle/spif_oracle.c) against outputs of the reference itself.

1. tests/golden/ref_cpu_golden.npz -- outputs of the UNMODIFIED reference CPU ops (made by tests/golden/make_golden.py
   from oracle/_ref/libspif_ref.so); checked everywhere, including the GPU box where /root/reference is absent.
2. oracle/_ref live (when it was built in this container): more shapes, thread counts, neuron_mask, biases.
The reference ships no golden vectors for this path; these are the pin.
"""
import os
import zlib

import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, BF16, F16, Q4_0, Q8_0, TYPE_NAMES

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_cpu_golden.npz")


def crc(a):
    return np.uint32(zlib.crc32(np.ascontiguousarray(a).tobytes()))


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


# BF16: the reference's mul+add may or may not be contracted by the compiler -> ~1e-6, everything else bit-exact
def same(a, b, t):
    return np.array_equal(a, b) if t != BF16 else rel_err(a, b) < 2e-6


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("t", [F16, BF16, Q8_0, Q4_0])
@pytest.mark.parametrize("ci", range(4))
def test_oracle_matches_golden_ffn(oracle, golden, t, ci):
    k = f"{TYPE_NAMES[t]}_{ci}"
    tt, E, F, T, seed = (int(v) for v in golden[k + "_meta"])
    rho = float(golden[k + "_rho"])
    Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
    qu, qg, qd = (oracle.quantize(t, w) for w in (Wu, Wg, Wd))
    # the oracle's weight quantisers reproduce the reference's byte for byte
    assert np.array_equal(golden[k + "_wcrc"], np.array([crc(qu), crc(qg), crc(qd)], dtype=np.uint32))
    x = synth.activations(E, T, seed + 1)
    p = synth.direct_mask(F, T, rho, seed + 2, near=0.01)
    assert same(oracle.mul_mat_sparse(t, qu, x, p, E, F), golden[k + "_up"], t)
    if t == Q4_0:
        return  # no reference axpy for Q4_0 (ggml-cpu.c:2225-2226): parity unpinned, see test_q4_0_axpy_bound
    for arch in (ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP):
        y, _, _, hid = oracle.sparse_ffn(arch, t, qu, qg, qd, x, p, E, F)
        assert same(hid, golden[f"{k}_hid{arch}"], t)
        assert same(y, golden[f"{k}_y{arch}"], t), rel_err(y, golden[f"{k}_y{arch}"])


@pytest.mark.parametrize("pt", [F16, BF16])
@pytest.mark.parametrize("ci", range(2))
def test_oracle_matches_golden_predictor(oracle, golden, pt, ci):
    k = f"pred_{TYPE_NAMES[pt]}_{ci}"
    _, E, R, F, T, seed = (int(v) for v in golden[k + "_meta"])
    wpu, wpd, bpd = synth.predictor_weights(E, R, F, 0.1, seed)
    bpu = np.random.default_rng(seed + 5).normal(0, 0.05, R).astype(np.float32)
    qpu, qpd = oracle.quantize(pt, wpu), oracle.quantize(pt, wpd)
    assert np.array_equal(golden[k + "_wcrc"], np.array([crc(qpu), crc(qpd)], dtype=np.uint32))
    x = synth.activations(E, T, seed + 1)
    p, _ = oracle.predictor(pt, qpu, bpu, qpd, bpd, x, E, R, F)
    assert same(p, golden[k + "_p"], pt)
    # the index lists the two produce are identical
    for t in range(T):
        assert np.array_equal(oracle.mask_compact(p[t]), np.nonzero(~(golden[k + "_p"][t] < 0.5))[0])


# ---- live reference (build container only) -------------------------------------------------------------------

@pytest.mark.parametrize("t", [F16, BF16, Q8_0, Q4_0])
@pytest.mark.parametrize("E,F,T", [(512, 1024, 1), (1376, 640, 5), (4096, 512, 2), (64, 33, 1)])
@pytest.mark.parametrize("n_threads", [1, 4])
def test_oracle_vs_live_reference(oracle, ref, t, E, F, T, n_threads):
    seed = synth.SEED0 + E + F + T
    Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
    qu, qg, qd = (oracle.quantize(t, w) for w in (Wu, Wg, Wd))
    assert np.array_equal(qu, ref.quantize(t, Wu))
    x = synth.activations(E, T, seed + 1)
    p = synth.direct_mask(F, T, 0.2, seed + 2, near=0.02)
    p[0, :3] = [np.nan, 0.5, np.nextafter(np.float32(0.5), np.float32(0))]  # NaN and == thr are active
    nm = (np.random.default_rng(seed).random(F) < 0.3).astype(np.int32)     # rows "on the GPU": skipped on the CPU
    for neuron_mask in (None, nm):
        a = oracle.mul_mat_sparse(t, qu, x, p, E, F, neuron_mask=neuron_mask)
        b = ref.mul_mat_sparse(t, qu, x, p, E, F, neuron_mask=neuron_mask, n_threads=n_threads)
        assert same(a, b, t)  # per-row dots do not depend on the thread count
        assert np.all(a[0, :2] != 0) or neuron_mask is not None
    if t == Q4_0:
        return
    hid = a.copy()
    hid[:, ::5] = 0.0
    ya = oracle.axpy_sparse(t, qd, hid, p, E, F)
    yb = ref.axpy_sparse(t, qd, hid, p, E, F, n_threads=n_threads)
    if n_threads == 1:
        assert same(ya, yb, t)
        return
    # n_threads > 1: summation order depends on the thread count (ggml-cpu.c:2298-2330).  The reference also has a
    # bug there: when ceil(n_ff / (nth*K)) * (nth*K - 1) >= n_ff some chunks are empty and the empty-chunk branch
    # (ggml-cpu.c:2306-2311) writes the thread's buffer -- possibly holding another token's partial sum -- into the
    # current token's dst without the lock (a lost-update race even for T == 1).  The multi-threaded reference is then
    # simply wrong, so the 1-thread order is the parity authority and those cases are not compared.
    cpt = n_threads * (64 if t == Q8_0 else 8)
    has_empty_chunks = -(-F // cpt) * (cpt - 1) >= F
    if not has_empty_chunks:
        assert rel_err(ya, yb) < 1e-5


@pytest.mark.parametrize("arch", [ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP])
def test_oracle_vs_live_reference_biases(oracle, ref, arch):
    E, F, T, t = 256, 512, 2, F16
    seed = synth.SEED0 + 99 + arch
    Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
    qu, qg, qd = (oracle.quantize(t, w) for w in (Wu, Wg, Wd))
    rng = np.random.default_rng(seed)
    bu, bg, bd = (rng.normal(0, 0.05, n).astype(np.float32) for n in (F, F, E))
    x = synth.activations(E, T, seed + 1)
    p = synth.direct_mask(F, T, 0.2, seed + 2)
    ya = oracle.sparse_ffn(arch, t, qu, qg, qd, x, p, E, F, b_up=bu, b_gate=bg, b_down=bd)
    yb = ref.sparse_ffn(arch, t, qu, qg, qd, x, p, E, F, b_up=bu, b_gate=bg, b_down=bd)
    for a, b in zip(ya, yb):
        assert np.array_equal(a, b)


# ---- restatements with no reference code: checked against the FP64 restatement / their definition ---------------

def test_q4_0_axpy_bound(oracle):
    """Q4_0 axpy: restated from the Q8_0 axpy with Q4_0 dequant (parity unpinned); must equal the dequantised-weights
    FP64 sum to F32 rounding."""
    E, F, T = 256, 512, 2
    seed = synth.SEED0 + 4
    _, _, Wd = synth.ffn_weights(E, F, seed)
    qd = oracle.quantize(Q4_0, Wd)
    hid = np.random.default_rng(seed).normal(0, 1, (T, F)).astype(np.float32)
    p = synth.direct_mask(F, T, 0.3, seed + 2)
    y = oracle.axpy_sparse(Q4_0, qd, hid, p, E, F)
    W = oracle.dequantize(Q4_0, qd, F, E).astype(np.float64)
    act = ~(p < 0.5)
    y64 = (hid.astype(np.float64) * act) @ W
    assert rel_err(y, y64) < 1e-5
    assert np.abs(W - Wd).max() <= np.abs(Wd).max() / 8 * 1.01 + 1e-9  # |dw| <= d (d/2 away from the clamp at q = 15)


def test_topk_mask_definition(oracle):
    rng = np.random.default_rng(5)
    s = rng.normal(0, 1, 1000).astype(np.float32)
    s[10] = s[20] = s[30] = np.float32(0.77)  # ties -> lower index first
    for k in (0, 1, 17, 500, 1000, 2000):
        m = oracle.topk_mask(s, k)
        order = np.lexsort((np.arange(s.size), -s.astype(np.float64)))
        want = np.zeros(s.size, np.float32)
        want[order[:min(k, s.size)]] = 1.0
        assert np.array_equal(m, want)


def test_mask_compact_definition(oracle):
    p = synth.direct_mask(4097, 1, 0.1, 3, near=0.05)[0]
    p[:4] = [np.nan, 0.5, 0.49999997, -0.0]
    idx = oracle.mask_compact(p)
    assert np.array_equal(idx, np.nonzero(~(p < np.float32(0.5)))[0])
    assert idx[0] == 0 and idx[1] == 1 and 2 not in idx


def test_f64_restatement_close(oracle):
    E, F, T = 512, 1024, 2
    seed = synth.SEED0 + 11
    Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
    x = synth.activations(E, T, seed + 1)
    p = synth.direct_mask(F, T, 0.15, seed + 2)
    for t, tol in ((F16, 1e-3), (BF16, 8e-3), (Q8_0, 2e-2)):
        qu, qg, qd = (oracle.quantize(t, w) for w in (Wu, Wg, Wd))
        y = oracle.sparse_ffn(ARCH_PROSPARSE, t, qu, qg, qd, x, p, E, F)[0]
        y64, _ = oracle.sparse_ffn_f64(ARCH_PROSPARSE, t, qu, qg, qd, x, p, E, F)
        assert rel_err(y, y64) < tol


def test_rms_norm_restatement_matches_reference():
    """the numpy restatement of build_norm(LLM_NORM_RMS) (rms_norm + mul) is bit-exact against the reference's CPU ops"""
    import oracle_lib
    if not oracle_lib.Ref.available():
        pytest.skip("oracle/_ref not built")
    ref = oracle_lib.Ref()
    if ref._rms is None:
        pytest.skip("oracle/_ref predates spif_ref_rms_norm_mul: rebuild it")
    rng = np.random.default_rng(41)
    for n_embd, n_tok, eps in ((512, 3, 1e-5), (5120, 2, 1e-6), (1376, 1, 1e-5), (8192, 1, 0.0)):
        x = (rng.normal(0, 1, (n_tok, n_embd)) * rng.choice([0.01, 1.0, 30.0])).astype(np.float32)
        w = rng.normal(1, 0.2, n_embd).astype(np.float32)
        assert np.array_equal(oracle_lib.rms_norm_mul(x, w, eps), ref.rms_norm_mul(x, w, eps)), (n_embd, n_tok)
