"""GPU parity of the batched (n_tok >= 2) sparse FFN: union-of-masks gathered GEMM on tcgen05 tensor cores, against the
oracle (reference CPU semantics: x and hidden rounded to the weight float type, FP32 accumulation) and against the
per-token kernel on the same inputs.  -m gpu."""
import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, BF16, F16, TYPE_NAMES

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


# same bars as test_gpu_parity.py: 1e-3 of max|y| for F16 (north_star), BF16's own mantissa bound
TOL = {F16: 1e-3, BF16: 8e-3}
# (n_embd, n_ff, n_tok, rho): one tile / several tiles per CTA, 16- and 32-token accumulators, ragged union sizes
SHAPES = [(512, 1024, 2, 0.12), (1024, 2048, 8, 0.12), (5120, 13824, 8, 0.10), (5120, 13824, 32, 0.10),
          (4096, 1536, 17, 0.3), (8192, 640, 16, 0.5), (1024, 40000, 5, 0.9), (512, 384, 32, 0.02)]


@pytest.mark.parametrize("arch", [ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP])
@pytest.mark.parametrize("wtype", [F16, BF16])
@pytest.mark.parametrize("n_embd,n_ff,n_tok,rho", SHAPES)
def test_batched_sparse_ffn(oracle, arch, wtype, n_embd, n_ff, n_tok, rho):
    import sparkinfer_b200 as sp
    if n_ff > 20000 and (arch != ARCH_PROSPARSE or wtype != F16):
        pytest.skip("large case once")
    seed = synth.SEED0 + 11 * n_embd + n_ff + arch + n_tok
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, n_tok, seed + 1)
    p = synth.direct_mask(n_ff, n_tok, rho, seed + 2)
    rng = np.random.default_rng(seed + 5)
    b_up = rng.normal(0, 0.05, n_ff).astype(np.float32) if arch == ARCH_RELU_UP else None
    b_down = rng.normal(0, 0.05, n_embd).astype(np.float32) if arch == ARCH_RELU_UP else None
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    y_ref, _, _, hid_ref = oracle.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff, b_up=b_up, b_down=b_down)
    gate = None if arch == ARCH_RELU_UP else dev(qg)
    kw = dict(up_b=None if b_up is None else dev(b_up), down_b=None if b_down is None else dev(b_down), ws=ws)
    y, hid, stats = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff, want_hidden=True,
                                  want_stats=True, **kw)
    y, hid, stats = y.cpu().numpy(), hid.cpu().numpy(), stats.cpu().numpy()
    act = ~(p < 0.5)
    assert rel_err(y, y_ref) < TOL[wtype], (TYPE_NAMES[wtype], arch, rel_err(y, y_ref))
    assert rel_err(hid * act, hid_ref * act) < TOL[wtype]
    assert np.all(hid[~act] == 0.0)
    assert np.array_equal(stats[:, 0], act.sum(axis=1))          # per-token active counts
    assert np.all(stats[:, 2] == int(act.any(axis=0).sum()))     # union size
    # the per-token schedule (the reference's, mm-sparse.cu:107-210) on the same inputs
    if n_ff <= 20000:  # (the per-token kernel keeps its mask bits in registers: m <= 4096 x warps)
        y_tok = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff, flags=sp.FLAG_PER_TOKEN, **kw)
        assert rel_err(y, y_tok.cpu().numpy()) < TOL[wtype]
    # run-to-run bit stability (fixed stream-K partition and summation order)
    y2 = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff, **kw)
    assert np.array_equal(y, y2.cpu().numpy())


def test_batched_shard_rows_and_empty_union(oracle):
    """a shard (m < n_ff rows + neuron_idx, the reference's GPU cache layout) and an all-inactive batch"""
    import sparkinfer_b200 as sp
    n_embd, n_ff, n_tok = 1024, 4096, 8
    seed = synth.SEED0 + 99
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    x = synth.activations(n_embd, n_tok, seed + 1)
    p = synth.direct_mask(n_ff, n_tok, 0.2, seed + 2)
    rows = np.sort(np.random.default_rng(seed).permutation(n_ff)[: n_ff // 2]).astype(np.int32)
    qu, qg, qd = (oracle.quantize(F16, w[rows]) for w in (Wu, Wg, Wd))
    # oracle on the full matrices with the mask restricted to the shard's neurons
    p_sh = np.full_like(p, 0.1)
    p_sh[:, rows] = p[:, rows]
    fu, fg, fd = (oracle.quantize(F16, w) for w in (Wu, Wg, Wd))
    y_ref = oracle.sparse_ffn(ARCH_PROSPARSE, F16, fu, fg, fd, x, p_sh, n_embd, n_ff)[0]
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    y = sp.sparse_ffn(ARCH_PROSPARSE, F16, dev(qu), dev(qg), dev(qd), dev(x), dev(p), n_ff, neuron_idx=dev(rows), ws=ws)
    assert rel_err(y.cpu().numpy(), y_ref) < TOL[F16]
    y0 = sp.sparse_ffn(ARCH_PROSPARSE, F16, dev(fu), dev(fg), dev(fd), dev(x), dev(np.full_like(p, 0.1)), n_ff, ws=ws)
    assert float(y0.abs().max()) == 0.0


QSHAPES = [(512, 1024, 2, 0.12), (1024, 2048, 8, 0.12), (4096, 1536, 17, 0.3), (5120, 13824, 32, 0.10), (1024, 30000, 5, 0.9)]


@pytest.mark.parametrize("arch", [ARCH_PROSPARSE, ARCH_RELU_UP])
@pytest.mark.parametrize("wtype", [8, 2])  # Q8_0, Q4_0 (ggml type ids)
@pytest.mark.parametrize("n_embd,n_ff,n_tok,rho", QSHAPES)
def test_batched_quant_sparse_ffn(oracle, arch, wtype, n_embd, n_ff, n_tok, rho):
    """Q8_0 / Q4_0 union GEMM: raw blocks dequantised to F16 operand tiles in shared memory (w16 = RN(q * d)), F16 x.
    The reference quantises x to 8 bit instead (mmq-sparse.cu, ggml-cpu.c:1832-1856), so the bar is the quant format's
    own bound against both the oracle and the FP64 restatement on the same stored weights (Q4_0 axpy: parity unpinned,
    the reference aborts for it, ggml-cpu.c:2225-2226)."""
    import sparkinfer_b200 as sp
    if n_ff > 20000 and arch != ARCH_PROSPARSE:
        pytest.skip("large case once")
    seed = synth.SEED0 + 13 * n_embd + n_ff + arch + n_tok + wtype
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, n_tok, seed + 1)
    p = synth.direct_mask(n_ff, n_tok, rho, seed + 2)
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    y_ref = oracle.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff)[0]
    y64 = oracle.sparse_ffn_f64(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff)[0].astype(np.float32)
    gate = None if arch == ARCH_RELU_UP else dev(qg)
    y, stats = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff, want_stats=True, ws=ws)
    y, stats = y.cpu().numpy(), stats.cpu().numpy()
    act = ~(p < 0.5)
    assert rel_err(y, y64) < 2e-2, (wtype, arch, rel_err(y, y64))
    assert rel_err(y, y_ref) < 2e-2, (wtype, arch, rel_err(y, y_ref))
    assert np.array_equal(stats[:, 0], act.sum(axis=1))
    if n_tok > 12:  # the union GEMM (<= 12 tokens of quantised weights take the per-token warp-per-neuron kernel: 8-bit x)
        assert rel_err(y, y64) <= rel_err(y_ref, y64) + 1e-3  # F16 activations are no worse than the 8-bit ones
        assert np.all(stats[:, 2] == int(act.any(axis=0).sum()))
    y2 = sp.sparse_ffn(arch, wtype, dev(qu), gate, dev(qd), dev(x), dev(p), n_ff, ws=ws)
    assert np.array_equal(y, y2.cpu().numpy())


def test_shard_exchange_single_rank_self_exchange(oracle):
    """world = 1 exercises the whole peer protocol against this GPU's own buffers (slots, flags, sequence numbers, parity
    reuse over many calls): the fused partial-sum + exchange kernel and the stand-alone form must reproduce the plain
    layer bit for bit.  The 2-GPU run is tools/test_peer_exchange.py under torchrun."""
    import sparkinfer_b200 as sp
    n_embd, n_ff = 1024, 2048
    seed = synth.SEED0 + 5
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (dev(oracle.quantize(F16, w)) for w in (Wu, Wg, Wd))
    comm = sp.ShardComm(8 * n_embd)
    for n_tok in (1, 8):
        x = dev(synth.activations(n_embd, n_tok, seed + 1))
        p = dev(synth.direct_mask(n_ff, n_tok, 0.15, seed + 2))
        ws = sp.Workspace(n_embd, n_ff, n_tok)
        y0 = sp.sparse_ffn(ARCH_PROSPARSE, F16, qu, qg, qd, x, p, n_ff, ws=ws)
        for _ in range(5):  # odd and even sequence numbers
            y1 = sp.sparse_ffn(ARCH_PROSPARSE, F16, qu, qg, qd, x, p, n_ff, ws=ws, comm=comm)
            assert torch.equal(y0, y1)
        y2 = sp.shard_allreduce(y0.clone(), comm)
        assert torch.equal(y0, y2)


def test_batched_f16_hidden_saturates_instead_of_inf(oracle):
    """union GEMM, F16: hidden values beyond the F16 range must saturate at +-65504, not turn y into inf / NaN (the
    reference CUDA kernels keep hidden in F32, axpy-sparse.cu:44).  Expected y from the kernel's own F32 hidden_out."""
    import sparkinfer_b200 as sp
    n_embd, n_ff, n_tok = 512, 1024, 4
    seed = synth.SEED0 + 991
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(F16, w) for w in (Wu * 40, Wg * 40, Wd))   # big gate / up values: |h| reaches 1e5 and more
    x = (synth.activations(n_embd, n_tok, seed + 1) * 20).astype(np.float32)
    p = synth.direct_mask(n_ff, n_tok, 0.2, seed + 2)
    ws = sp.Workspace(n_embd, n_ff, n_tok)
    y, hid = sp.sparse_ffn(ARCH_PROSPARSE, F16, dev(qu), dev(qg), dev(qd), dev(x), dev(p), n_ff, want_hidden=True, ws=ws)
    y, hid = y.cpu().numpy(), hid.cpu().numpy()
    assert np.abs(hid).max() > 65504.0, "the case must overflow F16"
    assert np.all(np.isfinite(y))
    h16 = np.clip(hid, -65504.0, 65504.0).astype(np.float16).astype(np.float64)
    want = h16 @ Wd.astype(np.float16).astype(np.float64)
    assert rel_err(y, want.astype(np.float32)) < 2e-3
