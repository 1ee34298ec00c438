"""GPU parity of the look-ahead layer kernel (spif_sparse_ffn_lookahead: sparse FFN of layer l + predictor of layer l+1 +
the next layer's active-row list + the fixed-order sum over CTAs, ONE launch) against the oracle, through the C-ABI.
Includes the exact BASELINE.json shapes: 5120x13824 (13B, configs[1]), 8192x28672 (70B, configs[4]), 4096x11008 (7B)."""
import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, BF16, F16, TYPE_NAMES

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = {F16: 1e-3, BF16: 8e-3}                  # north_star: 1e-3 relative for FP16 (F32 activations, reference CUDA semantics)
TOL_CPU_ROUNDING = {F16: 4e-4, BF16: 3e-3}     # hidden rounded to the weight type: one ulp on a neuron's contribution


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("m", [33, 1024, 13824, 28672])
def test_mask_to_list_bit_exact(oracle, m):
    import sparkinfer_b200 as sp
    p = synth.direct_mask(m, 1, 0.1, synth.SEED0 + m, near=0.01)
    p[0, :5] = [np.nan, 0.5, np.nextafter(np.float32(0.5), np.float32(0)), 0.0, 1.0][:min(5, m)]
    lst = sp.mask_to_list(dev(p[0]))
    assert lst.rows() == oracle.mask_compact(p[0]).tolist()


FFN_SHAPES = [(512, 1024), (1376, 2048), (4096, 11008), (5120, 13824), (8192, 28672), (3584, 18944)]


@pytest.mark.parametrize("arch", [ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP])
@pytest.mark.parametrize("wtype", [F16, BF16])
@pytest.mark.parametrize("n_embd,n_ff", FFN_SHAPES)
def test_lookahead_ffn_only(oracle, arch, wtype, n_embd, n_ff):
    """pred == NULL: the sparse FFN alone, list from a direct mask; y, hidden, stats against the oracle; bit-stable"""
    import sparkinfer_b200 as sp
    if (n_embd, n_ff) in ((8192, 28672), (3584, 18944)) and (arch != ARCH_PROSPARSE):
        pytest.skip("big shapes: one arch is enough")
    seed = synth.SEED0 + 13 * n_embd + n_ff + arch
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(wtype, w) for w in (Wu, Wg, Wd))
    x = synth.activations(n_embd, 1, seed + 1)
    p = synth.direct_mask(n_ff, 1, 0.10, seed + 2)
    b_down = np.random.default_rng(seed + 4).normal(0, 0.1, n_embd).astype(np.float32)
    y_ref, _, _, hid_ref = oracle.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff, b_down=b_down)
    comm = sp.ShardComm(n_embd + 1024)
    ws = sp.Workspace(n_embd, n_ff, 1)
    du, dg, dd, dx = dev(qu), (None if arch == ARCH_RELU_UP else dev(qg)), dev(qd), dev(x[0])
    lst = sp.mask_to_list(dev(p[0]))
    act = ~(p < 0.5)
    for flags, tol in ((0, TOL[wtype]), (sp.FLAG_CPU_ROUNDING, TOL_CPU_ROUNDING[wtype])):
        y = torch.full((n_embd,), float("nan"), device="cuda")
        hid = torch.empty((n_ff,), device="cuda")
        stats = torch.zeros((4,), dtype=torch.int32, device="cuda")
        sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=arch, w_type=wtype, up=du, gate=dg, down_t=dd, cur=dx, n_ff=n_ff,
                                                   down_b=dev(b_down), flags=flags, out=y, hidden_out=hid, stats=stats),
                                list_in=lst)
        y, hid, stats = y.cpu().numpy(), hid.cpu().numpy(), stats.cpu().numpy()
        if not rel_err(y, y_ref[0]) < tol:  # diagnosis of a failure: is it the list, a hidden value or the sum over CTAs?
            wrong_h = np.nonzero(np.abs(hid - hid_ref[0] * act[0]) > 1e-3 * np.abs(hid_ref).max())[0]
            raise AssertionError((TYPE_NAMES[wtype], arch, flags, rel_err(y, y_ref[0]), "stats", stats.tolist(), "active",
                                  int(act.sum()), "list ok", lst.rows() == np.nonzero(act[0])[0].tolist(), "hidden wrong at",
                                  wrong_h[:8].tolist(), "columns off", int((np.abs(y - y_ref[0]) > tol * np.abs(y_ref).max()).sum())))
        assert rel_err(hid * act[0], hid_ref[0] * act[0]) < (2e-5 if flags else TOL[wtype])
        assert np.all(hid[~act[0]] == 0.0)
        assert stats[0] == act.sum()
        assert stats[1] == np.count_nonzero(hid)
    ya = torch.empty((n_embd,), device="cuda")
    yb = torch.empty((n_embd,), device="cuda")
    for out in (ya, yb):
        sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=arch, w_type=wtype, up=du, gate=dg, down_t=dd, cur=dx, n_ff=n_ff, out=out),
                                list_in=lst)
    assert torch.equal(ya, yb)
    assert comm.status() == 0 and comm.epoch() == 4


@pytest.mark.parametrize("ptype", [F16, BF16])
@pytest.mark.parametrize("n_embd,r,m", [(512, 256, 1024), (5120, 1024, 13824), (4096, 1024, 11008), (8192, 1024, 28672),
                                        (5120, 2048, 2048), (4544, 1024, 4096)])
def test_lookahead_predictor_only(oracle, ptype, n_embd, r, m):
    """ffn == NULL: build_predictor + the active-row list; flips against the oracle only where |logit| < 1e-4"""
    import sparkinfer_b200 as sp
    seed = synth.SEED0 + n_embd + r + m
    wpu, wpd, bpd = synth.predictor_weights(n_embd, r, m, 0.1, seed)
    bpu = np.random.default_rng(seed + 5).normal(0, 0.05, r).astype(np.float32)
    qpu, qpd = oracle.quantize(ptype, wpu), oracle.quantize(ptype, wpd)
    x = synth.activations(n_embd, 1, seed + 1)
    p_ref, z_ref = oracle.predictor(ptype, qpu, bpu, qpd, bpd, x, n_embd, r, m)
    comm = sp.ShardComm(n_embd + r)
    ws = sp.Workspace(n_embd, m, 1)
    p_gpu = torch.empty((m,), device="cuda")
    z_gpu = torch.empty((m,), device="cuda")
    lst = sp.ActiveList(m)
    sp.sparse_ffn_lookahead(comm, ws, pred=dict(pred_up=dev(qpu), pred_up_b=dev(bpu), pred_down=dev(qpd), pred_down_b=dev(bpd),
                                                cur=dev(x[0]), p_type=ptype, r=r, m=m, flags=sp.FLAG_CPU_ROUNDING, out=p_gpu,
                                                logits_out=z_gpu), list_out=lst)
    p_gpu, z_gpu = p_gpu.cpu().numpy(), z_gpu.cpu().numpy()
    assert np.abs(z_gpu - z_ref[0]).max() < 1e-4 * max(1.0, np.abs(z_ref).max())
    flips = (p_gpu < 0.5) != (p_ref[0] < 0.5)
    assert np.all(np.abs(z_ref[0][flips]) < 1e-4), f"{flips.sum()} flips, worst |logit| {np.abs(z_ref[0][flips]).max()}"
    # (a last-bit difference in an h1 element can cross a rounding boundary of the weight type: 2^-11 F16, 2^-8 BF16)
    assert np.abs(p_gpu - p_ref[0]).max() < (1e-5 if ptype == F16 else 2e-4)
    # the list is bit-exact for the kernel's OWN sigmoid output (mask convention !(p < 0.5))
    assert lst.rows() == np.nonzero(~(p_gpu < 0.5))[0].tolist()
    assert comm.status() == 0


@pytest.mark.parametrize("n_embd,r,n_ff,L", [(1024, 256, 2048, 3), (5120, 1024, 13824, 2)])
def test_lookahead_chain(oracle, n_embd, r, n_ff, L):
    """the decode order of the reference (llama-graph.cpp:932-946): predictor(0); then per layer FFN(l) + predictor(l+1)
    in one launch, the list handed from launch to launch.  FFN outputs are checked against the oracle run on the GPU's
    own masks (mask flips near the threshold are the predictor test's business)."""
    import sparkinfer_b200 as sp
    wt = F16
    seed = synth.SEED0 + 99 + n_embd
    layers = []
    for l in range(L):
        Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed + 10 * l)
        wpu, wpd, bpd = synth.predictor_weights(n_embd, r, n_ff, 0.1, seed + 10 * l + 5)
        layers.append(dict(q=[oracle.quantize(wt, w) for w in (Wu, Wg, Wd)], qp=[oracle.quantize(wt, wpu), oracle.quantize(wt, wpd)],
                           bpd=bpd))
    xs = [synth.activations(n_embd, 1, seed + 1000 + l) for l in range(L)]
    comm = sp.ShardComm(n_embd + r)
    ws = sp.Workspace(n_embd, n_ff, 1)
    dw = [[dev(q) for q in ly["q"]] + [dev(q) for q in ly["qp"]] + [dev(ly["bpd"])] for ly in layers]
    dx = [dev(x[0]) for x in xs]
    lists = [sp.ActiveList(n_ff), sp.ActiveList(n_ff)]
    p_out = [torch.empty((n_ff,), device="cuda") for _ in range(L)]
    ys = [torch.empty((n_embd,), device="cuda") for _ in range(L)]

    def pred(l, x_from):
        return dict(pred_up=dw[l][3], pred_down=dw[l][4], pred_down_b=dw[l][5], cur=dx[x_from], p_type=wt, r=r, m=n_ff,
                    out=p_out[l])

    def step():
        sp.sparse_ffn_lookahead(comm, ws, pred=pred(0, 0), list_out=lists[0])
        for l in range(L):
            sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=ARCH_PROSPARSE, w_type=wt, up=dw[l][0], gate=dw[l][1], down_t=dw[l][2],
                                                       cur=dx[l], n_ff=n_ff, out=ys[l]),
                                    pred=pred(l + 1, l) if l + 1 < L else None,
                                    list_in=lists[l % 2], list_out=lists[(l + 1) % 2] if l + 1 < L else None)

    step()
    torch.cuda.synchronize()
    first = [y.clone() for y in ys]
    for l in range(L):
        p = p_out[l].cpu().numpy()[None]
        y_ref = oracle.sparse_ffn(ARCH_PROSPARSE, wt, *layers[l]["q"], xs[l], p, n_embd, n_ff)[0]
        assert rel_err(ys[l].cpu().numpy(), y_ref[0]) < 1e-3, l
        p_ref, z_ref = oracle.predictor(wt, layers[l]["qp"][0], None, layers[l]["qp"][1], layers[l]["bpd"], xs[max(l - 1, 0)],
                                        n_embd, r, n_ff)
        flips = (p[0] < 0.5) != (p_ref[0] < 0.5)
        assert np.all(np.abs(z_ref[0][flips]) < 1e-3)  # F32 activations here (no CPU_ROUNDING): a slightly wider band
    # the same chain captured in a CUDA graph and replayed: bit-identical, epoch counts every launch
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
        s.synchronize()
        with torch.cuda.graph(g, stream=s):
            step()
    torch.cuda.synchronize()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for l in range(L):
        assert torch.equal(ys[l], first[l])
    assert comm.status() == 0 and comm.epoch() == (L + 1) * 5


def test_lookahead_neuron_idx(oracle):
    """rows held here are a subset of the neurons (GPU cache / shard): list in ROW space, hidden_out / biases by neuron id"""
    import sparkinfer_b200 as sp
    n_embd, n_ff, m, wt = 1024, 4096, 1536, F16
    seed = synth.SEED0 + 4242
    rng = np.random.default_rng(seed)
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    nidx = np.sort(rng.choice(n_ff, m, replace=False)).astype(np.int32)
    b_up = rng.normal(0, 0.05, n_ff).astype(np.float32)
    b_gate = rng.normal(0, 0.05, n_ff).astype(np.float32)
    x = synth.activations(n_embd, 1, seed + 1)
    p_row = synth.direct_mask(m, 1, 0.15, seed + 2)      # mask indexed by row
    p_full = np.full((1, n_ff), 0.1, np.float32)
    p_full[0, nidx] = p_row[0]
    qu, qg, qd = (oracle.quantize(wt, w) for w in (Wu, Wg, Wd))
    y_ref, _, _, hid_ref = oracle.sparse_ffn(ARCH_RELU_GATED, wt, qu, qg, qd, x, p_full, n_embd, n_ff, b_up=b_up, b_gate=b_gate)
    # inactive rows of the oracle's dense hidden hold act(bias): compare on active rows only
    qu_s, qg_s, qd_s = (oracle.quantize(wt, w[nidx]) for w in (Wu, Wg, Wd))
    # the oracle adds biases to inactive neurons too (llama-graph.cpp:1049-1059) and their hidden feeds the down axpy only
    # if active: y is unaffected
    comm = sp.ShardComm(n_embd + 256)
    ws = sp.Workspace(n_embd, n_ff, 1)
    y = torch.empty((n_embd,), device="cuda")
    hid = torch.empty((n_ff,), device="cuda")
    sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=ARCH_RELU_GATED, w_type=wt, up=dev(qu_s), gate=dev(qg_s), down_t=dev(qd_s),
                                               cur=dev(x[0]), n_ff=n_ff, neuron_idx=dev(nidx), up_b=dev(b_up), gate_b=dev(b_gate),
                                               out=y, hidden_out=hid), list_in=sp.mask_to_list(dev(p_row[0])))
    act = ~(p_full[0] < 0.5)
    assert rel_err(y.cpu().numpy(), y_ref[0]) < 1e-3
    assert rel_err(hid.cpu().numpy() * act, hid_ref[0] * act) < 1e-3


def test_lookahead_rejects(oracle):
    import sparkinfer_b200 as sp
    from sparkinfer_b200._lib import SpifError
    comm = sp.ShardComm(512 + 256)
    ws = sp.Workspace(512, 1024, 1)
    z = torch.zeros(512 * 1024 * 2, dtype=torch.uint8, device="cuda")
    x = torch.zeros(512, device="cuda")
    with pytest.raises(SpifError):  # r must be a multiple of 256
        sp.sparse_ffn_lookahead(comm, ws, pred=dict(pred_up=z, pred_down=z, cur=x, p_type=F16, r=128, m=1024))
    with pytest.raises(SpifError):  # nothing to do
        sp.sparse_ffn_lookahead(comm, ws)


@pytest.mark.parametrize("flags_name", ["f32", "cpu_rounding"])
@pytest.mark.parametrize("n_embd,r,n_ff", [(512, 256, 1024), (1376, 256, 2048), (5120, 1024, 13824)])
def test_lookahead_fused_rmsnorm_and_residual(oracle, n_embd, r, n_ff, flags_name):
    """SURVEY 8f2 (src/models/llama.cpp:98-101,144): build_norm(ffn_inp, ffn_norm, RMS) before and ggml_add(cur, ffn_inp) after
    the sparse FFN, fused into the layer kernel (spif_sparse_ffn_lookahead_ex).  Oracle = the reference's rms_norm + mul
    arithmetic (oracle_lib.rms_norm_mul, pinned bit-exact against oracle/_ref) -> oracle predictor / sparse FFN -> + residual."""
    import oracle_lib
    import sparkinfer_b200 as sp
    seed = synth.SEED0 + 7 * n_embd + n_ff + r
    rng = np.random.default_rng(seed)
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    qu, qg, qd = (oracle.quantize(F16, w) for w in (Wu, Wg, Wd))
    wpu, wpd, bpd = synth.predictor_weights(n_embd, r, n_ff, 0.1, seed + 3)
    qpu, qpd = oracle.quantize(F16, wpu), oracle.quantize(F16, wpd)
    x = (synth.activations(n_embd, 1, seed + 1) * 3.0).astype(np.float32)          # raw ffn_inp, not unit variance
    w_norm = rng.normal(1.0, 0.2, n_embd).astype(np.float32)
    eps = 1e-5
    b_down = rng.normal(0, 0.1, n_embd).astype(np.float32)
    p = synth.direct_mask(n_ff, 1, 0.10, seed + 2)
    x_n = oracle_lib.rms_norm_mul(x, w_norm, eps)
    y_ffn = oracle.sparse_ffn(ARCH_PROSPARSE, F16, qu, qg, qd, x_n, p, n_embd, n_ff, b_down=b_down)[0]
    y_ref = (y_ffn[0] + x[0]).astype(np.float32)
    p_ref, z_ref = oracle.predictor(F16, qpu, None, qpd, bpd, x_n, n_embd, r, n_ff)
    flags = sp.FLAG_CPU_ROUNDING if flags_name == "cpu_rounding" else 0
    comm = sp.ShardComm(n_embd + r)
    ws = sp.Workspace(n_embd, n_ff, 1)
    dx = dev(x[0])
    y = torch.full((n_embd,), float("nan"), device="cuda")
    z_gpu = torch.empty((n_ff,), device="cuda")
    p_gpu = torch.empty((n_ff,), device="cuda")
    lst_out = sp.ActiveList(n_ff)
    sp.sparse_ffn_lookahead(comm, ws,
                            ffn=dict(arch=ARCH_PROSPARSE, w_type=F16, up=dev(qu), gate=dev(qg), down_t=dev(qd), cur=dx, n_ff=n_ff,
                                     down_b=dev(b_down), flags=flags, out=y),
                            pred=dict(pred_up=dev(qpu), pred_down=dev(qpd), pred_down_b=dev(bpd), cur=dx, p_type=F16, r=r, m=n_ff,
                                      flags=sp.FLAG_CPU_ROUNDING, out=p_gpu, logits_out=z_gpu),
                            list_in=sp.mask_to_list(dev(p[0])), list_out=lst_out,
                            io=dict(norm_w=dev(w_norm), norm_eps=eps, residual=dx))
    y, z_gpu, p_gpu = y.cpu().numpy(), z_gpu.cpu().numpy(), p_gpu.cpu().numpy()
    tol = TOL_CPU_ROUNDING[F16] if flags else TOL[F16]
    assert rel_err(y - x[0], y_ffn[0]) < tol, rel_err(y - x[0], y_ffn[0])      # the FFN part alone (the residual dominates y)
    assert rel_err(y, y_ref) < tol
    if flags:  # (the predictor follows the FFN's rounding mode: one codec per launch)
        assert np.abs(z_gpu - z_ref[0]).max() < 1e-4 * max(1.0, np.abs(z_ref).max())
        flips = (p_gpu < 0.5) != (p_ref[0] < 0.5)
        assert np.all(np.abs(z_ref[0][flips]) < 1e-4)
    assert lst_out.rows() == np.nonzero(~(p_gpu < 0.5))[0].tolist()
    assert comm.status() == 0
    # without io the same call must still treat `cur` as the already-normalised input
    y2 = torch.empty((n_embd,), device="cuda")
    sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=ARCH_PROSPARSE, w_type=F16, up=dev(qu), gate=dev(qg), down_t=dev(qd),
                                               cur=dev(x_n[0]), n_ff=n_ff, down_b=dev(b_down), flags=flags, out=y2),
                            list_in=sp.mask_to_list(dev(p[0])))
    assert rel_err(y2.cpu().numpy(), y_ffn[0]) < tol
