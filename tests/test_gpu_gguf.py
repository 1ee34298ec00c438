"""GPU: checkpoint -> sparkinfer_b200/convert.py -> model GGUF -> FfnStack.from_gguf -> one decode step, against the oracle run
on the checkpoint's own values (the data formats either side of the path).  Covers what only the file round
trip can break: ffn_down transposition, predictor renames and biases, pred_lora, the F16 / Q8_0 row bytes, row staging."""
import json

import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, F16, Q8_0

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

E, F, R, L = 1024, 2048, 256, 3


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def checkpoint(tmp_path):
    from safetensors.torch import save_file
    seed = synth.SEED0 + 4242
    model, pred = tmp_path / "hf", tmp_path / "pred"
    model.mkdir()
    pred.mkdir()
    (model / "config.json").write_text(json.dumps({"architectures": ["ProSparseLlamaForCausalLM"], "num_hidden_layers": L,
                                                   "hidden_size": E, "intermediate_size": F, "rms_norm_eps": 1e-5}))
    sd, layers = {}, []
    for l in range(L):
        Wu, Wg, Wd = synth.ffn_weights(E, F, seed + 10 * l)          # all [n_ff, n_embd]: Wd is down^T
        wpu, wpd, bpd = synth.predictor_weights(E, R, F, 0.1, seed + 10 * l + 5)
        bpu = np.random.default_rng(seed + 10 * l + 7).normal(0, 0.05, R).astype(np.float32)
        sd[f"model.layers.{l}.mlp.up_proj.weight"] = torch.from_numpy(Wu)
        sd[f"model.layers.{l}.mlp.gate_proj.weight"] = torch.from_numpy(Wg)
        sd[f"model.layers.{l}.mlp.down_proj.weight"] = torch.from_numpy(np.ascontiguousarray(Wd.T))  # HF: [n_embd, n_ff]
        sd[f"model.layers.{l}.post_attention_layernorm.weight"] = torch.ones(E)
        torch.save({"fc1.weight": torch.from_numpy(wpu), "fc1.bias": torch.from_numpy(bpu), "fc2.weight": torch.from_numpy(wpd),
                    "fc2.bias": torch.from_numpy(bpd)}, str(pred / f"model_{l}.pt"))
        layers.append(dict(W=(Wu, Wg, Wd), wpu=wpu, bpu=bpu, wpd=wpd, bpd=bpd))
    save_file(sd, str(model / "model.safetensors"))
    return str(model), str(pred), layers


@pytest.mark.parametrize("outtype,wtype,tol", [("f16", F16, 1e-3), ("q8_0", Q8_0, 2e-2)])
def test_checkpoint_to_gguf_to_stack_matches_oracle(oracle, tmp_path, outtype, wtype, tol):
    from sparkinfer_b200.convert import convert
    from sparkinfer_b200.stack import FfnStack
    model, pred, layers = checkpoint(tmp_path)
    out = str(tmp_path / "m.gguf")
    convert(model, out, pred_path=pred, outtype=outtype, pred_bias=True)
    st, info = FfnStack.from_gguf(out, n_tok=1, device="cuda")
    assert (info["arch"], info["n_layer"], info["n_embd"], info["n_ff"], info["r"]) == ("prosparse-llama", L, E, F, R)
    assert st.arch == ARCH_PROSPARSE and st.w_type == wtype and st.p_type == F16 and st.m == F
    assert info["ffn_norm"].shape == (L, E) and abs(info["rms_eps"] - 1e-5) < 1e-12
    xs = np.stack([synth.activations(E, 1, synth.SEED0 + 77 + l) for l in range(L)])
    st.x.copy_(torch.from_numpy(xs).cuda())
    fused = st.fused_supported()
    assert fused == (outtype == "f16")
    (st.step_fused if fused else st.step_eager)()
    torch.cuda.synchronize()
    p = st.p.cpu().numpy()
    y = st.y.cpu().numpy()
    for l in range(L):
        ly = layers[l]
        q = [oracle.quantize(wtype, w) for w in ly["W"]]
        # predictor of layer l reads x_{l-1} (look-ahead, llama-graph.cpp:939-946), layer 0 its own input; with BOTH biases
        p_ref, z_ref = oracle.predictor(F16, oracle.quantize(F16, ly["wpu"]), ly["bpu"], oracle.quantize(F16, ly["wpd"]), ly["bpd"],
                                        xs[max(l - 1, 0)], E, R, F)
        flips = (p[l, 0] < 0.5) != (p_ref[0] < 0.5)
        assert np.all(np.abs(z_ref[0][flips]) < 1e-3), (l, int(flips.sum()))
        assert 0.02 < float((~(p[l, 0] < 0.5)).mean()) < 0.3
        y_ref = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, *q, xs[l], p[l], E, F)[0]
        assert rel_err(y[l, 0], y_ref[0]) < tol, (l, rel_err(y[l, 0], y_ref[0]))
    assert st.comm.status() == 0


def test_sharded_rows_from_gguf(oracle, tmp_path):
    """rows_of_layer: a shard stages only its neuron rows; two half-stacks on one GPU sum to the full stack's y"""
    from sparkinfer_b200.convert import convert
    from sparkinfer_b200.stack import FfnStack
    model, pred, layers = checkpoint(tmp_path)
    out = str(tmp_path / "m.gguf")
    convert(model, out, pred_path=pred, outtype="f16", pred_bias=True)
    full, _ = FfnStack.from_gguf(out, device="cuda", layers=[1])
    g = np.arange(F // 16)
    halves = [np.sort(np.concatenate([np.arange(16 * i, 16 * i + 16) for i in g[g % 2 == s]])) for s in (0, 1)]
    x = synth.activations(E, 1, synth.SEED0 + 5)
    ys = []
    for st in [full] + [FfnStack.from_gguf(out, device="cuda", layers=[1], rows_of_layer=[h])[0] for h in halves]:
        st.x[0].copy_(torch.from_numpy(x).cuda())
        st.step_eager()
        torch.cuda.synchronize()
        ys.append((st.y[0, 0].cpu().numpy(), st.p[0, 0].cpu().numpy()))
    assert ys[1][1].shape == (F // 2,) and np.array_equal(ys[1][1], ys[0][1][halves[0]]) and np.array_equal(ys[2][1], ys[0][1][halves[1]])
    assert rel_err(ys[1][0] + ys[2][0], ys[0][0]) < 1e-5


def test_model_split_order_from_gguf(oracle, tmp_path):
    """profile -> model-split producer -> load in the split's order: masks are the unsplit masks permuted, y is unchanged"""
    from sparkinfer_b200 import model_split as ms
    from sparkinfer_b200.convert import convert
    from sparkinfer_b200.stack import FfnStack
    model, pred, _ = checkpoint(tmp_path)
    out = str(tmp_path / "m.gguf")
    convert(model, out, pred_path=pred, outtype="f16", pred_bias=True)
    plain, _ = FfnStack.from_gguf(out, device="cuda")
    xs = np.stack([np.stack([synth.activations(E, 1, synth.SEED0 + 900 + 7 * s + l)[0] for l in range(L)]) for s in range(6)])
    rates = ms.profile_rates(plain, xs)
    assert rates.shape == (L, F) and 0.02 < rates.mean() < 0.3
    pattern, perms = ms.build_split(rates, g=16, n_shards=2)
    split = str(tmp_path / "ms.gguf")
    ms.write_split(split, 16, pattern, perms)
    reordered, _ = FfnStack.from_gguf(out, device="cuda", model_split=split)
    x = torch.from_numpy(xs[0][:, None, :]).cuda()
    res = []
    for st in (plain, reordered):
        st.x.copy_(x)
        st.step_fused()
        torch.cuda.synchronize()
        res.append((st.p.cpu().numpy(), st.y.cpu().numpy()))
    for l in range(L):
        p_new, p_old = res[1][0][l, 0], res[0][0][l, 0][perms[l]]
        assert np.abs(p_new - p_old).max() < 1e-6, l                         # a row's dot does not depend on where the row sits
        clear = np.abs(p_old - 0.5) > 1e-5
        assert np.array_equal((p_new < 0.5)[clear], (p_old < 0.5)[clear]), l
        assert rel_err(res[1][1][l, 0], res[0][1][l, 0]) < 1e-4, l          # same neurons, another summation order
