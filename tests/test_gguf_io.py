"""CPU: the model file of the path -- the library's GGUF consumer (spif_gguf_*), the converter
(sparkinfer_b200/convert.py, counterpart of convert_hf_to_gguf.py:4594-4653) and its row quantisers.

  * the parser against a file written by the REFERENCE's gguf-py writer (tests/golden/model_tiny_ref.gguf, committed with its
    generating script): hyper-parameters, pred_lora, every tensor's type / shape / bytes, row-range reads;
  * the converter's output read back by an independent reader (the `gguf` package of this image) and by the library, and
    compared byte for byte with the golden file's tensors (same checkpoint values, same layout: F16 matrices, F32 vectors,
    ffn_down transposed, predictor renames, pred_lora);
  * the quantisers bit-exact against the reference's ggml functions (through the oracle library);
  * malformed files are error codes."""
import json
import os
import sys

import numpy as np
import pytest

from oracle_lib import BF16, F16, F32, Q4_0, Q8_0

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "model_tiny_ref.gguf")
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_model_gguf as tiny  # noqa: E402  (numpy only at import time)


def hf_dir(tmp_path, bias_free_arch="ProSparseLlamaForCausalLM", fmt="safetensors"):
    """the tiny checkpoint as a HF directory + a predictor directory, the converter's two inputs"""
    torch = pytest.importorskip("torch")
    t = tiny.tiny_model()
    model = tmp_path / "hf"
    pred = tmp_path / "pred"
    model.mkdir()
    pred.mkdir()
    (model / "config.json").write_text(json.dumps({
        "architectures": [bias_free_arch], "num_hidden_layers": tiny.N_LAYER, "hidden_size": tiny.N_EMBD,
        "intermediate_size": tiny.N_FF, "rms_norm_eps": tiny.EPS}))
    sd = {k: torch.from_numpy(v) for k, v in t.items() if k.startswith("model.")}
    sd["model.layers.0.self_attn.q_proj.weight"] = torch.zeros(4, 4)  # outside the path: must be skipped
    if fmt == "safetensors":
        from safetensors.torch import save_file
        save_file(sd, str(model / "model.safetensors"))
    else:
        torch.save(sd, str(model / "pytorch_model.bin"))
    for l in range(tiny.N_LAYER):
        torch.save({k.split(".", 2)[2]: torch.from_numpy(v) for k, v in t.items() if k.startswith(f"pred.{l}.")},
                   str(pred / f"model_{l}.pt"))
    return str(model), str(pred)


def test_parser_reads_reference_written_file():
    from sparkinfer_b200.gguf_io import ModelGguf
    t = tiny.tiny_model()
    with ModelGguf(GOLDEN) as g:
        assert (g.arch, g.n_layer, g.n_embd, g.n_ff) == ("prosparse-llama", tiny.N_LAYER, tiny.N_EMBD, tiny.N_FF)
        assert abs(g.rms_eps - tiny.EPS) < 1e-12 and g.pred_rank.tolist() == [tiny.RANK] * tiny.N_LAYER
        assert len(g.names()) == 8 * tiny.N_LAYER and g.find("blk.0.ffn_gate.bias") is None
        for l in range(tiny.N_LAYER):
            up = g.find(f"blk.{l}.ffn_up.weight")
            assert (up.type, up.n_dims, up.ne[0], up.ne[1]) == (F16, 2, tiny.N_EMBD, tiny.N_FF) and up.nbytes == tiny.N_FF * tiny.N_EMBD * 2
            dn = g.find(f"blk.{l}.ffn_down.weight")
            assert (dn.ne[0], dn.ne[1]) == (tiny.N_EMBD, tiny.N_FF), "transposed: one row per neuron"
            want = t[f"model.layers.{l}.mlp.down_proj.weight"].T.astype(np.float16)
            got = g.read_rows(f"blk.{l}.ffn_down.weight").view(np.float16)
            assert np.array_equal(got, want)
            # a shard's rows only
            part = g.read_rows(f"blk.{l}.ffn_down.weight", 16, 32).view(np.float16)
            assert np.array_equal(part, want[16:48])
            idx = np.random.default_rng(l).permutation(tiny.N_FF)[:40]       # any order, e.g. a model-split permutation
            assert np.array_equal(g.gather_rows(f"blk.{l}.ffn_down.weight", idx).view(np.float16), want[idx])
            assert np.array_equal(g.gather_rows(f"blk.{l}.ffn_pred_down.bias", idx).view(np.float32).reshape(-1),
                                  t[f"pred.{l}.fc2.bias"][idx])
            pu = g.find(f"blk.{l}.ffn_pred_up.weight")
            pd = g.find(f"blk.{l}.ffn_pred_down.weight")
            assert (pu.ne[0], pu.ne[1], pd.ne[0], pd.ne[1]) == (tiny.N_EMBD, tiny.RANK, tiny.RANK, tiny.N_FF)
            b = g.read_rows(f"blk.{l}.ffn_pred_down.bias", 8, 16).view(np.float32).reshape(-1)  # 1-D: rows are elements
            assert np.array_equal(b, t[f"pred.{l}.fc2.bias"][8:24])
            nw = g.read_rows(f"blk.{l}.ffn_norm.weight").view(np.float32).reshape(-1)
            assert np.array_equal(nw, t[f"model.layers.{l}.post_attention_layernorm.weight"])


@pytest.mark.parametrize("fmt", ["safetensors", "bin"])
def test_converter_output_equals_reference_layout(tmp_path, fmt):
    """same checkpoint values -> the converter's file holds, tensor by tensor, the bytes of the reference-written golden file"""
    from sparkinfer_b200.convert import convert
    from sparkinfer_b200.gguf_io import ModelGguf
    model, pred = hf_dir(tmp_path, fmt=fmt)
    out = str(tmp_path / "m.gguf")
    info = convert(model, out, pred_path=pred, outtype="f16", pred_bias=True)
    assert info["arch"] == "prosparse-llama" and info["pred_lora"] == [tiny.RANK] * tiny.N_LAYER
    with ModelGguf(out) as a, ModelGguf(GOLDEN) as b:
        assert (a.arch, a.n_layer, a.n_embd, a.n_ff, a.rms_eps) == (b.arch, b.n_layer, b.n_embd, b.n_ff, b.rms_eps)
        assert a.pred_rank.tolist() == b.pred_rank.tolist()
        assert sorted(a.names()) == sorted(b.names())
        for n in b.names():
            ta, tb = a.find(n), b.find(n)
            assert (ta.type, ta.n_dims, list(ta.ne), ta.nbytes) == (tb.type, tb.n_dims, list(tb.ne), tb.nbytes), n
            assert np.array_equal(a.read_rows(n), b.read_rows(n)), n
    # without --pred-bias the predictor biases are dropped (ReluMLP(bias=False), convert_hf_to_gguf.py:4588-4589)
    info = convert(model, out, pred_path=pred, outtype="f16", pred_bias=False)
    assert not any(n.endswith("ffn_pred_up.bias") or n.endswith("ffn_pred_down.bias") for n in info["tensors"])
    # without predictors: pred_lora all zero and ffn_down NOT transposed (:4606-4608, :4646)
    info = convert(model, out, pred_path=None)
    with ModelGguf(out) as a:
        assert a.pred_rank.tolist() == [0, 0]
        dn = a.find("blk.0.ffn_down.weight")
        assert (dn.ne[0], dn.ne[1]) == (tiny.N_FF, tiny.N_EMBD)


def test_converter_output_read_by_independent_reader(tmp_path):
    gguf = pytest.importorskip("gguf")
    from sparkinfer_b200.convert import convert
    model, pred = hf_dir(tmp_path)
    out = str(tmp_path / "q.gguf")
    convert(model, out, pred_path=pred, outtype="q8_0", pred_bias=True)
    r = gguf.GGUFReader(out)
    arch = bytes(r.fields["general.architecture"].parts[-1]).decode()
    assert arch == "prosparse-llama"
    assert [int(r.fields["prosparse-llama.pred_lora"].parts[i][0]) for i in r.fields["prosparse-llama.pred_lora"].data] == [tiny.RANK] * 2
    by = {t.name: t for t in r.tensors}
    t = tiny.tiny_model()
    up = by["blk.1.ffn_up.weight"]
    assert up.tensor_type == gguf.GGMLQuantizationType.Q8_0 and [int(v) for v in up.shape] == [tiny.N_EMBD, tiny.N_FF]
    from sparkinfer_b200.gguf_io import quantize_rows
    assert np.array_equal(np.asarray(up.data).reshape(tiny.N_FF, -1), quantize_rows(t["model.layers.1.mlp.up_proj.weight"], Q8_0))
    # block-format FFN matrices: the predictor stays F16 (the library's predictor computes in F16 / BF16), vectors F32
    assert by["blk.1.ffn_pred_up.weight"].tensor_type == gguf.GGMLQuantizationType.F16
    assert by["blk.1.ffn_pred_down.bias"].tensor_type == gguf.GGMLQuantizationType.F32
    assert by["blk.1.ffn_norm.weight"].tensor_type == gguf.GGMLQuantizationType.F32
    # and the dequantised rows are the reference's dequantisation of them
    deq = gguf.quants.dequantize(np.asarray(up.data), up.tensor_type).reshape(tiny.N_FF, tiny.N_EMBD)
    assert np.abs(deq - t["model.layers.1.mlp.up_proj.weight"]).max() < 0.05 * 2 / 127 * 4


@pytest.mark.parametrize("wtype", [F32, F16, BF16, Q8_0, Q4_0])
def test_row_quantisers_bit_exact(oracle, wtype):
    """quantize_rows == the reference's ggml_fp32_to_fp16_row / _bf16_row / quantize_row_q8_0_ref / quantize_row_q4_0_ref"""
    from sparkinfer_b200.gguf_io import quantize_rows
    rng = np.random.default_rng(5 + wtype)
    a = rng.normal(0, 1, (48, 256)).astype(np.float32)
    a[0, :32] = 0                                 # an all-zero block: d = 0
    a[1, 3], a[1, 7] = -5.0, 5.0                  # equal magnitudes: the first one sets the Q4_0 scale
    a[2, :64] = np.round(a[2, :64] * 4) / 4       # exact .5 ties after scaling
    a[3, :32] = np.linspace(-127, 127, 32)        # Q8_0 values landing on integers
    a[4, 0], a[4, 1] = 1e-30, -1e-30              # denormal-range scale
    a[5, 5] = 70000.0                             # F16 overflow -> inf, like ggml
    want = np.asarray(oracle.quantize(wtype, a)).view(np.uint8).reshape(48, -1)
    got = quantize_rows(a, wtype)
    assert got.shape == want.shape and np.array_equal(got, want)


def test_malformed_files_are_error_codes(tmp_path):
    from sparkinfer_b200 import _lib
    from sparkinfer_b200.gguf_io import ModelGguf
    raw = open(GOLDEN, "rb").read()
    cases = {"truncated_data": raw[:len(raw) - 4096], "truncated_header": raw[:200], "bad_magic": b"GGUX" + raw[4:],
             "version_1": raw[:4] + (1).to_bytes(4, "little") + raw[8:], "tiny": raw[:10]}
    for name, blob in cases.items():
        p = tmp_path / f"{name}.gguf"
        p.write_bytes(blob)
        with pytest.raises(_lib.SpifError) as e:
            ModelGguf(str(p))
        assert e.value.code in (-2, -5), (name, e.value.code)
    # a tensor whose dimensions multiply past 2^64: rejected, not wrapped into a small size
    import struct
    name = b"blk.0.ffn_up.weight"
    kv = struct.pack("<Q", 20) + b"general.architecture" + struct.pack("<IQ", 8, 5) + b"llama"
    info = struct.pack("<Q", len(name)) + name + struct.pack("<I4QIQ", 4, 1 << 40, 1 << 40, 1 << 40, 1 << 40, 0, 0)
    p = tmp_path / "wraps.gguf"
    p.write_bytes(b"GGUF" + struct.pack("<IQQ", 3, 1, 1) + kv + info + b"\0" * 64)
    with pytest.raises(_lib.SpifError) as e:
        ModelGguf(str(p))
    assert e.value.code == -2
    with pytest.raises(_lib.SpifError):
        ModelGguf(str(tmp_path / "does_not_exist.gguf"))
    with ModelGguf(GOLDEN) as g:
        with pytest.raises(KeyError):
            g.read_rows("blk.9.ffn_up.weight")
        with pytest.raises(_lib.SpifError) as e:  # rows past the end
            g.read_rows("blk.0.ffn_up.weight", tiny.N_FF - 4, 8)
        assert e.value.code == -2
        with pytest.raises(_lib.SpifError) as e:  # a row index outside the tensor
            g.gather_rows("blk.0.ffn_up.weight", [0, tiny.N_FF])
        assert e.value.code == -2
        with pytest.raises(_lib.SpifError) as e:  # destination too small
            g.read_rows("blk.0.ffn_up.weight", 0, 8, out=np.empty((8, 2 * tiny.N_EMBD), dtype=np.uint8)[:4])
        assert e.value.code == -4


def test_from_gguf_stages_the_rows_it_should(tmp_path, monkeypatch):
    """FfnStack.from_gguf's host logic without a GPU: device = cpu, pinning and the stack constructor stubbed.  Whole layers,
    a shard's rows, and a model-split order must hand the constructor exactly the file's rows in that order."""
    torch = pytest.importorskip("torch")
    from sparkinfer_b200 import model_split as ms
    from sparkinfer_b200.convert import convert
    from sparkinfer_b200.gguf_io import quantize_rows
    from sparkinfer_b200.stack import FfnStack
    model, pred = hf_dir(tmp_path)
    out = str(tmp_path / "m.gguf")
    convert(model, out, pred_path=pred, outtype="q8_0", pred_bias=True)
    got = {}

    def fake_init(self, arch, w_type, p_type, n_embd, m, r, layers, **kw):
        got.update(arch=arch, w_type=w_type, p_type=p_type, n_embd=n_embd, m=m, r=r, layers=layers, kw=kw)

    monkeypatch.setattr(FfnStack, "__init__", fake_init)
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self: self)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    t = tiny.tiny_model()
    F, E, R = tiny.N_FF, tiny.N_EMBD, tiny.RANK

    def check(rows_per_layer, which):
        assert (got["arch"], got["w_type"], got["p_type"], got["n_embd"], got["r"]) == (0, Q8_0, F16, E, R)
        for i, l in enumerate(which):
            rows = rows_per_layer[i]
            w = got["layers"][i]
            assert got["m"] == len(rows) == w.m
            up = quantize_rows(t[f"model.layers.{l}.mlp.up_proj.weight"], Q8_0)[rows]
            dn = quantize_rows(t[f"model.layers.{l}.mlp.down_proj.weight"].T, Q8_0)[rows]
            pd = quantize_rows(t[f"pred.{l}.fc2.weight"], F16)[rows]
            assert np.array_equal(w.up.numpy().reshape(len(rows), -1), up)
            assert np.array_equal(w.down_t.numpy().reshape(len(rows), -1), dn)
            assert np.array_equal(w.pred_down.numpy().reshape(len(rows), -1), pd)
            assert np.array_equal(w.pred_down_b.numpy(), t[f"pred.{l}.fc2.bias"][rows])
            assert w.pred_up.numel() == R * E * 2 and np.array_equal(w.pred_up_b.numpy(), t[f"pred.{l}.fc1.bias"])  # replicated
            assert w.up_b is None and w.gate_b is None and w.down_b is None and w.gate is not None

    st, info = FfnStack.from_gguf(out, device="cpu")
    check([np.arange(F)] * 2, [0, 1])
    assert info["ffn_norm"].shape == (2, E) and info["r"] == R
    mine = np.concatenate([np.arange(16 * gi, 16 * gi + 16) for gi in range(1, F // 16, 2)])  # odd groups: this shard's rows
    FfnStack.from_gguf(out, device="cpu", layers=[1], rows_of_layer=[mine])
    check([mine], [1])
    rates = np.random.default_rng(3).random((2, F))
    pattern, perms = ms.build_split(rates, g=16)
    split = str(tmp_path / "ms.gguf")
    ms.write_split(split, 16, pattern, perms)
    FfnStack.from_gguf(out, device="cpu", model_split=split)
    check([perms[0], perms[1]], [0, 1])


def test_parser_survives_corrupted_headers(tmp_path):
    """test-gguf.cpp style robustness: bytes of the header region of a valid file overwritten at random -> spif_gguf_open returns
    (0 or an error code) and, when it opens, every indexed tensor can be read or is refused; never a crash or an overrun."""
    from sparkinfer_b200 import _lib
    from sparkinfer_b200.gguf_io import ModelGguf
    raw = bytearray(open(GOLDEN, "rb").read())
    head = 2600  # KV + tensor infos of the golden file end before this
    rng = np.random.default_rng(99)
    opened = 0
    for it in range(400):
        b = bytearray(raw)
        for _ in range(int(rng.integers(1, 4))):
            pos = int(rng.integers(4, head))
            b[pos] = int(rng.integers(0, 256)) if rng.random() < 0.7 else 0xFF
        if it % 5 == 0:
            b = b[:int(rng.integers(8, len(b)))]
        p = tmp_path / "fuzz.gguf"
        p.write_bytes(bytes(b))
        try:
            g = ModelGguf(str(p))
        except _lib.SpifError:
            continue
        opened += 1
        for n in g.names():
            try:
                g.read_rows(n)
            except (_lib.SpifError, ValueError, KeyError):
                pass
        g.close()
    assert opened > 20  # most single-byte damage to names / values still parses
