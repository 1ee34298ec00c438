"""Checkpoint -> model GGUF for the sparse-FFN path: the producer of what spif_gguf_* / FfnStack.from_gguf consume.

Reference: convert_hf_to_gguf.py -- ProSparseLlamaModel / BambooModel (:4594-4653; the OPT, Falcon and Qwen2 classes carry the
same three overrides, :2154-2240, :3468-3540):
  * the predictor of layer i comes from `<pred_path>/model_<i>.pt`, a torch state dict with fc1.weight {r, n_embd} and
    fc2.weight {n_ff, r} (+ fc1.bias / fc2.bias with --pred-bias), renamed blk.i.ffn_pred_up / ffn_pred_down (:4582-4590, :4617-4634);
  * `{arch}.pred_lora` = the rank r of every layer's predictor, 0 without predictors (:4604-4609);
  * ffn_down.weight is stored TRANSPOSED (one row per neuron) when predictors are present (:4640-4648);
  * type policy of prepare_tensors (:548-611): 1-D tensors and *_norm.weight stay F32, matrices take --outtype.
Differences, on purpose: only the tensors of the path are written (attention, embeddings and the tokenizer are out of scope,
SURVEY.md 8); `q4_0` is an --outtype here (the reference needs llama-quantize for it; same quantize_row_q4_0_ref arithmetic);
predictor matrices are written as F16 when the FFN matrices are a block format, because the library's predictor computes in
F16 / BF16 (`pred_type` overrides).  torch is used to read checkpoints only.

CLI:  python -m sparkinfer_b200.convert <hf_dir> --pred-path <dir> --outfile model.gguf [--outtype f16] [--pred-bias]
"""
import argparse
import json
import os
import re

import numpy as np

from . import gguf_io as G

# config.json "architectures" -> general.architecture (src/llama-arch.cpp:10-11 and the stock names)
ARCH_OF_HF = {
    "ProSparseLlamaForCausalLM": "prosparse-llama", "BambooForCausalLM": "bamboo", "OPTForCausalLM": "opt",
    "FalconForCausalLM": "falcon", "RWForCausalLM": "falcon", "Qwen2ForCausalLM": "qwen2", "Qwen2Model": "qwen2",
    "LlamaForCausalLM": "llama",
}

# HF tensor name -> (GGUF base name, is ffn_down); gguf-py/gguf/tensor_mapping.py (llama / qwen2 / bamboo, opt, falcon spellings)
_HF_PATTERNS = [
    (r"(?:model\.)?layers\.(\d+)\.mlp\.gate_proj\.(weight|bias)", "ffn_gate", False),
    (r"(?:model\.)?layers\.(\d+)\.mlp\.up_proj\.(weight|bias)", "ffn_up", False),
    (r"(?:model\.)?layers\.(\d+)\.mlp\.down_proj\.(weight|bias)", "ffn_down", True),
    (r"(?:model\.)?layers\.(\d+)\.post_attention_layernorm\.(weight)", "ffn_norm", False),
    (r"(?:model\.)?decoder\.layers\.(\d+)\.fc1\.(weight|bias)", "ffn_up", False),
    (r"(?:model\.)?decoder\.layers\.(\d+)\.fc2\.(weight|bias)", "ffn_down", True),
    (r"(?:model\.)?decoder\.layers\.(\d+)\.final_layer_norm\.(weight|bias)", "ffn_norm", False),
    (r"(?:transformer\.)?h\.(\d+)\.mlp\.dense_h_to_4h\.(weight|bias)", "ffn_up", False),
    (r"(?:transformer\.)?h\.(\d+)\.mlp\.dense_4h_to_h\.(weight|bias)", "ffn_down", True),
]


def map_name(hf_name):
    """-> (gguf name, layer, is_down_weight) or None for a tensor outside the path"""
    for pat, base, is_down in _HF_PATTERNS:
        m = re.fullmatch(pat, hf_name)
        if m:
            return f"blk.{int(m.group(1))}.{base}.{m.group(2)}", int(m.group(1)), is_down and m.group(2) == "weight"
    return None


def _checkpoint_tensors(model_dir):
    """name -> (shape, zero-argument loader returning a float32 numpy array); safetensors shards first, then pytorch_model*.bin.
    Nothing is read until a loader is called (safetensors: lazy slices; torch: mmap)."""
    import torch
    files = sorted(f for f in os.listdir(model_dir) if f.endswith(".safetensors"))
    out = {}
    if files:
        from safetensors import safe_open
        for fn in files:
            path = os.path.join(model_dir, fn)
            with safe_open(path, framework="pt") as f:
                for k in f.keys():
                    def load(path=path, k=k):
                        with safe_open(path, framework="pt") as g:
                            return g.get_tensor(k).float().numpy()
                    out[k] = (tuple(f.get_slice(k).get_shape()), load)
        return out
    files = sorted(f for f in os.listdir(model_dir) if re.fullmatch(r"pytorch_model.*\.bin|.*\.pt", f))
    for fn in files:
        sd = torch.load(os.path.join(model_dir, fn), map_location="cpu", weights_only=True, mmap=True)
        for k, v in sd.items():
            out[k] = (tuple(v.shape), (lambda v=v: v.float().numpy()))
    return out


def _predictor_files(pred_path):
    found = []
    for f in os.listdir(pred_path):
        m = re.fullmatch(r"model_(\d+)\.pt", f)
        if m:
            found.append((int(m.group(1)), f))
    return [f for _, f in sorted(found)]  # sorted by index, then enumerated like the reference (:4619-4628)


def convert(model_dir, outfile, pred_path=None, outtype="f16", pred_bias=False, pred_type=None):
    """writes the GGUF; returns {"arch", "n_layer", "n_embd", "n_ff", "pred_lora", "tensors": [names]}"""
    import torch
    cfg = json.load(open(os.path.join(model_dir, "config.json")))
    hf_arch = (cfg.get("architectures") or ["LlamaForCausalLM"])[0]
    if hf_arch not in ARCH_OF_HF:
        raise ValueError(f"architecture {hf_arch} has no sparse-FFN path in the reference")
    arch = ARCH_OF_HF[hf_arch]
    n_layer = int(cfg.get("num_hidden_layers", cfg.get("n_layer", 0)))
    n_embd = int(cfg.get("hidden_size", cfg.get("n_embd", 0)))
    n_ff = int(cfg.get("intermediate_size", cfg.get("ffn_dim", 4 * n_embd)))
    wtype = G.TYPE_NAMES[outtype]
    if pred_type is None:
        pred_type = outtype if outtype in ("f16", "bf16") else "f16"
    ptype = G.TYPE_NAMES[pred_type]

    w = G.GgufWriter(outfile)
    w.add_str("general.architecture", arch)
    w.add_u32(f"{arch}.block_count", n_layer)
    w.add_u32(f"{arch}.embedding_length", n_embd)
    w.add_u32(f"{arch}.feed_forward_length", n_ff)
    eps = cfg.get("rms_norm_eps", cfg.get("layer_norm_epsilon", cfg.get("layer_norm_eps")))
    if eps is not None:
        w.add_f32(f"{arch}.attention.layer_norm_rms_epsilon", float(eps))

    # predictors: the rank of every layer first (a KV), tensors after the model's own
    pred_files = _predictor_files(pred_path) if pred_path else []
    pred_lora, pred_tensors = [], []
    for i, fn in enumerate(pred_files):
        sd = torch.load(os.path.join(pred_path, fn), map_location="cpu", weights_only=True)
        sd = {k.replace("fc1", "ffn_pred_up").replace("fc2", "ffn_pred_down"): v for k, v in sd.items()}
        r, e = sd["ffn_pred_up.weight"].shape
        f, r2 = sd["ffn_pred_down.weight"].shape
        if (e, r2, f) != (n_embd, r, n_ff):
            raise ValueError(f"{fn}: predictor shapes {tuple(sd['ffn_pred_up.weight'].shape)}, "
                             f"{tuple(sd['ffn_pred_down.weight'].shape)} do not fit n_embd {n_embd}, n_ff {n_ff}")
        pred_lora.append(int(r))
        for k in ("ffn_pred_up.weight", "ffn_pred_up.bias", "ffn_pred_down.weight", "ffn_pred_down.bias"):
            if k.endswith("bias") and not pred_bias:  # ReluMLP(bias=False) + load_state_dict(strict=False) drops them (:4588-4589)
                continue
            if k in sd:
                pred_tensors.append((f"blk.{i}.{k}", sd[k].float().numpy()))
    if not pred_files:
        pred_lora = [0] * n_layer
    w.add_array_i32(f"{arch}.pred_lora", pred_lora)

    names = []

    def add(name, loader, shape, transpose, ttype):
        shape = tuple(int(s) for s in shape)
        if len(shape) == 1 or name.endswith("_norm.weight"):
            ttype = G.GGML_F32
        if len(shape) == 1:
            w.add_tensor(name, ttype, shape, lambda: np.ascontiguousarray(loader(), dtype=np.float32))
        else:
            rows, cols = (shape[1], shape[0]) if transpose else shape
            w.add_tensor(name, ttype, (cols, rows),
                         lambda: G.quantize_rows(loader().T if transpose else loader(), ttype))
        names.append(name)

    for k, (shape, load) in _checkpoint_tensors(model_dir).items():
        mapped = map_name(k)
        if mapped is None:
            continue
        name, _, is_down = mapped
        add(name, load, shape, transpose=is_down and bool(pred_files), ttype=wtype)
    for name, arr in pred_tensors:
        add(name, (lambda arr=arr: arr), arr.shape, transpose=False, ttype=ptype)
    w.write()
    return {"arch": arch, "n_layer": n_layer, "n_embd": n_embd, "n_ff": n_ff, "pred_lora": pred_lora, "tensors": names}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("model", help="directory with config.json and *.safetensors / pytorch_model*.bin")
    ap.add_argument("--pred-path", default=None, help="directory with the predictors model_<i>.pt (reference: --pred-path)")
    ap.add_argument("--pred-bias", action="store_true", help="keep fc1.bias / fc2.bias of the predictors")
    ap.add_argument("--outfile", required=True)
    ap.add_argument("--outtype", default="f16", choices=sorted(G.TYPE_NAMES))
    ap.add_argument("--pred-type", default=None, choices=["f16", "bf16", "f32"])
    a = ap.parse_args(argv)
    info = convert(a.model, a.outfile, a.pred_path, a.outtype, a.pred_bias, a.pred_type)
    print(json.dumps({k: v for k, v in info.items() if k != "tensors"} | {"n_tensors": len(info["tensors"])}))


if __name__ == "__main__":
    main()
