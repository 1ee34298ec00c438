"""Writes tests/golden/model_tiny_ref.gguf with the REFERENCE's own writer (/root/reference/gguf-py: GGUFWriter.add_pred_lora,
gguf_writer.py:1168; MODEL_ARCH.PROSPARSE_LLAMA), laid out the way convert_hf_to_gguf.py:4594-4653 lays out a ProSparse model:
F16 matrices, F32 vectors, ffn_down transposed, predictors as blk.N.ffn_pred_up / ffn_pred_down.  Run in the build container
only (the reference is not on the GPU box):  python tests/golden/make_model_gguf.py
tests/test_gguf_io.py reads the committed file with the library's parser and regenerates the values from tiny_model() below."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
N_LAYER, N_EMBD, N_FF, RANK, EPS, SEED = 2, 64, 128, 32, 1e-5, 20261105


def tiny_model():
    """name -> float32 array in checkpoint (HF) orientation; predictors under 'pred.<layer>.<fc1|fc2>.<weight|bias>'"""
    rng = np.random.default_rng(SEED)
    t = {}
    for l in range(N_LAYER):
        t[f"model.layers.{l}.mlp.gate_proj.weight"] = rng.normal(0, 0.05, (N_FF, N_EMBD)).astype(np.float32)
        t[f"model.layers.{l}.mlp.up_proj.weight"] = rng.normal(0, 0.05, (N_FF, N_EMBD)).astype(np.float32)
        t[f"model.layers.{l}.mlp.down_proj.weight"] = rng.normal(0, 0.05, (N_EMBD, N_FF)).astype(np.float32)
        t[f"model.layers.{l}.post_attention_layernorm.weight"] = rng.uniform(0.5, 1.5, (N_EMBD,)).astype(np.float32)
        t[f"pred.{l}.fc1.weight"] = rng.normal(0, N_EMBD ** -0.5, (RANK, N_EMBD)).astype(np.float32)
        t[f"pred.{l}.fc2.weight"] = rng.normal(0, RANK ** -0.5, (N_FF, RANK)).astype(np.float32)
        t[f"pred.{l}.fc1.bias"] = rng.normal(0, 0.1, (RANK,)).astype(np.float32)
        t[f"pred.{l}.fc2.bias"] = rng.normal(-0.5, 0.5, (N_FF,)).astype(np.float32)
    return t


def main():
    sys.path.insert(0, "/root/reference/gguf-py")
    import gguf
    assert gguf.__file__.startswith("/root/reference/"), gguf.__file__
    t = tiny_model()
    out = os.path.join(HERE, "model_tiny_ref.gguf")
    w = gguf.GGUFWriter(out, arch=gguf.MODEL_ARCH_NAMES[gguf.MODEL_ARCH.PROSPARSE_LLAMA])
    w.add_block_count(N_LAYER)
    w.add_embedding_length(N_EMBD)
    w.add_feed_forward_length(N_FF)
    w.add_layer_norm_rms_eps(EPS)
    w.add_pred_lora([RANK] * N_LAYER)
    f16 = lambda a: np.ascontiguousarray(a).astype(np.float16)
    for l in range(N_LAYER):
        w.add_tensor(f"blk.{l}.ffn_gate.weight", f16(t[f"model.layers.{l}.mlp.gate_proj.weight"]))
        w.add_tensor(f"blk.{l}.ffn_up.weight", f16(t[f"model.layers.{l}.mlp.up_proj.weight"]))
        w.add_tensor(f"blk.{l}.ffn_down.weight", f16(t[f"model.layers.{l}.mlp.down_proj.weight"].T))  # transposed, :4646
        w.add_tensor(f"blk.{l}.ffn_norm.weight", t[f"model.layers.{l}.post_attention_layernorm.weight"])
        w.add_tensor(f"blk.{l}.ffn_pred_up.weight", f16(t[f"pred.{l}.fc1.weight"]))
        w.add_tensor(f"blk.{l}.ffn_pred_up.bias", t[f"pred.{l}.fc1.bias"])
        w.add_tensor(f"blk.{l}.ffn_pred_down.weight", f16(t[f"pred.{l}.fc2.weight"]))
        w.add_tensor(f"blk.{l}.ffn_pred_down.bias", t[f"pred.{l}.fc2.bias"])
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    print(out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
