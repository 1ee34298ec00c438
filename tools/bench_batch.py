"""Batched sparse-FFN timing on the GPU box (BASELINE.json config 3: decode at batch 1/8/32).

For each batch size T: L distinct layers (weights >> L2), per-token masks from the Zipf hotness model at mean activity rho,
one spif_sparse_ffn call per layer captured in a CUDA graph, CUDA-event timing.  Reports us/layer, tokens/s of the FFN
stack, the union size, and effective GB/s on the bytes the algorithm has to touch: (2 A_u + Z_u) * row_bytes with A_u the
union of active rows (gated arch), next to the per-token schedule (the reference's, T x the row traffic).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkinfer_b200 as sp  # noqa: E402

ROW_BYTES = {sp.GGML_TYPE_F16: lambda e: 2 * e, sp.GGML_TYPE_BF16: lambda e: 2 * e,
             sp.GGML_TYPE_Q8_0: lambda e: e // 32 * 34, sp.GGML_TYPE_Q4_0: lambda e: e // 32 * 18}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-embd", type=int, default=5120)
    ap.add_argument("--n-ff", type=int, default=13824)
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--rho", type=float, default=0.10)
    ap.add_argument("--type", default="f16", choices=["f16", "bf16", "q8_0", "q4_0"])
    ap.add_argument("--batches", default="2,8,16,32")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--per-token", action="store_true", help="also time the per-token schedule")
    a = ap.parse_args()
    E, F, L = a.n_embd, a.n_ff, a.layers
    wtype = {"f16": sp.GGML_TYPE_F16, "bf16": sp.GGML_TYPE_BF16, "q8_0": sp.GGML_TYPE_Q8_0, "q4_0": sp.GGML_TYPE_Q4_0}[a.type]
    rb = ROW_BYTES[wtype](E)
    gen = torch.Generator(device="cuda").manual_seed(11)

    def mk():
        if a.type in ("f16", "bf16"):
            dt = torch.float16 if a.type == "f16" else torch.bfloat16
            return torch.empty((F, E), dtype=dt, device="cuda").normal_(0, 0.02, generator=gen).view(torch.uint8)
        # quantised rows: random quants, one F16 scale per 32-element block (valid block_q8_0 / block_q4_0 bytes)
        nb = E // 32
        qbytes = 32 if a.type == "q8_0" else 16
        blk = torch.randint(0, 256, (F, nb, 2 + qbytes), dtype=torch.uint8, device="cuda", generator=gen)
        d = torch.full((F, nb, 1), 0.02 / (100 if a.type == "q8_0" else 6), dtype=torch.float16, device="cuda")
        blk[:, :, :2] = d.view(torch.uint8).reshape(F, nb, 2)
        return blk.reshape(F, nb * (2 + qbytes)).contiguous()

    up, gate, down = [mk() for _ in range(L)], [mk() for _ in range(L)], [mk() for _ in range(L)]
    n_g = F // 16
    out = []
    for T in [int(v) for v in a.batches.split(",")]:
        x = torch.randn((T, E), device="cuda", generator=gen)
        masks, stats = [], []
        for l in range(L):
            ranks = torch.randperm(n_g, device="cuda", generator=gen).float() + 1
            w = 1.0 / ranks ** 0.7
            p_g = torch.clamp(w / w.mean() * a.rho, 0, 1).repeat_interleave(16)
            masks.append(torch.where(torch.rand((T, F), device="cuda", generator=gen) < p_g[None, :], 0.9, 0.1).float())
            stats.append(torch.zeros((T, 4), dtype=torch.int32, device="cuda"))
        ws = sp.Workspace(E, F, T)
        y = torch.empty((T, E), device="cuda")

        def run(flags):
            for l in range(L):
                sp.sparse_ffn(sp.ARCH_PROSPARSE, wtype, up[l], gate[l], down[l], x, masks[l], F, flags=flags, ws=ws, out=y,
                              stats=stats[l], want_stats=True)

        res = {"T": T}
        for name, flags in (("batched", 0),) + ((("per_token", sp.FLAG_PER_TOKEN),) if a.per_token else ()):
            run(flags)
            torch.cuda.synchronize()
            st = torch.stack(stats).cpu().numpy()
            g = torch.cuda.CUDAGraph()
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                run(flags)
                s.synchronize()
                with torch.cuda.graph(g, stream=s):
                    run(flags)
            torch.cuda.synchronize()
            for _ in range(3):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / a.iters / L
            act = np.stack([(m >= 0.5).cpu().numpy() for m in masks])  # [L, T, F]
            A_u = act.any(axis=1).sum(axis=1).mean()
            A_tok = act.sum(axis=2).mean()
            if name == "batched":
                Z_u = float(st[:, 0, 3].mean()) if st[:, 0, 2].max() > 0 else A_u
                byt = (2 * A_u + Z_u) * rb
            else:
                Z_u = None
                byt = float((2 * st[:, :, 0] + st[:, :, 1]).sum(axis=1).mean()) * rb
            res[name] = {"us_per_layer": round(us, 2), "tok_s_stack40": round(T / (us * 40) * 1e6, 1),
                         "union_rows": float(A_u), "rows_per_token": float(A_tok), "algorithmic_MB": round(byt / 1e6, 2),
                         "GBps": round(byt / us / 1e3, 1)}
        out.append(res)
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
