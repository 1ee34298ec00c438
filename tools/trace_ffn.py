"""Device-side phase timing of the fused sparse-FFN kernel (globaltimer stamps per CTA). Debug tool for the GPU box."""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkinfer_b200 as sp  # noqa: E402
from sparkinfer_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-embd", type=int, default=5120)
    ap.add_argument("--n-ff", type=int, default=13824)
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--rho", type=float, default=0.10)
    ap.add_argument("--mode", default="fused")
    ap.add_argument("--wtype", default="f16", choices=["f16", "q4_0", "q8_0"])
    ap.add_argument("--stable", action="store_true", help="SPIF_FLAG_MASK_STABLE")
    a = ap.parse_args()
    E, F, L = a.n_embd, a.n_ff, a.layers
    gen = torch.Generator(device="cuda").manual_seed(7)
    mk = lambda: torch.empty((F, E), dtype=torch.float16, device="cuda").normal_(0, 0.02, generator=gen).view(torch.uint8)
    wtype, flags = sp.GGML_TYPE_F16, (sp.FLAG_MASK_STABLE if a.stable else 0)
    if a.wtype != "f16":
        bb = 18 if a.wtype == "q4_0" else 34
        wtype = sp.GGML_TYPE_Q4_0 if a.wtype == "q4_0" else sp.GGML_TYPE_Q8_0

        def mk():
            blk = torch.randint(0, 256, (F, E // 32, bb), dtype=torch.uint8, device="cuda", generator=gen)
            d = torch.full((F, E // 32, 1), 0.02 / 6, dtype=torch.float16, device="cuda")
            blk[:, :, :2] = d.view(torch.uint8).reshape(F, E // 32, 2)
            return blk.reshape(F, E // 32 * bb).contiguous()
    up, gate, down = [mk() for _ in range(L)], [mk() for _ in range(L)], [mk() for _ in range(L)]
    x = torch.randn((1, E), device="cuda", generator=gen)
    mask = [torch.where(torch.rand((1, F), device="cuda", generator=gen) < a.rho, 0.9, 0.1).float() for _ in range(L)]
    ws = sp.Workspace(E, F, 1)
    lib = _lib.load()
    P = torch.cuda.get_device_properties(0).multi_processor_count
    traces = [torch.zeros((P, 8), dtype=torch.int64, device="cuda") for _ in range(L)]
    lib.lib if False else None
    setter = lib.spif_debug_set_trace
    setter.argtypes = [ctypes.c_void_p]
    setter.restype = None
    y = torch.empty((1, E), device="cuda")

    def run(l):
        if a.mode == "fused":
            sp.sparse_ffn(sp.ARCH_PROSPARSE, wtype, up[l], gate[l], down[l], x, mask[l], F, flags=flags, ws=ws, out=y)
        elif a.mode == "gemv":
            sp.mul_mat_sparse(up[l], sp.GGML_TYPE_F16, x, mask[l], ws=ws)
        else:
            sp.axpy_sparse(down[l], sp.GGML_TYPE_F16, mask[l], mask[l], n_embd=E, ws=ws)

    for l in range(L):  # warm-up (also pages in code)
        run(l)
    torch.cuda.synchronize()
    for rep in range(2):
        for l in range(L):
            setter(ctypes.c_void_p(traces[l].data_ptr()))
            run(l)
        torch.cuda.synchronize()
    setter(None)
    names = ["entry->pdlwait", "pdlwait->list", "list->first_batch", "first_batch->loop_done", "loop_done->end"]
    prev0 = None
    for l in range(L):
        t = traces[l].cpu().numpy()
        t0 = t[:, 0].min()
        if prev0 is not None:
            print(f"   entry of layer {l} - entry of layer {l - 1}: {(t0 - prev0) / 1e3:6.2f} us; phase A done (quant kernel, stamp 7) "
                  f"med {np.median((t[:, 7] - t0) / 1e3):6.2f}")
        prev0 = t0
        rel = (t[:, :6] - t0) / 1e3
        seg = np.diff(rel, axis=1)
        line = " ".join(f"{n}: med {np.median(seg[:, i]):5.2f} max {seg[:, i].max():5.2f} |" for i, n in enumerate(names))
        print(f"layer {l}: cta_end med {np.median(rel[:, 5]):6.2f} max {rel[:, 5].max():6.2f} us; start spread {rel[:, 0].max():5.2f}; "
              f"items {t[:, 6].min()}..{t[:, 6].max()} :: {line}")


if __name__ == "__main__":
    main()
