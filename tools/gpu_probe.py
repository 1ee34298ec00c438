"""Exploratory timing of the kernels at model dimensions (run on the GPU box).  Not the bench: see bench.py."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkinfer_b200 as sp  # noqa: E402


def rand_half_bytes(rows, cols, std, gen):
    w = torch.empty((rows, cols), dtype=torch.float16, device="cuda")
    w.normal_(0.0, std, generator=gen)
    return w


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-embd", type=int, default=5120)
    ap.add_argument("--n-ff", type=int, default=13824)
    ap.add_argument("--layers", type=int, default=40)
    ap.add_argument("--rank", type=int, default=1024)
    ap.add_argument("--rho", type=float, default=0.10)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    gen = torch.Generator(device="cuda").manual_seed(1234)
    E, F, L, R = a.n_embd, a.n_ff, a.layers, a.rank
    up = [rand_half_bytes(F, E, 0.02, gen) for _ in range(L)]
    gate = [rand_half_bytes(F, E, 0.02, gen) for _ in range(L)]
    down = [rand_half_bytes(F, E, 0.02, gen) for _ in range(L)]
    wpu = [rand_half_bytes(R, E, E ** -0.5, gen) for _ in range(L)]
    wpd = [rand_half_bytes(F, R, R ** -0.5, gen) for _ in range(L)]
    x = [torch.randn((1, E), device="cuda", generator=gen) for _ in range(L)]
    mask = [torch.where(torch.rand((1, F), device="cuda", generator=gen) < a.rho, 0.9, 0.1).float() for _ in range(L)]
    ws = sp.Workspace(E, F, 1)
    y = [torch.empty((1, E), device="cuda") for _ in range(L)]
    pout = [torch.empty((1, F), device="cuda") for _ in range(L)]
    u8 = lambda t: t.view(torch.uint8)

    def run_ffn():
        for l in range(L):
            sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, u8(up[l]), u8(gate[l]), u8(down[l]), x[l], mask[l], F,
                          ws=ws, out=y[l])

    def run_pred():
        for l in range(L):
            sp.predictor(u8(wpu[l]), None, u8(wpd[l]), None, x[l], sp.GGML_TYPE_F16, R, F, ws=ws, out=pout[l])

    def run_gemv():
        for l in range(L):
            sp.mul_mat_sparse(u8(up[l]), sp.GGML_TYPE_F16, x[l], mask[l], ws=ws)

    def run_axpy():
        for l in range(L):
            sp.axpy_sparse(u8(down[l]), sp.GGML_TYPE_F16, mask[l], mask[l], n_embd=E, ws=ws)

    if a.check:
        l = 0
        act = ~(mask[l] < 0.5)
        u = (x[l] @ up[l].float().T) * act
        g = (x[l] @ gate[l].float().T) * act
        h = torch.where(g > 0.01, g, torch.zeros_like(g)) * u
        yr = h @ down[l].float()
        run_ffn()
        torch.cuda.synchronize()
        err = (y[l] - yr).abs().max() / yr.abs().max()
        print("check fused vs torch fp32:", float(err))
        z = torch.relu(x[l] @ wpu[l].float().T) @ wpd[l].float().T
        run_pred()
        torch.cuda.synchronize()
        print("check predictor:", float((pout[l] - torch.sigmoid(z)).abs().max()))

    res = {}
    for name, fn, nbytes in (
        ("ffn_fused", run_ffn, None), ("predictor", run_pred, (R * E + F * R) * 2), ("gemv_up", run_gemv, None),
            ("axpy_down", run_axpy, None)):
        fn()
        torch.cuda.synchronize()
        g = None
        if a.graph:
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                fn()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    fn()
            torch.cuda.synchronize()
        ts = []
        for _ in range(a.iters):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if g is not None:
                g.replay()
            else:
                fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / L)
        us = sorted(ts)[len(ts) // 2]
        if nbytes is None:
            A = sum(int((~(m < 0.5)).sum()) for m in mask) / L
            nb = {"ffn_fused": 3 * A * E * 2, "gemv_up": A * E * 2, "axpy_down": A * E * 2}[name]
        else:
            nb = nbytes
        res[name] = {"us_per_layer": round(us, 2), "GBps": round(nb / us / 1e3, 1), "MB": round(nb / 1e6, 2)}
        print(name, res[name], flush=True)
    print(json.dumps({"cfg": vars(a), "res": res}))


if __name__ == "__main__":
    main()
