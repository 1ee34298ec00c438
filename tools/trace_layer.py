"""Device-side phase timing of the look-ahead layer kernel (globaltimer stamps per CTA, spif_layer.cu LAYER_STAMP).
Debug tool for the GPU box: python tools/trace_layer.py [--n-embd 5120 --n-ff 13824 --r 1024 --layers 8]"""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sparkinfer_b200 as sp  # noqa: E402
from sparkinfer_b200 import _lib  # noqa: E402
from sparkinfer_b200.stack import FfnStack, LayerWeights  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-embd", type=int, default=5120)
    ap.add_argument("--n-ff", type=int, default=13824)
    ap.add_argument("--r", type=int, default=1024)
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--rho", type=float, default=0.10)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    sys.argv = sys.argv[:1]
    import bench
    E, F, L, R = a.n_embd, a.n_ff, a.layers, a.r
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(7)
    u8 = lambda t: t.view(torch.uint8).reshape(-1)
    half = lambda rows, cols, std: u8(torch.empty((rows, cols), dtype=torch.float16, device=dev).normal_(0.0, std, generator=gen))
    layers = []
    for l in range(L):
        p_n = np.repeat(bench.synth_hotness(F, a.rho, 1234 + l), bench.GROUP)
        bpd = torch.from_numpy(bench.bias_for(p_n)).to(dev)
        layers.append(LayerWeights(up=half(F, E, 0.02), gate=half(F, E, 0.02), down_t=half(F, E, 0.02),
                                   pred_up=half(R, E, E ** -0.5), pred_down=half(F, R, R ** -0.5), pred_down_b=bpd, m=F))
    st = FfnStack(0, sp.GGML_TYPE_F16, sp.GGML_TYPE_F16, E, F, R, layers, n_tok=1, device=dev)
    st.x.normal_(0, 1, generator=gen)
    lib = _lib.load()
    setter = lib.spif_debug_set_trace
    setter.argtypes = [ctypes.c_void_p]
    setter.restype = None
    P = torch.cuda.get_device_properties(0).multi_processor_count
    st.step_fused()
    torch.cuda.synchronize()
    g = st.capture(st.step_fused)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"graph replay: {e0.elapsed_time(e1) / 20 * 1e3 / L:.2f} us per layer ({L} layers + 1 predictor-only launch)")
    # traced eager steps: one trace buffer per launch (launch 0 = predictor only)
    traces = [torch.zeros((P, 512), dtype=torch.int64, device=dev) for _ in range(L + 1)]
    for rep in range(a.reps):
        for t in traces:
            t.zero_()
        setter(ctypes.c_void_p(traces[0].data_ptr()))
        st._lookahead(None, 0, 0)
        for l in range(L):
            setter(ctypes.c_void_p(traces[l + 1].data_ptr()))
            st._lookahead(l, l + 1 if l + 1 < L else None, l)
        torch.cuda.synchronize()
    setter(None)
    # producer stamps: 0 pre-wait prefetch issued, 1 wait done, 2 list built, 3 first GU issued, 4 flush retired, 5 loop done, 6 list out
    # consumer stamps: 8 first item, 9 h1 loaded, 10 END seen, 11 partials complete, 12 y written
    names = {7: "entry", 0: "prewait", 1: "wait", 14: "poll0", 2: "list", 8: "c.first", 13: "h1.sent", 9: "h1.got", 4: "flush", 10: "c.end", 6: "l.out", 11: "c.parts", 3: "s.ld", 5: "s.l2", 12: "c.y"}
    prev_end = None
    for i, t in enumerate(traces):
        t = t.cpu().numpy().astype(np.float64)
        t0 = t[:, 7][t[:, 7] > 0].min()
        line = []
        for k, nm in names.items():
            col = t[:, k]
            ok = col > 0
            if ok.any():
                rel = (col[ok] - t0) / 1e3
                line.append(f"{nm} {rel.min():5.2f}/{np.median(rel):5.2f}/{rel.max():5.2f}")
        end = max(t[:, 12].max(), t[:, 6].max(), t[:, 10].max())
        gap = "" if prev_end is None else f" gap_from_prev_end {(t0 - prev_end) / 1e3:5.2f}"
        print(f"launch {i}: total {(end - t0) / 1e3:6.2f} us start_spread {(t[:, 7].max() - t0) / 1e3:5.2f}{gap} | min/med/max since first CTA entry: " + " | ".join(line))
        prev_end = end
        if i in (0, 2, L):  # per-item timeline of the median CTA and CTA 0
            names_t = {1: "WPU", 2: "GU", 3: "DN", 4: "WPD", 5: "FLUSH", 6: "END"}
            slow = int(np.argmax(t[:, 13])) if t[:, 13].max() > 0 else int(np.argmax(t[:, 4]))
            for cta in (0, 77, slow):
                row = t[cta]
                print(f"   cta {cta}: " + " ".join(f"{nm}={(row[k] - t0) / 1e3:.2f}" for k, nm in names.items() if row[k] > 0))
                items = []
                for q in range(120):
                    ty = int(row[16 + 4 * q])
                    if ty == 0:
                        break
                    iss, lnd, ret = [(row[16 + 4 * q + j] - row[15]) / 1965.0 if row[16 + 4 * q + j] > 0 else float("nan") for j in (1, 2, 3)]
                    items.append(f"{names_t.get(ty, ty)}:{iss:.1f}/{lnd:.1f}/{ret:.1f}")
                print(f"   cta {cta} items (issued/landed/retired, us since CTA entry at 1965 MHz): " + " ".join(items))
                ex = {500: "epi.lastGU", 501: "epi.final", 502: "prod.final", 504: "help.flushbar"}
                print(f"   cta {cta} extra: " + " ".join(f"{nm}={(row[k] - row[15]) / 1965.0:.2f}" for k, nm in ex.items() if row[k] > 0))


if __name__ == "__main__":
    main()
