"""Stress: the look-ahead layer kernel (FFN only and FFN + predictor chains) launched many times with the workspace poisoned
(all bytes 0xFF = NaN partials) before every call; any CTA that reads a partial y before its owner wrote it shows up as NaN,
any other race as a result that differs from the first one of the same configuration (the kernel is bit-stable).
    python tools/stress_layer.py [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np, torch
import sparkinfer_b200 as sp, synth
from oracle_lib import Oracle, F16

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
o = Oracle()
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
bad_total = 0
for (n_embd, n_ff, rho) in ((1376, 2048, 0.10), (512, 1024, 0.10), (1376, 2048, 0.03), (5120, 13824, 0.10), (4096, 11008, 0.02)):
    for arch in (0, 1, 2):
        seed = synth.SEED0 + 13 * n_embd + n_ff + arch
        Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
        qu, qg, qd = (dev(o.quantize(F16, w)) for w in (Wu, Wg, Wd))
        x = dev(synth.activations(n_embd, 1, seed + 1)[0])
        first, nan_cnt, diff_cnt = None, 0, 0
        for it in range(iters):
            p = synth.direct_mask(n_ff, 1, rho, seed + 2)
            if it % 4 == 0:  # fresh objects like a new test does
                comm = sp.ShardComm(n_embd + 1024)
                ws = sp.Workspace(n_embd, n_ff, 1)
                lst = sp.mask_to_list(dev(p[0]))
            ws.buf.fill_(0xFF)
            y = torch.full((n_embd,), float("nan"), device="cuda")
            sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=arch, w_type=F16, up=qu, gate=(None if arch == 2 else qg), down_t=qd,
                                                       cur=x, n_ff=n_ff, flags=0, out=y), list_in=lst)
            if it % 3 == 0:  # a second launch right behind the first (PDL overlap), same inputs
                y2 = torch.full((n_embd,), float("nan"), device="cuda")
                sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=arch, w_type=F16, up=qu, gate=(None if arch == 2 else qg),
                                                           down_t=qd, cur=x, n_ff=n_ff, flags=0, out=y2), list_in=lst)
                if not torch.equal(y, y2):
                    diff_cnt += 1
            if torch.isnan(y).any():
                nan_cnt += 1
            elif first is None:
                first = y.clone()
            elif not torch.equal(first, y):
                diff_cnt += 1
                d = (first - y).abs()
                if diff_cnt <= 3:
                    print("    diff: cols", torch.nonzero(d > 0).flatten()[:12].tolist(), "n", int((d > 0).sum()), "max", float(d.max() / first.abs().max()))
        assert comm.status() == 0
        print(f"shape {n_embd}x{n_ff} rho {rho} arch {arch}: {iters} launches, NaN {nan_cnt}, differing {diff_cnt}", flush=True)
        bad_total += nan_cnt + diff_cnt
print("STRESS", "FAILED" if bad_total else "ok", bad_total)
