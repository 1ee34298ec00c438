"""Debug: the look-ahead layer kernel with pred == NULL on one shape/arch; which neurons' contributions differ from the oracle.
    python tools/dbg_layer4.py n_embd n_ff arch wtype [flags]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np, torch
import sparkinfer_b200 as sp, synth
from oracle_lib import Oracle

n_embd, n_ff, arch, wtype = (int(v) for v in sys.argv[1:5])
flags = int(sys.argv[5]) if len(sys.argv) > 5 else 0
o = Oracle()
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
seed = synth.SEED0 + 13 * n_embd + n_ff + arch
Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
qu, qg, qd = (o.quantize(wtype, w) for w in (Wu, Wg, Wd))
x = synth.activations(n_embd, 1, seed + 1)
p = synth.direct_mask(n_ff, 1, 0.10, seed + 2)
b_down = np.random.default_rng(seed + 4).normal(0, 0.1, n_embd).astype(np.float32)
y_ref, _, _, hid_ref = o.sparse_ffn(arch, wtype, qu, qg, qd, x, p, n_embd, n_ff, b_down=b_down)
comm = sp.ShardComm(n_embd + 1024)
ws = sp.Workspace(n_embd, n_ff, 1)
lst = sp.mask_to_list(dev(p[0]))
y = torch.full((n_embd,), float("nan"), device="cuda")
hid = torch.empty((n_ff,), device="cuda")
stats = torch.zeros((4,), dtype=torch.int32, device="cuda")
sp.sparse_ffn_lookahead(comm, ws, ffn=dict(arch=arch, w_type=wtype, up=dev(qu), gate=(None if arch == 2 else dev(qg)), down_t=dev(qd),
                                           cur=dev(x[0]), n_ff=n_ff, down_b=dev(b_down), flags=flags, out=y, hidden_out=hid, stats=stats),
                        list_in=lst)
y, hid, stats = y.cpu().numpy(), hid.cpu().numpy(), stats.cpu().numpy()
act = ~(p[0] < 0.5)
rows = np.nonzero(act)[0]
print("impl", os.environ.get("SPIF_LAYER_IMPL", "default"), "shape", n_embd, n_ff, "arch", arch, "flags", flags)
print(" rel_err y", np.abs(y - y_ref[0]).max() / np.abs(y_ref[0]).max(), "stats", stats, "ref active", act.sum(), "ref nonzero", np.count_nonzero(hid_ref[0] * act))
dh = np.abs(hid - hid_ref[0] * act)
print(" hidden: max abs diff", dh.max(), "at", int(dh.argmax()), "n wrong", int((dh > 1e-3 * np.abs(hid_ref).max()).sum()))
Wd16 = Wd.astype(np.float16).astype(np.float32)
nz = rows[hid_ref[0][rows] != 0]
c, *_ = np.linalg.lstsq(Wd16[nz].T.astype(np.float64), (y - y_ref[0]).astype(np.float64), rcond=None)
bad = np.nonzero(np.abs(c) > 1e-3 * np.abs(hid_ref[0][nz]).max())[0]
P = 148
blk = lambda r: int((np.searchsorted([(b * n_ff) // P for b in range(P + 1)], r, side="right") - 1))
print(" neurons whose contribution differs:", len(bad), "of", len(nz))
for b in bad[:20]:
    n = int(nz[b])
    print("   neuron", n, "coef", c[b], "h_ref", hid_ref[0][n], "h_gpu", hid[n], "rank", int(np.searchsorted(rows, n)), "block", blk(n),
          "rank in block", int((rows[(rows >= (blk(n) * n_ff) // P)] < n).sum()))
err = np.abs(y - y_ref[0])
print(" columns wrong", int((err > 1e-3 * np.abs(y_ref).max()).sum()), "of", n_embd, "first", np.nonzero(err > 1e-3 * np.abs(y_ref).max())[0][:10])
