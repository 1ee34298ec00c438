"""Generate tests/golden/*.npz from the UNMODIFIED reference CPU ops (oracle/_ref/libspif_ref.so, built by
oracle/Makefile from /root/reference).  Run in the build container only:  python tests/golden/make_golden.py

Each fixture holds the seeded inputs' recipe (sizes + seeds; inputs are re-derived with tests/synth.py, and a CRC of
the stored weight bytes guards against generator drift) and the reference's outputs.  The reference has no golden
vectors of its own for this path, so these are "outputs of the reference itself run here".
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402
from oracle_lib import BF16, F16, Q4_0, Q8_0, TYPE_NAMES, Ref  # noqa: E402

CASES = [  # (n_embd, n_ff, n_tok, rho)
    (256, 512, 1, 0.15), (512, 1024, 3, 0.10), (1024, 768, 8, 0.20), (160, 96, 2, 0.5),
]
PRED_CASES = [(256, 64, 512, 1), (512, 128, 1024, 2)]


def crc(a):
    return np.uint32(zlib.crc32(np.ascontiguousarray(a).tobytes()))


def main():
    ref = Ref()
    out = {}
    for t in (F16, BF16, Q8_0, Q4_0):
        for ci, (E, F, T, rho) in enumerate(CASES):
            seed = synth.SEED0 + 1000 * t + ci
            Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
            qu, qg, qd = (ref.quantize(t, w) for w in (Wu, Wg, Wd))
            x = synth.activations(E, T, seed + 1)
            p = synth.direct_mask(F, T, rho, seed + 2, near=0.01)
            k = f"{TYPE_NAMES[t]}_{ci}"
            out[k + "_meta"] = np.array([t, E, F, T, seed], dtype=np.int64)
            out[k + "_rho"] = np.float64(rho)
            out[k + "_wcrc"] = np.array([crc(qu), crc(qg), crc(qd)], dtype=np.uint32)
            out[k + "_up"] = ref.mul_mat_sparse(t, qu, x, p, E, F)
            if t != Q4_0:  # the reference CPU axpy aborts for Q4_0 (ggml-cpu.c:2225-2226)
                for arch in (0, 1, 2):
                    y, up, gate, hid = ref.sparse_ffn(arch, t, qu, qg, qd, x, p, E, F)
                    out[f"{k}_y{arch}"] = y
                    out[f"{k}_hid{arch}"] = hid
    for pt in (F16, BF16):
        for ci, (E, R, F, T) in enumerate(PRED_CASES):
            seed = synth.SEED0 + 77 * pt + ci
            wpu, wpd, bpd = synth.predictor_weights(E, R, F, 0.1, seed)
            bpu = np.random.default_rng(seed + 5).normal(0, 0.05, R).astype(np.float32)
            qpu, qpd = ref.quantize(pt, wpu), ref.quantize(pt, wpd)
            x = synth.activations(E, T, seed + 1)
            k = f"pred_{TYPE_NAMES[pt]}_{ci}"
            out[k + "_meta"] = np.array([pt, E, R, F, T, seed], dtype=np.int64)
            out[k + "_wcrc"] = np.array([crc(qpu), crc(qpd)], dtype=np.uint32)
            out[k + "_p"] = ref.predictor(pt, qpu, bpu, qpd, bpd, x, E, R, F)
    path = os.path.join(HERE, "ref_cpu_golden.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays;", ref.describe())


if __name__ == "__main__":
    main()
