"""ctypes bindings for the CHECKERS: oracle/libspif_oracle.so (our C restatement) and, when present,
oracle/_ref/libspif_ref.so (the unmodified reference CPU ops, built by oracle/Makefile).

Test infrastructure only -- the product package (sparkinfer_b200/) must never import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

F32, F16, Q4_0, Q8_0, BF16 = 0, 1, 2, 8, 30
ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, ARCH_SILU_GATED = 0, 1, 2, 3
TYPE_NAMES = {F16: "f16", BF16: "bf16", Q8_0: "q8_0", Q4_0: "q4_0", F32: "f32"}

c_vp, c_i, c_i64, c_f, c_sz, c_d = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float,
                                    ctypes.c_size_t, ctypes.c_double)


def _ptr(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def build_oracle():
    """(re)build oracle/libspif_oracle.so and, where /root/reference exists, oracle/_ref/libspif_ref.so"""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "-j8", "all"], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


class _Lib:
    def __init__(self, path, prefix):
        self.lib = ctypes.CDLL(path)
        self.p = prefix

    def fn(self, name, restype, argtypes):
        f = getattr(self.lib, self.p + name)
        f.restype = restype
        f.argtypes = argtypes
        return f


class Oracle:
    """our C restatement (O-cpu order, plus O-f64)"""

    def __init__(self):
        path = os.path.join(ORACLE_DIR, "libspif_oracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = _Lib(path, "spifo_")
        self.L = L
        self._row_size = L.fn("row_size", c_sz, [c_i, c_i64])
        self._quant = L.fn("quantize_rows", c_i, [c_i, c_vp, c_vp, c_i64, c_i64])
        self._dequant = L.fn("dequantize_row", c_i, [c_i, c_vp, c_vp, c_i64])
        self._q8act = L.fn("quantize_row_q8_0_act", None, [c_vp, c_vp, c_i64])
        self._compact = L.fn("mask_compact", c_i, [c_vp, c_i, c_f, c_vp])
        self._topk = L.fn("topk_mask", c_i, [c_vp, c_i, c_i, c_vp])
        self._mms = L.fn("mul_mat_sparse", c_i, [c_i, c_vp, c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_f])
        self._axpy = L.fn("axpy_sparse", c_i, [c_i, c_vp, c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_f])
        self._pred = L.fn("predictor", c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i])
        self._ffn = L.fn("sparse_ffn", c_i, [c_i, c_i] + [c_vp] * 12 + [c_i, c_i, c_i, c_f])
        self._ffn64 = L.fn("sparse_ffn_f64", c_i, [c_i, c_i] + [c_vp] * 10 + [c_i, c_i, c_i, c_f])
        self._stack = L.fn("ffn_stack_time", c_d, [c_i, c_i, c_i, c_i] + [c_vp] * 8 + [c_i] * 7)

    def row_size(self, t, n):
        return self._row_size(t, n)

    def quantize(self, t, a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        rows, n = a.shape
        out = np.empty(rows * self.row_size(t, n), dtype=np.uint8)
        assert self._quant(t, _ptr(a), _ptr(out), rows, n) == 0
        return out

    def dequantize(self, t, buf, rows, n):
        out = np.empty((rows, n), dtype=np.float32)
        rs = self.row_size(t, n)
        buf = np.ascontiguousarray(buf).view(np.uint8)
        for r in range(rows):
            assert self._dequant(t, ctypes.c_void_p(buf.ctypes.data + r * rs), ctypes.c_void_p(out[r].ctypes.data), n) == 0
        return out

    def quantize_act_q8_0(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        out = np.empty(self.row_size(Q8_0, x.size), dtype=np.uint8)
        self._q8act(_ptr(x), _ptr(out), x.size)
        return out

    def mask_compact(self, p, thr=0.5):
        p = np.ascontiguousarray(p, dtype=np.float32)
        idx = np.empty(p.size, dtype=np.int32)
        c = self._compact(_ptr(p), p.size, thr, _ptr(idx))
        return idx[:c].copy()

    def topk_mask(self, score, k):
        score = np.ascontiguousarray(score, dtype=np.float32)
        out = np.empty(score.size, dtype=np.float32)
        self._topk(_ptr(score), score.size, k, _ptr(out))
        return out

    def mul_mat_sparse(self, t, w, x, p, n_embd, n_ff, neuron_idx=None, neuron_mask=None, thr=0.5):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        m = n_ff if neuron_idx is None else len(neuron_idx)
        dst = np.empty((T, n_ff), dtype=np.float32)
        rc = self._mms(t, _ptr(w), m, _ptr(neuron_idx), _ptr(neuron_mask), _ptr(x), _ptr(p), _ptr(dst), n_embd, n_ff, T, thr)
        assert rc == 0
        return dst

    def axpy_sparse(self, t, wt, hid, p, n_embd, n_ff, neuron_idx=None, neuron_mask=None, thr=0.5):
        hid = np.ascontiguousarray(hid, dtype=np.float32).reshape(-1, n_ff)
        T = hid.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        m = n_ff if neuron_idx is None else len(neuron_idx)
        dst = np.empty((T, n_embd), dtype=np.float32)
        rc = self._axpy(t, _ptr(wt), m, _ptr(neuron_idx), _ptr(neuron_mask), _ptr(hid), _ptr(p), _ptr(dst), n_embd, n_ff, T, thr)
        assert rc == 0
        return dst

    def predictor(self, pt, wpu, bpu, wpd, bpd, x, n_embd, r, n_ff):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        out = np.empty((T, n_ff), dtype=np.float32)
        logits = np.empty((T, n_ff), dtype=np.float32)
        rc = self._pred(pt, _ptr(wpu), _ptr(bpu), _ptr(wpd), _ptr(bpd), _ptr(x), _ptr(out), _ptr(logits), n_embd, r, n_ff, T)
        assert rc == 0
        return out, logits

    def sparse_ffn(self, arch, t, wu, wg, wd, x, p, n_embd, n_ff, b_up=None, b_gate=None, b_down=None, thr=0.5):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        y = np.empty((T, n_embd), np.float32)
        up, gate, hid = (np.zeros((T, n_ff), np.float32) for _ in range(3))
        rc = self._ffn(arch, t, _ptr(wu), _ptr(wg), _ptr(wd), _ptr(b_up), _ptr(b_gate), _ptr(b_down), _ptr(x), _ptr(p),
                       _ptr(y), _ptr(up), _ptr(gate), _ptr(hid), n_embd, n_ff, T, thr)
        assert rc == 0
        return y, up, gate, hid

    def sparse_ffn_f64(self, arch, t, wu, wg, wd, x, p, n_embd, n_ff, b_up=None, b_gate=None, b_down=None, thr=0.5):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        y = np.empty((T, n_embd), np.float64)
        hid = np.empty((T, n_ff), np.float64)
        rc = self._ffn64(arch, t, _ptr(wu), _ptr(wg), _ptr(wd), _ptr(b_up), _ptr(b_gate), _ptr(b_down), _ptr(x), _ptr(p),
                         _ptr(y), _ptr(hid), n_embd, n_ff, T, thr)
        assert rc == 0
        return y, hid


class Ref:
    """the unmodified reference CPU ops (oracle/_ref). available() is False where it was not built."""

    PATH = os.path.join(ORACLE_DIR, "_ref", "libspif_ref.so")

    @classmethod
    def available(cls):
        return os.path.exists(cls.PATH)

    def __init__(self):
        L = _Lib(self.PATH, "spif_ref_")
        self.L = L
        self._describe = L.fn("describe", ctypes.c_char_p, [])
        self._row_size = L.fn("row_size", c_sz, [c_i, c_i64])
        self._quant = L.fn("quantize", c_sz, [c_i, c_vp, c_vp, c_i64, c_i64])
        self._dequant = L.fn("dequantize", c_i, [c_i, c_vp, c_vp, c_i64])
        self._mms = L.fn("mul_mat_sparse", c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i])
        self._axpy = L.fn("axpy_sparse", c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i])
        self._pred = L.fn("predictor", c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i])
        self._ffn = L.fn("sparse_ffn", c_i, [c_i, c_i] + [c_vp] * 12 + [c_i, c_i, c_i, c_i])
        self._stack = L.fn("ffn_stack_time", c_d, [c_i, c_i, c_i, c_i] + [c_vp] * 8 + [c_i] * 7)
        try:
            self._rms = L.fn("rms_norm_mul", c_i, [c_vp, c_vp, ctypes.c_float, c_i, c_i, c_vp, c_i])
        except AttributeError:  # an oracle/_ref built before round 2's f2 row
            self._rms = None

    def describe(self):
        return self._describe().decode()

    def rms_norm_mul(self, x, w, eps, n_threads=1):
        """the reference's build_norm(x, w, LLM_NORM_RMS): ggml_rms_norm + ggml_mul on its CPU backend"""
        x = np.ascontiguousarray(x, dtype=np.float32)
        w = np.ascontiguousarray(w, dtype=np.float32)
        n_tok, n_embd = x.shape
        out = np.empty_like(x)
        assert self._rms(x.ctypes.data, w.ctypes.data, float(eps), n_embd, n_tok, out.ctypes.data, n_threads) == 0
        return out

    def row_size(self, t, n):
        return self._row_size(t, n)

    def quantize(self, t, a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        rows, n = a.shape
        out = np.empty(rows * self.row_size(t, n), dtype=np.uint8)
        self._quant(t, _ptr(a), _ptr(out), rows, n)
        return out

    def dequantize(self, t, buf, rows, n):
        out = np.empty((rows, n), dtype=np.float32)
        assert self._dequant(t, _ptr(np.ascontiguousarray(buf)), _ptr(out), rows * n) == 0
        return out

    def mul_mat_sparse(self, t, w, x, p, n_embd, n_ff, neuron_mask=None, n_threads=1):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        if neuron_mask is None:
            neuron_mask = np.zeros(n_ff, np.int32)
        dst = np.empty((T, n_ff), dtype=np.float32)
        assert self._mms(t, _ptr(w), _ptr(x), _ptr(p), _ptr(neuron_mask), _ptr(dst), n_embd, n_ff, T, n_threads) == 0
        return dst

    def axpy_sparse(self, t, wt, hid, p, n_embd, n_ff, neuron_mask=None, n_threads=1):
        hid = np.ascontiguousarray(hid, dtype=np.float32).reshape(-1, n_ff)
        T = hid.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        if neuron_mask is None:
            neuron_mask = np.zeros(n_ff, np.int32)
        dst = np.empty((T, n_embd), dtype=np.float32)
        assert self._axpy(t, _ptr(wt), _ptr(hid), _ptr(p), _ptr(neuron_mask), _ptr(dst), n_embd, n_ff, T, n_threads) == 0
        return dst

    def predictor(self, pt, wpu, bpu, wpd, bpd, x, n_embd, r, n_ff, n_threads=1):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        out = np.empty((T, n_ff), dtype=np.float32)
        assert self._pred(pt, _ptr(wpu), _ptr(bpu), _ptr(wpd), _ptr(bpd), _ptr(x), _ptr(out), n_embd, r, n_ff, T, n_threads) == 0
        return out

    def sparse_ffn(self, arch, t, wu, wg, wd, x, p, n_embd, n_ff, b_up=None, b_gate=None, b_down=None, n_threads=1):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, n_embd)
        T = x.shape[0]
        p = np.ascontiguousarray(p, dtype=np.float32).reshape(T, n_ff)
        y = np.empty((T, n_embd), np.float32)
        up, gate, hid = (np.zeros((T, n_ff), np.float32) for _ in range(3))
        rc = self._ffn(arch, t, _ptr(wu), _ptr(wg), _ptr(wd), _ptr(b_up), _ptr(b_gate), _ptr(b_down), _ptr(x), _ptr(p),
                       _ptr(y), _ptr(up), _ptr(gate), _ptr(hid), n_embd, n_ff, T, n_threads)
        assert rc == 0
        return y, up, gate, hid


class RefCuda:
    """the reference's own CUDA kernels for the path, compiled unmodified for sm_100 (oracle/_ref/libspif_refcuda.so, built by
    oracle/Makefile from /root/reference).  GPU baseline of bench.py and a third, reported comparison; never the product."""

    PATH = os.path.join(ORACLE_DIR, "_ref", "libspif_refcuda.so")

    @classmethod
    def available(cls):
        return os.path.exists(cls.PATH)

    def __init__(self):
        self.lib = ctypes.CDLL(self.PATH)
        self.lib.spif_refcuda_describe.restype = ctypes.c_char_p
        self._stack = self.lib.spif_refcuda_stack
        self._stack.restype = c_i
        self._stack.argtypes = [c_i, c_i, c_i, c_i] + [c_vp] * 9 + [c_i, c_i, c_i, c_i, c_i, c_i, c_vp]

    def describe(self):
        return self.lib.spif_refcuda_describe().decode()

    def stack(self, arch, wtype, ptype, w_up, w_gate, w_down, wpu, wpd, bpd, x, y, p_io, n_embd, n_ff, r, with_pred=True,
              warmup=3, iters=10):
        """all tensors are torch CUDA tensors (lists per layer for the weights; x, y [L, n_embd], p_io [L, n_ff] float32).
        -> (us per step through a CUDA graph, us per step eager)"""
        nl = len(w_up)

        def arr(lst):
            if lst is None:
                return None
            a = (ctypes.c_void_p * nl)(*[ctypes.c_void_p(t.data_ptr()) for t in lst])
            return a

        keep = [arr(w_up), arr(w_gate), arr(w_down), arr(wpu), arr(wpd), arr(bpd)]
        out = (ctypes.c_double * 2)()
        rc = self._stack(arch, wtype, ptype, nl, *[ctypes.cast(k, c_vp) if k is not None else None for k in keep],
                         ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(y.data_ptr()), ctypes.c_void_p(p_io.data_ptr()),
                         n_embd, n_ff, r, 1 if with_pred else 0, warmup, iters, ctypes.cast(out, c_vp))
        if rc != 0:
            raise RuntimeError(f"spif_refcuda_stack failed: {rc}")
        return float(out[0]), float(out[1])


def stack_time(lib, arch, wtype, ptype, w_up, w_gate, w_down, wpu, wpd, bpd, x, n_embd, n_ff, r, n_tok,
               n_threads, warmup, iters, want_y=False):
    """time one decode step over a stack of layers on the CPU (lib = Oracle() or Ref()). returns (median_us, y|None)"""
    nl = len(w_up)

    def arr(lst):
        if lst is None:
            return None
        a = (ctypes.c_void_p * nl)(*[ctypes.c_void_p(v.ctypes.data) for v in lst])
        return ctypes.cast(a, ctypes.c_void_p)

    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x) if want_y else None
    keep = [arr(w_up), arr(w_gate), arr(w_down), arr(wpu), arr(wpd), arr(bpd)]
    us = lib._stack(arch, wtype, ptype, nl, keep[0], keep[1], keep[2], keep[3], keep[4], keep[5], _ptr(x), _ptr(y),
                    n_embd, n_ff, r, n_tok, n_threads, warmup, iters)
    return us, y


def rms_norm_mul(x, w, eps):
    """numpy restatement of build_norm(x, w, LLM_NORM_RMS) as the reference CPU backend computes it
    (ggml_compute_forward_rms_norm_f32, ggml-cpu/ops.cpp:3730-3779, then ggml_mul): squares rounded to F32 and summed
    sequentially in double, mean and scale in F32, y = (x * scale) * w with a rounding after each product.
    Pinned bit-exact against oracle/_ref in tests/test_oracle_golden.py."""
    x = np.asarray(x, dtype=np.float32)
    w = np.asarray(w, dtype=np.float32)
    out = np.empty_like(x)
    for t in range(x.shape[0]):
        sq = (x[t] * x[t]).astype(np.float32)
        total = 0.0
        for v in sq.astype(np.float64):  # sequential double sum (the order matters only in the last bit of the double)
            total += v
        mean = np.float32(total / x.shape[1])
        scale = np.float32(1.0) / np.sqrt(np.float32(mean + np.float32(eps)), dtype=np.float32)
        out[t] = ((x[t] * np.float32(scale)).astype(np.float32) * w).astype(np.float32)
    return out
