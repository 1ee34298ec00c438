"""The model file of the path (GGUF): reader over the C-ABI (spif_gguf_*), a writer, and the row quantisers the converter needs.

Reference: the loader side is src/llama-model.cpp:661, 2730-2768 (pred_lora + the ffn / predictor tensors), the writer side is
gguf-py/gguf/gguf_writer.py (container) and ggml/src/ggml-quants.c:66-96, 204-230 (quantize_row_q4_0_ref / quantize_row_q8_0_ref)
plus ggml/src/ggml-impl.h fp32 -> fp16 / bf16 conversions.  Host code: numpy for bytes, the C library for parsing.  The
quantisers are bit-exact restatements (tests/test_gguf_io.py compares them with the reference's functions).
"""
import ctypes
import struct

import numpy as np

from . import _lib
from ._lib import check

GGML_F32, GGML_F16, GGML_Q4_0, GGML_Q8_0, GGML_I32, GGML_BF16 = 0, 1, 2, 8, 26, 30
TYPE_NAMES = {"f32": GGML_F32, "f16": GGML_F16, "bf16": GGML_BF16, "q8_0": GGML_Q8_0, "q4_0": GGML_Q4_0}
QK = 32


def row_bytes(ggml_type, n):
    if ggml_type == GGML_I32:
        return 4 * n
    rb = _lib.load().spif_row_size(ggml_type, n)
    if rb == 0 or (ggml_type in (GGML_Q4_0, GGML_Q8_0) and n % QK):
        raise ValueError(f"type {ggml_type} with rows of {n} elements is not supported by the path")
    return rb


# ---- consumer -----------------------------------------------------------------------------------------------------------
class ModelGguf:
    """hyper-parameters and tensor index of a model GGUF; read_rows copies a row range of one tensor into host memory"""

    def __init__(self, path):
        self.L = _lib.load()
        self.h = ctypes.c_void_p()
        check(self.L.spif_gguf_open(str(path).encode(), ctypes.byref(self.h)), f"spif_gguf_open({path})")
        arch = ctypes.create_string_buffer(64)
        nl, ne, nf, eps = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32(), ctypes.c_float()
        check(self.L.spif_gguf_hparams(self.h, arch, 64, ctypes.byref(nl), ctypes.byref(ne), ctypes.byref(nf), ctypes.byref(eps)),
              "spif_gguf_hparams")
        self.arch, self.n_layer, self.n_embd, self.n_ff, self.rms_eps = arch.value.decode(errors="surrogateescape"), nl.value, ne.value, nf.value, eps.value
        self.pred_rank = np.zeros(self.n_layer, dtype=np.int32)
        check(self.L.spif_gguf_pred_rank(self.h, self.pred_rank.ctypes.data_as(ctypes.c_void_p), self.n_layer), "spif_gguf_pred_rank")

    def close(self):
        if self.h:
            self.L.spif_gguf_close(self.h)
            self.h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def names(self):
        out, buf = [], ctypes.create_string_buffer(256)
        for i in range(self.L.spif_gguf_n_tensors(self.h)):
            check(self.L.spif_gguf_tensor_at(self.h, i, buf, 256, None), "spif_gguf_tensor_at")
            out.append(buf.value.decode(errors="surrogateescape"))  # names are bytes in the file
        return out

    def find(self, name):
        """-> _lib.GgufTensor or None (optional tensors: biases, ffn_gate of the ungated architectures)"""
        t = _lib.GgufTensor()
        rc = self.L.spif_gguf_find(self.h, name.encode(errors="surrogateescape"), ctypes.byref(t))
        if rc == -2:  # SPIF_E_SHAPE: no tensor of that name
            return None
        check(rc, f"spif_gguf_find({name})")
        return t

    def read_rows(self, name, row0=0, n_rows=None, out=None):
        """rows [row0, row0 + n_rows) of tensor `name` as bytes [n_rows, row_bytes] (1-D tensors: one element per row);
        `out`: a writable uint8 numpy view to fill (e.g. of a pinned torch tensor) instead of a fresh array"""
        t = self.find(name)
        if t is None:
            raise KeyError(name)
        vec = t.n_dims <= 1
        rows = t.ne[0] if vec else t.ne[1] * t.ne[2] * t.ne[3]
        n_rows = rows - row0 if n_rows is None else n_rows
        rb = row_bytes(t.type, 1 if vec else t.ne[0])
        if out is None:
            out = np.empty((n_rows, rb), dtype=np.uint8)
        assert out.dtype == np.uint8 and out.flags.c_contiguous  # too small a destination: SPIF_E_WORKSPACE from the library
        check(self.L.spif_gguf_read(self.h, ctypes.byref(t), row0, n_rows, out.ctypes.data_as(ctypes.c_void_p), out.nbytes),
              f"spif_gguf_read({name})")
        return out

    def gather_rows(self, name, rows, out=None):
        """rows[i] of tensor `name` -> row i of the result, any order (a model-split permutation, a shard's neuron list)"""
        t = self.find(name)
        if t is None:
            raise KeyError(name)
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        rb = row_bytes(t.type, 1 if t.n_dims <= 1 else t.ne[0])
        if out is None:
            out = np.empty((rows.size, rb), dtype=np.uint8)
        assert out.dtype == np.uint8 and out.flags.c_contiguous
        check(self.L.spif_gguf_gather(self.h, ctypes.byref(t), rows.ctypes.data_as(ctypes.c_void_p), rows.size,
                                      out.ctypes.data_as(ctypes.c_void_p), out.nbytes), f"spif_gguf_gather({name})")
        return out


# ---- row quantisers (the converter's --outtype) ---------------------------------------------------------------------------
def _roundf(v):
    """C roundf on float32 arrays: half away from zero, exact (no v + 0.5 double rounding)"""
    t = np.trunc(v)
    return t + np.where(np.abs(v - t) >= np.float32(0.5), np.copysign(np.float32(1), v), np.float32(0)).astype(np.float32)


def _bf16_bits(a):
    """ggml_compute_fp32_to_bf16 (ggml-impl.h): NaN -> quiet NaN with the payload's top bits, else round to nearest even"""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    nan = (u & 0x7FFFFFFF) > 0x7F800000
    r = (u + (np.uint32(0x7FFF) + ((u >> 16) & 1))) >> 16
    return np.where(nan, (u >> 16) | 64, r).astype(np.uint16)


def quantize_rows(a, ggml_type):
    """float32 [rows, n] -> bytes [rows, row_bytes(ggml_type, n)] in ggml row layout"""
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2
    rows, n = a.shape
    if ggml_type == GGML_F32:
        return a.view(np.uint8).reshape(rows, -1)
    if ggml_type == GGML_F16:
        with np.errstate(over="ignore"):
            return a.astype(np.float16).view(np.uint8).reshape(rows, -1)
    if ggml_type == GGML_BF16:
        return _bf16_bits(a).view(np.uint8).reshape(rows, -1)
    if n % QK:
        raise ValueError("block formats need rows of whole 32-element blocks")
    b = a.reshape(rows, n // QK, QK)
    if ggml_type == GGML_Q8_0:  # quantize_row_q8_0_ref: d = amax / 127, q = roundf(x / d)
        amax = np.abs(b).max(axis=2)
        d = amax / np.float32(127)
        with np.errstate(divide="ignore"):
            inv = np.where(d != 0, np.float32(1) / d, np.float32(0)).astype(np.float32)
        q = _roundf(b * inv[:, :, None]).astype(np.int8)
        out = np.empty((rows, n // QK, 34), dtype=np.uint8)
        out[:, :, :2] = d.astype(np.float16).view(np.uint8).reshape(rows, n // QK, 2)
        out[:, :, 2:] = q.view(np.uint8)
        return out.reshape(rows, -1)
    if ggml_type == GGML_Q4_0:  # quantize_row_q4_0_ref: d = (value of largest magnitude, first on ties) / -8
        k = np.abs(b).argmax(axis=2)  # first occurrence of the maximum, like the strict `amax < fabsf(v)`
        mx = np.take_along_axis(b, k[:, :, None], axis=2)[:, :, 0]
        d = mx / np.float32(-8)
        with np.errstate(divide="ignore"):
            inv = np.where(d != 0, np.float32(1) / d, np.float32(0)).astype(np.float32)
        x = b * inv[:, :, None] + np.float32(8.5)
        q = np.minimum(x.astype(np.int8), 15).astype(np.uint8)  # (int8_t) truncates toward zero
        out = np.empty((rows, n // QK, 18), dtype=np.uint8)
        out[:, :, :2] = d.astype(np.float16).view(np.uint8).reshape(rows, n // QK, 2)
        out[:, :, 2:] = q[:, :, :16] | (q[:, :, 16:] << 4)
        return out.reshape(rows, -1)
    raise ValueError(f"unsupported ggml type {ggml_type}")


# ---- producer ---------------------------------------------------------------------------------------------------------------
_KV_U32, _KV_I32, _KV_F32, _KV_BOOL, _KV_STR, _KV_ARR = 4, 5, 6, 7, 8, 9


class GgufWriter:
    """GGUF v3, little endian.  Tensors are registered with their type and shape first and produced one at a time while the
    file is written, so a 26 GB model never sits in memory twice."""

    def __init__(self, path, alignment=32):
        self.path, self.alignment = path, alignment
        self.kv, self.tensors = [], []

    @staticmethod
    def _str(s):
        b = s.encode()
        return struct.pack("<Q", len(b)) + b

    def add_str(self, key, v):
        self.kv.append(self._str(key) + struct.pack("<I", _KV_STR) + self._str(v))

    def add_u32(self, key, v):
        self.kv.append(self._str(key) + struct.pack("<II", _KV_U32, v))

    def add_i32(self, key, v):
        self.kv.append(self._str(key) + struct.pack("<Ii", _KV_I32, v))

    def add_f32(self, key, v):
        self.kv.append(self._str(key) + struct.pack("<If", _KV_F32, v))

    def add_array_i32(self, key, vs):
        vs = [int(v) for v in vs]
        self.kv.append(self._str(key) + struct.pack("<IIQ", _KV_ARR, _KV_I32, len(vs)) + struct.pack(f"<{len(vs)}i", *vs))

    def add_array_f32(self, key, vs):
        vs = [float(v) for v in vs]
        self.kv.append(self._str(key) + struct.pack("<IIQ", _KV_ARR, _KV_F32, len(vs)) + struct.pack(f"<{len(vs)}f", *vs))

    def add_tensor(self, name, ggml_type, ne, producer):
        """ne: ggml order (ne[0] = row length); producer() -> array of exactly the tensor's bytes, rows in order"""
        ne = tuple(int(v) for v in ne)
        rows = int(np.prod(ne[1:])) if len(ne) > 1 else 1
        self.tensors.append((name, ggml_type, ne, row_bytes(ggml_type, ne[0]) * rows, producer))

    def write(self):
        al = self.alignment
        pad = lambda n: (n + al - 1) // al * al
        head = [b"GGUF", struct.pack("<IQQ", 3, len(self.tensors), len(self.kv))] + self.kv
        off = 0
        for name, t, ne, nbytes, _ in self.tensors:
            head.append(self._str(name) + struct.pack("<I", len(ne)) + struct.pack(f"<{len(ne)}Q", *ne) + struct.pack("<IQ", t, off))
            off += pad(nbytes)
        blob = b"".join(head)
        with open(self.path, "wb") as f:
            f.write(blob)
            f.write(b"\0" * (pad(len(blob)) - len(blob)))
            for name, _, _, nbytes, producer in self.tensors:
                data = np.ascontiguousarray(producer())
                if data.nbytes != nbytes:
                    raise ValueError(f"{name}: producer gave {data.nbytes} bytes, the declared shape needs {nbytes}")
                f.write(data.view(np.uint8).reshape(-1).data)
                f.write(b"\0" * (pad(nbytes) - nbytes))
