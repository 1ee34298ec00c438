"""CPU: the model-split file (`-ms`) consumer / producer of the C-ABI (spif_model_split_*) against the reference's OWN GGUF
code (gguf_init_from_file / gguf_write_to_file, compiled unmodified into oracle/_ref/libspif_ref.so):
  ours-write -> reference-read, reference-write -> ours-read, the keys and tensor names the reference's cache manager loads
  (src/llama-sparkinfer.cpp:150-158, 269-276), then the load-time pipeline budget -> reorder on the parsed values."""
import ctypes
import os

import numpy as np
import pytest

from sparkinfer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "libspif_ref.so")
c_vp = ctypes.c_void_p


def make_split(n_layer=5, n_ff=1024, g=16, seed=3):
    rng = np.random.default_rng(seed)
    pattern = rng.random(n_layer).astype(np.float32)
    pattern /= pattern.sum()
    perms = np.stack([rng.permutation(n_ff) for _ in range(n_layer)]).astype(np.int32)
    return g, pattern, perms


def ours_write(path, g, pattern, perms):
    L = _lib.load()
    rc = L.spif_model_split_write(path.encode(), g, pattern.ctypes.data, len(pattern), perms.ctypes.data, perms.shape[1])
    assert rc == 0


def ours_read(path):
    L = _lib.load()
    h = c_vp()
    rc = L.spif_model_split_open(path.encode(), ctypes.byref(h))
    if rc:
        return rc, None
    g, nl, nf = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    assert L.spif_model_split_info(h, ctypes.byref(g), ctypes.byref(nl), ctypes.byref(nf)) == 0
    pattern = np.empty(nl.value, np.float32)
    assert L.spif_model_split_pattern(h, pattern.ctypes.data, nl.value) == 0
    perms = np.empty((nl.value, nf.value), np.int32)
    for l in range(nl.value):
        assert L.spif_model_split_perm(h, l, perms[l].ctypes.data, nf.value) == 0
    assert L.spif_model_split_perm(h, nl.value, perms[0].ctypes.data, nf.value) != 0
    L.spif_model_split_close(h)
    return 0, (g.value, pattern, perms)


def test_roundtrip_and_errors(tmp_path):
    g, pattern, perms = make_split()
    path = str(tmp_path / "ms.gguf")
    ours_write(path, g, pattern, perms)
    rc, (g2, pattern2, perms2) = ours_read(path)
    assert rc == 0 and g2 == g and np.array_equal(pattern2, pattern) and np.array_equal(perms2, perms)
    raw = open(path, "rb").read()
    for cut in (3, 20, 60, len(raw) // 2, len(raw) - 5):  # truncated anywhere: an error code, never a crash
        p2 = str(tmp_path / f"cut{cut}.gguf")
        open(p2, "wb").write(raw[:cut])
        assert ours_read(p2)[0] != 0
    p3 = str(tmp_path / "bad_magic.gguf")
    open(p3, "wb").write(b"GGUX" + raw[4:])
    assert ours_read(p3)[0] != 0
    # a tensor that is not a permutation is rejected by spif_model_split_perm
    bad = perms.copy()
    bad[2, 7] = bad[2, 8]
    p4 = str(tmp_path / "notperm.gguf")
    ours_write(p4, g, pattern, bad)
    L = _lib.load()
    h = c_vp()
    assert L.spif_model_split_open(p4.encode(), ctypes.byref(h)) == 0
    out = np.empty(perms.shape[1], np.int32)
    assert L.spif_model_split_perm(h, 1, out.ctypes.data, perms.shape[1]) == 0
    assert L.spif_model_split_perm(h, 2, out.ctypes.data, perms.shape[1]) != 0
    L.spif_model_split_close(h)
    assert L.spif_model_split_write(p4.encode(), 16, pattern.ctypes.data, len(pattern), perms.ctypes.data, 1000) != 0  # n_ff % g


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/libspif_ref.so not built (needs /root/reference)")
    lib = ctypes.CDLL(REF)

    class InitParams(ctypes.Structure):
        _fields_ = [("no_alloc", ctypes.c_bool), ("ctx", ctypes.POINTER(c_vp))]

    lib.gguf_init_from_file.restype = c_vp
    lib.gguf_init_from_file.argtypes = [ctypes.c_char_p, InitParams]
    lib.gguf_find_key.restype = ctypes.c_int64
    lib.gguf_find_key.argtypes = [c_vp, ctypes.c_char_p]
    lib.gguf_get_val_i32.restype = ctypes.c_int32
    lib.gguf_get_val_i32.argtypes = [c_vp, ctypes.c_int64]
    lib.gguf_get_arr_data.restype = c_vp
    lib.gguf_get_arr_data.argtypes = [c_vp, ctypes.c_int64]
    lib.gguf_get_arr_n.restype = ctypes.c_size_t
    lib.gguf_get_arr_n.argtypes = [c_vp, ctypes.c_int64]
    lib.gguf_get_n_tensors.restype = ctypes.c_int64
    lib.gguf_get_n_tensors.argtypes = [c_vp]
    lib.gguf_get_tensor_name.restype = ctypes.c_char_p
    lib.gguf_get_tensor_name.argtypes = [c_vp, ctypes.c_int64]
    lib.ggml_get_tensor.restype = c_vp
    lib.ggml_get_tensor.argtypes = [c_vp, ctypes.c_char_p]
    lib.ggml_get_data.restype = c_vp
    lib.ggml_get_data.argtypes = [c_vp]
    lib.ggml_nbytes.restype = ctypes.c_size_t
    lib.ggml_nbytes.argtypes = [c_vp]
    lib.gguf_free.argtypes = [c_vp]
    lib.ggml_free.argtypes = [c_vp]
    lib.gguf_init_empty.restype = c_vp
    lib.gguf_set_val_i32.argtypes = [c_vp, ctypes.c_char_p, ctypes.c_int32]
    lib.gguf_set_val_str.argtypes = [c_vp, ctypes.c_char_p, ctypes.c_char_p]
    lib.gguf_set_arr_data.argtypes = [c_vp, ctypes.c_char_p, ctypes.c_int, c_vp, ctypes.c_size_t]
    lib.gguf_add_tensor.argtypes = [c_vp, c_vp]
    lib.gguf_write_to_file.restype = ctypes.c_bool
    lib.gguf_write_to_file.argtypes = [c_vp, ctypes.c_char_p, ctypes.c_bool]

    class GgmlInit(ctypes.Structure):
        _fields_ = [("mem_size", ctypes.c_size_t), ("mem_buffer", c_vp), ("no_alloc", ctypes.c_bool)]

    lib.ggml_init.restype = c_vp
    lib.ggml_init.argtypes = [GgmlInit]
    lib.ggml_new_tensor_1d.restype = c_vp
    lib.ggml_new_tensor_1d.argtypes = [c_vp, ctypes.c_int, ctypes.c_int64]
    lib.ggml_set_name.restype = c_vp
    lib.ggml_set_name.argtypes = [c_vp, ctypes.c_char_p]
    lib._InitParams, lib._GgmlInit = InitParams, GgmlInit
    return lib


def test_reference_reads_our_file(ref, tmp_path):
    """the consumer's exact calls (src/llama-sparkinfer.cpp:150-158, 269-276) on a file written by spif_model_split_write"""
    g, pattern, perms = make_split(n_layer=7, n_ff=2048, g=32, seed=11)
    path = str(tmp_path / "ours.gguf")
    ours_write(path, g, pattern, perms)
    meta = c_vp()
    ctx = ref.gguf_init_from_file(path.encode(), ref._InitParams(False, ctypes.pointer(meta)))
    assert ctx, "the reference's gguf reader rejected the file"
    assert ref.gguf_get_val_i32(ctx, ref.gguf_find_key(ctx, b"ffn_group_size")) == g
    kid = ref.gguf_find_key(ctx, b"ffn_normalized_pattern")
    assert ref.gguf_get_arr_n(ctx, kid) == len(pattern)
    got = np.ctypeslib.as_array(ctypes.cast(ref.gguf_get_arr_data(ctx, kid), ctypes.POINTER(ctypes.c_float)), (len(pattern),))
    assert np.array_equal(got, pattern)
    assert ref.gguf_get_n_tensors(ctx) == perms.shape[0]
    for i in range(perms.shape[0]):
        name = ref.gguf_get_tensor_name(ctx, i)
        assert name == f"blk.{i}.ffn_reorder_perms".encode()
        t = ref.ggml_get_tensor(meta, name)
        assert t and ref.ggml_nbytes(t) == perms.shape[1] * 4
        data = np.ctypeslib.as_array(ctypes.cast(ref.ggml_get_data(t), ctypes.POINTER(ctypes.c_int32)), (perms.shape[1],))
        assert np.array_equal(data, perms[i])
    ref.gguf_free(ctx)
    ref.ggml_free(meta)


def test_we_read_the_reference_writers_file(ref, tmp_path):
    """a model-split file produced with the reference's gguf writer (plus keys / tensors we must skip) parses to the same values"""
    g, pattern, perms = make_split(n_layer=4, n_ff=512, g=16, seed=5)
    gctx = ref.gguf_init_empty()
    ref.gguf_set_val_str(gctx, b"general.name", b"model split for a synthetic model")   # unknown key, string
    ref.gguf_set_val_i32(gctx, b"ffn_group_size", g)
    extra = np.arange(9, dtype=np.int32)
    ref.gguf_set_arr_data(gctx, b"some.other.array", 5, extra.ctypes.data, len(extra))      # GGUF_TYPE_INT32 array, skipped
    ref.gguf_set_arr_data(gctx, b"ffn_normalized_pattern", 6, pattern.ctypes.data, len(pattern))
    ctx = ref.ggml_init(ref._GgmlInit(16 << 20, None, False))
    keep = []
    for l in range(perms.shape[0]):
        t = ref.ggml_new_tensor_1d(ctx, 26, perms.shape[1])  # GGML_TYPE_I32
        ref.ggml_set_name(t, f"blk.{l}.ffn_reorder_perms".encode())
        ctypes.memmove(ref.ggml_get_data(t), perms[l].ctypes.data, perms[l].nbytes)
        ref.gguf_add_tensor(gctx, t)
        keep.append(t)
    other = ref.ggml_new_tensor_1d(ctx, 0, 33)  # an F32 tensor with another name: skipped
    ref.ggml_set_name(other, b"blk.0.something_else")
    ref.gguf_add_tensor(gctx, other)
    path = str(tmp_path / "ref.gguf")
    assert ref.gguf_write_to_file(gctx, path.encode(), False)
    ref.gguf_free(gctx)
    ref.ggml_free(ctx)
    rc, (g2, pattern2, perms2) = ours_read(path)
    assert rc == 0 and g2 == g and np.array_equal(pattern2, pattern) and np.array_equal(perms2, perms)


def test_load_time_pipeline(tmp_path):
    """file -> per-layer cache budget (ctor :179-202) -> row reorder (:283-304), all through the C-ABI"""
    L = _lib.load()
    g, pattern, perms = make_split(n_layer=6, n_ff=256, g=16, seed=9)
    path = str(tmp_path / "ms.gguf")
    ours_write(path, g, pattern, perms)
    rc, (g2, pattern2, perms2) = ours_read(path)
    assert rc == 0
    n_embd, n_ff = 64, perms.shape[1]
    group_bytes = n_embd * 2 * g
    groups = np.zeros(len(pattern2), np.int32)
    budget = group_bytes * 3 * 40
    assert L.spif_cache_group_budget(budget, group_bytes, 1, pattern2.ctypes.data, len(pattern2), n_ff // g, groups.ctypes.data) == 0
    assert groups.sum() == min(40, len(pattern2) * (n_ff // g)) and groups.max() <= n_ff // g
    W = np.random.default_rng(1).integers(0, 65535, (n_ff, n_embd), dtype=np.uint16)
    out = np.empty_like(W)
    assert L.spif_reorder_rows(W.ctypes.data, out.ctypes.data, perms2[3].ctypes.data, n_ff, n_embd * 2) == 0
    assert np.array_equal(out, W[perms2[3]])
