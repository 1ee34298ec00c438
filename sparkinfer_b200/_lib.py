"""ctypes loader for libspif_b200.so -- the C-ABI declared in include/spif_b200.h.

There is no fallback: if the shared library is missing the import of any op fails loudly.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libspif_b200.so")

c_vp, c_i, c_i64, c_f, c_sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t


class FfnArgs(ctypes.Structure):
    """struct spif_ffn_args (include/spif_b200.h)"""
    _fields_ = [
        ("arch", ctypes.c_int32), ("type", ctypes.c_int32), ("n_embd", ctypes.c_int32), ("n_ff", ctypes.c_int32),
        ("m", ctypes.c_int32), ("n_tok", ctypes.c_int32), ("flags", ctypes.c_int32), ("thr", ctypes.c_float),
        ("w_up", c_vp), ("w_gate", c_vp), ("w_down_t", c_vp), ("b_up", c_vp), ("b_gate", c_vp), ("b_down", c_vp),
        ("x", c_vp), ("sparse_idx", c_vp), ("neuron_idx", c_vp), ("y", c_vp), ("hidden_out", c_vp), ("stats_out", c_vp),
    ]


class PredArgs(ctypes.Structure):
    """struct spif_pred_args (include/spif_b200.h)"""
    _fields_ = [
        ("ptype", ctypes.c_int32), ("n_embd", ctypes.c_int32), ("r", ctypes.c_int32), ("m", ctypes.c_int32),
        ("flags", ctypes.c_int32), ("thr", ctypes.c_float),
        ("wpu", c_vp), ("bpu", c_vp), ("wpd", c_vp), ("bpd", c_vp), ("x", c_vp), ("sparse_idx_out", c_vp), ("logits_out", c_vp),
    ]


class LayerIo(ctypes.Structure):
    """struct spif_layer_io (include/spif_b200.h)"""
    _fields_ = [("norm_w", c_vp), ("norm_eps", ctypes.c_float), ("reserved", ctypes.c_int32), ("residual", c_vp)]


class GgufTensor(ctypes.Structure):
    """struct spif_gguf_tensor (include/spif_b200.h)"""
    _fields_ = [("type", ctypes.c_int32), ("n_dims", ctypes.c_int32), ("ne", ctypes.c_int64 * 4), ("offset", ctypes.c_uint64),
                ("nbytes", ctypes.c_uint64)]


class ShardCommStruct(ctypes.Structure):
    """struct spif_shard_comm (include/spif_b200.h)"""
    _fields_ = [("rank", ctypes.c_int32), ("world", ctypes.c_int32), ("cap_floats", ctypes.c_int32),
                ("reserved", ctypes.c_int32), ("data", c_vp * 8), ("seq", c_vp), ("mc", c_vp)]


# name -> (restype, argtypes); every symbol include/spif_b200.h declares
SYMBOLS = {
    "spif_abi_version": (c_i, []),
    "spif_error_string": (ctypes.c_char_p, [c_i]),
    "spif_device_sm_count": (c_i, [ctypes.POINTER(c_i)]),
    "spif_row_size": (c_sz, [c_i, c_i64]),
    "spif_workspace_size": (c_sz, [c_i, c_i, c_i]),
    "spif_mask_compact": (c_i, [c_vp, c_vp, c_i, c_i, c_i, c_f, c_i, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "spif_topk_mask": (c_i, [c_vp, c_i, c_i, c_i, c_vp, c_vp]),
    "spif_mul_mat_sparse": (c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_f, c_i, c_vp, c_sz, c_vp]),
    "spif_axpy_sparse": (c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_f, c_i, c_vp, c_sz, c_vp]),
    "spif_fatrelu": (c_i, [c_vp, c_vp, c_i64, c_f, c_vp]),
    "spif_shifted_step": (c_i, [c_vp, c_vp, c_i64, c_f, c_vp]),
    "spif_predictor": (c_i, [c_i, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i, c_i, c_i, c_i, c_i, c_vp, c_sz, c_vp]),
    "spif_sparse_ffn": (c_i, [ctypes.POINTER(FfnArgs), c_vp, c_sz, c_vp]),
    "spif_dfr_update": (c_i, [c_vp, c_i, c_i, c_i, c_f, c_f, c_i, c_vp, c_vp]),
    "spif_shard_comm_data_bytes": (c_sz, [c_i, c_i]),
    "spif_shard_allreduce": (c_i, [c_vp, c_i64, ctypes.POINTER(ShardCommStruct), c_vp]),
    "spif_sparse_ffn_sharded": (c_i, [ctypes.POINTER(FfnArgs), ctypes.POINTER(ShardCommStruct), c_vp, c_sz, c_vp]),
    "spif_shard_comm_status": (c_i, [ctypes.POINTER(ShardCommStruct), ctypes.POINTER(c_i), c_vp]),
    "spif_active_list_bytes": (c_sz, [c_i]),
    "spif_mask_to_list": (c_i, [c_vp, c_i, c_f, c_vp, c_vp]),
    "spif_sparse_ffn_lookahead": (c_i, [ctypes.POINTER(FfnArgs), ctypes.POINTER(PredArgs), c_vp, c_vp,
                                        ctypes.POINTER(ShardCommStruct), c_vp, c_sz, c_vp]),
    "spif_sparse_ffn_lookahead_ex": (c_i, [ctypes.POINTER(FfnArgs), ctypes.POINTER(PredArgs), ctypes.POINTER(LayerIo), c_vp, c_vp,
                                           ctypes.POINTER(ShardCommStruct), c_vp, c_sz, c_vp]),
    "spif_noalloc_max_bytes": (c_sz, []),
    "spif_sparse_ffn_noalloc": (c_i, [ctypes.POINTER(FfnArgs), c_vp]),
    "spif_initial_placement": (c_i, [c_vp, c_i, c_i, c_vp]),
    "spif_rebalance_plan": (c_i, [c_vp, c_vp, c_i, c_i, c_i, c_f, c_vp, ctypes.POINTER(c_i)]),
    "spif_cache_group_budget": (c_i, [c_i64, c_i64, c_i, c_vp, c_i, c_i, c_vp]),
    "spif_reorder_rows": (c_i, [c_vp, c_vp, c_vp, c_i64, c_i64]),
    "spif_balancer_create": (c_i, [c_i, c_i, c_i, c_i, c_vp, c_f, c_f, c_f, ctypes.POINTER(c_vp)]),
    "spif_balancer_destroy": (None, [c_vp]),
    "spif_balancer_slots": (c_i, [c_vp, c_i, c_vp]),
    "spif_balancer_state": (c_i, [c_vp, ctypes.POINTER(c_f), ctypes.POINTER(c_i64), ctypes.POINTER(c_i)]),
    "spif_balancer_plan": (c_i, [c_vp, c_vp, c_i, c_f, c_vp, c_i, ctypes.POINTER(c_i), ctypes.POINTER(c_f), ctypes.POINTER(c_f),
                                 ctypes.POINTER(c_i)]),
    "spif_balancer_commit": (c_i, [c_vp]),
    "spif_balancer_abort": (c_i, [c_vp]),
    "spif_model_split_open": (c_i, [ctypes.c_char_p, ctypes.POINTER(c_vp)]),
    "spif_model_split_close": (None, [c_vp]),
    "spif_model_split_info": (c_i, [c_vp, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    "spif_model_split_pattern": (c_i, [c_vp, c_vp, c_i]),
    "spif_model_split_perm": (c_i, [c_vp, c_i, c_vp, c_i]),
    "spif_model_split_write": (c_i, [ctypes.c_char_p, c_i, c_vp, c_i, c_vp, c_i]),
    "spif_gguf_open": (c_i, [ctypes.c_char_p, ctypes.POINTER(c_vp)]),
    "spif_gguf_close": (None, [c_vp]),
    "spif_gguf_hparams": (c_i, [c_vp, ctypes.c_char_p, c_i, ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32),
                                ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(c_f)]),
    "spif_gguf_pred_rank": (c_i, [c_vp, c_vp, c_i]),
    "spif_gguf_n_tensors": (c_i, [c_vp]),
    "spif_gguf_tensor_at": (c_i, [c_vp, c_i, ctypes.c_char_p, c_i, ctypes.POINTER(GgufTensor)]),
    "spif_gguf_find": (c_i, [c_vp, ctypes.c_char_p, ctypes.POINTER(GgufTensor)]),
    "spif_gguf_read": (c_i, [c_vp, ctypes.POINTER(GgufTensor), c_i64, c_i64, c_vp, c_sz]),
    "spif_gguf_gather": (c_i, [c_vp, ctypes.POINTER(GgufTensor), c_vp, c_i64, c_vp, c_sz]),
}

_lib = None


def load():
    """dlopen the product library; raises if it has not been built (python __graft_entry__.py build)"""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA extension has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class SpifError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = load().spif_error_string(code)
        super().__init__(f"{where}: spif error {code}: {msg.decode() if msg else '?'}")


def check(code, where):
    if code != 0:
        raise SpifError(code, where)
