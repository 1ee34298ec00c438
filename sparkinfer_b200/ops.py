"""Host-side mirror of the reference's ggml operator surface for the sparse-FFN path.

Names and argument meaning follow the reference's op constructors (ggml/include/ggml.h:964-980,1159-1165,1443-1455;
ggml/src/ggml.c:3310-3356) and graph recipe (src/llama-graph.cpp:865-1142):

    mul_mat_sparse(a, b, sparse_idx, neu_info)      GGML_OP_MUL_MAT_SPARSE
    axpy_sparse(a, b, sparse_idx, neu_info)         GGML_OP_AXPY_SPARSE
    fatrelu(a, threshold) / shifted_step(a, t)      GGML_OP_FATRELU / GGML_OP_SHIFTED_STEP
    predictor(...)                                  llm_graph_context::build_predictor
    sparse_ffn(...)                                 llm_graph_context::build_sparse_ffn (GPU-resident rows)

Weights are passed as raw byte tensors (uint8, CUDA) in ggml row layout plus a ggml type id; activations are F32
CUDA tensors, token-major.  Errors the reference turns into GGML_ASSERT/GGML_ABORT surface as SpifError.
torch is used for device memory and streams only.
"""
import ctypes

import torch

from . import _lib
from ._lib import FfnArgs, LayerIo, PredArgs, ShardCommStruct, check

GGML_TYPE_F32, GGML_TYPE_F16, GGML_TYPE_Q4_0, GGML_TYPE_Q8_0, GGML_TYPE_BF16 = 0, 1, 2, 8, 30
ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, ARCH_SILU_GATED = 0, 1, 2, 3
FLAG_CPU_ROUNDING, FLAG_MASK_BY_ROW, FLAG_MASK_STABLE, FLAG_PER_TOKEN = 1, 2, 8, 16
SPARSE_THRESHOLD = 0.5  # SPIF_SPARSE_THRESHOLD, mm-sparse.cu:5-7


def _p(t):
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "device-resident contiguous tensor required"
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def row_size(ggml_type, n_embd):
    return _lib.load().spif_row_size(ggml_type, n_embd)


class Workspace:
    """Caller-owned scratch (the analogue of ctx.pool() allocations, mmq-sparse.cu:339): allocated once, reused."""

    def __init__(self, n_embd, n_ff, n_tok=1, device="cuda"):
        self.nbytes = _lib.load().spif_workspace_size(n_embd, n_ff, n_tok)
        self.buf = torch.empty(self.nbytes + 256, dtype=torch.uint8, device=device)
        off = (-self.buf.data_ptr()) % 256
        self.ptr = ctypes.c_void_p(self.buf.data_ptr() + off)


class ShardComm:
    """Exchange buffers of the per-layer sum over GPU shards (struct spif_shard_comm).  The buffers are symmetric memory:
    every rank allocates the same layout and maps all peers' copies (torch.distributed._symmetric_memory is plumbing for
    the CUDA VMM / IPC handle exchange; the exchange itself is done by our kernels over NVLink peer stores).  world == 1
    works without torch.distributed (self-exchange; used by the single-GPU tests)."""

    SEQ_WORDS = 1024  # SPIF_COMM_SEQ_WORDS

    def __init__(self, cap_floats, group=None, device="cuda"):
        L = _lib.load()
        cap = (int(cap_floats) + 3) // 4 * 4
        world = 1 if group is None else torch.distributed.get_world_size(group)
        rank = 0 if group is None else torch.distributed.get_rank(group)
        data_b = (L.spif_shard_comm_data_bytes(world, cap) + 255) // 256 * 256
        if group is None:
            self.buf = torch.zeros(data_b, dtype=torch.uint8, device=device)
            ptrs = [self.buf.data_ptr()]
            self.handle = None
        else:
            import torch.distributed._symmetric_memory as symm
            self.buf = symm.empty(data_b, dtype=torch.uint8, device=device)
            self.buf.zero_()
            self.handle = symm.rendezvous(self.buf, group)
            off = int(getattr(self.handle, "offset", 0) or 0)
            ptrs = [int(p) + off for p in self.handle.buffer_ptrs]
            assert ptrs[rank] == self.buf.data_ptr(), "symmetric-memory offset"
            # NVLS multicast mapping of the same buffers (one store reaches every rank): SPIF_COMM_MULTICAST=0 keeps unicast stores
            import os
            try:
                mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
            except Exception:  # no NVLS on this box / driver: unicast stores
                mc = 0
            self.multicast = (mc + off) if (mc and os.environ.get("SPIF_COMM_MULTICAST", "1") != "0") else 0
            torch.cuda.synchronize()
            torch.distributed.barrier(group)  # everybody's buffers are zero before anyone stores into them
        self.seq = torch.zeros(self.SEQ_WORDS, dtype=torch.int32, device=device)
        c = ShardCommStruct()
        c.rank, c.world, c.cap_floats, c.reserved = rank, world, cap, 0
        for r in range(world):
            c.data[r] = ptrs[r]
        c.seq = self.seq.data_ptr()
        c.mc = getattr(self, "multicast", 0) or None
        self.c = c
        self.rank, self.world, self.cap = rank, world, cap

    def status(self):
        """synchronises the current stream; 0 = every exchange completed, 1 = a peer timed out (outputs were NaN)"""
        err = ctypes.c_int(0)
        check(_lib.load().spif_shard_comm_status(ctypes.byref(self.c), ctypes.byref(err), _stream()), "spif_shard_comm_status")
        return err.value

    def epoch(self):
        return int(self.seq[0].item())


def shard_allreduce(y, comm):
    """y = sum over shards of y, in place (spif_shard_allreduce: one-shot exchange over peer memory, rank-order sum)"""
    check(_lib.load().spif_shard_allreduce(_p(y), y.numel(), ctypes.byref(comm.c), _stream()), "spif_shard_allreduce")
    return y


def _ws(ws, n_embd, n_ff, n_tok, device):
    return ws if ws is not None else Workspace(n_embd, n_ff, n_tok, device)


def mask_compact(sparse_idx, neuron_idx=None, thr=SPARSE_THRESHOLD, flags=0, want_union=False):
    """-> (idx [n_tok, m] int32, count [n_tok] int32[, union_idx [m], union_count [1]])"""
    L = _lib.load()
    n_tok, n_mask = sparse_idx.shape
    m = n_mask if neuron_idx is None else neuron_idx.numel()
    n_ff = n_mask  # with FLAG_MASK_BY_ROW the mask is shard-local and n_mask == m
    idx = torch.empty((n_tok, m), dtype=torch.int32, device=sparse_idx.device)
    cnt = torch.empty((n_tok,), dtype=torch.int32, device=sparse_idx.device)
    uidx = torch.empty((m,), dtype=torch.int32, device=sparse_idx.device) if want_union else None
    ucnt = torch.empty((1,), dtype=torch.int32, device=sparse_idx.device) if want_union else None
    check(L.spif_mask_compact(_p(sparse_idx), _p(neuron_idx), m, n_ff, n_tok, thr, flags, _p(idx), _p(cnt), _p(uidx),
                              _p(ucnt), _stream()), "spif_mask_compact")
    return (idx, cnt, uidx, ucnt) if want_union else (idx, cnt)


def topk_mask(score, k, out=None):
    """sparse_idx[t][i] = 1.0 for the k largest score[t][i] (ties -> lower index), else 0.0 (Llama-3 / SiLU config)"""
    L = _lib.load()
    n_tok, n_ff = score.shape
    out = out if out is not None else torch.empty_like(score)
    check(L.spif_topk_mask(_p(score), n_ff, n_tok, k, _p(out), _stream()), "spif_topk_mask")
    return out


def mul_mat_sparse(a, a_type, b, sparse_idx, neu_info=None, n_ff=None, thr=SPARSE_THRESHOLD, flags=0, ws=None):
    """ggml_mul_mat_sparse(ctx, a, b, sparse_idx, neu_info): a = weights {n_embd, m} (bytes), b = x {n_embd, n_tok},
    result {n_ff, n_tok} F32, zero where inactive (ggml.c:3310-3331)."""
    L = _lib.load()
    n_tok, n_embd = b.shape
    m = a.numel() // max(1, row_size(a_type, n_embd)) if neu_info is None else neu_info.numel()
    n_ff = n_ff if n_ff is not None else (sparse_idx.shape[1] if not (flags & FLAG_MASK_BY_ROW) else m)
    dst = torch.empty((n_tok, n_ff), dtype=torch.float32, device=b.device)
    ws = _ws(ws, n_embd, n_ff, n_tok, b.device)
    check(L.spif_mul_mat_sparse(a_type, _p(a), _p(b), _p(sparse_idx), _p(neu_info), _p(dst), n_embd, n_ff, m, n_tok, thr,
                                flags, ws.ptr, ws.nbytes, _stream()), "spif_mul_mat_sparse")
    return dst


def axpy_sparse(a, a_type, b, sparse_idx, neu_info=None, n_embd=None, thr=SPARSE_THRESHOLD, flags=0, ws=None):
    """ggml_axpy_sparse(ctx, a, b, sparse_idx, neu_info): a = W_down^T {n_embd, m} (bytes), b = hidden {n_ff, n_tok},
    result {n_embd, n_tok} (ggml.c:3333-3356)."""
    L = _lib.load()
    n_tok, n_ff = b.shape
    assert n_embd is not None
    m = a.numel() // max(1, row_size(a_type, n_embd)) if neu_info is None else neu_info.numel()
    dst = torch.empty((n_tok, n_embd), dtype=torch.float32, device=b.device)
    ws = _ws(ws, n_embd, n_ff, n_tok, b.device)
    check(L.spif_axpy_sparse(a_type, _p(a), _p(b), _p(sparse_idx), _p(neu_info), _p(dst), n_embd, n_ff, m, n_tok, thr,
                             flags, ws.ptr, ws.nbytes, _stream()), "spif_axpy_sparse")
    return dst


def fatrelu(a, threshold):
    L = _lib.load()
    y = torch.empty_like(a)
    check(L.spif_fatrelu(_p(a), _p(y), a.numel(), threshold, _stream()), "spif_fatrelu")
    return y


def shifted_step(a, threshold):
    L = _lib.load()
    y = torch.empty_like(a)
    check(L.spif_shifted_step(_p(a), _p(y), a.numel(), threshold, _stream()), "spif_shifted_step")
    return y


def predictor(pred_up, pred_up_b, pred_down, pred_down_b, cur, p_type, r, n_ff, flags=0, want_logits=False, ws=None,
              out=None, logits_out=None):
    """build_predictor(cur, pred_up, pred_up_b, pred_down, pred_down_b): -> sparse_idx {n_ff, n_tok} (and logits)"""
    L = _lib.load()
    n_tok, n_embd = cur.shape
    out = out if out is not None else torch.empty((n_tok, n_ff), dtype=torch.float32, device=cur.device)
    logits = logits_out if logits_out is not None else (
        torch.empty((n_tok, n_ff), dtype=torch.float32, device=cur.device) if want_logits else None)
    ws = _ws(ws, n_embd, n_ff, n_tok, cur.device)
    check(L.spif_predictor(p_type, _p(pred_up), _p(pred_up_b), _p(pred_down), _p(pred_down_b), _p(cur), _p(out), _p(logits),
                           n_embd, r, n_ff, n_tok, flags, ws.ptr, ws.nbytes, _stream()), "spif_predictor")
    return (out, logits) if (want_logits or logits_out is not None) else out


def sparse_ffn(arch, w_type, up, gate, down_t, cur, sparse_idx, n_ff, neuron_idx=None, up_b=None, gate_b=None,
               down_b=None, thr=SPARSE_THRESHOLD, flags=0, want_hidden=False, want_stats=False, ws=None, out=None,
               stats=None, comm=None):
    """build_sparse_ffn for the rows resident on this device.  -> y {n_embd, n_tok} [, hidden {n_ff, n_tok}][, stats]"""
    L = _lib.load()
    n_tok, n_embd = cur.shape
    m = n_ff if neuron_idx is None else neuron_idx.numel()
    y = out if out is not None else torch.empty((n_tok, n_embd), dtype=torch.float32, device=cur.device)
    hid = torch.empty((n_tok, n_ff), dtype=torch.float32, device=cur.device) if want_hidden else None
    if stats is None and want_stats:
        stats = torch.zeros((n_tok, 4), dtype=torch.int32, device=cur.device)
    ws = _ws(ws, n_embd, n_ff, n_tok, cur.device)
    a = FfnArgs(arch=arch, type=w_type, n_embd=n_embd, n_ff=n_ff, m=m, n_tok=n_tok, flags=flags, thr=thr,
                w_up=_p(up), w_gate=_p(gate), w_down_t=_p(down_t), b_up=_p(up_b), b_gate=_p(gate_b), b_down=_p(down_b),
                x=_p(cur), sparse_idx=_p(sparse_idx), neuron_idx=_p(neuron_idx), y=_p(y), hidden_out=_p(hid),
                stats_out=_p(stats))
    if comm is not None:  # sharded layer: y is the sum over all shards on return (spif_sparse_ffn_sharded)
        check(L.spif_sparse_ffn_sharded(ctypes.byref(a), ctypes.byref(comm.c), ws.ptr, ws.nbytes, _stream()),
              "spif_sparse_ffn_sharded")
    else:
        check(L.spif_sparse_ffn(ctypes.byref(a), ws.ptr, ws.nbytes, _stream()), "spif_sparse_ffn")
    res = [y]
    if want_hidden:
        res.append(hid)
    if want_stats:
        res.append(stats)
    return res[0] if len(res) == 1 else tuple(res)


class ActiveList:
    """opaque device buffer holding one layer's active-row list (spif_active_list_bytes): produced by
    sparse_ffn_lookahead or mask_to_list, consumed by the next sparse_ffn_lookahead"""
    HDR, MAX_BLOCKS = 4, 1024

    def __init__(self, m, device="cuda"):
        self.m = int(m)
        n = _lib.load().spif_active_list_bytes(self.m)
        self.buf = torch.zeros(n // 4, dtype=torch.int32, device=device)

    def rows(self):
        """-> ascending active rows (host numpy), for tests"""
        h = self.buf.cpu().numpy()
        nblk = int(h[0])
        cnt = h[self.HDR:self.HDR + nblk]
        rows = h[self.HDR + self.MAX_BLOCKS:]
        out = []
        for b in range(nblk):
            s = (b * self.m) // nblk
            out.extend(rows[s:s + cnt[b]].tolist())
        return out


def mask_to_list(sparse_idx, thr=SPARSE_THRESHOLD, out=None):
    """float mask indexed by row {m} -> ActiveList (spif_mask_to_list: ballot + popc per block of rows)"""
    m = sparse_idx.numel()
    out = out if out is not None else ActiveList(m, sparse_idx.device)
    check(_lib.load().spif_mask_to_list(_p(sparse_idx), m, thr, _p(out.buf), _stream()), "spif_mask_to_list")
    return out


def sparse_ffn_lookahead(comm, ws, ffn=None, pred=None, list_in=None, list_out=None, io=None):
    """spif_sparse_ffn_lookahead: the sparse FFN of one layer (ffn: dict of sparse_ffn's arguments arch, w_type, up, gate,
    down_t, cur, n_ff, [neuron_idx, up_b, gate_b, down_b, flags, out, hidden_out, stats]) + the look-ahead predictor of
    the next (pred: dict pred_up, pred_up_b, pred_down, pred_down_b, cur, p_type, r, m, [flags, thr, out, logits_out]) in
    one launch; list_in / list_out are ActiveList objects.  io: dict(norm_w, norm_eps, residual) = the layer's RMSNorm before
    and residual add after, fused (spif_sparse_ffn_lookahead_ex); `cur` is then the RAW ffn_inp."""
    L = _lib.load()
    fa = pa = None
    if ffn is not None:
        cur = ffn["cur"]
        n_embd = cur.numel()
        n_ff = ffn["n_ff"]
        nidx = ffn.get("neuron_idx")
        m = n_ff if nidx is None else nidx.numel()
        fa = FfnArgs(arch=ffn["arch"], type=ffn["w_type"], n_embd=n_embd, n_ff=n_ff, m=m, n_tok=1, flags=ffn.get("flags", 0),
                     thr=SPARSE_THRESHOLD, w_up=_p(ffn["up"]), w_gate=_p(ffn.get("gate")), w_down_t=_p(ffn["down_t"]),
                     b_up=_p(ffn.get("up_b")), b_gate=_p(ffn.get("gate_b")), b_down=_p(ffn.get("down_b")), x=_p(cur),
                     sparse_idx=None, neuron_idx=_p(nidx), y=_p(ffn["out"]), hidden_out=_p(ffn.get("hidden_out")),
                     stats_out=_p(ffn.get("stats")))
    if pred is not None:
        cur = pred["cur"]
        pa = PredArgs(ptype=pred["p_type"], n_embd=cur.numel(), r=pred["r"], m=pred["m"], flags=pred.get("flags", 0),
                      thr=pred.get("thr", SPARSE_THRESHOLD), wpu=_p(pred["pred_up"]), bpu=_p(pred.get("pred_up_b")),
                      wpd=_p(pred["pred_down"]), bpd=_p(pred.get("pred_down_b")), x=_p(cur),
                      sparse_idx_out=_p(pred.get("out")), logits_out=_p(pred.get("logits_out")))
    if io is not None:
        lio = LayerIo(norm_w=_p(io.get("norm_w")), norm_eps=float(io.get("norm_eps", 0.0)), reserved=0, residual=_p(io.get("residual")))
        check(L.spif_sparse_ffn_lookahead_ex(ctypes.byref(fa) if fa is not None else None,
                                             ctypes.byref(pa) if pa is not None else None, ctypes.byref(lio),
                                             _p(list_in.buf) if list_in is not None else None,
                                             _p(list_out.buf) if list_out is not None else None, ctypes.byref(comm.c), ws.ptr,
                                             ws.nbytes, _stream()), "spif_sparse_ffn_lookahead_ex")
        return
    check(L.spif_sparse_ffn_lookahead(ctypes.byref(fa) if fa is not None else None, ctypes.byref(pa) if pa is not None else None,
                                      _p(list_in.buf) if list_in is not None else None,
                                      _p(list_out.buf) if list_out is not None else None, ctypes.byref(comm.c), ws.ptr,
                                      ws.nbytes, _stream()), "spif_sparse_ffn_lookahead")


def dfr_update(sparse_idx, dfr_scores, g, lam, ema=True, thr=SPARSE_THRESHOLD):
    """build_dfr's score update (src/llama-graph.cpp:911-918), in place on dfr_scores {n_ff / g}"""
    L = _lib.load()
    n_tok, n_ff = sparse_idx.shape
    check(L.spif_dfr_update(_p(sparse_idx), n_ff, n_tok, g, thr, lam, 1 if ema else 0, _p(dfr_scores), _stream()),
          "spif_dfr_update")
    return dfr_scores


# -- shard placement / swap planning (host code inside the library, no device work) -------------------------------------
def initial_placement(group_hotness, n_shards):
    """spif_initial_placement: snake-deal of the groups sorted by hotness -> int32 shard_of_group[n_g]"""
    import numpy as np
    L = _lib.load()
    h = np.ascontiguousarray(group_hotness, dtype=np.float32)
    out = np.empty(h.size, dtype=np.int32)
    check(L.spif_initial_placement(h.ctypes.data, h.size, n_shards, out.ctypes.data), "spif_initial_placement")
    return out


def rebalance_plan(dfr, shard_of_group, n_shards, max_swaps=8, min_gain=0.0):
    """spif_rebalance_plan (the re-targeted sparkinfer_reload_plan, src/llama-sparkinfer.cpp:45-91)
    -> (list of (group_from_hot_shard, group_from_cold_shard, hot_shard, cold_shard), new shard_of_group)"""
    import numpy as np
    L = _lib.load()
    d = np.ascontiguousarray(dfr, dtype=np.float32)
    shard = np.array(shard_of_group, dtype=np.int32, copy=True)
    plan = np.zeros((max(int(max_swaps), 1), 4), dtype=np.int32)
    n = ctypes.c_int(0)
    check(L.spif_rebalance_plan(d.ctypes.data, shard.ctypes.data, d.size, n_shards, int(max_swaps), float(min_gain),
                                plan.ctypes.data, ctypes.byref(n)), "spif_rebalance_plan")
    return [tuple(r) for r in plan[:n.value].tolist()], shard
