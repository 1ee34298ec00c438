"""Producer of the model-split file (`-ms`): offline firing rates -> hotness reorder + per-layer budget pattern -> GGUF.

The reference only CONSUMES this file (sparkinfer_cache_manager ctor, src/llama-sparkinfer.cpp:150-158, 257, 269-352): KV
`ffn_group_size`, `ffn_normalized_pattern` (each layer's share of the VRAM budget) and per layer `blk.N.ffn_reorder_perms`
(new_row -> old_row).  It ships no tool that makes one.  What the consumer assumes about the contents:
  * after the reorder, a GROUP is g consecutive rows and is cached / moved as a unit (:339-352), so rows that fire together
    should sit together: the permutation sorts neurons by firing rate, hottest first;
  * the pattern splits one budget over the layers (:179-202): a layer with more expected active rows gets a larger share.
Shard-aware variant (n_shards > 1): the hot-sorted groups are dealt over the GPU shards -- hottest group first,
each to the least-loaded shard that still has room (offline there is time for that; the online spif_initial_placement is the
cheaper snake deal) -- and every shard's groups are laid out contiguously, so shard s holds rows [s*m, (s+1)*m) of the
reordered matrices -- one contiguous row-range read per tensor (spif_gguf_read) -- with near-equal expected active rows.

Host code: numpy for the index arithmetic, the C library for the file and the row reorder.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import check


def _deal_groups(heat_g, n_shards):
    """groups in descending heat, each to the least-loaded shard with fewer than n_g / n_shards groups -> shard_of_group"""
    n_g = heat_g.size
    cap = n_g // n_shards
    shard = np.empty(n_g, dtype=np.int32)
    load = np.zeros(n_shards)
    count = np.zeros(n_shards, dtype=np.int64)
    for grp in np.argsort(-heat_g, kind="stable"):
        open_ = np.nonzero(count < cap)[0]
        s = open_[np.argmin(load[open_])]
        shard[grp] = s
        load[s] += heat_g[grp]
        count[s] += 1
    return shard


def build_split(rates, g=16, n_shards=1):
    """rates [n_layer, n_ff]: fraction of tokens on which each neuron is active (offline profile).
    -> (pattern float32 [n_layer] summing to 1, perms int32 [n_layer, n_ff] with perms[l][new_row] = old_row)"""
    rates = np.ascontiguousarray(rates, dtype=np.float64)
    n_layer, n_ff = rates.shape
    if n_ff % g or (n_ff // g) % n_shards:
        raise ValueError("n_ff must be whole groups of g rows, dealt evenly over the shards")
    n_g = n_ff // g
    perms = np.empty((n_layer, n_ff), dtype=np.int32)
    for l in range(n_layer):
        by_heat = np.argsort(-rates[l], kind="stable").astype(np.int32)  # hottest neuron first; ties keep file order
        if n_shards > 1:
            # groups of the sorted order, dealt like the online placement (equal count, near-equal load per shard) ...
            heat_g = rates[l][by_heat].reshape(n_g, g).sum(axis=1)
            shard = _deal_groups(heat_g, n_shards)
            # ... then laid out shard by shard, each shard's groups still hottest first
            order_g = np.argsort(shard, kind="stable")
            by_heat = by_heat.reshape(n_g, g)[order_g].reshape(-1)
        perms[l] = by_heat
    load = rates.sum(axis=1)
    pattern = (load / load.sum() if load.sum() > 0 else np.full(n_layer, 1.0 / n_layer)).astype(np.float32)
    return pattern, perms


def shard_rows_from_split(perms, layers, rank=0, world=1):
    """rows of the MODEL file (old numbering) that shard `rank` of `world` holds, in the order it holds them: rows
    [rank*m, (rank+1)*m) of the reordered matrices.  -> list over `layers` of int32 arrays (FfnStack.from_gguf rows_of_layer)"""
    n_ff = perms.shape[1]
    if n_ff % world:
        raise ValueError("n_ff must divide over the shards")
    m = n_ff // world
    return [np.ascontiguousarray(perms[l][rank * m:(rank + 1) * m], dtype=np.int32) for l in layers]


def write_split(path, g, pattern, perms):
    """spif_model_split_write: a GGUF the unmodified reference loads (tests/test_model_split.py)"""
    pattern = np.ascontiguousarray(pattern, dtype=np.float32)
    perms = np.ascontiguousarray(perms, dtype=np.int32)
    check(_lib.load().spif_model_split_write(str(path).encode(), int(g), pattern.ctypes.data, len(pattern), perms.ctypes.data,
                                             perms.shape[1]), "spif_model_split_write")


def read_split(path):
    """-> (g, pattern [n_layer], perms [n_layer, n_ff])"""
    L = _lib.load()
    h = ctypes.c_void_p()
    check(L.spif_model_split_open(str(path).encode(), ctypes.byref(h)), f"spif_model_split_open({path})")
    try:
        g, nl, nf = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        check(L.spif_model_split_info(h, ctypes.byref(g), ctypes.byref(nl), ctypes.byref(nf)), "spif_model_split_info")
        pattern = np.empty(nl.value, np.float32)
        check(L.spif_model_split_pattern(h, pattern.ctypes.data, nl.value), "spif_model_split_pattern")
        perms = np.empty((nl.value, nf.value), np.int32)
        for l in range(nl.value):
            check(L.spif_model_split_perm(h, l, perms[l].ctypes.data, nf.value), "spif_model_split_perm")
    finally:
        L.spif_model_split_close(h)
    return g.value, pattern, perms


def reorder_rows(rows, perm):
    """spif_reorder_rows on host bytes [n, row_bytes] (or a 1-D vector): out[new] = rows[perm[new]] (:283-323)"""
    a = np.ascontiguousarray(rows)
    n = a.shape[0]
    flat = a.reshape(n, -1).view(np.uint8)
    out = np.empty_like(flat)
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    check(_lib.load().spif_reorder_rows(flat.ctypes.data, out.ctypes.data, perm.ctypes.data, n, flat.shape[1]), "spif_reorder_rows")
    return out.view(a.dtype).reshape(a.shape)


def profile_rates(stack, xs):
    """offline profile on the GPU: run the stack's predictors over the sample xs [n_samples, n_layer, n_embd] (the FFN inputs of
    every layer) and count how often each neuron's mask is on.  -> float64 [n_layer, n_ff] in the stack's CURRENT row order."""
    import torch
    cnt = torch.zeros((stack.L, stack.m), dtype=torch.float64, device=stack.x.device)
    for x in xs:
        stack.x[:, 0].copy_(torch.as_tensor(x, device=stack.x.device))
        stack.predict(0, 0)
        for l in range(stack.L - 1):
            stack.predict(l + 1, l)  # look-ahead order: the mask of layer l+1 comes from the input of layer l
        cnt += (~(stack.p[:, 0] < 0.5)).to(torch.float64)
    return (cnt / max(len(xs), 1)).cpu().numpy()
