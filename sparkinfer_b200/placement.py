"""Neuron-group -> GPU-shard placement and the shard-level balancer (host logic; numpy only, no torch).

Reference: the cache manager places neuron GROUPS (g consecutive rows after the offline hotness reorder,
src/llama-sparkinfer.cpp:339-352) on the GPU by decayed firing rate, and rebalances online: build_dfr
(src/llama-graph.cpp:910-930) keeps an EMA per group, sparkinfer_reload_plan (src/llama-sparkinfer.cpp:45-91) diffs the
top-m_g groups against the resident set and emits swap pairs.  Re-targeted here: every group is
resident on exactly one of G GPU shards; the quantity to equalise is the expected number of ACTIVE rows per shard
(the slowest shard sets the start of the per-layer allreduce), i.e. minimise max_s sum_{grp in S_s} dfr[grp].
"""
import numpy as np


def initial_placement(group_hotness, n_shards):
    """snake-deal the groups sorted by hotness: equal group COUNT per shard and near-equal expected active rows.
    -> int32 array shard_of_group[n_g]"""
    h = np.asarray(group_hotness, dtype=np.float64)
    n_g = h.size
    assert n_g % n_shards == 0, "groups must divide evenly over the shards (equal rows per shard)"
    order = np.argsort(-h, kind="stable")
    shard = np.empty(n_g, dtype=np.int32)
    for k, grp in enumerate(order):
        rnd, pos = divmod(k, n_shards)
        shard[grp] = pos if rnd % 2 == 0 else n_shards - 1 - pos
    return shard


def shard_rows(shard_of_group, shard, g):
    """ascending neuron ids held by `shard` (this is the shard's neuron_idx table, mm-sparse.cu:20)"""
    groups = np.nonzero(np.asarray(shard_of_group) == shard)[0]
    return (groups[:, None] * g + np.arange(g)[None, :]).reshape(-1).astype(np.int32)


def shard_loads(dfr, shard_of_group, n_shards):
    return np.bincount(np.asarray(shard_of_group), weights=np.asarray(dfr, dtype=np.float64), minlength=n_shards)


def imbalance(dfr, shard_of_group, n_shards):
    """max_s load / mean_s load (1.0 = perfectly balanced)"""
    ld = shard_loads(dfr, shard_of_group, n_shards)
    return float(ld.max() / max(ld.mean(), 1e-30))


def rebalance_plan(dfr, shard_of_group, n_shards, max_swaps=8, min_gain=0.0):
    """Greedy swap plan, the analogue of sparkinfer_reload_plan's copy_pair list (ggml-sparkinfer.hpp:16-24): repeatedly
    take the most- and least-loaded shards and swap the pair of groups (one from each) that lowers the maximum most.
    Swaps keep the number of groups per shard constant, so buffers never resize: the slot of the outgoing group is
    reused by the incoming one (same as the reference's cache slots).
    -> (list of (group_from_hot_shard, group_from_cold_shard, hot_shard, cold_shard), new shard_of_group)"""
    d = np.asarray(dfr, dtype=np.float64)
    shard = np.array(shard_of_group, dtype=np.int32, copy=True)
    plan = []
    for _ in range(max_swaps):
        ld = shard_loads(d, shard, n_shards)
        hot, cold = int(ld.argmax()), int(ld.argmin())
        gap = ld[hot] - ld[cold]
        if hot == cold or gap <= 0:
            break
        gh = np.nonzero(shard == hot)[0]
        gc = np.nonzero(shard == cold)[0]
        # moving delta = d[a] - d[b] from hot to cold: new pair max = max(hot - delta, cold + delta) = mid + |delta - gap/2|,
        # so the best pair has delta > 0 closest to gap / 2.  For every a the two candidates are the neighbours of
        # d[a] - gap/2 in the sorted scores of the cold shard: O(n log n) per swap instead of the n x n table.
        oc = np.argsort(d[gc], kind="stable")
        dc = d[gc][oc]
        pos = np.searchsorted(dc, d[gh] - gap / 2)
        best = (np.inf, -1, -1)
        for cand in (np.clip(pos - 1, 0, gc.size - 1), np.clip(pos, 0, gc.size - 1)):
            delta = d[gh] - dc[cand]
            cost = np.where(delta > 0, ld[cold] + gap / 2 + np.abs(delta - gap / 2), np.inf)
            i = int(np.argmin(cost))
            if cost[i] < best[0]:
                best = (float(cost[i]), i, int(oc[cand[i]]))
        c_best, i, j = best
        if not np.isfinite(c_best) or ld[hot] - c_best <= min_gain:
            break
        a, b = int(gh[i]), int(gc[j])
        shard[a], shard[b] = cold, hot
        plan.append((a, b, hot, cold))
    return plan, shard
