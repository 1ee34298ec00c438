"""CPU: neuron-group -> shard placement, the shard-level balancer, and the N > 1 data path (world_size 2, gloo):
every rank computes ITS neurons with the oracle, one allreduce of y per layer, result == the unsharded oracle."""
import os
import sys

import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, F16, Q8_0, Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_initial_placement_is_balanced():
    from sparkinfer_b200 import placement
    rng = np.random.default_rng(1)
    hot = synth.hotness(13824, 16, 0.1, rng)[::16]           # per group
    for G in (2, 4, 8):
        shard = placement.initial_placement(hot, G)
        assert np.all(np.bincount(shard, minlength=G) == hot.size // G)      # equal rows per shard
        assert placement.imbalance(hot, shard, G) < 1.01
        # contiguous dealing (what "no placement" would give) is far worse on a Zipf profile sorted by hotness
        order = np.argsort(-hot)
        naive = np.empty_like(shard)
        naive[order] = np.arange(hot.size) // (hot.size // G)
        assert placement.imbalance(hot, naive, G) > 1.5
    rows = placement.shard_rows(placement.initial_placement(hot, 4), 2, 16)
    assert rows.size == 13824 // 4 and np.all(np.diff(rows) > 0)


def test_rebalance_plan_lowers_the_maximum_and_keeps_counts():
    from sparkinfer_b200 import placement
    rng = np.random.default_rng(2)
    n_g, G = 864, 8
    dfr = rng.random(n_g) ** 4
    # equal group counts but the worst possible load split: shard 0 holds the hottest n_g/G groups, and so on
    shard = np.empty(n_g, dtype=np.int32)
    shard[np.argsort(-dfr)] = np.arange(n_g) // (n_g // G)
    before = placement.imbalance(dfr, shard, G)
    counts = np.bincount(shard, minlength=G)
    cur = shard
    for _ in range(400):
        plan, cur = placement.rebalance_plan(dfr, cur, G, max_swaps=8)
        if not plan:
            break
        for a, b, hot, cold_s in plan:
            assert hot != cold_s
    after = placement.imbalance(dfr, cur, G)
    assert np.array_equal(np.bincount(cur, minlength=G), counts)             # swaps never change rows per shard
    assert after < before and after < 1.05, (before, after)


def _worker(rank, world, port, wtype, q):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
    import torch
    import torch.distributed as dist
    from sparkinfer_b200 import placement
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o = Oracle()
        E, F, T, g = 512, 1024, 2, 16
        Wu, Wg, Wd = synth.ffn_weights(E, F, 77)
        x = synth.activations(E, T, 78)
        p = synth.direct_mask(F, T, 0.15, 79)
        hot = synth.hotness(F, g, 0.15, np.random.default_rng(80))[::g]
        shard_of = placement.initial_placement(hot, world)
        rows = placement.shard_rows(shard_of, rank, g)
        # this rank's rows of every matrix, and the shard-local mask (SPIF_FLAG_MASK_BY_ROW semantics)
        qu, qg, qd = (o.quantize(wtype, w[rows]) for w in (Wu, Wg, Wd))
        y_part = o.sparse_ffn(ARCH_PROSPARSE, wtype, qu, qg, qd, x, p[:, rows], E, rows.size)[0]
        t = torch.from_numpy(y_part.copy())
        dist.all_reduce(t)                                   # the ONE exchange step of the path
        # the same through the op surface with a neuron_idx table: dense dst, disjoint support per shard
        u_part = o.mul_mat_sparse(wtype, qu, x, p, E, F, neuron_idx=rows)
        tu = torch.from_numpy(u_part.copy())
        dist.all_reduce(tu)
        if rank == 0:
            full = o.sparse_ffn(ARCH_PROSPARSE, wtype, *(o.quantize(wtype, w) for w in (Wu, Wg, Wd)), x, p, E, F)
            err = float(np.abs(t.numpy() - full[0]).max() / np.abs(full[0]).max())
            exact_u = bool(np.array_equal(tu.numpy(), full[1]))
            q.put((err, exact_u))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("wtype", [F16, Q8_0])
def test_sharded_path_world2_gloo(wtype):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + wtype
    procs = [ctx.Process(target=_worker, args=(r, 2, port, wtype, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(timeout=180)
        assert pr.exitcode == 0
    err, exact_u = q.get(timeout=5)
    assert exact_u, "per-shard gate/up results are disjoint: their sum must equal the unsharded result bit for bit"
    assert err < 2e-6, err                                   # only the summation order of y differs


# ---------------------------------------------------------------------------------------------------------------------
# balancer execution: swap plan -> rows exchanged between the two shards, slot tables stay in step on both ranks
def _balancer_worker(rank, world, port, q):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
    import torch
    import torch.distributed as dist
    from sparkinfer_b200 import placement
    from sparkinfer_b200.balancer import ShardBalancer
    from sparkinfer_b200.stack import LayerWeights
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g, n_g, E, R, L = 16, 32, 64, 8, 3
        m = n_g // world * g
        rng = np.random.default_rng(5)
        hots = [synth.hotness(n_g * g, g, 0.2, rng)[::g] for _ in range(L)]
        layers, sogs = [], []
        for l in range(L):
            order = np.argsort(-hots[l], kind="stable")
            sog = np.empty(n_g, dtype=np.int32)
            sog[order] = np.arange(n_g) // (n_g // world)     # worst case: the hottest groups all on shard 0
            sogs.append(sog)
            mine = np.nonzero(sog == rank)[0]

            def rows(tag, width, dtype):
                # every element of a group's rows encodes (layer, matrix, group): the content identifies the group
                v = (np.repeat(mine, g)[:, None] * 8 + tag + 1000 * l) * np.ones((1, width))
                return torch.from_numpy(v.astype(dtype))
            layers.append(LayerWeights(up=rows(0, E, np.float16).view(torch.uint8).reshape(-1),
                                       gate=rows(1, E, np.float16).view(torch.uint8).reshape(-1),
                                       down_t=rows(2, E, np.float16).view(torch.uint8).reshape(-1),
                                       pred_up=None, pred_down=rows(3, R, np.float16).view(torch.uint8).reshape(-1),
                                       pred_down_b=rows(4, 1, np.float32).reshape(-1), m=m))
        bal = ShardBalancer(layers, sogs, rank, world, group=None, g=g)
        for l in range(L):                                    # scores as the DFR kernel would leave them
            bal.dfr[l] = torch.from_numpy(hots[l][bal.local_groups(l)].astype(np.float32))
        res = bal.rebalance(max_swaps=6)
        ok = True
        for l in range(L):
            mine = bal.local_groups(l)
            w = layers[l]
            for tag, t, dt, width in ((0, w.up, torch.float16, E), (1, w.gate, torch.float16, E),
                                      (2, w.down_t, torch.float16, E), (3, w.pred_down, torch.float16, R)):
                got = t.view(dt).reshape(m, width).double().numpy()
                want = (np.repeat(mine, g)[:, None] * 8 + tag + 1000 * l) * np.ones((1, width))
                ok &= bool(np.array_equal(got, want.astype(np.float16).astype(np.float64)))
            ok &= bool(np.array_equal(w.pred_down_b.numpy(), (np.repeat(mine, g) * 8 + 4 + 1000 * l).astype(np.float32)))
            ok &= bool(np.array_equal(bal.dfr[l].numpy(), hots[l][mine].astype(np.float32)))   # scores moved with the rows
            ok &= sorted(np.concatenate(list(bal.slots[l])).tolist()) == list(range(n_g))        # still a partition
        again = bal.rebalance(max_swaps=6)                   # scores() after the move must reproduce the planned state
        lam_calm = bal.lam                                    # the promised balance held -> decay lowered (more reactive)
        for l in range(L):                                    # hotness changes completely: the promise cannot hold
            bal.dfr[l] = torch.from_numpy(np.sort(hots[(l + 1) % L])[::-1][rank::world].copy().astype(np.float32))
        bal.rebalance(max_swaps=6)
        lam_drift = bal.lam
        # the asynchronous form (plan -> staged exchange -> anchor): tables pending until finish, rows follow the tables
        for l in range(L):
            bal.dfr[l] = torch.from_numpy(hots[(l + 2) % L][bal.local_groups(l)].astype(np.float32))
        before_tabs = [s.copy() for s in bal.slots]
        info = bal.rebalance_begin(max_swaps=6)
        ok &= all(np.array_equal(a, b) for a, b in zip(before_tabs, bal.slots))       # not committed yet
        bal.rebalance_finish()
        ok &= info["swaps"] > 0 and not all(np.array_equal(a, b) for a, b in zip(before_tabs, bal.slots))
        for l in range(L):
            mine = bal.local_groups(l)
            got = layers[l].up.view(torch.float16).reshape(m, E).double().numpy()
            want = (np.repeat(mine, g)[:, None] * 8 + 0 + 1000 * l) * np.ones((1, E))
            ok &= bool(np.array_equal(got, want.astype(np.float16).astype(np.float64)))
        q.put((rank, ok, res, again["imbalance_before"], [s.tolist() for s in bal.slots], lam_calm, lam_drift))
    finally:
        dist.destroy_process_group()


def test_balancer_swaps_rows_world2_gloo():
    pytest.importorskip("torch")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_balancer_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    got = [q.get(timeout=180) for _ in range(2)]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    got.sort(key=lambda t: t[0])
    for _, ok, res, again_before, _, lam_calm, lam_drift in got:
        # adaptive decay (ggml-sparkinfer.hpp:169-173): 0.9 * 0.95 after a round whose promise held, * 1.05 after one that did not
        assert abs(lam_calm - 0.9 * 0.95) < 1e-6 and abs(lam_drift - 0.9 * 0.95 * 1.05) < 1e-6
        assert ok, "rows / bias / scores in a slot do not belong to the group the slot table names"
        assert res["swaps"] > 0 and res["bytes_sent_this_rank"] > 0
        assert res["imbalance_after"] < res["imbalance_before"] and res["imbalance_after"] < 1.1
        assert abs(again_before - res["imbalance_after"]) < 1e-6
    assert got[0][4] == got[1][4], "slot tables diverged between the ranks"


def test_rebalance_plan_matches_exhaustive_pair_search():
    """the sorted-search pair selection picks a pair with the same cost as the n x n table it replaced"""
    from sparkinfer_b200 import placement as P

    def exhaustive(d, shard, G, max_swaps):
        shard = shard.copy()
        for _ in range(max_swaps):
            ld = P.shard_loads(d, shard, G)
            hot, cold = int(ld.argmax()), int(ld.argmin())
            if ld[hot] - ld[cold] <= 0:
                break
            gh, gc = np.nonzero(shard == hot)[0], np.nonzero(shard == cold)[0]
            delta = d[gh][:, None] - d[gc][None, :]
            cost = np.maximum(ld[hot] - delta, ld[cold] + delta)
            cost[delta <= 0] = np.inf
            i, j = np.unravel_index(np.argmin(cost), cost.shape)
            if not np.isfinite(cost[i, j]) or ld[hot] - cost[i, j] <= 0:
                break
            shard[gh[i]], shard[gc[j]] = cold, hot
        return P.imbalance(d, shard, G)

    rng = np.random.default_rng(3)
    for G in (2, 4, 8):
        n = G * 48
        d = rng.random(n) ** 3
        shard = np.repeat(np.arange(G), n // G).astype(np.int32)[np.argsort(np.argsort(-d))]   # hottest block on shard 0
        plan, new = P.rebalance_plan(d, shard, G, max_swaps=24)
        assert np.all(np.bincount(new, minlength=G) == n // G)
        assert abs(P.imbalance(d, new, G) - exhaustive(d, shard, G, 24)) < 1e-9
        assert P.imbalance(d, new, G) < P.imbalance(d, shard, G)


def test_c_abi_planner_matches_host_logic():
    """spif_initial_placement / spif_rebalance_plan (host code in libspif_b200.so, no GPU needed) take the same decisions
    as placement.py on float32 scores, tie-breaks included"""
    import ctypes
    from sparkinfer_b200 import _lib, placement as P
    L = _lib.load()
    rng = np.random.default_rng(9)
    for G, n_g in ((2, 64), (4, 864), (8, 1792)):
        d = (rng.random(n_g) ** 3).astype(np.float32)
        d[::7] = d[3]                                           # plenty of exact ties
        got = np.empty(n_g, dtype=np.int32)
        assert L.spif_initial_placement(d.ctypes.data, n_g, G, got.ctypes.data) == 0
        assert np.array_equal(got, P.initial_placement(d, G))
        shard = np.repeat(np.arange(G), n_g // G).astype(np.int32)[np.argsort(np.argsort(-d, kind="stable"), kind="stable")]
        want_plan, want_shard = P.rebalance_plan(d, shard, G, max_swaps=32)
        mine = shard.copy()
        plan = np.zeros((32, 4), dtype=np.int32)
        n = ctypes.c_int(0)
        assert L.spif_rebalance_plan(d.ctypes.data, mine.ctypes.data, n_g, G, 32, 0.0, plan.ctypes.data, ctypes.byref(n)) == 0
        assert n.value == len(want_plan) and n.value > 0
        assert [tuple(r) for r in plan[:n.value].tolist()] == [tuple(int(v) for v in r) for r in want_plan]
        assert np.array_equal(mine, want_shard)
    bad = np.zeros(6, dtype=np.int32)
    assert L.spif_initial_placement(np.zeros(6, np.float32).ctypes.data, 6, 4, bad.ctypes.data) != 0   # 6 % 4 != 0


def test_c_abi_cache_budget_and_reorder():
    """load-time half of the cache manager (src/llama-sparkinfer.cpp:179-202, 283-311): groups per layer from the
    model-split file's normalized pattern, and the new_row -> old_row row reorder; host code, no GPU"""
    from sparkinfer_b200 import _lib
    L = _lib.load()
    rng = np.random.default_rng(21)

    def reference_budget(vram, group_bytes, gate, pattern, n_group):
        budget = int(vram // (group_bytes * (3 if gate else 2)))
        out = [min(int(np.float32(budget) * np.float32(p)), n_group) for p in pattern]
        left = budget - sum(out)
        while left > 0:
            before = left
            for il in range(len(out)):
                if left > 0 and out[il] < n_group:
                    out[il] += 1
                    left -= 1
            if left == before:
                break
        return out

    for n_layer, n_group, gate in ((40, 864, 1), (32, 896, 1), (24, 512, 0)):
        group_bytes = 16 * 2 * 5120
        pattern = rng.random(n_layer).astype(np.float32)
        pattern /= pattern.sum()
        for frac in (0.0, 0.13, 0.5, 0.97, 1.0, 3.0):
            vram = int(frac * n_layer * n_group * group_bytes * (3 if gate else 2))
            got = np.zeros(n_layer, dtype=np.int32)
            assert L.spif_cache_group_budget(vram, group_bytes, gate, pattern.ctypes.data, n_layer, n_group, got.ctypes.data) == 0
            want = reference_budget(vram, group_bytes, gate, pattern, n_group)
            if frac <= 1.0:
                assert got.tolist() == want, (n_layer, frac)
            assert got.max() <= n_group and (frac < 1.0 or np.all(got == n_group))       # full budget -> gpu_only everywhere

    n, rb = 1000, 34 * 5
    src = rng.integers(0, 256, (n, rb), dtype=np.uint8)
    perm = rng.permutation(n).astype(np.int32)
    dst = np.zeros_like(src)
    assert L.spif_reorder_rows(src.ctypes.data, dst.ctypes.data, perm.ctypes.data, n, rb) == 0
    assert np.array_equal(dst, src[perm])
    bias = rng.normal(size=n).astype(np.float32)                                          # 1-D tensor: rows of 4 bytes
    out = np.zeros_like(bias)
    assert L.spif_reorder_rows(bias.ctypes.data, out.ctypes.data, perm.ctypes.data, n, 4) == 0
    assert np.array_equal(out, bias[perm])
    perm[3] = n                                                                            # out of range -> error, no write
    assert L.spif_reorder_rows(src.ctypes.data, dst.ctypes.data, perm.ctypes.data, n, rb) != 0


def test_balancer_state_machine_c_abi():
    """spif_balancer (host C++): plan -> pending, commit / abort, the moves of both ends of every swap mirror each other, the
    tables stay a partition, lambda follows ggml-sparkinfer.hpp:169-173 re-targeted (promise held -> lower, drifted -> raise)"""
    import ctypes
    from sparkinfer_b200 import _lib, placement
    from sparkinfer_b200.balancer import Move
    L = _lib.load()
    n_layer, n_g, world = 3, 24, 4
    n_slot = n_g // world
    rng = np.random.default_rng(12)
    sog = np.stack([np.arange(n_g) // n_slot for _ in range(n_layer)]).astype(np.int32)       # contiguous = unbalanced
    hot = [np.sort(rng.random(n_g))[::-1].astype(np.float32) for _ in range(n_layer)]           # hottest groups on shard 0

    def make(rank):
        h = ctypes.c_void_p()
        assert L.spif_balancer_create(n_layer, n_g, world, rank, sog.ctypes.data, 0.9, 0.05, 0.02, ctypes.byref(h)) == 0
        return h

    def slots(h):
        out = []
        for l in range(n_layer):
            t = np.empty((world, n_slot), np.int32)
            assert L.spif_balancer_slots(h, l, t.ctypes.data) == 0
            out.append(t)
        return out

    def slot_scores(tabs, hotness):
        s = np.empty((world, n_layer, n_slot), np.float32)
        for l in range(n_layer):
            for r in range(world):
                s[r, l] = hotness[l][tabs[l][r]]
        return s

    def plan(h, scores, max_swaps, cap=256):
        mv = (Move * cap)()
        n, tot, b, a = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_float(0), ctypes.c_float(0)
        rc = L.spif_balancer_plan(h, scores.ctypes.data, max_swaps, 0.0, mv, cap, ctypes.byref(n), ctypes.byref(b), ctypes.byref(a),
                                  ctypes.byref(tot))
        return rc, [mv[i] for i in range(min(n.value, cap))], n.value, b.value, a.value, tot.value

    hs = [make(r) for r in range(world)]
    bad = sog.copy()
    bad[0, 0] = 1                                                                               # unequal split is rejected
    h_bad = ctypes.c_void_p()
    assert L.spif_balancer_create(n_layer, n_g, world, 0, bad.ctypes.data, 0.9, 0.05, 0.02, ctypes.byref(h_bad)) != 0
    sc = slot_scores(slots(hs[0]), hot)
    rc, _, need, _, _, _ = plan(hs[0], sc, 5, cap=1)
    assert rc == -4 and need > 1                                                                # SPIF_E_WORKSPACE, nothing pending
    res = [plan(h, sc, 5) for h in hs]
    assert all(r[0] == 0 for r in res)
    before, after, total = res[0][3], res[0][4], res[0][5]
    assert after < before and total > 0 and all((r[3], r[4], r[5]) == (before, after, total) for r in res)
    ref_before = np.mean([placement.imbalance(hot[l].astype(np.float64), sog[l], world) for l in range(n_layer)])
    assert abs(before - ref_before) < 1e-5
    # every swap appears once on each of its two ranks, mirrored, at the same (round, layer); ordering is (round, layer)
    seen = {}
    for rank, r in enumerate(res):
        keys = [(m.round, m.layer) for m in r[1]]
        assert keys == sorted(keys)
        for m in r[1]:
            seen.setdefault((m.round, m.layer), []).append((rank, m.peer, m.group_out, m.group_in))
    assert sum(len(v) for v in seen.values()) == 2 * total
    for v in seen.values():
        (r0, p0, o0, i0), (r1, p1, o1, i1) = v
        assert (r0, p0, o0, i0) == (p1, r1, i1, o1)
    # pending until commit; a second plan before commit / abort is refused
    assert all(np.array_equal(a, b) for a, b in zip(slots(hs[0]), [sog[l].reshape(world, n_slot) * 0 + np.arange(n_g).reshape(world, n_slot) for l in range(n_layer)]))
    assert plan(hs[0], sc, 5)[0] != 0
    assert L.spif_balancer_abort(hs[0]) == 0 and L.spif_balancer_commit(hs[0]) != 0
    assert plan(hs[0], sc, 5)[0] == 0                                                           # (lambda: promise held -> lowered)
    for h in hs:
        assert L.spif_balancer_commit(h) == 0
    tabs = slots(hs[1])
    for l in range(n_layer):
        assert sorted(tabs[l].reshape(-1).tolist()) == list(range(n_g))
        # the committed tables realise the planned balance
    got_after = np.mean([placement.imbalance(hot[l].astype(np.float64), np.argsort(tabs[l].reshape(-1)) // n_slot, world)
                         for l in range(n_layer)])
    assert abs(got_after - after) < 1e-5
    assert all(np.array_equal(a, b) for a, b in zip(slots(hs[0]), slots(hs[3])))
    lam = ctypes.c_float(0)
    assert L.spif_balancer_state(hs[1], ctypes.byref(lam), None, None) == 0 and abs(lam.value - 0.9) < 1e-7   # first plan: no history
    # the promise holds -> lambda lowered; then the hotness flips -> drifted -> raised
    rc = plan(hs[1], slot_scores(slots(hs[1]), hot), 5)
    assert rc[0] == 0 and L.spif_balancer_commit(hs[1]) == 0
    assert L.spif_balancer_state(hs[1], ctypes.byref(lam), None, None) == 0 and abs(lam.value - 0.9 * 0.95) < 1e-6
    flipped = [h_[::-1].copy() for h_ in hot]
    rc = plan(hs[1], slot_scores(slots(hs[1]), flipped), 5)
    assert rc[0] == 0 and L.spif_balancer_state(hs[1], ctypes.byref(lam), None, None) == 0
    assert abs(lam.value - 0.9 * 0.95 * 1.05) < 1e-6
    for h in hs:
        L.spif_balancer_destroy(h)
