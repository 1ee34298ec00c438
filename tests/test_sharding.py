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
