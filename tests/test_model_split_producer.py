"""CPU: the model-split PRODUCER (sparkinfer_b200/model_split.py; the reference ships only the consumer,
src/llama-sparkinfer.cpp:150-158, 257, 269-352): what build_split emits has the properties the consumer relies on, survives the
file, and reordering a layer by it changes nothing but the order of the neurons (predictor mask permuted exactly, FFN output
equal up to summation order) -- checked with the oracle."""
import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, F16, Q8_0


def zipf_rates(n_layer, n_ff, seed, rho=0.1):
    rng = np.random.default_rng(seed)
    out = []
    for l in range(n_layer):
        w = 1.0 / (rng.permutation(n_ff) + 1.0) ** 0.7
        out.append(np.clip(w / w.mean() * rho * (1 + 0.5 * l), 0, 1))
    return np.stack(out)


def test_single_shard_split_sorts_by_heat_and_budgets_by_load(tmp_path):
    from sparkinfer_b200 import model_split as ms
    rates = zipf_rates(4, 2048, 1)
    pattern, perms = ms.build_split(rates, g=16)
    assert pattern.dtype == np.float32 and abs(float(pattern.sum()) - 1) < 1e-6
    assert np.allclose(pattern, rates.sum(1) / rates.sum(), rtol=1e-6)          # a hotter layer gets a larger share
    for l in range(4):
        assert np.array_equal(np.sort(perms[l]), np.arange(2048))               # a permutation: new_row -> old_row
        assert np.all(np.diff(rates[l][perms[l]]) <= 0)                         # hottest first
    path = tmp_path / "ms.gguf"
    ms.write_split(path, 16, pattern, perms)
    g2, pattern2, perms2 = ms.read_split(path)
    assert g2 == 16 and np.array_equal(pattern2, pattern) and np.array_equal(perms2, perms)
    with pytest.raises(ValueError):
        ms.build_split(rates[:, :2040], g=16)


@pytest.mark.parametrize("n_shards", [2, 4, 8])
def test_shard_aware_split_balances_contiguous_blocks(n_shards):
    from sparkinfer_b200 import model_split as ms
    n_ff, g = 4096, 16
    rates = zipf_rates(3, n_ff, 7)
    _, perms = ms.build_split(rates, g=g, n_shards=n_shards)
    _, hot_sorted = ms.build_split(rates, g=g, n_shards=1)
    m = n_ff // n_shards
    for l in range(3):
        assert np.array_equal(np.sort(perms[l]), np.arange(n_ff))
        r = rates[l][perms[l]]
        loads = r.reshape(n_shards, m).sum(axis=1)                              # shard s = rows [s*m, (s+1)*m) after the reorder
        assert loads.max() / loads.mean() < 1.02, loads
        naive = rates[l][hot_sorted[l]].reshape(n_shards, m).sum(axis=1)        # contiguous blocks of the plain hot sort
        assert naive.max() / naive.mean() > 1.5
        heat_g = r.reshape(n_shards, m // g, g).sum(axis=2)
        assert np.all(np.diff(heat_g, axis=1) <= 1e-12)                         # inside a shard: groups hottest first
        # groups stay whole: every group of the reordered matrix is a group of the hot-sorted order
        a = {tuple(x) for x in perms[l].reshape(-1, g)}
        b = {tuple(x) for x in hot_sorted[l].reshape(-1, g)}
        assert a == b


@pytest.mark.parametrize("wtype", [F16, Q8_0])
def test_reordered_layer_computes_the_same_ffn(oracle, wtype):
    """the consumer's reorder_tensor_2d / _1d (:283-323) applied with the producer's permutation: up / gate / down^T /
    predictor-down rows + bias move together; the mask is the old mask permuted, y is unchanged"""
    from sparkinfer_b200 import model_split as ms
    E, R, F = 256, 64, 512
    seed = synth.SEED0 + 31 + wtype
    Wu, Wg, Wd = synth.ffn_weights(E, F, seed)
    wpu, wpd, bpd = synth.predictor_weights(E, R, F, 0.15, seed + 1)
    x = synth.activations(E, 2, seed + 2)
    qu, qg, qd = (oracle.quantize(wtype, w).reshape(F, -1) for w in (Wu, Wg, Wd))
    qpu, qpd = oracle.quantize(F16, wpu), oracle.quantize(F16, wpd).reshape(F, -1)
    p, _ = oracle.predictor(F16, qpu, None, qpd, bpd, x, E, R, F)
    y = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, qu, qg, qd, x, p, E, F)[0]
    rates = (~(p < 0.5)).mean(axis=0)[None]
    _, perms = ms.build_split(rates, g=16, n_shards=2)
    perm = perms[0]
    ru, rg, rd, rpd = (ms.reorder_rows(q, perm) for q in (qu, qg, qd, qpd))
    rb = ms.reorder_rows(bpd, perm)
    assert np.array_equal(rb, bpd[perm]) and np.array_equal(ru, qu[perm])
    p2, _ = oracle.predictor(F16, qpu, None, rpd, rb, x, E, R, F)
    assert np.array_equal(p2, p[:, perm])
    y2 = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, ru, rg, rd, x, p2, E, F)[0]
    assert np.abs(y2 - y).max() <= 1e-5 * np.abs(y).max()


def test_shard_rows_from_split_partition_the_layer():
    from sparkinfer_b200 import model_split as ms
    rates = zipf_rates(3, 1024, 11)
    _, perms = ms.build_split(rates, g=16, n_shards=4)
    parts = [ms.shard_rows_from_split(perms, [2, 0], rank=r, world=4) for r in range(4)]
    for i, l in enumerate([2, 0]):
        rows = np.concatenate([parts[r][i] for r in range(4)])
        assert np.array_equal(rows, perms[l]) and all(parts[r][i].size == 256 for r in range(4))
    whole = ms.shard_rows_from_split(perms, [1])
    assert np.array_equal(whole[0], perms[1])
    with pytest.raises(ValueError):
        ms.shard_rows_from_split(perms, [0], 0, 3)
