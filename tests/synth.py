"""Seeded synthetic inputs for the sparse-FFN path.  numpy only; shared by CPU and GPU tests."""
import numpy as np

SEED0 = 0x5BA7C0DE


def hotness(n_ff, g, rho, rng):
    """per-neuron firing probability: Zipf-like over groups of g neurons, normalised to mean rho, clipped to [0, 1]"""
    n_g = (n_ff + g - 1) // g
    ranks = rng.permutation(n_g) + 1
    w = 1.0 / ranks ** 0.7
    p_g = np.clip(w / w.mean() * rho, 0.0, 1.0)
    return np.repeat(p_g, g)[:n_ff]


def direct_mask(n_ff, n_tok, rho, seed, g=16, near=0.001):
    """sparse_idx in {0.1, 0.9} by Bernoulli(p_i), plus a fraction `near` of entries within 1e-5 of 0.5"""
    rng = np.random.default_rng(seed)
    p_i = hotness(n_ff, g, rho, rng)
    m = np.where(rng.random((n_tok, n_ff)) < p_i[None, :], 0.9, 0.1).astype(np.float32)
    k = int(near * m.size)
    if k:
        pos = rng.choice(m.size, k, replace=False)
        m.reshape(-1)[pos] = (0.5 + rng.uniform(-1e-5, 1e-5, k)).astype(np.float32)
    return m


def ffn_weights(n_embd, n_ff, seed, nmat=3):
    rng = np.random.default_rng(seed)
    return [rng.normal(0.0, 0.02, (n_ff, n_embd)).astype(np.float32) for _ in range(nmat)]


def activations(n_embd, n_tok, seed):
    return np.random.default_rng(seed).normal(0.0, 1.0, (n_tok, n_embd)).astype(np.float32)


def predictor_weights(n_embd, r, n_ff, rho, seed, g=16):
    """Wpu {r, n_embd}, Wpd {n_ff, r} ~ N(0, 1/fan_in); per-neuron bias at the Gaussian quantile that makes neuron i
    fire with probability p_i (logits are ~N(0, s^2) with s estimated analytically: relu halves the variance)"""
    from statistics import NormalDist
    rng = np.random.default_rng(seed)
    wpu = rng.normal(0.0, 1.0 / np.sqrt(n_embd), (r, n_embd)).astype(np.float32)
    wpd = rng.normal(0.0, 1.0 / np.sqrt(r), (n_ff, r)).astype(np.float32)
    p_i = np.clip(hotness(n_ff, g, rho, rng), 1e-4, 1 - 1e-4)
    s = np.sqrt(0.5)  # x ~ N(0,1): h = relu(N(0,1)), E[h^2] = 1/2 -> logit std ~ sqrt(1/2)
    nd = NormalDist()
    bpd = np.array([s * nd.inv_cdf(float(p)) for p in p_i], dtype=np.float32)
    return wpu, wpd, bpd
