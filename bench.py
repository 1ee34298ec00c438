#!/usr/bin/env python
"""bench.py -- decode tok/s of the activation-sparse FFN path (predictor -> mask -> gathered gate/up -> down axpy) on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run, one rank per GPU)
    python bench.py --impl reference ...                     (the reference's own CPU ops on the box's host cores)

A "step" is one decode token (bs = 1) through the sparse-FFN path of EVERY layer of the named architecture: per layer
the look-ahead predictor (src/llama-graph.cpp:932-946) and build_sparse_ffn (:896-1142).  Attention is out of scope
(SURVEY.md 8): the per-layer FFN inputs x_l are synthetic N(0,1) vectors, weights are random-init in the real shapes,
the predictor is synthetic with per-neuron biases that give a Zipf-like hotness at mean activity rho.

  value     whole-job tok/s with the x_l already resident in HBM (one D2D refresh of the x stack per step);
  e2e       the same metric with the x stack coming from pinned HOST memory every step and the y stack read back;
  roofline  the fused sparse-FFN kernel: algorithmic bytes (2A+Z)*row_bytes per layer over its live-measured duration;
  cpu_baseline  the reference CPU path (oracle/_ref, else the oracle port) on a bounded sample of the same workload.

At N > 1 the FFN neurons are sharded over the ranks and every layer ends in one NCCL allreduce of y (strong scaling).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

F16, Q4_0, Q8_0, BF16 = 1, 2, 8, 30
WORKLOADS = {
    # name: n_embd, n_ff, n_layer, activation recipe (spif_arch), label as in BASELINE.json configs
    "prosparse-13b": dict(n_embd=5120, n_ff=13824, n_layer=40, arch=0, label="ProSparse-Llama-2-13B FP16 bs=1 decode"),
    "prosparse-7b": dict(n_embd=4096, n_ff=11008, n_layer=32, arch=0, label="ProSparse-Llama-2-7B FP16 bs=1 decode"),
    "relullama-70b": dict(n_embd=8192, n_ff=28672, n_layer=80, arch=0, label="ReluLLaMA-70B FP16 bs=1 decode"),
    # configs[2] at batch 1 as a whole decode step: quantised FFN matrices (random valid blocks), F16 predictor
    "prosparse-13b-q4_0": dict(n_embd=5120, n_ff=13824, n_layer=40, arch=0, weights="q4_0",
                               label="ProSparse-Llama-2-13B Q4_0 bs=1 decode (F16 predictor)"),
    "prosparse-13b-q8_0": dict(n_embd=5120, n_ff=13824, n_layer=40, arch=0, weights="q8_0",
                               label="ProSparse-Llama-2-13B Q8_0 bs=1 decode (F16 predictor)"),
    # configs[3]: SiLU gate, mask = top-k of the predictor logits (k = rho * n_ff); no reference code path for it
    "llama3-8b-topk": dict(n_embd=4096, n_ff=14336, n_layer=32, arch=3, topk=True,
                           label="Llama-3-8B SiLU top-k FP16 bs=1 decode"),
}
GROUP = 16  # neurons per placement group


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="prosparse-13b", choices=sorted(WORKLOADS))
    ap.add_argument("--rho", type=float, default=0.10, help="mean fraction of neurons the predictor marks active")
    ap.add_argument("--rank", type=int, default=1024, help="predictor rank r (pred_lora)")
    ap.add_argument("--pool", type=int, default=8, help="distinct x stacks cycled through (distinct masks per token)")
    ap.add_argument("--cpu-layers", type=int, default=4, help="layers in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="per-layer sum over GPU shards: fused peer-memory exchange (default) or ncclAllReduce")
    ap.add_argument("--placement", default="snake", choices=["snake", "contiguous"],
                    help="N > 1: initial neuron-group placement (snake = placement.initial_placement; contiguous = none)")
    ap.add_argument("--balance-rounds", type=int, default=0,
                    help="N > 1: rounds of the online shard balancer (observe masks, plan, swap groups) before the timed region")
    ap.add_argument("--balance-tokens", type=int, default=32, help="decode steps observed per balancer round")
    ap.add_argument("--balance-async", action="store_true",
                    help="balancer rounds move the rows in windows between decode steps (rebalance_begin / _step / _finish) "
                         "instead of all at once; reports tok/s during the exchange")
    ap.add_argument("--balance-swaps", type=int, default=64, help="max group swaps per layer per balancer round")
    ap.add_argument("--no-batched", action="store_true", help="skip the batch 8/32 legs (BASELINE.json configs[2] shape)")
    ap.add_argument("--seed", type=int, default=0x5BA7C0DE)
    ap.add_argument("--schedule", default="fused", choices=["fused", "serial", "two-streams"],
                    help="fused (default): ONE launch per layer, FFN(l) + look-ahead predictor(l+1) (spif_sparse_ffn_lookahead), "
                         "true data edges of the reference graph; serial: the reference's node order, 4 launches per layer on one "
                         "stream; two-streams: round-1 schedule, free-running predictor chain next to the FFN chain (looser than "
                         "a real decode: kept for comparison only)")
    ap.add_argument("--parity-layers", type=int, default=4, help="layers of the timed stack checked against the CPU oracle")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the reference-CUDA-kernels leg (oracle/_ref/libspif_refcuda.so)")
    ap.add_argument("--no-scale-70b", action="store_true", help="N > 1: skip the relullama-70b scaling leg")
    ap.add_argument("--no-other-configs", action="store_true", help="N = 1: skip the prosparse-7b / llama3-8b-topk quick legs")
    ap.add_argument("--activations", default="f32", choices=["f32", "weight"],
                    help="f32: x / hidden values stay F32 (flags 0); weight: rounded to the weight type like the reference CPU "
                         "ops, ggml-cpu.c:1832-1856 (SPIF_FLAG_CPU_ROUNDING): exact f16 x f16 products, F32 accumulation")
    ap.add_argument("--no-early-mask", action="store_true", help="do not use SPIF_FLAG_MASK_STABLE")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons through NVML while the timed region runs"""

    def __init__(self, index, period=0.002):
        super().__init__(daemon=True)
        self.period, self.samples, self.reasons, self.max_mhz = period, [], set(), None
        self._stop_evt = threading.Event()
        self.h = None
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # NVML missing: clocks are reported as null, the run is still valid
            self.h = None

    def run(self):
        if self.h is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=1.0)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ---------------------------------------------------------------------------------------------------------------------
def synth_hotness(n_ff, rho, seed):
    """per-neuron firing probability, Zipf-like over groups of GROUP neurons, mean rho (same recipe as tests/synth.py)"""
    rng = np.random.default_rng(seed)
    n_g = n_ff // GROUP
    ranks = rng.permutation(n_g) + 1
    w = 1.0 / ranks ** 0.7
    p_g = np.clip(w / w.mean() * rho, 1e-4, 1.0 - 1e-4)
    return p_g  # per group


def bias_for(p_neuron):
    """logit ~ N(0, 1/2) for the synthetic predictor (x ~ N(0,1), relu halves the variance) -> bias at the p-quantile"""
    from statistics import NormalDist
    nd = NormalDist()
    s = float(np.sqrt(0.5))
    return np.array([s * nd.inv_cdf(float(p)) for p in p_neuron], dtype=np.float32)


# ---------------------------------------------------------------------------------------------------------------------
def build_stack(a, wl, rank, world, device):
    """random-init weights of the named architecture, resident on this rank's GPU (only this shard's rows)"""
    import torch
    import sparkinfer_b200 as sp
    from sparkinfer_b200 import placement
    from sparkinfer_b200.stack import FfnStack, LayerWeights

    E, F, L, R = wl["n_embd"], wl["n_ff"], wl["n_layer"], a.rank
    n_g = F // GROUP
    assert n_g % world == 0, "neuron groups must divide over the shards"
    m = F // world
    gen = torch.Generator(device=device).manual_seed(a.seed + 7919 * rank)
    u8 = lambda t: t.view(torch.uint8).reshape(-1)

    def half(rows, cols, std):
        return u8(torch.empty((rows, cols), dtype=torch.float16, device=device).normal_(0.0, std, generator=gen))

    wq = wl.get("weights", "f16")
    w_type = {"f16": sp.GGML_TYPE_F16, "q4_0": sp.GGML_TYPE_Q4_0, "q8_0": sp.GGML_TYPE_Q8_0}[wq]

    def ffn_rows(rows):
        """random-init rows of one FFN matrix in the workload's weight format (quantised: random valid blocks, d fixed)"""
        if wq == "f16":
            return half(rows, E, 0.02)
        bb, dval = (18, 0.02 / 6) if wq == "q4_0" else (34, 0.02 / 64)
        nb = E // 32
        blk = torch.randint(0, 256, (rows, nb, bb), dtype=torch.uint8, device=device, generator=gen)
        d = torch.full((rows, nb, 1), dval, dtype=torch.float16, device=device)
        blk[:, :, :2] = d.view(torch.uint8).reshape(rows, nb, 2)
        return blk.reshape(-1).contiguous()

    layers, expect, shard_maps = [], [], []
    for l in range(L):
        p_g = synth_hotness(F, a.rho, a.seed + l)                   # identical on every rank
        if a.placement == "contiguous" and world > 1:
            # no placement: the offline hotness reorder leaves the groups sorted, dealt in contiguous blocks the hottest
            # all land on shard 0 (what the online balancer then has to repair)
            shard_of = np.empty(n_g, dtype=np.int32)
            shard_of[np.argsort(-p_g, kind="stable")] = np.arange(n_g) // (n_g // world)
        else:
            shard_of = placement.initial_placement(p_g, world)
        shard_maps.append(shard_of)
        mine = np.nonzero(shard_of == rank)[0]
        p_n = np.repeat(p_g[mine], GROUP)                           # this shard's neurons, ascending
        expect.append(float(p_n.sum()))
        bpd = torch.from_numpy(bias_for(p_n)).to(device)
        # the replicated predictor up-projection must be the SAME on every rank: seed it by layer only
        gpu_l = torch.Generator(device=device).manual_seed(a.seed + 104729 * (l + 1))
        wpu = u8(torch.empty((R, E), dtype=torch.float16, device=device).normal_(0.0, E ** -0.5, generator=gpu_l))
        layers.append(LayerWeights(up=ffn_rows(m), gate=ffn_rows(m), down_t=ffn_rows(m), pred_up=wpu,
                                   pred_down=half(m, R, R ** -0.5), pred_down_b=bpd, m=m))
    group = torch.distributed.group.WORLD if world > 1 else None
    st = FfnStack(wl["arch"], w_type, sp.GGML_TYPE_F16, E, m, R, layers, n_tok=1, group=group, device=device,
                  exchange=a.exchange, topk=int(a.rho * m) if wl.get("topk") else 0,
                  flags=sp.FLAG_CPU_ROUNDING if a.activations == "weight" else 0)
    st.shard_maps = shard_maps
    return st, expect


def log(msg):
    if os.environ.get("SPIF_BENCH_VERBOSE"):
        print(f"[bench rank {os.environ.get('RANK', '0')} t={time.time() % 1000:.1f}] {msg}", file=sys.stderr, flush=True)


def run_ours(a):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the sparse-FFN path has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    import sparkinfer_b200 as sp  # noqa: F401  (raises if libspif_b200.so is missing)

    wl = WORKLOADS[a.workload]
    E, F, L, R = wl["n_embd"], wl["n_ff"], wl["n_layer"], a.rank
    log('init done')
    st, expect = build_stack(a, wl, rank, world, device)
    log('stack built')
    wq = wl.get("weights", "f16")
    row_bytes = {"f16": 2 * E, "q4_0": E // 32 * 18, "q8_0": E // 32 * 34}[wq]

    # x pool: identical on all ranks (x is replicated), pinned on the host for the e2e leg
    rng = np.random.default_rng(a.seed + 1)
    x_host = torch.from_numpy(rng.normal(0.0, 1.0, (a.pool, L, 1, E)).astype(np.float32)).pin_memory()
    y_host = torch.empty((L, 1, E), dtype=torch.float32).pin_memory()
    x_pool = x_host.to(device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    fused = a.schedule == "fused" and st.fused_supported()
    schedule = a.schedule if (fused or a.schedule != "fused") else "two-streams"
    step_stats = st.step_fused if fused else st.step_eager
    # eager warm-up (sets kernel attributes) + per-pool-entry activity statistics for the byte accounting
    A_pool, Z_pool = [], []
    for i in range(a.pool):
        st.x.copy_(x_pool[i])
        step_stats(want_stats=True)
        torch.cuda.synchronize()
        s = st.stats.cpu().numpy()[:, 0, :]
        A_pool.append(s[:, 0].astype(np.int64))
        Z_pool.append(s[:, 1].astype(np.int64))
    launches_per_step = st.launches_per_step
    log('eager warm-up done')
    st.early_mask = not a.no_early_mask
    step_fn = st.step_fused if fused else (st.step_two_streams if schedule == "two-streams" else st.step_eager)
    g_step = st.capture(step_fn)
    launches_per_step = st.launches_per_step
    log('step graph captured')
    g_ffn = st.capture(st.ffn_only_fused if fused else st.ffn_only_eager)
    g_pred = st.capture(st.predict_only_fused if fused else st.predict_only_eager)

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def step_resident(i):
        st.x.copy_(x_pool[i % a.pool])     # D2D, 4*L*E bytes: a new token's activations -> new masks every step
        g_step.replay()

    def step_e2e(i):
        st.x.copy_(x_host[i % a.pool], non_blocking=True)
        g_step.replay()
        y_host.copy_(st.y, non_blocking=True)
        torch.cuda.current_stream().synchronize()   # a decode step's result is needed on the host before the next

    log('all graphs captured')

    def shard_activity():
        """active rows per layer of every shard, averaged over the x pool: [world, L]; imbalance = max_s / mean_s"""
        mine = torch.tensor(np.mean(np.stack(A_pool), axis=0), dtype=torch.float64, device=device)
        if world == 1:
            return mine.cpu().numpy()[None]
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        return torch.stack(parts).cpu().numpy()

    def measure_activity():
        A_pool.clear(), Z_pool.clear()
        for i in range(a.pool):
            st.x.copy_(x_pool[i])
            step_stats(want_stats=True)
            torch.cuda.synchronize()
            s = st.stats.cpu().numpy()[:, 0, :]
            A_pool.append(s[:, 0].astype(np.int64))
            Z_pool.append(s[:, 1].astype(np.int64))

    act = shard_activity()
    balance = {"placement": a.placement if world > 1 else "single shard",
               "active_rows_imbalance": float((act.max(axis=0) / np.maximum(act.mean(axis=0), 1e-9)).mean()),
               "active_rows_per_shard": [float(v) for v in act.mean(axis=1)]}
    if world > 1 and a.balance_rounds > 0:
        # the online balancer: decayed firing rates from the masks of real steps, greedy swap plan, the
        # rows of the swapped groups exchanged between the two ranks.  Weights move in place, the graphs stay valid.
        from sparkinfer_b200.balancer import ShardBalancer
        bal = ShardBalancer(st.layers, st.shard_maps, rank, world, group=dist.group.WORLD, g=GROUP, lam=0.9)
        n_before = max(50, a.steps // 10)
        ms_before = timed(step_resident, n_before, 5) / n_before
        log('balancer: before-leg timed')
        rounds, t_obs, t_move, tok, ms_during = [], 0.0, 0.0, 0, []
        for rd in range(a.balance_rounds):
            t0 = time.perf_counter()
            for _ in range(a.balance_tokens):
                step_resident(tok)
                bal.observe(st.p)
                tok += 1
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            log(f'balancer round {rd}: observed')
            if a.balance_async:
                # plan -> one window of the exchange after every decode step -> anchor
                bal.rebalance_begin(max_swaps=a.balance_swaps)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                n_during, left = 0, 1
                while left > 0:
                    step_resident(tok + n_during)
                    left = bal.rebalance_step()
                    n_during += 1
                e1.record()
                e1.synchronize()
                ms_during.append(e0.elapsed_time(e1) / n_during)
                r = bal.rebalance_finish()
            else:
                r = bal.rebalance(max_swaps=a.balance_swaps)
            log(f'balancer round {rd}: moved {r["swaps"]} swaps')
            barrier()
            t2 = time.perf_counter()
            t_obs, t_move = t_obs + (t1 - t0), t_move + (t2 - t1)
            rounds.append({"swaps": r["swaps"], "lambda": r["lambda"], "dfr_imbalance_before": r["imbalance_before"],
                           "dfr_imbalance_planned": r["imbalance_after"],
                           "tok_s_during_exchange": (1e3 / ms_during[-1]) if (a.balance_async and ms_during) else None,
                           "steps_during_exchange": n_during if a.balance_async else None})
        measure_activity()
        act2 = shard_activity()
        balance.update({"active_rows_imbalance_before": balance["active_rows_imbalance"],
                        "active_rows_imbalance": float((act2.max(axis=0) / np.maximum(act2.mean(axis=0), 1e-9)).mean()),
                        "active_rows_per_shard_before": balance["active_rows_per_shard"],
                        "active_rows_per_shard": [float(v) for v in act2.mean(axis=1)],
                        "tok_s_before": 1e3 / ms_before, "rounds": rounds, "tokens_observed": tok,
                        "swaps_total": bal.swaps_done, "bytes_sent_rank0": bal.bytes_moved,
                        "observe_s": t_obs, "plan_and_move_s": t_move, "dfr_launches_per_token": L})
        log('balancer done')
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    ms_value = timed(step_resident, a.steps, max(a.warmup, 3))
    clocks = sampler.stop()
    log('value leg done')
    ms_e2e = timed(step_e2e, a.steps, max(a.warmup, 3))
    log('e2e leg done')

    # ---- parity leg: layers 0..k of the stack against the CPU oracle on the host (unsharded problem), every N
    parity = None
    if not a.no_parity and a.parity_layers > 0:
        parity = parity_check(a, st, wl, x_pool[0], g_step, step_stats, rank, world, device)
    log('parity leg done')

    # ---- roofline legs: (i) the dominant kernel of the step, (ii) the sparse FFNs alone, (iii) the predictors alone
    st.x.copy_(x_pool[0])
    g_step.replay()
    torch.cuda.synchronize()
    reps = max(10, min(200, a.steps // 10))
    ms_ffn = timed(lambda i: g_ffn.replay(), reps, 3) / reps
    ms_pred = timed(lambda i: g_pred.replay(), reps, 3) / reps
    ms_step0 = timed(lambda i: g_step.replay(), reps, 3) / reps      # the step on the masks of pool[0] (no x refresh)
    A0, Z0 = A_pool[0], Z_pool[0]
    ffn_bytes = float(((2 * A0 + Z0) * row_bytes).sum())          # this rank's shard, all layers of one step
    pred_bytes = float(L * (E * R // (world if fused else 1) + R * st.m) * 2)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback 6650"
    traffic = None
    if world == 1 and a.workload == "prosparse-13b" and a.rho == 0.10 and a.rank == 1024:  # the ncu capture is of exactly this launch
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_top_kernel.json")))
            if (tj.get("kernel", "").startswith("spif_layer")) == fused:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass

    def leave():
        # NCCL communicators that were captured into CUDA graphs can hang in destroy_process_group / interpreter
        # teardown: synchronise, flush, and leave without running destructors
        if world > 1:
            del_graphs = (g_step, g_ffn, g_pred)  # noqa: F841 (keep alive until the barrier has passed)
            barrier()
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)

    # N = 1: the other single-GPU configurations of BASELINE.json (configs[0] architecture and configs[3]) as quick legs, so
    # that the driver's default run records them too (same schedule, same timing rules, parity on 2 layers each)
    other_configs = []
    if world == 1 and not a.no_other_configs and a.workload == "prosparse-13b":
        for name in ("prosparse-7b", "llama3-8b-topk", "prosparse-13b-q4_0", "prosparse-13b-q8_0"):
            torch.cuda.empty_cache()
            other_configs.append(quick_leg(a, name, rank, world, device, barrier))
    scale_70b = None
    if world > 1 and not a.no_scale_70b and a.workload == "prosparse-13b":
        # the north-star's scaling target (BASELINE.json configs[4]): ReluLLaMA-70B neuron-sharded over the same ranks
        g_step = g_ffn = g_pred = None   # the 13B graphs and weights make room for the 70B shard
        st.layers = None
        torch.cuda.empty_cache()
        scale_70b = quick_leg(a, "relullama-70b", rank, world, device, barrier)
        log('70b leg done')
        if rank != 0:
            barrier()
            sys.stdout.flush()
            os._exit(0)
    elif rank != 0:
        leave()
        return
    ms_step = ms_value / a.steps
    A_mean = float(np.mean([x.mean() for x in A_pool]))
    Z_mean = float(np.mean([x.mean() for x in Z_pool]))
    if fused:
        # ONE launch per layer does FFN(l) + predictor(l+1): algorithmic bytes of a launch = (2A+Z)*row_bytes + the predictor's
        # two matrices (this shard's part); duration = the captured step / launches (L + 1: the extra one is predictor(0))
        n_launch = L + 1
        bytes_launch = (ffn_bytes + pred_bytes) / n_launch
        us_launch = ms_step0 * 1e3 / n_launch
        kname = "spif_layer4_kernel<f16> (sparse FFN of layer l + predictor of layer l+1 + list + sum over CTAs%s): one spif_sparse_ffn_lookahead call" % (
            " + shard exchange" if world > 1 else "")
    else:
        bytes_launch, us_launch = ffn_bytes / L, ms_ffn * 1e3 / L
        kname = ("spif_ffn_direct_kernel<f16> (+ spif_reduce_partials_kernel): one spif_sparse_ffn call" if wq == "f16" else
                 f"spif_ffn_qwarp_kernel<{wq}> (+ spif_reduce_partials_kernel): one spif_sparse_ffn call")
    achieved = bytes_launch / (us_launch * 1e-6) / 1e9
    out = {
        "metric": "decode tok/s ProSparse-13B bs1; sparse-FFN effective HBM GB/s vs B200 peak",
        "value": 1e3 / ms_step, "unit": "tok/s (sparse-FFN path of all layers: predictor + mask + gate/up + down)",
        "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f16 weights, f32 accumulate" if wq == "f16" else f"{wq} weights x 8-bit activations (dp4a), f32 accumulate; f16 predictor",
        "data": "synthetic: random-init weights, synthetic predictor (Zipf hotness over 16-neuron groups), N(0,1) x",
        "config": {"workload": wl["label"], "n_embd": E, "n_ff": F, "n_layer": L, "pred_rank": R, "rho": a.rho,
                   "active_rows_per_layer": A_mean * world if world > 1 else A_mean,
                   "nonzero_hidden_per_layer": Z_mean * world if world > 1 else Z_mean,
                   "sharding": f"neuron groups of {GROUP} over {world} GPU(s), one sum of y over shards per layer ({st.exchange})"
                   if world > 1 else "all neurons resident on 1 GPU",
                   "schedule": schedule + (": one launch per layer, FFN(l) + predictor(l+1); predictor(l+1) reads x_l, launch l+1 "
                                           "follows launch l on one stream (the reference graph's data edges)" if fused else ""),
                   "l2": "inputs larger than L2: distinct weights per layer, %.1f GB touched per step" %
                         ((ffn_bytes + pred_bytes) / 1e9), "x_pool": a.pool,
                   "activations": "rounded to the weight type like the reference CPU ops (ggml-cpu.c:1832-1856), exact products, "
                                  "F32 accumulation" if a.activations == "weight" else "F32 (weights converted to F32 per element)"},
        "e2e": {"value": 1e3 / (ms_e2e / a.steps), "unit": "tok/s", "h2d_bytes_per_step": 4 * L * E,
                "d2h_bytes_per_step": 4 * L * E},
        "gpu_launches": launches_per_step * a.steps,
        "clocks": clocks,
        "parity": parity,
        "shard_balance": balance,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "kernel": kname,
                     "bytes_per_launch": bytes_launch, "us_per_launch": us_launch,
                     "ffn": {"achieved": ffn_bytes / (ms_ffn * 1e-3) / 1e9, "frac": ffn_bytes / (ms_ffn * 1e-3) / 1e9 / peak,
                             "us_per_layer": ms_ffn * 1e3 / L, "bytes_per_layer": ffn_bytes / L,
                             "what": ("the layer kernel with pred == NULL + spif_mask_to_list (2 launches)" if fused else
                                      "spif_ffn_direct_kernel + spif_reduce_partials_kernel") +
                                     (", shard exchange included" if (fused and world > 1) else
                                      (", shard exchange NOT included" if world > 1 else ""))},
                     "predictor": {"achieved": pred_bytes / (ms_pred * 1e-3) / 1e9, "us_per_layer": ms_pred * 1e3 / L,
                                   "bytes_per_layer": pred_bytes / L},
                     "layer": {"achieved": (ffn_bytes + pred_bytes + 4 * L * (2 * E + st.m)) / (ms_step * 1e-3) / 1e9,
                               "us_per_layer": ms_step * 1e3 / L}},
    }
    if scale_70b is not None:
        out["scale_70b"] = scale_70b
    if other_configs:
        out["other_configs"] = other_configs
    if world == 1 and not a.no_gpu_baseline and wq == "f16":
        out["gpu_baseline"] = gpu_baseline(a, st, wl, x_pool[0], device)
    if world == 1 and not a.no_batched and wq == "f16":
        out["batched"] = batched_legs(a, st, wl, device)
    if world == 1 and not a.no_cpu_baseline and wq == "f16":  # (the CPU leg times F16 host weights)
        out["cpu_baseline"] = cpu_baseline(a, wl, iters=8, warmup=2)
    print(json.dumps(out), flush=True)
    leave()


def gpu_baseline(a, st, wl, x0, device):
    """the reference's OWN CUDA kernels (mm-sparse.cu, axpy-sparse.cu, mmvf.cu, unary.cu, binbcast.cu: compiled unmodified for
    sm_100 into oracle/_ref/libspif_refcuda.so) on the SAME device-resident weights and inputs, in the reference's node
    order on one stream, timed with CUDA events -- once through a CUDA graph (how a GGML_CUDA_GRAPHS build decodes) and once
    as eager launches.  A reported baseline: the only place outside tests/ that loads anything from oracle/."""
    import oracle_lib
    import torch
    if wl.get("topk") or not oracle_lib.RefCuda.available():
        return {"unavailable": "no reference code path for this config" if wl.get("topk") else
                "oracle/_ref/libspif_refcuda.so not built (needs /root/reference at build time)"}
    rc = oracle_lib.RefCuda()
    E, F, L, R = wl["n_embd"], wl["n_ff"], wl["n_layer"], a.rank
    x = x0[:, 0].contiguous()
    y = torch.zeros((L, E), dtype=torch.float32, device=device)
    p_io = torch.zeros((L, F), dtype=torch.float32, device=device)
    ly = st.layers
    args = (wl["arch"], F16, F16, [w.up for w in ly], [w.gate for w in ly], [w.down_t for w in ly], [w.pred_up for w in ly],
            [w.pred_down for w in ly], [w.pred_down_b for w in ly], x, y, p_io, E, F, R)
    us_graph, us_eager = rc.stack(*args, with_pred=True, warmup=3, iters=20)
    # the same kernels on OUR masks (no predictor): its y against ours, a reported third comparison
    st.x.copy_(x0)
    st.step_fused() if st.fused_supported() else st.step_eager()
    torch.cuda.synchronize()
    p_io.copy_(st.p[:, 0])
    us_ffn_graph, _ = rc.stack(*args, with_pred=False, warmup=1, iters=10)
    torch.cuda.synchronize()
    ours = st.y[:, 0]
    err = float(((y - ours).abs().amax(dim=1) / ours.abs().amax(dim=1).clamp_min(1e-30)).max().item())
    return {"value": 1e6 / us_graph, "unit": "tok/s", "us_per_layer": us_graph / L, "eager_tok_s": 1e6 / us_eager,
            "ffn_only_us_per_layer": us_ffn_graph / L, "kind": "reference CUDA kernels, sm_100 (unmodified sources), CUDA graph",
            "what": rc.describe(), "max_rel_err_vs_ours_same_masks": err,
            "note": "half2 accumulation (GGML_PREC_DEFAULT, mm-sparse.cu:355-361) and float atomics in the axpy (axpy-sparse.cu:84)"}


def quick_leg(a, workload, rank, world, device, barrier):
    """a second workload in the same process (N > 1: the 70B scaling leg): tok/s with x resident, same timing rules"""
    import copy

    import torch
    import torch.distributed as dist
    a2 = copy.copy(a)
    a2.workload = workload
    wl = WORKLOADS[workload]
    E, L = wl["n_embd"], wl["n_layer"]
    st, _ = build_stack(a2, wl, rank, world, device)
    rng = np.random.default_rng(a.seed + 1)
    pool = min(a.pool, 4)
    x_pool = torch.from_numpy(rng.normal(0.0, 1.0, (pool, L, 1, E)).astype(np.float32)).to(device)
    fused = a.schedule == "fused" and st.fused_supported()
    step_stats = st.step_fused if fused else st.step_eager
    A_all, Z_all = [], []
    for i in range(pool):
        st.x.copy_(x_pool[i])
        step_stats(want_stats=True)
        torch.cuda.synchronize()
        sn = st.stats.cpu().numpy()[:, 0, :]
        A_all.append(sn[:, 0].astype(np.int64)), Z_all.append(sn[:, 1].astype(np.int64))
    # not fusable (quantised FFN weights): the reference's data edges with the overlap they allow, on two streams --
    # predictor(l+1) beside FFN(l), never ahead of FFN(l-1) (FfnStack.step_overlap)
    g = st.capture(st.step_fused if fused else st.step_overlap)
    steps = max(20, min(a.steps, 200))

    def run(i):
        st.x.copy_(x_pool[i % pool])
        g.replay()
    for i in range(5):
        run(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        run(i)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    parity = parity_check(a2, st, wl, x_pool[0], g, step_stats, rank, world, device, n_layers=2) if not a.no_parity else None
    row_b = {"f16": 2 * E, "q4_0": E // 32 * 18, "q8_0": E // 32 * 34}[wl.get("weights", "f16")]
    ffn_bytes = float(((2 * A_all[0] + Z_all[0]) * row_b).sum())
    pred_bytes = float(L * (E * a.rank // (world if fused else 1) + a.rank * st.m) * 2)
    res = {"workload": wl["label"], "n_gpus": world, "value": 1e3 / ms, "unit": "tok/s", "ms_per_step": ms, "steps": steps,
           "us_per_layer": ms * 1e3 / L, "schedule": "fused" if fused else "overlap: predictor(l+1) beside FFN(l), after FFN(l-1); 4 launches per layer",
           "exchange": st.exchange,
           "active_rows_per_layer": float(np.mean([x.mean() for x in A_all])) * world,
           "per_gpu_algorithmic_GBps": (ffn_bytes + pred_bytes) / (ms * 1e-3) / 1e9, "parity": parity}
    del g
    return res


def parity_check(a, st, wl, x0, g_step, step_stats, rank, world, device, n_layers=None):
    """Runs one step on x0 and checks layers 0..k-1 against the CPU oracle on the host, on the UNSHARDED problem (the rows
    of all shards concatenated in rank order; a sum over neurons does not care about their order):
      index_flips   predictor: neurons whose active / inactive decision differs from the oracle's, with the CPU-rounding flag
                    (x and h1 rounded to F16 like the reference CPU backend); flips_outside_band counts those with
                    |oracle logit| >= 1e-4 -- must be 0 (BASELINE.json north_star)
      max_rel_err   sparse FFN: max |y - y_oracle| / max |y_oracle| over the checked layers, the oracle run on the GPU's own
                    masks; the captured graph's default mode (F32 activations, tolerance 1e-3)
      ranks_bit_identical   N > 1: y of every layer is bit-identical on all ranks"""
    import oracle_lib  # test infrastructure: the checker, never the product path
    import torch
    import torch.distributed as dist
    E, F, L, R = wl["n_embd"], wl["n_ff"], wl["n_layer"], a.rank
    k = min(n_layers if n_layers is not None else a.parity_layers, L)
    st.x.copy_(x0)
    g_step.replay()
    torch.cuda.synchronize()
    y_gpu = st.y[:k, 0].clone()
    p_gpu = st.p[:k, 0].clone()
    # CPU-rounding pass (eager): predictor flips within the 1e-4 band
    old_flags = st.flags
    st.flags = old_flags | 1
    step_stats(want_stats=False)
    torch.cuda.synchronize()
    p_cr = st.p[:k, 0].clone()
    st.flags = old_flags
    g_step.replay()          # leave the default-mode masks / lists in place
    torch.cuda.synchronize()

    def gather(t):
        """concatenate a per-shard tensor over ranks (rank order) -> host numpy on rank 0"""
        if world == 1:
            return t.cpu().numpy()
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous())
        return torch.cat(parts, dim=0).cpu().numpy() if rank == 0 else None

    same = None
    if world > 1:
        ys = [torch.empty_like(st.y) for _ in range(world)]
        dist.all_gather(ys, st.y)
        same = all(torch.equal(ys[0], v) for v in ys[1:])
    # F16: 1e-3 (north_star).  Quantised: the CUDA path quantises x like the reference CUDA kernels (roundf(x / d), quantize.cu),
    # the oracle like the reference CPU ops (round-half-even of x * id): the quant format's own bound, 2e-2 as in tests/
    tol = 1e-3 if st.w_type in (F16, BF16) else 2e-2
    res = {"layers_checked": k, "max_rel_err": 0.0, "index_flips": 0, "flips_outside_band": 0, "tolerance": tol,
           "oracle": None, "ranks_bit_identical": same}
    o = oracle_lib.Oracle() if rank == 0 else None
    if rank == 0:
        res["oracle"] = "oracle/libspif_oracle.so (C restatement pinned bit-exact to the reference CPU ops)"
    for l in range(k):
        w = st.layers[l]
        rows = lambda t, n: gather(t.view(st.m, -1))
        up, gate, down = rows(w.up, E), rows(w.gate, E), rows(w.down_t, E)
        wpd, bpd = rows(w.pred_down, R), gather(w.pred_down_b)
        pl, pcr = gather(p_gpu[l]), gather(p_cr[l])
        if rank != 0:
            continue
        x_l = x0[l].cpu().numpy()
        y_ref = o.sparse_ffn(wl["arch"], st.w_type, up.reshape(-1), gate.reshape(-1), down.reshape(-1), x_l, pl[None], E, F)[0]
        yg = y_gpu[l].cpu().numpy()
        res["max_rel_err"] = max(res["max_rel_err"], float(np.abs(yg - y_ref[0]).max() / max(np.abs(y_ref).max(), 1e-30)))
        if not wl.get("topk"):
            x_p = x0[max(l - 1, 0)].cpu().numpy()   # look-ahead: the mask of layer l comes from x_{l-1} (layer 0: its own x)
            p_ref, z_ref = o.predictor(F16, w.pred_up.cpu().numpy(), None, wpd.reshape(-1), bpd, x_p, E, R, F)
            flips = (pcr < 0.5) != (p_ref[0] < 0.5)
            res["index_flips"] += int(flips.sum())
            res["flips_outside_band"] += int((np.abs(z_ref[0][flips]) >= 1e-4).sum())
    if world > 1:
        t = torch.tensor([res["max_rel_err"], res["index_flips"], res["flips_outside_band"]], dtype=torch.float64, device=device)
        dist.broadcast(t, src=0)
        res["max_rel_err"], res["index_flips"], res["flips_outside_band"] = float(t[0]), int(t[1]), int(t[2])
    res["ok"] = bool(res["max_rel_err"] < tol and res["flips_outside_band"] == 0 and (same is None or same))
    return res


def batched_legs(a, st, wl, device, batches=(8, 32), n_layers=8, reps=20):
    """decode at batch T > 1 (BASELINE.json configs[2]: Q4_0 at batch 1/8/32; the F16 weights of this run as well):
    per-token masks from the synthetic predictor, then one spif_sparse_ffn call per layer -- the union-of-masks gathered GEMM
    on tcgen05 tensor cores (every union row read once) next to the reference's per-(row, token) schedule
    (mm-sparse.cu:107-210, T x the row traffic).  Bytes = (2 A_u + Z_u) * row_bytes with A_u / Z_u the union rows / union
    rows with a non-zero hidden value.  Q4_0 weights are random valid block_q4_0 bytes (random-init, like the F16 ones)."""
    import torch
    import sparkinfer_b200 as sp
    E, F = wl["n_embd"], st.m
    nl = min(n_layers, st.L)
    res = []
    gen = torch.Generator(device=device).manual_seed(a.seed + 77)

    def q4_rows():
        nb = E // 32
        blk = torch.randint(0, 256, (F, nb, 18), dtype=torch.uint8, device=device, generator=gen)
        d = torch.full((F, nb, 1), 0.02 / 6, dtype=torch.float16, device=device)
        blk[:, :, :2] = d.view(torch.uint8).reshape(F, nb, 2)
        return blk.reshape(F, nb * 18).contiguous()

    def q8_rows():  # block_q8_0 = {f16 d, int8 qs[32]} (the quant format the reference's sparse kernels ship: mmq-sparse.cu)
        nb = E // 32
        blk = torch.randint(0, 256, (F, nb, 34), dtype=torch.uint8, device=device, generator=gen)
        d = torch.full((F, nb, 1), 0.02 / 64, dtype=torch.float16, device=device)
        blk[:, :, :2] = d.view(torch.uint8).reshape(F, nb, 2)
        return blk.reshape(F, nb * 34).contiguous()

    f16 = [(w.up, w.gate, w.down_t) for w in st.layers[:nl]]
    q4 = [(q4_rows(), q4_rows(), q4_rows()) for _ in range(nl)] if E % 512 == 0 else None
    q8 = [(q8_rows(), q8_rows(), q8_rows()) for _ in range(nl)] if E % 512 == 0 else None
    for wname, wtype, rb, weights, Ts in (("f16", st.w_type, 2 * E, f16, batches),
                                          ("q4_0", sp.GGML_TYPE_Q4_0, E // 32 * 18, q4, (1,) + tuple(batches)),
                                          ("q8_0", sp.GGML_TYPE_Q8_0, E // 32 * 34, q8, (1,) + tuple(batches))):
        if weights is None:
            continue
        for T in Ts:
            x = torch.randn((T, E), device=device, generator=gen)
            ws = sp.Workspace(E, F, T)
            y = torch.empty((T, E), device=device)
            masks = [sp.predictor(w.pred_up, None, w.pred_down, w.pred_down_b, x, st.p_type, st.r, F, ws=ws).clone()
                     for w in st.layers[:nl]]
            stats = [torch.zeros((T, 4), dtype=torch.int32, device=device) for _ in range(nl)]
            leg = {"weights": wname, "n_tok": T, "layers_timed": nl}
            # T == 1: "per_token" = every launch waits for its predecessor before it looks at its mask; "mask_stable" = the mask
            # of layer l exists a whole layer early, as in the reference's look-ahead order (SPIF_FLAG_MASK_STABLE): the kernel
            # scans it and requests its first rows under the previous launch's tail
            # (quantised weights at <= 12 tokens: spif_sparse_ffn itself picks the per-token kernel, there is no union leg)
            for name, flags in ((("per_token", sp.FLAG_PER_TOKEN), ("mask_stable", sp.FLAG_PER_TOKEN | sp.FLAG_MASK_STABLE))
                                if T == 1 else (("per_token", sp.FLAG_PER_TOKEN),) if (wname != "f16" and T <= 12) else
                                (("union_gemm", 0), ("per_token", sp.FLAG_PER_TOKEN))):
                def run():
                    for l in range(nl):
                        u, g_, d_ = weights[l]
                        sp.sparse_ffn(st.arch, wtype, u, g_, d_, x, masks[l], F, flags=flags, ws=ws, out=y, stats=stats[l],
                                      want_stats=True)
                run()
                torch.cuda.synchronize()
                sn = torch.stack(stats).cpu().numpy()
                g, side = torch.cuda.CUDAGraph(), torch.cuda.Stream(device=device)
                with torch.cuda.stream(side):
                    run()
                    side.synchronize()
                    with torch.cuda.graph(g, stream=side):
                        run()
                torch.cuda.synchronize()
                for _ in range(3):
                    g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    g.replay()
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / reps / nl
                if name == "union_gemm":
                    rows = float((2 * sn[:, 0, 2] + sn[:, 0, 3]).mean())
                    leg["union_rows"] = float(sn[:, 0, 2].mean())
                else:
                    rows = float((2 * sn[:, :, 0] + sn[:, :, 1]).sum(axis=1).mean())
                leg[name] = {"us_per_layer": us, "ffn_tok_s": T / (us * st.L) * 1e6, "algorithmic_GBps": rows * rb / us / 1e3}
            if wname == "q8_0" and T == 1 and not a.no_gpu_baseline:
                # the quant format the reference's sparse CUDA kernels ship (mmq-sparse.cu + quantize.cu, axpyq-sparse.cu, compiled
                # unmodified for sm_100): the same Q8_0 rows, masks and x through them, CUDA graph, no predictor
                import oracle_lib
                if oracle_lib.RefCuda.available():
                    rc = oracle_lib.RefCuda()
                    ly = st.layers[:nl]
                    xr = x.expand(nl, E).contiguous()
                    yr = torch.zeros((nl, E), dtype=torch.float32, device=device)
                    pr = torch.stack([mk[0] for mk in masks]).contiguous()
                    us_g, _ = rc.stack(st.arch, sp.GGML_TYPE_Q8_0, F16, [w[0] for w in weights], [w[1] for w in weights],
                                       [w[2] for w in weights], [w.pred_up for w in ly], [w.pred_down for w in ly],
                                       [w.pred_down_b for w in ly], xr, yr, pr, E, F, st.r, with_pred=False, warmup=2, iters=10)
                    torch.cuda.synchronize()
                    run()
                    torch.cuda.synchronize()
                    err = float(((yr[nl - 1] - y[0]).abs().max() / y[0].abs().max().clamp_min(1e-30)).item())
                    leg["reference_cuda_kernels"] = {"us_per_layer": us_g / nl, "max_rel_err_vs_ours_last_layer": err,
                                                     "what": "mmq-sparse.cu / axpyq-sparse.cu / quantize.cu, unmodified, sm_100"}
            res.append(leg)
    return res


# ---------------------------------------------------------------------------------------------------------------------
def host_layers(a, wl, n_layers):
    """random-init host copies of the first n_layers layers (CPU legs only; values differ from the GPU's, shapes and
    activity statistics are the same)"""
    import torch
    E, F, R = wl["n_embd"], wl["n_ff"], a.rank
    g = torch.Generator().manual_seed(a.seed)

    def half(rows, cols, std):
        t = torch.empty((rows, cols), dtype=torch.float32).normal_(0.0, std, generator=g).to(torch.float16)
        return t.numpy().view(np.uint8).reshape(-1)

    up, gate, down, wpu, wpd, bpd = [], [], [], [], [], []
    for l in range(n_layers):
        up.append(half(F, E, 0.02)), gate.append(half(F, E, 0.02)), down.append(half(F, E, 0.02))
        wpu.append(half(R, E, E ** -0.5)), wpd.append(half(F, R, R ** -0.5))
        bpd.append(bias_for(np.repeat(synth_hotness(F, a.rho, a.seed + l), GROUP)))
    x = np.random.default_rng(a.seed + 1).normal(0.0, 1.0, (n_layers, 1, E)).astype(np.float32)
    return up, gate, down, wpu, wpd, bpd, x


def cpu_stack_tok_s(a, wl, iters, warmup):
    """time the reference CPU sparse-FFN path (all host threads) on a bounded sample: the first --cpu-layers layers of
    the workload, scaled to the full layer count.  -> (tok/s, kind, cores, sample description)"""
    import oracle_lib  # test infrastructure: the checker / CPU baseline, never the product path
    wl_L = wl["n_layer"]
    nl = min(a.cpu_layers, wl_L)
    cores = os.cpu_count() or 1
    if oracle_lib.Ref.available() and wl["arch"] != 3:
        lib, kind = oracle_lib.Ref(), "reference"
    else:  # (the reference has no SiLU / top-k path: the port times threshold masks with the SiLU recipe there)
        lib, kind = oracle_lib.Oracle(), "port"
    up, gate, down, wpu, wpd, bpd, x = host_layers(a, wl, nl)
    us, _ = oracle_lib.stack_time(lib, wl["arch"], F16, F16, up, gate, down, wpu, wpd, bpd, x, wl["n_embd"], wl["n_ff"],
                                  a.rank, 1, cores, warmup, iters)
    if us <= 0:
        raise RuntimeError("CPU stack timing failed")
    tok_s = 1e6 / (us * wl_L / nl)
    sample = (f"{nl} of {wl_L} layers (predictor + sparse FFN, look-ahead order), median of {iters} steps after {warmup} "
              f"warm-up, {cores} threads, scaled x{wl_L / nl:g} to the full stack"
              + ("; reference built -march=x86-64-v3 (AVX2, no AVX-512) so that the binary travels" if kind == "reference" else ""))
    return tok_s, kind, cores, sample


def cpu_baseline(a, wl, iters, warmup):
    tok_s, kind, cores, sample = cpu_stack_tok_s(a, wl, iters, warmup)
    return {"value": tok_s, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample}


def run_reference(a):
    """the reference's own CPU implementation of the path (oracle/_ref = its unmodified ggml CPU ops) on the host cores"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    wl = WORKLOADS[a.workload]
    iters = max(3, min(a.steps, 200))       # bounded: each step is ~10 ms of CPU work per sampled layer
    warmup = max(3, min(a.warmup, 10))
    t0 = time.time()
    tok_s, kind, cores, sample = cpu_stack_tok_s(a, wl, iters, warmup)
    E, F, L = wl["n_embd"], wl["n_ff"], wl["n_layer"]
    out = {
        "impl": "reference",
        "metric": "decode tok/s ProSparse-13B bs1; sparse-FFN effective HBM GB/s vs B200 peak",
        "value": tok_s, "unit": "tok/s (sparse-FFN path of all layers: predictor + mask + gate/up + down)",
        "n_gpus": a.gpus, "steps": iters, "warmup": warmup, "ms_per_step": 1e3 / tok_s, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f16 weights, f32 accumulate", "data": "synthetic",
        "config": {"workload": wl["label"], "n_embd": E, "n_ff": F, "n_layer": L, "pred_rank": a.rank, "rho": a.rho,
                   "sharding": "host cores only (reference CPU path, all neurons CPU-side)"},
        "cpu_baseline": {"value": tok_s, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": tok_s, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.time() - t0, 1),
    }
    print(json.dumps(out), flush=True)


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
