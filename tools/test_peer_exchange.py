"""Multi-GPU check + timing of the per-layer shard exchange (run under torchrun on a box with >= 2 GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/test_peer_exchange.py

Checks: spif_shard_allreduce == ncclAllReduce (to summation-order rounding), bit-identical on all ranks; the fused layer
(spif_sparse_ffn_sharded on neuron shards) == the unsharded layer; then times both exchanges on a 13B-sized y (20 KB) inside
CUDA graphs.  Prints one JSON line on rank 0.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparkinfer_b200 as sp  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    out = {"world": world}

    # ---- stand-alone exchange vs NCCL
    comm = sp.ShardComm(32 * 8192, group, dev)
    for n in (5120, 8192, 32 * 5120):
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for it in range(4):
            y = torch.randn(n, device=dev, generator=g)
            ref = y.clone()
            dist.all_reduce(ref)
            got = sp.shard_allreduce(y.clone(), comm)
            err = float((got - ref).abs().max() / ref.abs().max())
            assert err < 1e-6, (n, it, err)
            gathered = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(gathered, got)
            assert all(torch.equal(gathered[0], t) for t in gathered), "exchange result differs between ranks"
    out["allreduce_parity"] = "ok"

    # ---- fused layer on neuron shards vs the unsharded layer
    E, F, T = 1024, 4096, 1
    g = torch.Generator(device=dev).manual_seed(7)  # same weights on every rank
    mk = lambda: torch.empty((F, E), dtype=torch.float16, device=dev).normal_(0, 0.02, generator=g)
    up, gate, down = mk(), mk(), mk()
    x = torch.randn((T, E), device=dev, generator=g)
    p = torch.where(torch.rand((T, F), device=dev, generator=g) < 0.15, 0.9, 0.1).float()
    ws = sp.Workspace(E, F, T)
    y_full = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, up.view(torch.uint8), gate.view(torch.uint8),
                           down.view(torch.uint8), x, p, F, ws=ws)
    rows = torch.arange(rank, F, world, device=dev, dtype=torch.int32)  # this rank's neurons
    sh = lambda w: w[rows.long()].contiguous().view(torch.uint8)
    lcomm = sp.ShardComm(T * E, group, dev)
    for it in range(3):
        y = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, sh(up), sh(gate), sh(down), x, p, F, neuron_idx=rows, ws=ws,
                          comm=lcomm)
        err = float((y - y_full).abs().max() / y_full.abs().max())
        assert err < 1e-5, err
    out["fused_layer_parity"] = "ok"

    # ---- ONE comm, the three forms interleaved, odd call counts (the sequence number is per comm and per call,
    # not per block, so launch shapes may differ from call to call)
    T8 = 8
    x8 = torch.randn((T8, E), device=dev, generator=g)
    p8 = torch.where(torch.rand((T8, F), device=dev, generator=g) < 0.15, 0.9, 0.1).float()
    ws8 = sp.Workspace(E, F, T8)
    y8_full = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, up.view(torch.uint8), gate.view(torch.uint8),
                            down.view(torch.uint8), x8, p8, F, ws=ws8)
    mcomm = sp.ShardComm(T8 * E + 256, group, dev)
    calls = 0
    for it in range(5):
        v = torch.randn(4 * 1280, device=dev, generator=torch.Generator(device=dev).manual_seed(500 + rank + it))
        ref = v.clone()
        dist.all_reduce(ref)
        got = sp.shard_allreduce(v.clone(), mcomm)                                   # stand-alone form: 5 blocks
        assert float((got - ref).abs().max() / ref.abs().max()) < 1e-6
        calls += 1
        y = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, sh(up), sh(gate), sh(down), x, p, F, neuron_idx=rows, ws=ws,
                          comm=mcomm)                                                 # fused into the partial sum: n_tok = 1
        assert float((y - y_full).abs().max() / y_full.abs().max()) < 1e-5
        calls += 1
        if it % 2 == 0:
            y8 = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, sh(up), sh(gate), sh(down), x8, p8, F, neuron_idx=rows,
                               ws=ws8, comm=mcomm)                                    # union GEMM + stand-alone form: n_tok = 8
            assert float((y8 - y8_full).abs().max() / y8_full.abs().max()) < 2e-3     # x / hidden rounded to F16 there
            calls += 1
        # look-ahead layer kernel on the same comm: FFN on this rank's rows (list in row space) + exchange
        lst = sp.mask_to_list(p[0, rows.long()].contiguous())
        yl = torch.empty((E,), device=dev)
        sp.sparse_ffn_lookahead(mcomm, ws, ffn=dict(arch=sp.ARCH_PROSPARSE, w_type=sp.GGML_TYPE_F16, up=sh(up), gate=sh(gate),
                                                    down_t=sh(down), cur=x[0], n_ff=F, neuron_idx=rows, out=yl), list_in=lst)
        assert float((yl - y_full[0]).abs().max() / y_full.abs().max()) < 1e-5
        calls += 1
    assert mcomm.status() == 0 and mcomm.epoch() == calls, (mcomm.status(), mcomm.epoch(), calls)
    out["mixed_forms_one_comm"] = {"calls": calls, "status": 0}

    # ---- sharded look-ahead chain: every rank computes r / world rows of the predictor's hidden vector and its own Wpd rows;
    # the masks of all shards together == the unsharded predictor's mask, y == the unsharded layer
    R = 256
    wpu = torch.empty((R, E), dtype=torch.float16, device=dev).normal_(0, E ** -0.5, generator=g)
    wpd = torch.empty((F, R), dtype=torch.float16, device=dev).normal_(0, R ** -0.5, generator=g)
    bpd = torch.full((F,), -0.9, device=dev)
    ccomm = sp.ShardComm(E + R, group, dev)
    one = sp.ShardComm(E + R, None, dev)
    p_full = torch.empty((F,), device=dev)
    sp.sparse_ffn_lookahead(one, ws, pred=dict(pred_up=wpu.view(torch.uint8), pred_down=wpd.view(torch.uint8), pred_down_b=bpd,
                                               cur=x[0], p_type=sp.GGML_TYPE_F16, r=R, m=F, out=p_full), list_out=sp.ActiveList(F))
    y_un = sp.sparse_ffn(sp.ARCH_PROSPARSE, sp.GGML_TYPE_F16, up.view(torch.uint8), gate.view(torch.uint8),
                         down.view(torch.uint8), x, p_full[None], F, ws=ws)
    m_loc = rows.numel()
    lists = [sp.ActiveList(m_loc, dev), sp.ActiveList(m_loc, dev)]
    p_loc = torch.empty((m_loc,), device=dev)
    yl = torch.empty((E,), device=dev)
    pred = dict(pred_up=wpu.view(torch.uint8), pred_down=sh(wpd), pred_down_b=bpd[rows.long()].contiguous(), cur=x[0],
                p_type=sp.GGML_TYPE_F16, r=R, m=m_loc, out=p_loc)
    for it in range(3):
        sp.sparse_ffn_lookahead(ccomm, ws, pred=pred, list_out=lists[0])
        sp.sparse_ffn_lookahead(ccomm, ws, ffn=dict(arch=sp.ARCH_PROSPARSE, w_type=sp.GGML_TYPE_F16, up=sh(up), gate=sh(gate),
                                                    down_t=sh(down), cur=x[0], n_ff=F, neuron_idx=rows, out=yl),
                                pred=pred, list_in=lists[0], list_out=lists[1])
        torch.cuda.synchronize()
        flips = int(((p_loc < 0.5) != (p_full[rows.long()] < 0.5)).sum())
        assert flips == 0 or float((p_loc - p_full[rows.long()]).abs().max()) < 1e-4, flips
        assert float((yl - y_un[0]).abs().max() / y_un.abs().max()) < 1e-5
        gathered = [torch.empty_like(yl) for _ in range(world)]
        dist.all_gather(gathered, yl)
        assert all(torch.equal(gathered[0], t) for t in gathered), "layer output differs between ranks"
    assert ccomm.status() == 0
    out["sharded_lookahead_chain"] = "ok"

    # ---- timing: 40 exchanges of a 13B-sized y per graph (one decode step's worth)
    n, reps = 5120, 40
    ys = [torch.randn(n, device=dev) for _ in range(reps)]
    tcomm = sp.ShardComm(n, group, dev)

    def timed(fn):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            fn()
            s.synchronize()
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph, stream=s):
                fn()
        torch.cuda.synchronize()
        dist.barrier()
        for _ in range(5):
            gph.replay()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            gph.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / 50 / reps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), gph

    us_peer, g1 = timed(lambda: [sp.shard_allreduce(y, tcomm) for y in ys])
    us_nccl, g2 = timed(lambda: [dist.all_reduce(y) for y in ys])
    out["us_per_exchange_20KB"] = {"peer": round(us_peer, 2), "nccl": round(us_nccl, 2)}
    if rank == 0:
        print(json.dumps(out), flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
