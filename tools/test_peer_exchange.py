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
