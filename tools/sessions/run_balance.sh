#!/bin/bash
# N-GPU run of the online shard balancer: contiguous (unbalanced) placement, then balancer rounds, then the timed region.
N=${1:-2}
WL=${2:-prosparse-13b}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR bench.py --gpus $N --workload $WL --steps 1000 --warmup 10 --placement contiguous \
    > gpurun_out/balance_${WL}_n${N}_off.json 2> gpurun_out/balance_${WL}_n${N}_off.err
echo "off rc=$?"
timeout 400 $TR bench.py --gpus $N --workload $WL --steps 1000 --warmup 10 --placement contiguous --balance-rounds 4 \
    > gpurun_out/balance_${WL}_n${N}_on.json 2> gpurun_out/balance_${WL}_n${N}_on.err
echo "on rc=$?"
timeout 300 $TR bench.py --gpus $N --workload $WL --steps 1000 --warmup 10 \
    > gpurun_out/balance_${WL}_n${N}_snake.json 2> gpurun_out/balance_${WL}_n${N}_snake.err
echo "snake rc=$?"
tail -3 gpurun_out/balance_${WL}_n${N}_on.err
