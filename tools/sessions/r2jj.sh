N=${1:-2}
mkdir -p gpurun_out
for mode in "" "--balance-async"; do
tag=$([ -z "$mode" ] && echo sync || echo async)
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 200 --warmup 10 --placement contiguous --balance-rounds 3 --balance-swaps 32 $mode --no-scale-70b > gpurun_out/r2jj_balance_n${N}_$tag.json 2> gpurun_out/r2jj_balance_n${N}_$tag.err; echo "balance $tag rc=$?"; tail -2 gpurun_out/r2jj_balance_n${N}_$tag.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2jj_balance_n${N}_$tag.json').read().strip().splitlines()[-1])
b=d['shard_balance']
print('$tag value', round(d['value'],1), 'before', round(b.get('tok_s_before',0),1), 'imb', round(b['active_rows_imbalance_before'],3), '->', round(b['active_rows_imbalance'],3), 'move_s', round(b['plan_and_move_s'],2), 'parity', d['parity']['ok'])
for r in b['rounds']: print('   ', json.dumps(r))
PY
done
