# usage: bash tools/sessions/r2n2.sh N    (multi-GPU: parity test, then the bench exactly as the driver launches it)
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/r2_multi_tests_n$N.log 2>&1; echo "multi tests rc=$?"; tail -5 gpurun_out/r2_multi_tests_n$N.log
fi
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 300 --warmup 10 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench_n$N.json').read().strip().splitlines()[-1])
r=d['roofline']
print('N=$N value', d['value'], 'e2e', d['e2e']['value'], 'frac', r['frac'], 'us/launch', r.get('us_per_launch'), 'parity', d.get('parity'))
print('scale_70b', json.dumps(d.get('scale_70b')))
print('shard_balance', json.dumps(d.get('shard_balance')))
PY
