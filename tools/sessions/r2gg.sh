mkdir -p gpurun_out
SPIF_EXPERIMENTAL=1 SPIF_TOPK_CLUSTER=1 timeout 180 python -m pytest tests/test_gpu_topk_silu.py -x -q -k cluster > gpurun_out/r2gg_cluster.log 2>&1; echo "cluster test rc=$?"; tail -5 gpurun_out/r2gg_cluster.log
for c in 0 1; do
SPIF_TOPK_CLUSTER=$c timeout 300 python bench.py --workload llama3-8b-topk --steps 100 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2gg_8b_c$c.json 2> gpurun_out/r2gg_8b_c$c.err; echo "8b cluster=$c rc=$?"
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r2gg_8b_c$c.json')); print('cluster $c', d['value'], d['ms_per_step'], d['parity'])
except Exception as e: print('fail', e)
PY
done
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py tests/test_gpu_batched.py -x -q 2>&1 | tail -3
