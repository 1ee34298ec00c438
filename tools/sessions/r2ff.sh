mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py tests/test_gpu_batched.py -x -q > gpurun_out/r2ff_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r2ff_tests.log
for fs in 1 0; do
SPIF_QW_FUSED_SUM=$fs timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-parity --no-other-configs > gpurun_out/r2ff_bench_fs$fs.json 2> gpurun_out/r2ff_bench_fs$fs.err; echo "bench fused_sum=$fs rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2ff_bench_fs$fs.json'))
for b in d['batched']:
    if b['weights']!='f16': print(json.dumps(b)[:420])
PY
done
