mkdir -p gpurun_out
timeout 900 python tools/stress_layer.py 150 > gpurun_out/r2y_stress.log 2>&1; echo "stress rc=$?"; tail -30 gpurun_out/r2y_stress.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -q -x -k "quant or neuron_idx" > gpurun_out/r2y_quant.log 2>&1; echo "quant rc=$?"; tail -5 gpurun_out/r2y_quant.log
for half in 1 0; do
SPIF_QW_HALF=$half timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2y_bench_h$half.json 2> gpurun_out/r2y_bench_h$half.err; echo "bench half=$half rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2y_bench_h$half.json'))
for b in d['batched']:
    if b['n_tok']==1 or b['weights']!='f16': print(json.dumps(b))
PY
done
