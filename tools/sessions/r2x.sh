mkdir -p gpurun_out
for impl in 3 0; do
SPIF_LAYER_IMPL=$impl timeout 300 python tools/dbg_layer4.py 1376 2048 1 1 0 2>&1 | tail -30
done
SPIF_LAYER_IMPL=3 timeout 300 python tools/dbg_layer4.py 1376 2048 0 1 0 2>&1 | tail -8
timeout 1500 python -m pytest tests/test_gpu_layer.py -q > gpurun_out/r2x_layer.log 2>&1; echo "layer rc=$?"; tail -15 gpurun_out/r2x_layer.log
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -q -k "quant or neuron_idx" > gpurun_out/r2x_quant.log 2>&1; echo "quant rc=$?"; tail -15 gpurun_out/r2x_quant.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2x_bench.json'))
for b in d['batched']: print(json.dumps(b))
PY
