mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity2.py -x -q -k "quant or neuron_idx" 2>&1 | tail -3
timeout 900 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-other-configs > gpurun_out/r2ll_bench.json 2> gpurun_out/r2ll_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2ll_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2ll_bench.json'))
for b in d['batched']:
    if b['weights']!='f16' and b['n_tok']<=8: print(json.dumps(b)[:520])
PY
for args in "--wtype q4_0" "--wtype q8_0"; do
echo "== trace $args"; timeout 300 python tools/trace_ffn.py $args 2>&1 | tail -2
done
