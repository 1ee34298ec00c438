mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -q -x > gpurun_out/r2r_t1.log 2>&1; echo "t1 rc=$?"
tail -5 gpurun_out/r2r_t1.log
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2r_trace.log 2>&1; echo "trace rc=$?"
head -1 gpurun_out/r2r_trace.log
timeout 600 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2r_bench.json 2> gpurun_out/r2r_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2r_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2r_bench.json'))
r=d['roofline']; print('value',d['value'],'e2e',d['e2e']['value'],'frac',r['frac'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'], 'parity', d['parity']['ok'], d['parity']['max_rel_err'])
PY
