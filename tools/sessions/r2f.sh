mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -q > gpurun_out/r2f_t1.log 2>&1; echo "t1 rc=$?"
tail -4 gpurun_out/r2f_t1.log
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2f_trace.log 2>&1; echo "trace rc=$?"
grep -E "graph replay|launch 0|launch 2|launch 8|cta" gpurun_out/r2f_trace.log | cut -c1-1700
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2f_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2f_bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'parity',d['parity']['ok'], d['parity']['max_rel_err'])
r=d['roofline']; print('frac',r['frac'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],r['ffn']['frac'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'])
print('gpu_baseline', d.get('gpu_baseline'))
PY
