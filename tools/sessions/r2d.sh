mkdir -p gpurun_out
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2d_trace.log 2>&1; echo "trace rc=$?"
cat gpurun_out/r2d_trace.log | cut -c1-1500
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'parity',d['parity']['ok'], d['parity']['max_rel_err'])
r=d['roofline']; print('frac',r['frac'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],r['ffn']['frac'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'])
PY
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --schedule two-streams --no-parity > gpurun_out/r2d_bench_2s.json 2> gpurun_out/r2d_bench_2s.err; echo "bench2s rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench_2s.json'))
print('two-streams value',d['value'])
r=d['roofline']; print('ffn',r['ffn']['us_per_layer'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'])
PY
