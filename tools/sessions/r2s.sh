mkdir -p gpurun_out
timeout 300 tools/ubench/stream_items > gpurun_out/r2s_stream_items.txt 2>&1; echo "ubench rc=$?"
cat gpurun_out/r2s_stream_items.txt
timeout 600 python -m pytest tests/test_gpu_layer.py -q -x > gpurun_out/r2s_t1.log 2>&1; echo "t1 rc=$?"
tail -3 gpurun_out/r2s_t1.log
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2s_trace.log 2>&1; echo "trace rc=$?"
head -1 gpurun_out/r2s_trace.log
timeout 600 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2s_bench.json'))
r=d['roofline']; print('value',d['value'],'e2e',d['e2e']['value'],'frac',r['frac'],'us/launch',r['us_per_launch'],'layer',r['layer']['us_per_layer'], 'parity', d['parity']['ok'], d['parity']['max_rel_err'])
PY
