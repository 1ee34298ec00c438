mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -q > gpurun_out/r2g_t1.log 2>&1; echo "t1 rc=$?"
tail -4 gpurun_out/r2g_t1.log
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2g_trace.log 2>&1; echo "trace rc=$?"
grep -E "graph replay|launch 0|launch 2|launch 8|cta" gpurun_out/r2g_trace.log | cut -c1-1700
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2g_bench.err
SPIF_LAYER_DEBUG=1 timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2g_bench_dbg.json 2> gpurun_out/r2g_bench_dbg.err; echo "bench dbg rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/r2g_bench.json','gpurun_out/r2g_bench_dbg.json'):
    try:
        d=json.load(open(f))
        r=d['roofline']; print(f, 'value',d['value'],'frac',r['frac'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'], 'parity', d.get('parity',{}).get('ok'))
    except Exception as e: print(f, 'ERR', e)
PY
