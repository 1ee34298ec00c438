mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -q -x > gpurun_out/r2l_t1.log 2>&1; echo "t1 rc=$?"
tail -3 gpurun_out/r2l_t1.log
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2l_trace.log 2>&1; echo "trace rc=$?"
for v in 0 4; do
SPIF_LAYER_DEBUG=$v timeout 600 python bench.py --steps 50 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2l_$v.json 2> gpurun_out/r2l_$v.err; echo "dbg $v rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2l_$v.json'))
r=d['roofline']; print('dbg $v value',d['value'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'])
PY
done
