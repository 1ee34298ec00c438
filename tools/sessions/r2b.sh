mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -x -q > gpurun_out/r2b_t1.log 2>&1; echo "t1 rc=$?"
tail -4 gpurun_out/r2b_t1.log
for h in 0 1; do
  SPIF_LAYER_L2_HINT=$h timeout 300 python tools/trace_layer.py > gpurun_out/r2b_trace_h$h.log 2>&1; echo "trace hint=$h rc=$?"
  head -1 gpurun_out/r2b_trace_h$h.log; sed -n 2,5p gpurun_out/r2b_trace_h$h.log; tail -1 gpurun_out/r2b_trace_h$h.log
done
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2b_bench.json'))
print('value',d['value'],'e2e',d['e2e']['value'],'parity',d['parity'])
print(json.dumps(d['roofline'],indent=0))
PY
