mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_layer.py -x -q > gpurun_out/r2ii_layer.log 2>&1; echo "layer rc=$?"; tail -6 gpurun_out/r2ii_layer.log
timeout 600 python bench.py --steps 300 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline --no-other-configs > gpurun_out/r2ii_bench.json 2> gpurun_out/r2ii_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2ii_bench.json'))
r=d['roofline']; print('value',round(d['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity', d['parity']['ok'])
PY
