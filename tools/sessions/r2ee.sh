mkdir -p gpurun_out
for dbg in 0 2 4 6; do
SPIF_LAYER_DEBUG=$dbg timeout 600 python bench.py --steps 300 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline --no-other-configs > gpurun_out/r2ee_bench_d$dbg.json 2> gpurun_out/r2ee_bench_d$dbg.err; echo "bench dbg=$dbg rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2ee_bench_d$dbg.json'))
r=d['roofline']; print('dbg $dbg value',round(d['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity', d['parity']['ok'], d['parity']['max_rel_err'])
PY
done
SPIF_LAYER_DEBUG=6 timeout 900 python -m pytest tests/test_gpu_layer.py -x -q 2>&1 | tail -3
SPIF_LAYER_DEBUG=6 timeout 600 python tools/trace_layer.py --layers 8 > gpurun_out/r2ee_trace_d6.txt 2>&1
grep -n "^launch 2" -A8 gpurun_out/r2ee_trace_d6.txt | cut -c1-1000
