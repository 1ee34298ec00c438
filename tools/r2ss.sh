mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_layer.py -x -q 2>&1 | tail -4
B="python bench.py --steps 300 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline --no-other-configs"
for ll in 0 1; do
SPIF_LAYER_LL_PARTIALS=$ll timeout 600 $B > gpurun_out/r2ss_ll$ll.json 2> gpurun_out/r2ss_ll$ll.err; echo "ll=$ll rc=$?"; tail -2 gpurun_out/r2ss_ll$ll.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2ss_ll$ll.json'))
r=d['roofline']; print('   value',round(d['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity', d['parity']['ok'], d['parity']['max_rel_err'])
PY
done
timeout 600 python tools/stress_layer.py 64 2>&1 | tail -4
