mkdir -p gpurun_out
SPIF_LAYER_SPEC_DOWN=1 timeout 900 python -m pytest tests/test_gpu_layer.py -x -q 2>&1 | tail -4
B="python bench.py --steps 300 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline --no-other-configs"
for cfg in "0 2" "1 0" "1 1" "1 2" "1 3"; do
set -- $cfg
SPIF_LAYER_SPEC_DOWN=$1 SPIF_LAYER_SPEC_FILL=$2 timeout 600 $B > gpurun_out/r2nn_s$1_f$2.json 2> gpurun_out/r2nn_s$1_f$2.err; echo "spec=$1 fill=$2 rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2nn_s$1_f$2.json'))
r=d['roofline']; print('   value',round(d['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity', d['parity']['ok'], d['parity']['max_rel_err'])
PY
done
SPIF_LAYER_SPEC_DOWN=1 timeout 600 python tools/stress_layer.py 60 2>&1 | tail -3
