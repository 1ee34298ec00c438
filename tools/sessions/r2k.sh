mkdir -p gpurun_out
for v in 0 2 4 6; do
SPIF_LAYER_DEBUG=$v timeout 600 python bench.py --steps 50 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2k_$v.json 2> gpurun_out/r2k_$v.err; echo "dbg $v rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2k_$v.json'))
r=d['roofline']; print('dbg $v value',d['value'],'us/launch',r['us_per_launch'],'ffn',r['ffn']['us_per_layer'],'pred',r['predictor']['us_per_layer'],'layer',r['layer']['us_per_layer'])
PY
done
