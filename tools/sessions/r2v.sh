mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_layer.py -q -x > gpurun_out/r2v_t1.log 2>&1; echo "t1 rc=$?"
tail -3 gpurun_out/r2v_t1.log
for act in f32 weight; do
timeout 600 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline --activations $act > gpurun_out/r2v_bench_$act.json 2> gpurun_out/r2v_bench_$act.err; echo "bench $act rc=$?"
tail -2 gpurun_out/r2v_bench_$act.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2v_bench_$act.json'))
r=d['roofline']; print('$act value',d['value'],'e2e',d['e2e']['value'],'frac',r['frac'],'us/launch',r['us_per_launch'],'layer',r['layer']['us_per_layer'], 'parity', d['parity']['ok'], d['parity']['max_rel_err'], d['parity']['index_flips'])
PY
done
