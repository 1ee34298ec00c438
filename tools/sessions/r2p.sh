# round 2 re-entry: where does the tree stand?  full GPU suite, then the bench with each layer-kernel variant and the round-1 schedule
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2p_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/r2p_tests.log
for v in 1 0; do
SPIF_LAYER_IMPL=$v timeout 600 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline > gpurun_out/r2p_impl$v.json 2> gpurun_out/r2p_impl$v.err; echo "impl $v rc=$?"
tail -2 gpurun_out/r2p_impl$v.err
done
timeout 600 python bench.py --steps 200 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline --schedule two-streams > gpurun_out/r2p_2s.json 2> gpurun_out/r2p_2s.err; echo "2s rc=$?"
python - <<'PY'
import json
for f in ('r2p_impl1','r2p_impl0','r2p_2s'):
    try:
        d=json.load(open('gpurun_out/%s.json'%f))
        r=d['roofline']; print(f, 'value',d['value'],'e2e',d['e2e']['value'],'frac',r['frac'],'us/launch',r['us_per_launch'],'ffn',r.get('ffn',{}).get('us_per_layer'),'pred',r.get('predictor',{}).get('us_per_layer'),'layer',r.get('layer',{}).get('us_per_layer'), 'parity', d.get('parity',{}).get('ok'), d.get('parity',{}).get('max_rel_err'))
        print('  gpu_baseline', d.get('gpu_baseline'))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2p_trace.log 2>&1; echo "trace rc=$?"
