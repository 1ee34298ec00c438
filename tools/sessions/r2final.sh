# round-2 closing run on the final tree: the whole GPU suite, smoke(), the default bench line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.txt 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/r2_gpu_tests.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2_smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.txt
timeout 900 python bench.py > gpurun_out/r2_bench_n1_default.json 2> gpurun_out/r2_bench_n1_default.err; echo "bench rc=$?"; tail -2 gpurun_out/r2_bench_n1_default.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench_n1_default.json').read().strip().splitlines()[-1])
r=d['roofline']; print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity',d['parity']['ok'],'gpu_baseline',round(d['gpu_baseline']['value'],1),'cpu',round(d['cpu_baseline']['value'],1))
for o in d['other_configs']: print(o['workload'], round(o['value'],1), o.get('schedule'), o['parity']['ok'])
PY
