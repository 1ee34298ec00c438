mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r2kk_bench.json 2> gpurun_out/r2kk_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2kk_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2kk_bench.json'))
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline'].get('traffic'), d['gpu_baseline']['value'], d['cpu_baseline']['value'])
for b in d.get('other_configs', []): print(b['workload'], round(b['value'],1), b['parity']['ok'])
for b in d['batched']:
    print(json.dumps(b)[:500])
PY
