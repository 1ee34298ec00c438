mkdir -p gpurun_out
timeout 900 python bench.py --steps 200 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2rr_bench.json 2> gpurun_out/r2rr_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2rr_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2rr_bench.json'))
print(round(d['value'],1), d['parity']['ok'])
for b in d.get('other_configs', []): print(b['workload'], round(b['value'],1), round(b['us_per_layer'],2), b['schedule'][:20], b['parity']['ok'], b['parity']['max_rel_err'])
PY
