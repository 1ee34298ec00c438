mkdir -p gpurun_out
timeout 900 python bench.py --steps 200 --warmup 10 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2qq_bench.json 2> gpurun_out/r2qq_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2qq_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2qq_bench.json'))
print(round(d['value'],1), d['parity']['ok'])
for b in d.get('other_configs', []): print(b['workload'], round(b['value'],1), round(b['us_per_layer'],2), b['schedule'], b['parity'])
PY
timeout 600 python bench.py --workload prosparse-13b-q4_0 --steps 100 --warmup 5 > gpurun_out/r2qq_bench_q4.json 2> gpurun_out/r2qq_bench_q4.err; echo "q4 main rc=$?"; tail -3 gpurun_out/r2qq_bench_q4.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2qq_bench_q4.json'))
print('q4_0 main', round(d['value'],1), d['config'].get('schedule'), json.dumps(d['roofline'])[:400], d['parity'])
PY
