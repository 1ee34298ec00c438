mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2bb_bench.json 2> gpurun_out/r2bb_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2bb_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2bb_bench.json'))
print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline'].get('traffic'))
for b in d.get('other_configs', []): print(json.dumps(b)[:600])
for b in d['batched']:
    if b['n_tok']==1: print(json.dumps(b))
PY
timeout 900 python bench.py --workload relullama-70b --steps 100 --warmup 5 --no-batched --no-cpu-baseline --no-gpu-baseline > gpurun_out/r2bb_bench_70b_n1.json 2> gpurun_out/r2bb_bench_70b.err; echo "70b rc=$?"; tail -2 gpurun_out/r2bb_bench_70b.err
python - <<PY
import json
d=json.load(open('gpurun_out/r2bb_bench_70b_n1.json'))
print('70B N=1', d['value'], d['roofline']['frac'], d['roofline'].get('us_per_launch'), d['parity'])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spif_ffn_qwarp -s 8 -c 3 -o gpurun_out/r2bb_qwarp -f python tools/trace_ffn.py --wtype q4_0 --layers 4 > gpurun_out/r2bb_ncu.log 2>&1; echo "ncu rc=$?"
