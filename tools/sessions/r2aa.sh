mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2aa_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2aa_tests.log
for args in "--wtype q4_0" "--wtype q4_0 --stable"; do
echo "== trace $args"; timeout 300 python tools/trace_ffn.py $args 2>&1 | tail -4
done
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2aa_bench.json 2> gpurun_out/r2aa_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2aa_bench.json'))
print(d['value'], d['roofline']['frac'])
for b in d['batched']:
    if b['n_tok']==1: print(json.dumps(b))
PY
