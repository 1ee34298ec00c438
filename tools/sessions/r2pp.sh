# final round-2 evidence: default bench line, reference arm, ncu launch list, one full capture of the layer kernel
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2pp_bench.json 2> gpurun_out/r2pp_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2pp_bench.err
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2pp_bench_ref.json 2> gpurun_out/r2pp_bench_ref.err; echo "ref rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity --no-other-configs"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2pp_launches.csv $B > gpurun_out/r2pp_ncu1.log 2>&1; echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spif_layer4 -s 200 -c 3 -o gpurun_out/r2pp_layer4 -f $B > gpurun_out/r2pp_ncu2.log 2>&1; echo "ncu2 rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2pp_bench.json'))
r=d['roofline']; print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'frac',round(r['frac'],4),'us/launch',round(r['us_per_launch'],2),'parity',d['parity']['ok'],'gpu_baseline',round(d['gpu_baseline']['value'],1),'cpu',round(d['cpu_baseline']['value'],1))
PY
