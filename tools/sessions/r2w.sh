# round-2 re-entry: full GPU suite, default bench line, ncu launch list + one full capture of the layer kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2w_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r2w_tests.log
timeout 900 python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2w_bench.err
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2w_bench_ref.json 2> gpurun_out/r2w_bench_ref.err; echo "ref rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2w_launches.csv $B > gpurun_out/r2w_ncu1.log 2>&1; echo "ncu1 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spif_layer4 -s 200 -c 3 -o gpurun_out/r2w_layer4 -f $B > gpurun_out/r2w_ncu2.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out | tail
