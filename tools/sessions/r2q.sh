mkdir -p gpurun_out
SPIF_LAYER_IMPL=0 timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2q_trace_impl0.log 2>&1; echo "trace rc=$?"
SPIF_LAYER_IMPL=0 timeout 600 ncu --set full --import-source on --clock-control none -k regex:spif_layer_kernel -s 50 -c 2 -o gpurun_out/r2q_layer0 -f python bench.py --steps 3 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2q_ncu0.log 2>&1; echo "ncu0 rc=$?"
SPIF_LAYER_IMPL=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:spif_layer2_kernel -s 50 -c 2 -o gpurun_out/r2q_layer1 -f python bench.py --steps 3 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2q_ncu1.log 2>&1; echo "ncu1 rc=$?"
ls -la gpurun_out/*.ncu-rep
