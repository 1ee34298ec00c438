mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:spif_layer2_kernel -s 50 -c 2 -o gpurun_out/r2n_layer2 -f python bench.py --steps 3 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2n_ncu.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/r2n_ncu.log
ls -la gpurun_out/r2n_layer2.ncu-rep
