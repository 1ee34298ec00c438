mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:spif_layer3_kernel -s 50 -c 1 -o gpurun_out/r2t_layer3 -f python bench.py --steps 3 --warmup 3 --no-batched --no-cpu-baseline --no-gpu-baseline --no-parity > gpurun_out/r2t_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r2t_layer3.ncu-rep
