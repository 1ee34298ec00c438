timeout 300 python -m pytest tests/test_gpu_batched.py -x -q 2>&1 | tail -15 > gpurun_out/s19_pytest6.log
for d in 0 1; do
SPIF_BATCH_DEBUG=$d timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:spif_batch_gateup -c 8 --csv --log-file gpurun_out/dbg_$d.csv python tools/bench_batch.py --layers 1 --iters 1 --batches 2,32 > /dev/null 2>&1
done
timeout 300 python tools/bench_batch.py --batches 2,8,32 > gpurun_out/s19_batch_f16_v4.txt 2>&1
