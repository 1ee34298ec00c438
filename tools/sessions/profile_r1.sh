# run on the GPU box: plain runs first, then the ncu passes of the same commands (see B200_PROFILING.md)
set -e
B="python bench.py --steps 3 --warmup 3 --pool 2 --no-cpu-baseline --no-batched"
C="python tools/bench_batch.py --layers 2 --iters 1 --batches 8,32"
$B > gpurun_out/r1c_plain.log 2>&1
$C > gpurun_out/r1c_plain_batch.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:spif_ -s 700 -c 800 --csv --log-file gpurun_out/r1c_launches.csv $B > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:spif_ffn_direct -s 400 -c 3 -f -o gpurun_out/r1c_ffn_direct $B > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:spif_batch -c 60 --csv --log-file gpurun_out/r1c_launches_batch.csv $C > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:spif_batch_gateup -s 4 -c 2 -f -o gpurun_out/r1c_batch_gateup $C > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:spif_batch_down -s 4 -c 2 -f -o gpurun_out/r1c_batch_down $C > /dev/null 2>&1
