mkdir -p gpurun_out
timeout 600 python tools/trace_layer.py --layers 8 > gpurun_out/r2dd_trace.txt 2>&1; echo "trace rc=$?"; head -1 gpurun_out/r2dd_trace.txt
grep -A0 "extra" gpurun_out/r2dd_trace.txt | head -12
timeout 900 python -m pytest tests/test_gpu_layer.py -x -q 2>&1 | tail -3
