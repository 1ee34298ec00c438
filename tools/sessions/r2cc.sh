mkdir -p gpurun_out
timeout 600 python tools/trace_layer.py --layers 8 > gpurun_out/r2cc_trace.txt 2>&1; echo "trace rc=$?"; head -3 gpurun_out/r2cc_trace.txt | cut -c1-600
