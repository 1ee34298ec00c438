mkdir -p gpurun_out
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2u_trace.log 2>&1; echo "trace rc=$?"
grep -E "^launch [2-3]" gpurun_out/r2u_trace.log | cut -c1-800
