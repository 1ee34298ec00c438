mkdir -p gpurun_out
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2c_trace.log 2>&1; echo "trace rc=$?"
cat gpurun_out/r2c_trace.log
