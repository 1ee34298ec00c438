mkdir -p gpurun_out
timeout 300 python tools/trace_layer.py --reps 2 > gpurun_out/r2j_trace.log 2>&1; echo "trace rc=$?"
grep -E "graph replay|launch 0|launch 2|launch 3|launch 8|cta" gpurun_out/r2j_trace.log | cut -c1-1900
