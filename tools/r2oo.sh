mkdir -p gpurun_out
SPIF_LAYER_SPEC_DOWN=1 SPIF_LAYER_SPEC_FILL=1 timeout 600 python tools/trace_layer.py --layers 8 > gpurun_out/r2oo_trace_spec.txt 2>&1; echo "trace rc=$?"
grep -n "^launch 2" -A8 gpurun_out/r2oo_trace_spec.txt | cut -c1-1100
