# round 2, GPU call A: the new layer kernel -- smallest test first, then the file, trace, bench
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_layer.py -x -q -k "mask_to_list or (ffn_only and 512)" > gpurun_out/r2a_t0.log 2>&1; echo "t0 rc=$?"
tail -5 gpurun_out/r2a_t0.log
timeout 900 python -m pytest tests/test_gpu_layer.py -x -q > gpurun_out/r2a_t1.log 2>&1; echo "t1 rc=$?"
tail -15 gpurun_out/r2a_t1.log
timeout 300 python tools/trace_layer.py > gpurun_out/r2a_trace.log 2>&1; echo "trace rc=$?"
tail -12 gpurun_out/r2a_trace.log
timeout 900 python bench.py --steps 200 --warmup 5 --no-batched > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/r2a_bench.json; tail -5 gpurun_out/r2a_bench.err
