mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2mm_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r2mm_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2mm_all.log 2>&1; echo "all rc=$?"; tail -7 gpurun_out/r2mm_all.log
