mkdir -p gpurun_out
for args in "--wtype q4_0" "--wtype q4_0 --stable" "--wtype q8_0 --stable"; do
echo "== trace $args"; timeout 300 python tools/trace_ffn.py $args 2>&1 | tail -12
done
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2z_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/r2z_tests.log
