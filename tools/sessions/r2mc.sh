# usage: bash tools/sessions/r2mc.sh N   (NVLS multicast stores for the shard exchange: parity at world 2, then the bench with and without)
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/r2mc_tests_n$N.log 2>&1; echo "multi tests rc=$?"; tail -3 gpurun_out/r2mc_tests_n$N.log
for MC in 1 0 1 0; do
SPIF_COMM_MULTICAST=$MC timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$MC bench.py --gpus $N --steps 200 --warmup 10 --no-scale-70b > gpurun_out/r2mc_n${N}_mc$MC.json 2> gpurun_out/r2mc_n${N}_mc$MC.err; echo "bench N=$N mc=$MC rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2mc_n${N}_mc$MC.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print("N=$N mc=$MC value", d["value"], "e2e", d["e2e"]["value"], "us/launch", r.get("us_per_launch"), "parity", d.get("parity"), d["config"]["sharding"][-40:])
except Exception as e:
    print('no line', e)
PY
tail -2 gpurun_out/r2mc_n${N}_mc$MC.err
done
