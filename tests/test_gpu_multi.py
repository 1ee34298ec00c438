"""Multi-GPU parity of the shard exchange and the sharded layers (skipped with fewer than 2 GPUs): runs
tools/test_peer_exchange.py under torchrun with 2 ranks -- stand-alone exchange vs ncclAllReduce, sharded layer vs unsharded,
the three exchange forms interleaved on ONE comm with odd call counts, and the sharded look-ahead chain.  -m gpu."""
import json
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2])
def test_peer_exchange_multi_gpu(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tools", "test_peer_exchange.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["allreduce_parity"] == "ok" and out["fused_layer_parity"] == "ok"
    assert out["mixed_forms_one_comm"]["status"] == 0 and out["sharded_lookahead_chain"] == "ok"
