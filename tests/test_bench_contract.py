"""CPU: the reference arm of bench.py (`--impl reference`: the reference's own CPU ops from oracle/_ref, else the oracle port)
prints ONE JSON line with the keys the driver reads; runs on a bounded sample (one layer of the 7B shape) in seconds."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "prosparse-7b",
                        "--cpu-layers", "1", "--steps", "3", "--warmup", "3"], capture_output=True, text=True, timeout=300,
                       cwd=ROOT, env={**os.environ, "RANK": "0", "WORLD_SIZE": "1"})
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["value"] > 0 and d["unit"].startswith("tok/s")
    assert d["steps"] == 3 and d["warmup"] >= 3 and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and "layers" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("ProSparse-Llama-2-7B") and d["gpu_launches"] == 0


def test_reference_arm_other_ranks_do_nothing():
    """under torchrun (N > 1) only rank 0 runs the CPU arm; the other ranks exit 0 without output"""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT,
                       env={**os.environ, "RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""
