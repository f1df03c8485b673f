"""bench.py --impl reference: the JSON line the driver parses, run here on a tiny workload.

Only the reference arm can run without a GPU (it times oracle/ on the host cores); the b200 arm
must refuse to run without the CUDA library rather than fall back (checked in test_abi.py).
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra, env=None):
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
           "--pods", "64", "--nodes", "200"] + extra
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    return lines


def test_reference_arm_line():
    lines = _run([])
    assert len(lines) == 1
    r = json.loads(lines[0])
    assert r["impl"] == "reference"
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in r, k
    assert r["steps"] == 2 and r["warmup"] == 1 and r["n_gpus"] == 1
    assert r["higher_is_better"] is True and r["vs_baseline"] is None
    assert "workload" in r["config"]
    cb = r["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == r["value"] and cb["sample"]
    e = r["e2e"]
    assert e["value"] == r["value"] and e["unit"] == r["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert r["value"] > 0


def test_reference_arm_other_ranks_idle():
    """Under torchrun only rank 0 runs and prints; the others exit 0 without work."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    assert _run(["--gpus", "2"], env=env) == []
