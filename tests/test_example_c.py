"""examples/score_one_pod.c: the C ABI used from plain C (what a cgo binding compiles against).

CPU: the example compiles as C99 against include/gpusched.h, links against the built library and,
with no CUDA device, stops at gs_create (no CPU fallback).  GPU: its printed scores equal the
oracle's for the same tables.
"""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "examples", "score_one_pod.c")


def _build(pkg, tmp_path):
    libdir = os.path.dirname(pkg.LIB_PATH)
    exe = str(tmp_path / "score_one_pod")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-L", libdir,
           "-lgpusched", "-Wl,-rpath," + libdir, "-o", exe]
    env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
    subprocess.run(cmd, check=True, capture_output=True, env=env)
    return exe


def test_example_builds_and_refuses_without_gpu(pkg, tmp_path):
    import torch

    exe = _build(pkg, tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present; covered by the gpu test")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 2 and "gs_create failed" in out.stderr


@pytest.mark.gpu
def test_example_matches_oracle(pkg, oracle_mod, tmp_path):
    exe = _build(pkg, tmp_path)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    rows = re.findall(r"node (\d+): Score\s+(\d+)\s+NormalizeScore\s+(\d+)\s+(no GPU selected|GPU #(\d+))\s+"
                      r"\(eligible (\d), F6 (\d), F7 (\d)\)", out.stdout)
    assert len(rows) == 2
    best = int(re.search(r"best node: (-?\d+)", out.stdout).group(1))

    # the same tables as the example
    synth = pkg.synth
    col_of = np.full((2, synth.MAX_GPUS_PER_NODE + 1), -1, dtype=np.int32)
    col_of[0, 1], col_of[0, 2], col_of[1, 1], col_of[1, 2] = 0, 1, 2, 3
    preds = synth.Predictions(
        conf=np.array([[100, 60, 140, 90], [200, 120, 260, 150]], dtype=np.float32),
        interf=(5.0 + np.arange(8, dtype=np.float32)).reshape(2, 2, 2), col_of=col_of)
    cluster = synth.Cluster(node_model=np.array([1, 2], dtype=np.uint8), node_ngpu=np.array([2, 2], dtype=np.uint8),
                            gpu_off=np.array([0, 1, 1, 1, 3], dtype=np.uint32), res_app=np.array([1, 0, 1], dtype=np.uint32),
                            res_slo=np.array([300, 150, 250], dtype=np.float32), max_gpus=2)
    pods = synth.Pods(app=np.array([0], dtype=np.uint32), slo=np.array([180], dtype=np.float32))
    ora = oracle_mod.Oracle(preds, cluster)
    raw, sel, _ = ora.score_batch(pods, normalize=False)
    norm, _, obest = ora.score_batch(pods, normalize=True)
    for node, score, nscore, _txt, gpu, elig, f6, f7 in rows:
        n = int(node)
        b = int(sel[0, n])
        assert int(score) == int(raw[0, n]) and int(nscore) == int(norm[0, n])
        assert (31 if gpu == "" else int(gpu)) == (b & 31)
        assert (int(elig), int(f6), int(f7)) == (b >> 7 & 1, b >> 5 & 1, b >> 6 & 1)
    assert best == int(obest[0])
