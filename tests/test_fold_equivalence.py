"""The unguarded set-up arithmetic of score_fast.cu (fold_fast) equals the guarded form bit for bit.

CPU only: tests/c/fold_check.c restates both forms with IEEE division and compares them over every
count pair of the streamlined domain (see its header)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_unguarded_fold_equals_guarded_fold(tmp_path):
    exe = str(tmp_path / "fold_check")
    env = {k: v for k, v in os.environ.items() if k not in ("CC", "CXX")}
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tests", "c", "fold_check.c"), "-lm"],
                   check=True, capture_output=True, env=env)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "0 differ" in out.stdout
