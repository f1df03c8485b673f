"""Builds libgpusched.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python k8s-gpu-scheduler_b200/build.py [--force] [--verbose]
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgpusched.so")
SOURCES = ["context.cu", "score_kernels.cu", "score_fast.cu", "score_pods.cu", "complete.cu", "eig_jacobi.cu", "tridiag.cu", "multi.cu", "noalloc.cu"]
HOST_SOURCES = ["host/gpu_plugin.cpp", "host/dcgm.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # no FMA contraction: the reference rounds after every operation
    "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-Wall,-fvisibility=hidden",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(b, f) for b, _, fs in os.walk(CSRC) for f in fs] + [
        os.path.join(HERE, "..", "include", "gpusched.h"), os.path.join(HERE, "..", "include", "gpusched_host.h"),
        os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, tag: str = "", defines=()) -> str:
    """tag/defines: tuning variants (libgpusched_<tag>.so built with extra -D flags; pick one at run
    time with GS_LIB_PATH)."""
    global LIB
    if tag:
        LIB = os.path.join(HERE, f"libgpusched_{tag}.so")
        force = True
    if not force and not needs_build():
        return LIB
    env = dict(os.environ)
    env.pop("CC", None), env.pop("CXX", None)
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, "build", (tag + "_" if tag else "") + src.replace(".cu", ".o"))
        cmd = [_nvcc(), *NVCC_FLAGS, *defines, "-ccbin", "/usr/bin/g++", "-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)))
        objs.append(obj)
    for src in HOST_SOURCES:
        obj = os.path.join(HERE, "build", os.path.basename(src).replace(".cpp", ".o"))
        cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-fvisibility=hidden", "-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)))
        objs.append(obj)
    log = []
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"== {src}\n{out}")
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
    cmd = [_nvcc(), "-shared", "-ccbin", "/usr/bin/g++", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs,
           "-Xcompiler", "-fPIC", "-cudart", "static"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    with open(os.path.join(HERE, "build", "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return LIB


if __name__ == "__main__":
    _tag = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--tag=")), "")
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, tag=_tag,
                defines=[a for a in sys.argv if a.startswith("-D")]))
