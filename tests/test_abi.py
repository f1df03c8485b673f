"""The C-ABI library loads without a GPU and exports every symbol include/gpusched.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpusched.h")).read()
    return sorted(set(re.findall(r"GS_API\s+[\w\s\*]+?\b(gs_\w+)\s*\(", text)))


def test_header_and_python_binding_agree(pkg):
    assert declared_symbols() == sorted(pkg.EXPORTS)


def test_library_exports_every_declared_symbol(pkg):
    import __graft_entry__ as entry

    entry.build()
    lib = ctypes.CDLL(pkg.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert b"sm_100a" in pkg.load_library().gs_version()


def test_header_compiles_as_c():
    import subprocess
    import tempfile

    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        open(src, "w").write('#include "gpusched.h"\nint main(void){gs_cluster c; gs_predictions p; (void)c; (void)p; return GS_OK;}\n')
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", src,
                        "-o", os.path.join(d, "t.o")], check=True)


def test_no_cpu_fallback_without_gpu(pkg):
    """Without a CUDA device the product path must fail loudly, not compute on the CPU."""
    import torch

    if torch.cuda.is_available():
        return
    try:
        pkg.Scheduler(0)
    except pkg.GsError as e:
        assert e.code == -2
    else:
        raise AssertionError("Scheduler() succeeded without a CUDA device")


def test_product_never_touches_the_oracle():
    """Nothing under the package directory may import, link or execute oracle/."""
    pkg_dir = os.path.join(ROOT, "k8s-gpu-scheduler_b200")
    for base, _, files in os.walk(pkg_dir):
        if os.path.basename(base) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".go")):
                text = open(os.path.join(base, f), errors="ignore").read()
                assert "liboracle" not in text and "score_oracle" not in text and "from oracle" not in text, f
