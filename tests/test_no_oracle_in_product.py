"""The oracle is the checker, never the product: nothing shipped may import, load or exec oracle/."""
import os
import re

from conftest import ROOT

PRODUCT_DIRS = ["pytorch_distributed_matmul_benchmark_b200", "include"]
PRODUCT_FILES = ["matmul_benchmark.py", "matmul_scaling_benchmark.py", "matmul_overlap_benchmark.py",
                 "matmul_distributed_benchmark.py", "compare_benchmarks.py", "run_benchmark.sh", "run_scaling_benchmark.sh", "run_overlap_benchmark.sh",
                 "run_distributed_benchmark.sh", "run_common.sh"]
PAT = re.compile(r"^\s*(from|import)\s+oracle\b|liboracle|oracle/_build|oracle\.gemm|gemm_oracle", re.M)


def _files():
    for d in PRODUCT_DIRS:
        for dirpath, _, names in os.walk(os.path.join(ROOT, d)):
            if "build" in dirpath.split(os.sep):
                continue
            for n in names:
                if n.endswith((".py", ".cu", ".cuh", ".h", ".sh", "Makefile")):
                    yield os.path.join(dirpath, n)
    for f in PRODUCT_FILES:
        p = os.path.join(ROOT, f)
        if os.path.exists(p):
            yield p


def test_product_never_touches_the_oracle():
    offenders = []
    for path in _files():
        text = open(path, errors="ignore").read()
        if PAT.search(text):
            offenders.append(os.path.relpath(path, ROOT))
    assert not offenders, f"product files referencing the oracle: {offenders}"


def test_product_has_no_library_gemm_fallback():
    """No torch.matmul / cuBLAS on the product path (hostlib.parity_error is the --check reference, not a fallback)."""
    bad = []
    for path in _files():
        if not path.endswith(".py"):
            continue
        rel = os.path.relpath(path, ROOT)
        for i, line in enumerate(open(path).read().splitlines(), 1):
            code = line.split("#")[0]
            if re.search(r"torch\.(matmul|bmm|mm)\(|cublas", code) and not rel.endswith("hostlib.py"):
                bad.append(f"{rel}:{i}")
    assert not bad, bad
    for path in _files():
        if path.endswith((".cu", ".cuh", "Makefile")):
            text = open(path).read()
            # comments may quote the library's measured numbers; what must not appear is a use of it
            assert not re.search(r"#include\s*[<\"]cublas|cublas(Lt)?[A-Z]\w*\s*\(|-lcublas|libcublas", text), path


def test_tma_loads_are_never_multicast():
    """TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
    memory hung the GPU.  The library's loads (operands A and B, always local) are plain per-CTA loads, so the rule
    holds by construction: no `cp.async.bulk` load with `.multicast::cluster` exists in the sources or in the SASS."""
    import subprocess
    from pytorch_distributed_matmul_benchmark_b200 import _lib
    for path in _files():
        if path.endswith((".cu", ".cuh")):
            for line in open(path).read().splitlines():
                assert not ("cp.async.bulk" in line and "multicast::cluster" in line), path
    sass = subprocess.check_output(["cuobjdump", "-sass", _lib.LIB_PATH], text=True, stderr=subprocess.DEVNULL)
    assert not re.search(r"(UTMALDG|UBLKCP)\S*MULTICAST", sass)
