"""The C-ABI library loads on a GPU-less host and exports exactly what include/b2mm.h declares.
No compute is launched here; only argument validation that returns before any CUDA call."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT
from pytorch_distributed_matmul_benchmark_b200 import _lib

HEADER = os.path.join(ROOT, "include", "b2mm.h")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"B2MM_API\s+[\w\s\*]+?\b(b2mm_\w+)\s*\(", text))


def test_header_python_and_library_agree():
    declared = _declared()
    assert declared, "no declarations parsed from include/b2mm.h"
    assert declared == set(_lib.SIGNATURES), "include/b2mm.h and _lib.SIGNATURES list different entry points"
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    assert declared <= exported, f"missing from libb2mm.so: {declared - exported}"
    assert {s for s in exported if s.startswith("b2mm_")} == declared, "undeclared b2mm_* symbols exported"


def test_every_declaration_cites_the_reference_or_says_it_has_no_counterpart():
    text = open(HEADER).read()
    for name in _declared():
        idx = text.index(name + "(")
        comment = text[text.rfind("/*", 0, idx):idx]
        cites = re.search(r"matmul_(scaling_)?benchmark\.py:\d+|matmul_overlap_benchmark\.py:\d+", comment)
        housekeeping = name in {"b2mm_version", "b2mm_last_error", "b2mm_launch_count", "b2mm_gemm_workspace_bytes",
                                "b2mm_comm_init"}
        assert cites or housekeeping or "No counterpart in the reference" in comment or "no counterpart in the reference" in comment, name


def test_library_loads_and_reports_version():
    lib = _lib.load()
    assert lib.b2mm_version() == 100
    assert lib.b2mm_last_error() in (b"", None) or isinstance(lib.b2mm_last_error(), bytes)
    assert _lib.launch_count() >= 0


def test_argument_errors_are_status_codes_not_crashes():
    lib = _lib.load()
    one = ctypes.c_void_p(16)  # never dereferenced: validation fails first
    rc = lib.b2mm_gemm(one, one, one, 8, 8, 8, 8, 8, 8, 1, 0, 0, 0, 7, None, 0, None)
    assert rc == _lib.ERR_BAD_ARG and b"dtype" in lib.b2mm_last_error()
    rc = lib.b2mm_gemm(one, one, one, -1, 8, 8, 8, 8, 8, 1, 0, 0, 0, _lib.BF16, None, 0, None)
    assert rc == _lib.ERR_BAD_ARG
    rc = lib.b2mm_gemm(None, one, one, 8, 8, 8, 8, 8, 8, 1, 0, 0, 0, _lib.BF16, None, 0, None)
    assert rc == _lib.ERR_BAD_ARG
    rc = lib.b2mm_gemm(one, one, one, 8, 8, 8, 4, 8, 8, 1, 0, 0, 0, _lib.BF16, None, 0, None)  # lda < K
    assert rc == _lib.ERR_BAD_ARG
    rc = lib.b2mm_gemm(one, one, one, 8, 12, 8, 8, 12, 12, 1, 0, 0, 0, _lib.BF16, None, 0, None)  # 24-byte rows
    assert rc == _lib.ERR_UNSUPPORTED and b"16 bytes" in lib.b2mm_last_error()
    with pytest.raises(_lib.B2mmError) as ei:
        _lib.check(rc)
    assert ei.value.status == _lib.ERR_UNSUPPORTED
    # empty problems are no-ops (torch.bmm on a 0-batch tensor; matmul_scaling_benchmark.py:111 at ws=8)
    assert lib.b2mm_gemm(None, None, None, 8, 8, 8, 8, 8, 8, 0, 0, 0, 0, _lib.BF16, None, 0, None) == _lib.OK
    assert lib.b2mm_all_reduce_sum(None, None, 0, _lib.BF16, None) == _lib.ERR_BAD_ARG  # no communicator


def test_workspace_query():
    lib = _lib.load()
    assert lib.b2mm_gemm_workspace_bytes(4096, 4096, 4096, 1, _lib.BF16) == 0
    assert lib.b2mm_gemm_workspace_bytes(4096, 4096, 4096, 1, _lib.FP16) == 0
    need = lib.b2mm_gemm_workspace_bytes(512, 256, 128, 2, _lib.FP32)
    assert need >= 2 * 4 * 2 * (512 * 128 + 128 * 256)  # hi + lo of both operands


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libb2mm.so"))
    with pytest.raises(ImportError, match="no fallback"):
        _lib.load()


def test_cpu_tensors_are_rejected():
    import torch
    import pytorch_distributed_matmul_benchmark_b200 as b2
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        b2.matmul(torch.zeros(8, 8), torch.zeros(8, 8))


def test_multicast_helpers_without_a_gpu():
    """multicast.py is host plumbing over the cuda-python driver bindings; what can be checked without a device:
    the raw-memory wrapper torch consumes, and that a driver error becomes a RuntimeError naming the call."""
    import pytest
    from pytorch_distributed_matmul_benchmark_b200 import multicast
    blob = multicast._Blob(0x7f0000000000, 4096)
    iface = blob.__cuda_array_interface__
    assert iface["shape"] == (4096,) and iface["typestr"] == "|u1" and iface["data"] == (0x7f0000000000, False)
    assert multicast._ok((0, 7), "x") == 7 and multicast._ok((0, 1, 2), "x") == (1, 2)
    drv = multicast._drv()
    with pytest.raises(RuntimeError, match="cuMemCreate failed"):
        multicast._ok((drv.CUresult.CUDA_ERROR_OUT_OF_MEMORY, None), "cuMemCreate")
