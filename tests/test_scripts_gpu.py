"""The drop-in scripts end to end on the GPU (single process), and the multi-rank modes when >= 2 GPUs are present."""
import re
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _run(cmd, timeout=900):
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-4000:]
    return p.stdout


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_matmul_benchmark_script():
    out = _run([sys.executable, "matmul_benchmark.py", "--sizes", "1024", "2048", "--iterations", "5", "--warmup", "2",
                "--dtype", "bfloat16", "--check"])
    assert "Running in single GPU mode" in out
    assert "Results for 1024x1024:" in out and "Results for 2048x2048:" in out
    assert out.count("-> OK") == 2 and "ERROR" not in out
    assert re.search(r"  - TFLOPS per GPU: \d+\.\d\d", out)
    assert "% of B200 theoretical peak" in out


@pytest.mark.parametrize("dtype", ["float16", "float32"])
def test_scaling_script_single_gpu_all_modes(dtype):
    for mode in ("independent", "batch_parallel", "matrix_parallel"):
        out = _run([sys.executable, "matmul_scaling_benchmark.py", "--sizes", "1024", "--iterations", "3", "--warmup",
                    "1", "--dtype", dtype, "--mode", mode, "--check"])
        assert f"  - Mode: {mode}" in out and "Results for 1024x1024:" in out
        assert "-> OK" in out and "ERROR" not in out and "FAIL" not in out


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode,extra", [("independent", []), ("batch_parallel", []), ("matrix_parallel", []),
                                        ("batch_parallel", ["--overlap", "--panels", "2"]),
                                        ("matrix_parallel", ["--overlap", "--panels", "4"]),
                                        ("matrix_parallel", ["--fused"]),
                                        ("batch_parallel", ["--fused", "--panels", "2"]),
                                        ("batch_parallel", ["--fused", "--panels", "2", "--dtype", "float32"]),
                                        ("matrix_parallel", ["--fused", "--dtype", "float32"])])
def test_scaling_script_two_gpus(mode, extra):
    dtype = [] if "--dtype" in extra else ["--dtype", "bfloat16"]
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                "--master-addr=127.0.0.1", f"--master-port={_free_port()}", "matmul_scaling_benchmark.py", "--sizes",
                "2048", "--iterations", "3", "--warmup", "2", "--mode", mode, "--check"] + dtype + extra)
    assert "Collective operations verified successfully across 2 GPUs" in out
    assert "Results for 2048x2048:" in out and "ERROR" not in out and "FAIL" not in out
    assert "-> OK" in out
    if "--overlap" in extra:
        assert "Overlapped end-to-end:" in out
    if "--fused" in extra:
        assert "Fused GEMM+all_" in out
        assert out.count("-> OK") >= (2 if mode == "batch_parallel" else 1)  # local bmm parity + fused-vs-NCCL parity


@pytest.mark.parametrize("mode", ["no_overlap", "overlap", "pipeline"])
def test_overlap_script_single_gpu(mode):
    out = _run([sys.executable, "matmul_overlap_benchmark.py", "--sizes", "1024", "--iterations", "4", "--warmup", "2",
                "--mode", mode])
    assert f"  - Mode: {mode}" in out and "Results for 1024x1024:" in out and "ERROR" not in out
    assert re.search(r"  - Actual TFLOPS: \d+\.\d\d \(FLOPs/Time\)", out)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode", ["no_overlap", "overlap", "pipeline"])
def test_overlap_script_two_gpus(mode):
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                "--master-addr=127.0.0.1", f"--master-port={_free_port()}", "matmul_overlap_benchmark.py", "--sizes",
                "2048", "--iterations", "5", "--warmup", "2", "--mode", mode, "--check"])
    assert "Results for 2048x2048:" in out and "ERROR" not in out and "FAIL" not in out
    assert "Parity of the reduced product" in out and "-> OK" in out
    assert "  - Note: In data parallel, each GPU does full matrix multiply" in out


@pytest.mark.parametrize("mode", ["independent", "data_parallel", "model_parallel"])
def test_distributed_draft_script_single_gpu(mode):
    out = _run([sys.executable, "matmul_distributed_benchmark.py", "--sizes", "1024", "--iterations", "3", "--warmup", "1",
                "--mode", mode, "--check"])
    assert f"  - Mode: {mode}" in out and "Results for 1024x1024:" in out
    assert "-> OK" in out and "ERROR" not in out and "FAIL" not in out


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode,size", [("data_parallel", "2048"), ("model_parallel", "2048"), ("model_parallel", "2056")])
def test_distributed_draft_script_two_gpus(mode, size):
    """2056 / 2 = 1028-wide bf16 slabs: a row pitch of 2056 bytes is not a multiple of 16, so this takes the padded-pitch path."""
    out = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                "--master-addr=127.0.0.1", f"--master-port={_free_port()}", "matmul_distributed_benchmark.py", "--sizes",
                size, "--iterations", "3", "--warmup", "2", "--mode", mode, "--check"])
    assert f"Results for {size}x{size}:" in out and "ERROR" not in out and "FAIL" not in out and "-> OK" in out
    assert "  - Communication time:" in out and "  - Scaling efficiency:" in out
