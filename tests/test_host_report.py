"""Host layer parity: CLI surface and report text against what the reference itself printed
(tests/golden/host_logic.json, captured from the reference's run_benchmarks with stub timings). CPU only."""
import contextlib
import io
import os
import re
import subprocess

import pytest
import torch

from conftest import ROOT
from pytorch_distributed_matmul_benchmark_b200 import hostlib
from pytorch_distributed_matmul_benchmark_b200 import matmul_benchmark as mb
from pytorch_distributed_matmul_benchmark_b200 import matmul_scaling_benchmark as ms


def _capture(fn):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        fn()
    return buf.getvalue()


def _normalise(text: str) -> str:
    # the one deliberate difference: the efficiency line names B200 and its peak instead of "RTX 6000 Ada"
    return re.sub(r"  - GPU Efficiency: .*\n", "  - GPU Efficiency: <peak-specific>\n", text)


def test_cli_defaults_match_reference(host_logic):
    a = mb.build_parser().parse_args([])
    assert [0, 1, a.sizes, str(hostlib.DTYPE_MAP[a.dtype]), a.iterations, a.warmup] == host_logic["cli_mb_default"]
    a = ms.build_parser().parse_args([])
    assert [0, 1, a.sizes, str(hostlib.DTYPE_MAP[a.dtype]), a.mode, a.iterations, a.warmup] == host_logic["cli_ms_default"]
    a = ms.build_parser().parse_args(["--sizes", "512", "1024", "--mode", "matrix_parallel", "--dtype", "float16",
                                      "--iterations", "7", "--warmup", "2"])
    assert [0, 1, a.sizes, str(hostlib.DTYPE_MAP[a.dtype]), a.mode, a.iterations, a.warmup] == host_logic["cli_ms_custom"]
    assert [m.value for m in ms.ScalingMode] == host_logic["modes"]
    with pytest.raises(SystemExit):
        ms.build_parser().parse_args(["--mode", "data_parallel"])
    with pytest.raises(SystemExit):
        mb.build_parser().parse_args(["--dtype", "float64"])


def test_basic_report_text_matches_reference(host_logic):
    t = host_logic["fake_time"]
    stub = lambda size, dtype, device, it, wu: (t, hostlib.calculate_tflops(size, t))  # noqa: E731
    out = _capture(lambda: mb.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, 50, 10, bench_fn=stub, device="cpu"))
    assert _normalise(out) == _normalise(host_logic["report_mb_bf16"])
    out = _capture(lambda: mb.run_benchmarks(0, 1, [4096], torch.float32, 50, 10, bench_fn=stub, device="cpu"))
    assert _normalise(out) == _normalise(host_logic["report_mb_fp32"])
    assert "% of B200 theoretical peak" in out or "B200" in out


@pytest.mark.parametrize("mode", list(ms.ScalingMode))
def test_scaling_report_text_matches_reference(host_logic, mode):
    t = host_logic["fake_time"]
    stubs = {
        ms.ScalingMode.INDEPENDENT: lambda size: (t, hostlib.calculate_tflops(size, t, 1)),
        ms.ScalingMode.BATCH_PARALLEL: lambda size: (t, hostlib.calculate_tflops(size, t, 4)),
        ms.ScalingMode.MATRIX_PARALLEL: lambda size: (t, hostlib.calculate_tflops(size, t, 1)),
    }
    out = _capture(lambda: ms.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, mode, 50, 10, bench_fns=stubs,
                                              device="cpu"))
    assert out == host_logic[f"report_ms_{mode.value}"]


def test_report_keeps_the_substrings_compare_benchmarks_greps(host_logic):
    """backup/compare_benchmarks.py:22-26 filters stdout on these substrings."""
    out = host_logic["report_ms_independent"]
    for needle in ("16384x16384", "Results for", "TFLOPS"):
        assert needle in out
    t = host_logic["fake_time"]
    mine = _capture(lambda: mb.run_benchmarks(0, 1, [16384], torch.bfloat16, 50, 10, device="cpu",
                                               bench_fn=lambda *a: (t, hostlib.calculate_tflops(16384, t))))
    for needle in ("16384x16384", "Results for", "Average time", "TFLOPS"):
        assert needle in mine


def test_per_size_error_net_prints_error_and_continues():
    """matmul_benchmark.py:143-148: an exception in one size becomes an ERROR: line; the loop goes on."""
    calls = []

    def flaky(size, dtype, device, it, wu):
        calls.append(size)
        if size == 8192:
            raise RuntimeError("boom")
        return 0.001, 1.0

    out = _capture(lambda: mb.run_benchmarks(0, 1, [4096, 8192, 16384], torch.bfloat16, 5, 1, bench_fn=flaky, device="cpu"))
    assert calls == [4096, 8192, 16384]
    assert "\n  ERROR: boom" in out and "Results for 16384x16384" in out and "Results for 8192x8192" not in out


def test_calculate_tflops_and_split_match_reference(host_logic):
    for n, t, v in host_logic["calculate_tflops_mb"]:
        assert hostlib.calculate_tflops(n, t) == pytest.approx(v, rel=1e-12)
    for n, t, k, v in host_logic["calculate_tflops_ms"]:
        assert hostlib.calculate_tflops(n, t, k) == pytest.approx(v, rel=1e-12)
    for N, ws, rank, s, e in host_logic["col_split"]:
        assert hostlib.split_columns(N, ws, rank) == (s, e)


def _dry_run(script, *args):
    """Run a launcher with python3 / torchrun replaced by an echo stub; returns what it would have executed."""
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        stub = os.path.join(d, "python3")
        open(stub, "w").write("#!/bin/bash\necho CMD: python3 \"$@\"\n")
        os.chmod(stub, 0o755)
        env = dict(os.environ, PATH=d + os.pathsep + os.environ["PATH"])
        return subprocess.run(["bash", os.path.join(ROOT, script), *args], env=env, capture_output=True, text=True, check=True).stdout


def test_launchers_keep_the_reference_positionals():
    """run_benchmark.sh: NUM_GPUS [DTYPE] (defaults 1 / bfloat16, port 29500); run_scaling_benchmark.sh:
    NUM_GPUS [MODE] [DTYPE] (2 / independent / bfloat16, port 29503) — same sweep, same banner lines."""
    out = _dry_run("run_benchmark.sh")
    assert "Starting distributed matrix multiplication benchmark with 1 GPU(s)\nData type: bfloat16\n" in out
    assert "Running in single GPU mode..." in out
    assert "CMD: python3 matmul_benchmark.py --sizes 4096 8192 16384 --iterations 50 --warmup 10 --dtype bfloat16" in out
    out = _dry_run("run_benchmark.sh", "4", "float16", "--check")
    assert "Running in distributed mode with 4 GPUs..." in out
    assert "torch.distributed.run" in out and "--nproc_per_node=4" in out and "--master_port=29500" in out
    assert out.strip().endswith("--dtype float16 --check")
    out = _dry_run("run_scaling_benchmark.sh")
    assert "Matrix Multiplication Scaling Benchmark\n  GPUs: 2\n  Mode: independent (independent, batch_parallel, matrix_parallel)\n  Data type: bfloat16\n" in out
    assert "--nproc_per_node=2" in out and "--master_port=29503" in out
    assert "matmul_scaling_benchmark.py --sizes 4096 8192 16384 --iterations 50 --warmup 10 --mode independent --dtype bfloat16" in out
    out = _dry_run("run_scaling_benchmark.sh", "1", "matrix_parallel", "float32")
    assert "CMD: python3 matmul_scaling_benchmark.py" in out and "--mode matrix_parallel --dtype float32" in out
    for script in ("run_benchmark.sh", "run_scaling_benchmark.sh", "run_overlap_benchmark.sh", "run_common.sh"):
        subprocess.check_call(["bash", "-n", os.path.join(ROOT, script)])


def test_cpu_device_is_refused_by_the_product_benchmark():
    with pytest.raises(RuntimeError, match="no CPU path"):
        mb.benchmark_matmul(64, torch.float32, "cpu", 1, 0)


# ---- backup/matmul_overlap_benchmark.py: its three modes are first-class here ----------------------------
from pytorch_distributed_matmul_benchmark_b200 import matmul_overlap_benchmark as ov  # noqa: E402


def test_overlap_cli_defaults_match_reference(host_logic):
    a = ov.build_parser().parse_args([])
    assert [0, 1, a.sizes, str(hostlib.DTYPE_MAP[a.dtype]), a.mode, a.iterations, a.warmup] == host_logic["cli_ov_default"]
    assert [m.value for m in ov.BenchmarkMode] == host_logic["modes_ov"]
    assert ov.PIPELINE_DEPTH == 3  # backup/matmul_overlap_benchmark.py:184
    with pytest.raises(SystemExit):
        ov.build_parser().parse_args(["--mode", "independent"])


@pytest.mark.parametrize("mode", list(ov.BenchmarkMode))
def test_overlap_report_text_matches_reference(host_logic, mode):
    t = host_logic["fake_time"]
    stubs = {m: (lambda size: (t, hostlib.calculate_tflops(size, t), 0.0)) for m in ov.BenchmarkMode}
    out = _capture(lambda: ov.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, mode, 50, 10, bench_fns=stubs, device="cpu"))
    assert out == host_logic[f"report_ov_{mode.value}"]


def test_overlap_launcher_keeps_the_reference_positionals():
    out = _dry_run("run_overlap_benchmark.sh")
    assert "Starting overlapped communication/computation benchmark\n  GPUs: 2\n  Mode: overlap (no_overlap, overlap, pipeline)\n  Data type: bfloat16\n" in out
    assert "--master_port=29502" in out
    assert "matmul_overlap_benchmark.py --sizes 4096 8192 16384 --iterations 50 --warmup 10 --mode overlap --dtype bfloat16" in out


def test_banner_runs_without_a_gpu():
    out = _capture(hostlib.print_banner)
    assert out.startswith("PyTorch version: ") and "CUDA available: " in out


# ---- backup/matmul_distributed_benchmark.py: independent / data_parallel / model_parallel -------
from pytorch_distributed_matmul_benchmark_b200 import matmul_distributed_benchmark as md  # noqa: E402


def test_distributed_cli_defaults_match_reference(host_logic):
    a = md.build_parser().parse_args([])
    assert [0, 1, a.sizes, str(hostlib.DTYPE_MAP[a.dtype]), a.mode, a.iterations, a.warmup] == host_logic["cli_md_default"]
    assert [m.value for m in md.BenchmarkMode] == host_logic["modes_md"]
    with pytest.raises(SystemExit):
        md.build_parser().parse_args(["--mode", "batch_parallel"])


@pytest.mark.parametrize("ws", [1, 2])
@pytest.mark.parametrize("mode", list(md.BenchmarkMode))
def test_distributed_report_text_matches_reference(host_logic, mode, ws):
    t, c = host_logic["fake_time"], host_logic["fake_comm"]
    stubs = {
        md.BenchmarkMode.INDEPENDENT: lambda size: (t, hostlib.calculate_tflops(size, t), 0.0),
        md.BenchmarkMode.DATA_PARALLEL: lambda size: (t + c, hostlib.calculate_tflops(size, t), c),
        md.BenchmarkMode.MODEL_PARALLEL: lambda size: (t + c, hostlib.calculate_tflops(size, t + c), c),
    }
    out = _capture(lambda: md.run_benchmarks(0, ws, [4096, 16384], torch.bfloat16, mode, 50, 10, bench_fns=stubs, device="cpu"))
    assert out == host_logic[f"report_md_{mode.value}_ws{ws}"]


def test_distributed_launcher_keeps_the_reference_positionals():
    out = _dry_run("run_distributed_benchmark.sh")
    assert "Starting distributed matrix multiplication benchmark\n  GPUs: 2\n  Mode: data_parallel\n  Data type: bfloat16\n" in out
    assert "--master_port=29501" in out
    assert "matmul_distributed_benchmark.py --sizes 4096 8192 16384 --iterations 50 --warmup 10 --mode data_parallel --dtype bfloat16" in out
    out = _dry_run("run_distributed_benchmark.sh", "1", "model_parallel", "float16")
    assert "CMD: python3 matmul_distributed_benchmark.py" in out and "--mode model_parallel --dtype float16" in out


def test_model_parallel_slab_storage():
    """Unequal slabs (N % ws != 0, the reference's :123-125 split) travel through equal-pitch storage."""
    for N, ws in ((200, 3), (33, 2), (4096, 8), (10, 4)):
        pitch = hostlib.slab_pitch(N, ws)
        widths = [e - s for s, e in (hostlib.split_columns(N, ws, r) for r in range(ws))]
        assert pitch % 8 == 0 and pitch >= max(widths) and sum(widths) == N
        slabs = torch.zeros(ws, N, pitch)
        full = torch.arange(N * N, dtype=torch.float32).view(N, N)
        for r in range(ws):
            s, e = hostlib.split_columns(N, ws, r)
            slabs[r, :, :e - s] = full[:, s:e]
        assert torch.equal(hostlib.assemble_slabs(slabs, N, ws), full)


# ---- backup/compare_benchmarks.py: the four-run comparison --------------------------------------------------------
def test_compare_benchmarks_filters_the_reference_reports(host_logic):
    """The block filter of backup/compare_benchmarks.py:22-26, applied to report text the reference itself printed."""
    import compare_benchmarks as cb
    blk = cb.block_for_size(host_logic["report_md_data_parallel_ws2"], 16384)
    assert blk[0] == "Results for 16384x16384:" and any("Communication overhead" in l for l in blk)
    assert cb.headline(blk) == (7.0, 1599.29)  # "Total time per operation: 7.000 ms", "Effective TFLOPS: 1599.29"
    blk = cb.block_for_size(host_logic["report_mb_bf16"], 16384)
    assert cb.headline(blk) == (5.5, 1599.29) and not any("4096" in l for l in blk)
    blk = cb.block_for_size(host_logic["report_ov_overlap"], 4096)
    assert blk and cb.headline(blk)[0] == 5.5
    assert cb.block_for_size(host_logic["report_ov_overlap"], 8192) == []
    assert [r[1:] for r in cb.RUNS] == [("run_benchmark.sh", None), ("run_distributed_benchmark.sh", "data_parallel"),
                                       ("run_overlap_benchmark.sh", "no_overlap"), ("run_overlap_benchmark.sh", "overlap")]


def test_compare_benchmarks_end_to_end_against_a_stub(host_logic, tmp_path):
    """All four launchers run (python3 replaced by a stub that prints a recorded reference report); the summary table
    carries the parsed time and TFLOPS of each."""
    import sys
    report = tmp_path / "report.txt"
    report.write_text(host_logic["report_md_data_parallel_ws2"])
    stub = tmp_path / "python3"
    stub.write_text(f"#!/bin/bash\ncat {report}\n")
    stub.chmod(0o755)
    env = dict(os.environ, PATH=str(tmp_path) + os.pathsep + os.environ["PATH"])
    p = subprocess.run([sys.executable, os.path.join(ROOT, "compare_benchmarks.py"), "--gpus", "2"], env=env, capture_output=True,
                       text=True, check=True)
    assert p.stdout.count("Results for 16384x16384:") == 4 and "### TEST 4: Overlap Benchmark - With Overlap" in p.stdout
    assert p.stdout.count("    7.000 ms    1599.29 TFLOPS") == 4
    dry = subprocess.run([sys.executable, os.path.join(ROOT, "compare_benchmarks.py"), "--dry-run", "--check"], capture_output=True,
                         text=True, check=True).stdout
    assert "run_distributed_benchmark.sh 2 data_parallel bfloat16 --check" in dry and "run_benchmark.sh 2 bfloat16 --check" in dry
