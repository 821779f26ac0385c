"""Generate the golden fixtures in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (it imports /root/reference, which does not
exist on the GPU box):

    python tests/golden/make_golden.py

What is recorded
  matmul_*.npz      inputs drawn with the reference's own recipes
                    (matmul_scaling_benchmark.py:73-75 independent, :111-116
                    batch_parallel, :176-183 matrix_parallel) and the output of the
                    reference's own operator call on them — torch.matmul / torch.bmm
                    on CPU, the branch matmul_benchmark.benchmark_matmul takes for
                    device='cpu' (:72).  16-bit tensors are stored as uint16 bit
                    patterns.
  host_logic.json   calculate_tflops of both scripts on a grid of arguments, the
                    column split and local-batch arithmetic evaluated by the
                    reference's own expressions, the argparse defaults/choices of
                    both main()s, and the reference's verbatim report text for
                    fixed fake timings (captured by running the reference's
                    run_benchmarks with its benchmark_* functions stubbed out).
"""
from __future__ import annotations

import contextlib
import io
import json
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)
import matmul_benchmark as mb  # noqa: E402
import matmul_scaling_benchmark as ms  # noqa: E402
sys.path.insert(0, os.path.join(REF, "backup"))
import matmul_overlap_benchmark as ov  # noqa: E402
import matmul_distributed_benchmark as md  # noqa: E402

DT = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}


def to_np(t: torch.Tensor) -> np.ndarray:
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    return t.contiguous().numpy()


def save(name: str, **tensors):
    np.savez_compressed(os.path.join(HERE, name), **{k: to_np(v) if isinstance(v, torch.Tensor) else v for k, v in tensors.items()})


def gen_matmul_fixtures():
    torch.set_num_threads(1)  # one deterministic summation order for the record
    for dname, dt in DT.items():
        # independent mode, ranks 0 and 1: matmul_scaling_benchmark.py:73-75, then torch.matmul (:92)
        for rank in (0, 1):
            torch.manual_seed(rank)
            A = torch.randn(160, 160, dtype=dt, device="cpu")
            B = torch.randn(160, 160, dtype=dt, device="cpu")
            save(f"matmul_independent_{dname}_r{rank}.npz", A=A, B=B, C=torch.matmul(A, B))
        # batch_parallel, ws=2 -> local batch 2: :111-116, then torch.bmm (:142)
        for rank in (0, 1):
            local_batch_size = 4 // 2
            torch.manual_seed(rank)
            A = torch.randn(local_batch_size, 96, 96, dtype=dt, device="cpu")
            B = torch.randn(local_batch_size, 96, 96, dtype=dt, device="cpu")
            save(f"bmm_batch_parallel_{dname}_ws2_r{rank}.npz", A=A, B=B, C=torch.bmm(A, B))
        # matrix_parallel, ws=3 on N=200 (remainder slab on the last rank): :176-183, matmul :211
        torch.manual_seed(0)
        N, ws = 136, 3
        A = torch.randn(N, N, dtype=dt, device="cpu")
        Bfull = torch.randn(N, N, dtype=dt, device="cpu")
        Cfull = torch.matmul(A, Bfull)
        slabs = {}
        for rank in range(ws):
            cols_per_gpu = N // ws
            start_col = rank * cols_per_gpu
            end_col = start_col + cols_per_gpu if rank < ws - 1 else N
            B_local = Bfull[:, start_col:end_col].contiguous()
            slabs[f"C_local_r{rank}"] = torch.matmul(A, B_local)
            slabs[f"cols_r{rank}"] = np.array([start_col, end_col])
        save(f"matmul_matrix_parallel_{dname}_ws3.npz", A=A, B=Bfull, C=Cfull, **slabs)
        # a long-K, ragged case (accumulation length matters for rounding)
        torch.manual_seed(7)
        A = torch.randn(40, 1536, dtype=dt, device="cpu")
        B = torch.randn(1536, 72, dtype=dt, device="cpu")
        save(f"matmul_longk_{dname}.npz", A=A, B=B, C=torch.matmul(A, B))


class _FakeCuda(types.SimpleNamespace):
    OutOfMemoryError = torch.cuda.OutOfMemoryError


class _TorchShim:
    """Stands in for the `torch` name inside a reference module so that its report code runs without a GPU."""

    def __init__(self, gpu_name: str):
        self.cuda = _FakeCuda(device_count=lambda: 1, get_device_name=lambda i=0: gpu_name, empty_cache=lambda: None,
                              is_available=lambda: False)

    def tensor(self, data, device=None, **kw):
        return torch.tensor(data, **kw)

    def __getattr__(self, name):
        return getattr(torch, name)


def capture_report(mod, call):
    shim = _TorchShim("NVIDIA B200")
    real = mod.torch
    mod.torch = shim
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            call()
    finally:
        mod.torch = real
    return buf.getvalue()


def gen_host_logic():
    out = {}
    grid = [(4096, 0.0001), (8192, 0.000673), (16384, 0.0055), (16384, 6.137e-3), (10, 1.0)]
    out["calculate_tflops_mb"] = [[n, t, mb.calculate_tflops(n, t)] for n, t in grid]
    out["calculate_tflops_ms"] = [[n, t, k, ms.calculate_tflops(n, t, num_ops=k)] for n, t in grid for k in (1, 2, 4)]
    # sharding arithmetic, evaluated with the reference's own expressions
    out["local_batch"] = [[4, ws, 4 // ws] for ws in (1, 2, 3, 4, 8)]  # matmul_scaling_benchmark.py:111
    splits = []
    for N in (4096, 16384, 200, 10):
        for ws in (1, 2, 3, 4, 8):
            for rank in range(ws):
                cols_per_gpu = N // ws
                start_col = rank * cols_per_gpu
                end_col = start_col + cols_per_gpu if rank < ws - 1 else N  # :179-181
                splits.append([N, ws, rank, start_col, end_col])
    out["col_split"] = splits

    # CLI defaults: run each main() with everything below argparse stubbed, record what it parsed
    def cli(mod, argv):
        seen = {}
        saved = (mod.run_benchmarks, mod.setup_distributed, sys.argv, torch.cuda.device_count)
        mod.run_benchmarks = lambda *a, **k: seen.setdefault("args", a)
        mod.setup_distributed = lambda: (0, 1)
        torch.cuda.device_count = lambda: 1
        sys.argv = ["prog"] + argv
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                mod.main()
        finally:
            mod.run_benchmarks, mod.setup_distributed, sys.argv, torch.cuda.device_count = saved
        return [str(a.value) if hasattr(a, "value") else (str(a) if isinstance(a, torch.dtype) else a) for a in seen["args"]]

    out["cli_mb_default"] = cli(mb, [])  # rank, world, sizes, dtype, iters, warmup
    out["cli_ms_default"] = cli(ms, [])  # rank, world, sizes, dtype, mode, iters, warmup
    out["cli_ms_custom"] = cli(ms, ["--sizes", "512", "1024", "--mode", "matrix_parallel", "--dtype", "float16",
                                    "--iterations", "7", "--warmup", "2"])
    out["modes"] = [m.value for m in ms.ScalingMode]

    # Report text for fixed fake timings
    fake_time = 0.0055
    saved = mb.benchmark_matmul
    mb.benchmark_matmul = lambda size, dtype, device, it, wu: (fake_time, mb.calculate_tflops(size, fake_time))
    try:
        out["report_mb_bf16"] = capture_report(mb, lambda: mb.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, 50, 10))
        out["report_mb_fp32"] = capture_report(mb, lambda: mb.run_benchmarks(0, 1, [4096], torch.float32, 50, 10))
    finally:
        mb.benchmark_matmul = saved

    saved = (ms.benchmark_independent, ms.benchmark_batch_parallel, ms.benchmark_matrix_parallel)
    ms.benchmark_independent = lambda size, dtype, device, rank, it, wu: (fake_time, ms.calculate_tflops(size, fake_time, 1))
    ms.benchmark_batch_parallel = lambda size, bs, dtype, device, rank, ws, it, wu: (
        fake_time, ms.calculate_tflops(size, fake_time, bs // ws))
    ms.benchmark_matrix_parallel = lambda size, dtype, device, rank, ws, it, wu: (
        fake_time, ms.calculate_tflops(size, fake_time, 1) / ws)
    try:
        for mode in ms.ScalingMode:
            out[f"report_ms_{mode.value}"] = capture_report(
                ms, lambda mode=mode: ms.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, mode, 50, 10))
    finally:
        ms.benchmark_independent, ms.benchmark_batch_parallel, ms.benchmark_matrix_parallel = saved
    out["fake_time"] = fake_time

    # backup/matmul_overlap_benchmark.py: CLI and report text (its three modes are first-class here)
    out["cli_ov_default"] = cli(ov, [])  # rank, world, sizes, dtype, mode, iters, warmup
    out["modes_ov"] = [m.value for m in ov.BenchmarkMode]
    saved = (ov.benchmark_no_overlap, ov.benchmark_overlap, ov.benchmark_pipeline)
    stub = lambda size, dtype, device, rank, it, wu: (fake_time, ov.calculate_tflops(size, fake_time), 0.0)  # noqa: E731
    ov.benchmark_no_overlap = ov.benchmark_overlap = ov.benchmark_pipeline = stub
    try:
        for mode in ov.BenchmarkMode:
            out[f"report_ov_{mode.value}"] = capture_report(
                ov, lambda mode=mode: ov.run_benchmarks(0, 1, [4096, 16384], torch.bfloat16, mode, 50, 10))
    finally:
        ov.benchmark_no_overlap, ov.benchmark_overlap, ov.benchmark_pipeline = saved

    # backup/matmul_distributed_benchmark.py (the earlier draft: independent / data_parallel / model_parallel)
    out["cli_md_default"] = cli(md, [])  # rank, world, sizes, dtype, mode, iters, warmup
    out["modes_md"] = [m.value for m in md.BenchmarkMode]
    fake_comm = 0.0015
    out["fake_comm"] = fake_comm
    saved = (md.benchmark_independent, md.benchmark_data_parallel, md.benchmark_model_parallel)
    md.benchmark_independent = lambda size, dtype, device, it, wu: (fake_time, md.calculate_tflops(size, fake_time), 0.0)
    md.benchmark_data_parallel = lambda size, dtype, device, rank, it, wu: (
        fake_time + fake_comm, md.calculate_tflops(size, fake_time), fake_comm)
    md.benchmark_model_parallel = lambda size, dtype, device, rank, ws, it, wu: (
        fake_time + fake_comm, md.calculate_tflops(size, fake_time + fake_comm), fake_comm)
    try:
        for mode in md.BenchmarkMode:
            for ws in (1, 2):  # no process group here, so ws=2 exercises the report branches only (:251-257)
                out[f"report_md_{mode.value}_ws{ws}"] = capture_report(
                    md, lambda mode=mode, ws=ws: md.run_benchmarks(0, ws, [4096, 16384], torch.bfloat16, mode, 50, 10))
    finally:
        md.benchmark_independent, md.benchmark_data_parallel, md.benchmark_model_parallel = saved

    # The reference's own CPU path, executed here once as a record (config 1 of BASELINE.json, reduced iterations)
    with contextlib.redirect_stdout(io.StringIO()):
        avg, tf = mb.benchmark_matmul(1024, torch.float32, "cpu", 3, 1)
    out["cpu_path_smoke"] = {"size": 1024, "dtype": "float32", "avg_time_s": avg, "tflops": tf,
                             "threads": torch.get_num_threads(), "note": "timing is machine dependent; only tflops == 2N^3/t/1e12 is asserted"}
    with open(os.path.join(HERE, "host_logic.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    if "--host-only" not in sys.argv:  # the matrix fixtures are deterministic; skip them when only the report goldens change
        gen_matmul_fixtures()
    torch.set_num_threads(os.cpu_count() or 1)
    gen_host_logic()
    total = sum(os.path.getsize(os.path.join(HERE, f)) for f in os.listdir(HERE))
    print(f"wrote fixtures to {HERE}: {total/1e6:.2f} MB")
