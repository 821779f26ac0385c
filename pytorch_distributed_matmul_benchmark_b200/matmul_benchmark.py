"""Drop-in for the reference's matmul_benchmark.py: one N x N GEMM per GPU, timed back to back.

Same CLI, same report text, same timing method (one CUDA-event pair around all
iterations, matmul_benchmark.py:55-67); the GEMM is libb2mm.so's tcgen05 kernel
instead of torch.matmul -> cuBLAS.  There is no CPU branch here: the reference's
CPU path is restated (as the timed baseline) in oracle/reference_port.py.
"""
from __future__ import annotations

import argparse
import json
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist

from . import hostlib
from .hostlib import calculate_tflops, cleanup_distributed, setup_distributed  # noqa: F401  (reference names)


def benchmark_matmul(matrix_size: int, dtype: torch.dtype, device: str, num_iterations: int = 50,
                     warmup_iterations: int = 10, ops: Optional[hostlib.Ops] = None,
                     check: bool = False) -> Tuple[float, float]:
    """Mirror of matmul_benchmark.py:39-79 (CUDA branch).  Returns (avg_time_s, tflops).

    Inputs: torch.randn, row-major, contiguous (:41-42).  The reference draws them
    unseeded; so does this.  C is allocated per call, as torch.matmul does.
    """
    if not device.startswith("cuda"):
        raise RuntimeError("b2mm has no CPU path; the reference's CPU branch is restated in oracle/reference_port.py")
    ops = ops or hostlib.device_ops(0, 1, device)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)

    print(f"  Performing {warmup_iterations} warmup iterations...")
    C = None
    for _ in range(warmup_iterations):
        C = ops.matmul(A, B)
    ops.synchronize()

    print(f"  Running {num_iterations} benchmark iterations...")
    start_event, end_event = ops.event(), ops.event()
    ops.synchronize()
    start_event.record()
    for _ in range(num_iterations):
        C = ops.matmul(A, B)
    end_event.record()
    ops.synchronize()

    total_time = start_event.elapsed_time(end_event) / 1000.0
    avg_time = total_time / num_iterations
    if check and C is not None:
        hostlib.report_check(dist.get_rank() if dist.is_initialized() else 0, hostlib.parity_error(C, A, B),
                             hostlib.CHECK_TOL[dtype], "vs float64 reference", indent="  ", device=device)
    return avg_time, calculate_tflops(matrix_size, avg_time)


def efficiency_line(tflops: float, dtype: torch.dtype) -> str:
    """The reference's last result line (matmul_benchmark.py:130-141), against B200 dense peak."""
    peak = hostlib.B200_PEAK_TFLOPS[dtype]
    return f"  - GPU Efficiency: {(tflops / peak) * 100:.1f}% of {hostlib.B200_PEAK_NOTE[dtype]} theoretical peak"


def run_benchmarks(rank: int, world_size: int, matrix_sizes: List[int], dtype: torch.dtype, num_iterations: int,
                   warmup_iterations: int, bench_fn: Optional[Callable] = None, device: Optional[str] = None,
                   check: bool = False, results: Optional[list] = None) -> None:
    """Size loop, cross-rank aggregation and report of matmul_benchmark.py:81-153.

    `bench_fn(size, dtype, device, iters, warmup) -> (avg_s, tflops)` defaults to
    benchmark_matmul; tests inject a stub to pin the report text.
    """
    device = device or hostlib.rank_device(rank)
    ops = None
    if bench_fn is None:
        ops = hostlib.device_ops(rank, 1, device)  # no payload collectives in this script
        bench_fn = lambda s, d, dev, it, wu: benchmark_matmul(s, d, dev, it, wu, ops=ops, check=check)  # noqa: E731

    if rank == 0:
        bar = "=" * 60
        print(f"\n{bar}\nMatrix Multiplication Benchmark\n{bar}")
        print("Configuration:")
        print(f"  - Number of GPUs: {world_size}")
        print(f"  - Data type: {dtype}")
        print("  - Device: CUDA")
        print(f"  - Iterations per test: {num_iterations}")
        print(f"  - Warmup iterations: {warmup_iterations}")
        print(f"{bar}\n")

    for size in matrix_sizes:
        if rank == 0:
            gib = size * size * hostlib.bytes_per_element(dtype) / (1024 ** 3)
            print(f"\nBenchmarking {size}x{size} matrix multiplication:")
            print(f"  - Memory per matrix: {gib:.2f} GB ({hostlib.dtype_name(dtype)})")
            print(f"  - Total memory for A, B, C: {3 * gib:.2f} GB")
        try:
            avg_time, tflops = bench_fn(size, dtype, device, num_iterations, warmup_iterations)
            total_tflops = hostlib.metric_all_reduce(tflops, device, "sum")
            avg_time_global = hostlib.metric_all_reduce(avg_time, device, "avg")
            if rank == 0:
                print(f"\nResults for {size}x{size}:")
                print(f"  - Average time per multiplication: {avg_time_global * 1000:.3f} ms")
                print(f"  - TFLOPS per GPU: {tflops:.2f}")
                print(f"  - Total TFLOPS (all GPUs): {total_tflops:.2f}")
                print(f"  - Required FLOPs per operation: {2.0 * size ** 3 / 1e12:.2f} TFLOPs")
                print(efficiency_line(tflops, dtype))
                if results is not None:
                    results.append({"size": size, "avg_time_ms": avg_time_global * 1000, "tflops_per_gpu": tflops,
                                    "total_tflops": total_tflops, "n_gpus": world_size})
        except torch.cuda.OutOfMemoryError:
            if rank == 0:
                print(f"\n  ERROR: Out of memory for {size}x{size} matrices")
        except Exception as e:  # the reference's per-size net: report and go on to the next size
            if rank == 0:
                print(f"\n  ERROR: {str(e)}")
        if torch.cuda.is_available():
            from . import ops as b2
            b2.release_workspaces()
            torch.cuda.empty_cache()
        if dist.is_initialized():
            dist.barrier()


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Distributed PyTorch Matrix Multiplication Benchmark")
    hostlib.add_common_args(parser)
    return parser


def main(argv: Optional[List[str]] = None) -> None:
    args = build_parser().parse_args(argv)
    dtype = hostlib.DTYPE_MAP[args.dtype]
    rank, world_size = setup_distributed()
    if rank == 0:
        hostlib.print_banner()
    results: list = []
    try:
        run_benchmarks(rank, world_size, args.sizes, dtype, args.iterations, args.warmup, check=args.check,
                       results=results)
    finally:
        cleanup_distributed()
    if rank == 0:
        if args.json:
            with open(args.json, "w") as f:
                json.dump({"script": "matmul_benchmark", "dtype": args.dtype, "results": results}, f, indent=1)
        bar = "=" * 60
        print(f"\n{bar}\nBenchmark completed!\n{bar}\n")


if __name__ == "__main__":
    main()
