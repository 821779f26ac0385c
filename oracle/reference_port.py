"""Python restatement of the reference's CPU timing path and input recipes.

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.

The reference's CPU implementation of the hot path is `torch.matmul` on CPU
tensors, timed by the `else` branch of `benchmark_matmul`
(matmul_benchmark.py:39-79, CPU branch :69-74).  /root/reference does not exist
on the GPU box, so that branch is restated here; torch itself (the reference's
only dependency, unpinned) is present on the box and is what the reference
would run there.  bench.py times this as `cpu_baseline` / `--impl reference`
("kind": "port").
"""
from __future__ import annotations

import time
from typing import Tuple

import torch


def calculate_tflops(matrix_size: int, time_seconds: float, num_ops: int = 1) -> float:
    """matmul_benchmark.py:34-37; matmul_scaling_benchmark.py:63-67."""
    flops = 2.0 * (matrix_size ** 3) * num_ops
    return (flops / time_seconds) / 1e12


def benchmark_matmul_cpu(matrix_size: int, dtype: torch.dtype, num_iterations: int = 50,
                         warmup_iterations: int = 10, seed: int | None = None) -> Tuple[float, float]:
    """matmul_benchmark.py:39-79 with device='cpu' (the only reference function with a CPU branch).

    `seed` is additive (the reference draws unseeded, :41-42); bench.py passes one
    so that repeated runs time the same data.
    """
    if seed is not None:
        torch.manual_seed(seed)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device="cpu")
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device="cpu")
    for _ in range(warmup_iterations):
        C = torch.matmul(A, B)  # noqa: F841
    start_time = time.perf_counter()
    for _ in range(num_iterations):
        C = torch.matmul(A, B)  # noqa: F841
    end_time = time.perf_counter()
    avg_time = (end_time - start_time) / num_iterations
    return avg_time, calculate_tflops(matrix_size, avg_time)


def independent_inputs(rank: int, matrix_size: int, dtype: torch.dtype, device: str = "cpu"):
    """Input recipe of benchmark_independent: matmul_scaling_benchmark.py:73-75."""
    torch.manual_seed(rank)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    return A, B


def batch_parallel_inputs(rank: int, world_size: int, matrix_size: int, dtype: torch.dtype, batch_size: int = 4,
                          device: str = "cpu"):
    """Input recipe of benchmark_batch_parallel: matmul_scaling_benchmark.py:111-116."""
    local_batch_size = batch_size // world_size
    torch.manual_seed(rank)
    A = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
    return A, B


def matrix_parallel_inputs(rank: int, world_size: int, matrix_size: int, dtype: torch.dtype, device: str = "cpu",
                           seed: int = 0):
    """Column-split recipe of benchmark_matrix_parallel (matmul_scaling_benchmark.py:176-183), made global.

    The reference draws A and B_local unseeded per rank, so no global B (or C) exists
    there.  For checking, every rank seeds identically, draws A and the FULL B, and
    takes its own column slab with the reference's split rule (:179-181).
    """
    torch.manual_seed(seed)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    cols_per_gpu = matrix_size // world_size
    start_col = rank * cols_per_gpu
    end_col = start_col + cols_per_gpu if rank < world_size - 1 else matrix_size
    return A, B, B[:, start_col:end_col].contiguous(), (start_col, end_col)
