"""Drop-in for the reference's backup/matmul_overlap_benchmark.py: no_overlap / overlap / pipeline.

Each iteration is a full N x N GEMM followed by an all_reduce(SUM) of its output (a data-parallel step).
Same CLI, timing method (one CUDA-event pair around the whole loop, :121-160) and report text.  What differs
is underneath, and the races are gone:

* the GEMM is libb2mm.so's tcgen05 kernel;
* `no_overlap` reduces with the library's NCCL all_reduce, serialised exactly like :51-69;
* `overlap` / `pipeline` reduce with the library's own self-synchronising all_reduce — inside the NVSwitch
  (b2mm_all_reduce_nvls_sync: multimem.ld_reduce / multimem.st on multicast buffers) where the hardware has it,
  else pull / add / push over peer mappings (b2mm_all_reduce_p2p_sync) — whose
  CTAs are co-resident with the GEMM's: the reduction of C_i runs on a second stream while GEMM i+1 runs on
  the whole device.  Every edge is an event: the reduction waits for its GEMM (the reference never makes
  comm_stream wait for compute_stream, :131-137), and a buffer is rewritten only after its reduction has
  finished (the reference's pipeline recomputes into the buffer it has just started reducing, :233-246).
"""
from __future__ import annotations

import argparse
import enum
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import hostlib
from .hostlib import cleanup_distributed, setup_distributed  # noqa: F401  (reference names)

PIPELINE_DEPTH = 3  # backup/matmul_overlap_benchmark.py:184


class BenchmarkMode(enum.Enum):
    NO_OVERLAP = "no_overlap"  # compute, wait, communicate, wait
    OVERLAP = "overlap"  # double-buffered: reduce C_i while computing C_{i+1}
    PIPELINE = "pipeline"  # depth-3 ring of buffers


def calculate_tflops(matrix_size: int, time_seconds: float) -> float:
    return hostlib.calculate_tflops(matrix_size, time_seconds)


def _compute_only_tflops(ops: hostlib.Ops, A, B, C, matrix_size: int) -> float:
    """TFLOPS from 10 extra compute-only GEMMs, as the reference reports (:76-88, :167-178, :265-276)."""
    start, end = ops.event(), ops.event()
    ops.synchronize()
    start.record()
    for _ in range(10):
        ops.matmul(A, B, out=C)
    end.record()
    ops.synchronize()
    return calculate_tflops(matrix_size, (start.elapsed_time(end) / 1000.0) / 10)


def benchmark_no_overlap(matrix_size: int, dtype: torch.dtype, device: str, rank: int, num_iterations: int = 50,
                         warmup_iterations: int = 10, ops: Optional[hostlib.Ops] = None,
                         check: bool = False) -> Tuple[float, float, float]:
    """Mirror of benchmark_no_overlap (:36-91)."""
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    C = torch.empty(matrix_size, matrix_size, dtype=dtype, device=device)
    comm_on = dist.is_initialized()
    for _ in range(warmup_iterations):
        ops.matmul(A, B, out=C)
        if comm_on:
            ops.all_reduce(C)
    ops.synchronize()
    start, end = ops.event(), ops.event()
    ops.synchronize()
    start.record()
    for _ in range(num_iterations):
        ops.matmul(A, B, out=C)
        ops.synchronize()
        if comm_on:
            ops.all_reduce(C)
        ops.synchronize()
    end.record()
    ops.synchronize()
    avg_time = (start.elapsed_time(end) / 1000.0) / max(num_iterations, 1)
    if check and comm_on:
        # NCCL: one rounding inside the switch (NVLS) or one per ring step
        _check_reduced_product(ops, lambda: ops.all_reduce(C), C, matrix_size, dtype, device, rank, ring=True)
    return avg_time, _compute_only_tflops(ops, A, B, C, matrix_size), 0.0


def _check_reduced_product(ops: hostlib.Ops, reduce_fn, out: torch.Tensor, matrix_size: int, dtype: torch.dtype,
                           device: str, rank: int, roundings: int = 2, ring: bool = False) -> None:
    """--check: every rank multiplies matrices seeded by its rank, reduces, and compares with the sum of all ranks'
    products recomputed locally (the reference reduces the rounded C tensors, so does the expectation).

    `roundings`: output-ulp budget of the reduction under test.  The library's own reductions add in fp32 and round
    ONCE whatever the world size; the NVSwitch (measured) truncates instead of rounding to nearest, so that one
    rounding may cost a full ulp: 2 x tolerance.  `ring`: the NCCL arm of `no_overlap`, which may round once per ring /
    tree step — each a PARTIAL sum, bounded by the ranks' terms: hostlib.reduction_pair_error is the bar there."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    want = mag = None
    for r in range(world):
        torch.manual_seed(4000 + r)
        Ar = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
        Br = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
        Cr = ops.matmul(Ar, Br)
        want = Cr.double() if want is None else want + Cr.double()
        m = hostlib.term_magnitudes([Cr])
        mag = m if mag is None else mag + m
        if r == rank:
            out.copy_(Cr)
    ops.synchronize()
    if dist.is_initialized():
        dist.barrier()
    reduce_fn()
    ops.synchronize()
    if ring:
        # NCCL rounds partial sums, which the ranks' terms bound (not the final, possibly cancelled, sum)
        err = hostlib.reduction_pair_error(out, want, mag, world)
        hostlib.report_check(rank, err, hostlib.CHECK_TOL[dtype], "of the reduced product vs float64 sum", indent="  ", device=device)
        return
    rms = want.pow(2).mean().sqrt()
    err = ((out.double() - want).abs() / (want.abs() + rms)).max().item()
    hostlib.report_check(rank, err, hostlib.CHECK_TOL[dtype] * roundings, "of the reduced product vs float64 sum", indent="  ",
                         device=device)


def _ring_benchmark(depth: int, matrix_size: int, dtype: torch.dtype, device: str, num_iterations: int,
                    warmup_iterations: int, ops: hostlib.Ops, rank: int = 0, check: bool = False) -> Tuple[float, float, float]:
    """`depth` (A, B, C) sets; GEMM i on the current stream, reduction of C_i on the comm stream, race-free."""
    As = [torch.randn(matrix_size, matrix_size, dtype=dtype, device=device) for _ in range(depth)]
    Bs = [torch.randn(matrix_size, matrix_size, dtype=dtype, device=device) for _ in range(depth)]
    comm_on = dist.is_initialized() and ops.all_reduce_sync is not None
    if comm_on:
        Cs = [(ops.reduce_buffer_like or ops.shared_like)(As[0]) for _ in range(depth)]  # NVLS or peer-visible outputs
        outs = [c.local for c in Cs]
    else:
        Cs, outs = [None] * depth, [torch.empty_like(As[0]) for _ in range(depth)]
    main = torch.cuda.current_stream()
    comm_stream = torch.cuda.Stream(priority=-1)
    computed = [torch.cuda.Event() for _ in range(depth)]
    reduced = [torch.cuda.Event() for _ in range(depth)]

    def step(i: int) -> None:
        k = i % depth
        if comm_on and i >= depth:
            main.wait_event(reduced[k])  # C_k is rewritten only after its previous reduction has finished
        ops.matmul(As[k], Bs[k], out=outs[k])
        if comm_on:
            computed[k].record(main)
            comm_stream.wait_event(computed[k])  # the edge the reference forgets
            ops.all_reduce_sync(Cs[k], comm_stream)
            reduced[k].record(comm_stream)

    for i in range(warmup_iterations):
        step(i)
    main.wait_stream(comm_stream)
    ops.synchronize()
    start, end = ops.event(), ops.event()
    start.record()
    for i in range(num_iterations):
        step(i)
    main.wait_stream(comm_stream)
    end.record()
    ops.synchronize()
    avg_time = (start.elapsed_time(end) / 1000.0) / max(num_iterations, 1)
    if check and comm_on:
        _check_reduced_product(ops, lambda: ops.all_reduce_sync(Cs[0]), outs[0], matrix_size, dtype, device, rank)
    return avg_time, _compute_only_tflops(ops, As[0], Bs[0], outs[0], matrix_size), 0.0


def benchmark_overlap(matrix_size: int, dtype: torch.dtype, device: str, rank: int, num_iterations: int = 50,
                      warmup_iterations: int = 10, ops: Optional[hostlib.Ops] = None,
                      check: bool = False) -> Tuple[float, float, float]:
    """The double-buffered overlap of benchmark_overlap (:93-180), with every dependency as an event."""
    return _ring_benchmark(2, matrix_size, dtype, device, num_iterations, warmup_iterations, ops, rank, check)


def benchmark_pipeline(matrix_size: int, dtype: torch.dtype, device: str, rank: int, num_iterations: int = 50,
                       warmup_iterations: int = 10, pipeline_depth: int = PIPELINE_DEPTH,
                       ops: Optional[hostlib.Ops] = None, check: bool = False) -> Tuple[float, float, float]:
    """The depth-3 pipeline of benchmark_pipeline (:182-278), with every dependency as an event."""
    return _ring_benchmark(pipeline_depth, matrix_size, dtype, device, num_iterations, warmup_iterations, ops, rank, check)


def run_benchmarks(rank: int, world_size: int, matrix_sizes: List[int], dtype: torch.dtype, mode: BenchmarkMode,
                   num_iterations: int, warmup_iterations: int, ops: Optional[hostlib.Ops] = None,
                   device: Optional[str] = None, bench_fns: Optional[dict] = None, check: bool = False) -> None:
    """Size loop, aggregation and report of run_benchmarks (:280-356)."""
    device = device or hostlib.rank_device(rank)
    if rank == 0:
        bar = "=" * 70
        print(f"\n{bar}\nOverlapped Communication/Computation Benchmark\n{bar}")
        print("Configuration:")
        print(f"  - Mode: {mode.value}")
        print(f"  - Number of GPUs: {world_size}")
        print(f"  - Data type: {dtype}")
        print(f"  - Iterations per test: {num_iterations}")
        print(f"  - Warmup iterations: {warmup_iterations}")
        print(f"{bar}\n")
    fns = {BenchmarkMode.NO_OVERLAP: benchmark_no_overlap, BenchmarkMode.OVERLAP: benchmark_overlap,
           BenchmarkMode.PIPELINE: benchmark_pipeline}
    for size in matrix_sizes:
        if rank == 0:
            gib = size * size * hostlib.bytes_per_element(dtype) / (1024 ** 3)
            print(f"\nBenchmarking {size}x{size} matrix multiplication:")
            print(f"  - Memory per matrix: {gib:.2f} GB ({hostlib.dtype_name(dtype)})")
            print(f"  - Mode: {mode.value}")
            print("  - Running warmup and benchmark...")
        try:
            if bench_fns is not None:
                avg_time, tflops, _ = bench_fns[mode](size)
            else:
                avg_time, tflops, _ = fns[mode](size, dtype, device, rank, num_iterations, warmup_iterations, ops=ops,
                                                check=check)
            avg_time_all = hostlib.metric_all_reduce(avg_time, device, "avg")
            hostlib.metric_all_reduce(tflops, device, "avg")
            if rank == 0:
                print(f"\nResults for {size}x{size}:")
                print(f"  - Average time per operation: {avg_time_all*1000:.3f} ms")
                flops_per_operation = 2.0 * size ** 3
                actual_tflops = (flops_per_operation / avg_time_all) / 1e12
                print(f"  - Actual TFLOPS: {actual_tflops:.2f} (FLOPs/Time)")
                if world_size > 1:
                    print("  - Note: In data parallel, each GPU does full matrix multiply")
                    print(f"  - System is achieving {actual_tflops:.2f} TFLOPS effectively")
                print(f"  - Required FLOPs per operation: {flops_per_operation / 1e12:.2f} TFLOPs")
        except torch.cuda.OutOfMemoryError:
            if rank == 0:
                print(f"\n  ERROR: Out of memory for {size}x{size} matrices")
        except Exception as e:
            if rank == 0:
                print(f"\n  ERROR: {str(e)}")
        if ops is not None:
            ops.release_shared()  # peers unmap this size's shared buffers before empty_cache() frees them
        if torch.cuda.is_available():
            torch.cuda.empty_cache()
        if dist.is_initialized():
            dist.barrier()


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Overlapped Communication/Computation Benchmark")
    hostlib.add_common_args(parser)
    parser.add_argument("--mode", type=str, default="overlap", choices=["no_overlap", "overlap", "pipeline"],
                        help="Benchmark mode (default: overlap)")
    return parser


def main(argv: Optional[List[str]] = None) -> None:
    args = build_parser().parse_args(argv)
    dtype = hostlib.DTYPE_MAP[args.dtype]
    mode = BenchmarkMode(args.mode)
    rank, world_size = setup_distributed()
    if rank == 0:
        hostlib.print_banner()
        if torch.cuda.is_available() and torch.cuda.device_count() > 1:
            print("\nChecking GPU interconnect...")
            print("  Multi-GPU system detected. Use nvidia-smi topo -m for interconnect details.")
    device = hostlib.rank_device(rank)
    ops = hostlib.device_ops(rank, world_size, device)
    try:
        run_benchmarks(rank, world_size, args.sizes, dtype, mode, args.iterations, args.warmup, ops=ops, device=device,
                       check=args.check)
    finally:
        ops.destroy()
        cleanup_distributed()
    if rank == 0:
        bar = "=" * 70
        print(f"\n{bar}\nBenchmark completed!\n{bar}\n")


if __name__ == "__main__":
    main()
