"""Drop-in for the reference's matmul_scaling_benchmark.py: independent / batch_parallel / matrix_parallel.

Same CLI, mode names, timing method and report text.  Underneath, every
torch.matmul / torch.bmm / dist.all_reduce / dist.all_gather on the timed path
(matmul_scaling_benchmark.py:92,142,150,211,221) goes through libb2mm.so.

Additive behaviour (flags the reference does not have):
  --overlap        pipeline the collective behind compute on a second CUDA stream
                   (after backup/matmul_overlap_benchmark.py:93-180, without its races);
                   both the reference-style serial split and the end-to-end time are reported
  --panels P       row panels per GEMM in --overlap mode
  --fused          matrix_parallel: ONE kernel whose epilogue stores each C tile into every rank's gather
                   buffer over NVLink (no collective kernel).  batch_parallel: GEMM row panels on the
                   whole device, each followed by the library's own peer-memory all-reduce (pull, fp32
                   add, push; self-synchronising, co-resident with the next panel's GEMM).  No NCCL kernel
                   on either path.  Reported like --overlap
  --batch-size B   total batch for batch_parallel (reference: 4, hard-coded at :283)
  --per-gpu-batch  fixed local batch per GPU (weak scaling; the reference's 4 // ws is 0 at ws=8)
"""
from __future__ import annotations

import argparse
import enum
import json
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import hostlib
from .hostlib import calculate_tflops, cleanup_distributed, setup_distributed  # noqa: F401  (reference names)


class ScalingMode(enum.Enum):
    INDEPENDENT = "independent"  # every GPU runs its own full GEMM
    BATCH_PARALLEL = "batch_parallel"  # batch dimension sharded, outputs all-reduced
    MATRIX_PARALLEL = "matrix_parallel"  # B column-split, C slabs all-gathered


def verify_collectives(rank: int, world_size: int, device: str, ops: hostlib.Ops) -> bool:
    """Collective smoke test of matmul_scaling_benchmark.py:26-57, through the library's own communicator."""
    if not dist.is_initialized() or world_size == 1:
        return True
    try:
        probe = torch.tensor([rank + 1.0], device=device)
        ops.all_reduce(probe)
        ops.synchronize()
        expected_sum = world_size * (world_size + 1) / 2
        if abs(probe.item() - expected_sum) > 0.001:
            print(f"Rank {rank}: all_reduce failed. Expected {expected_sum}, got {probe.item()}")
            return False
        mine = torch.tensor([rank * 2.0], device=device)
        slabs = torch.zeros(world_size, 1, device=device)
        ops.all_gather_into(slabs, mine)
        ops.synchronize()
        for i in range(world_size):
            if abs(slabs[i].item() - i * 2.0) > 0.001:
                print(f"Rank {rank}: all_gather failed for rank {i}. Expected {i * 2.0}, got {slabs[i].item()}")
                return False
        dist.barrier()
        if rank == 0:
            print(f"✓ Collective operations verified successfully across {world_size} GPUs")
        return True
    except Exception as e:
        print(f"Rank {rank}: Collective verification failed with error: {e}")
        return False


def _timed(ops: hostlib.Ops, fn) -> float:
    """One synchronised CUDA-event measurement of fn(), in ms (the reference's per-phase pattern, :137-145)."""
    start, end = ops.event(), ops.event()
    ops.synchronize()
    start.record()
    fn()
    end.record()
    ops.synchronize()
    return start.elapsed_time(end)


def _timed_loop(ops: hostlib.Ops, fn, iterations: int) -> float:
    """One event pair around `iterations` back-to-back calls; returns seconds per call (:84-98)."""
    start, end = ops.event(), ops.event()
    ops.synchronize()
    start.record()
    for _ in range(iterations):
        fn()
    end.record()
    ops.synchronize()
    return (start.elapsed_time(end) / 1000.0) / max(iterations, 1)


def benchmark_independent(matrix_size: int, dtype: torch.dtype, device: str, rank: int, num_iterations: int = 50,
                          warmup_iterations: int = 10, ops: Optional[hostlib.Ops] = None,
                          check: bool = False) -> Tuple[float, float]:
    """Mirror of matmul_scaling_benchmark.py:69-104: seeded per rank, no communication in the timed region."""
    torch.manual_seed(rank)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    C = None
    for _ in range(warmup_iterations):
        C = ops.matmul(A, B)
    ops.synchronize()
    if dist.is_initialized():
        dist.barrier()

    def step():
        nonlocal C
        C = ops.matmul(A, B)

    avg_time = _timed_loop(ops, step, num_iterations)
    if check and C is not None:
        _report_check(rank, hostlib.parity_error(C, A, B), dtype, device)
    return avg_time, calculate_tflops(matrix_size, avg_time, num_ops=1)


def benchmark_batch_parallel(matrix_size: int, batch_size: int, dtype: torch.dtype, device: str, rank: int,
                             world_size: int, num_iterations: int = 50, warmup_iterations: int = 10,
                             ops: Optional[hostlib.Ops] = None, overlap: bool = False, panels: int = 4,
                             local_batch_size: Optional[int] = None, check: bool = False,
                             fused: bool = False) -> Tuple[float, float]:
    """Mirror of matmul_scaling_benchmark.py:106-165: bmm on the local batch, then all_reduce(C, SUM)."""
    if local_batch_size is None:
        local_batch_size = batch_size // world_size  # :111 (0 at ws=8 with the default total of 4)
    torch.manual_seed(rank)
    A = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
    B = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
    C = torch.empty_like(A)
    comm_on = dist.is_initialized()

    for _ in range(warmup_iterations):
        ops.bmm(A, B, out=C)
        if comm_on:
            ops.all_reduce(C)
    ops.synchronize()
    if comm_on:
        dist.barrier()

    compute_ms = comm_ms = 0.0
    for _ in range(num_iterations):
        compute_ms += _timed(ops, lambda: ops.bmm(A, B, out=C))
        if comm_on:
            comm_ms += _timed(ops, lambda: ops.all_reduce(C))
    avg_compute_time = (compute_ms / 1000.0) / max(num_iterations, 1)
    avg_comm_time = (comm_ms / 1000.0) / max(num_iterations, 1)
    avg_total_time = avg_compute_time + avg_comm_time
    if rank == 0 and num_iterations > 0:
        print(f"    Compute time: {avg_compute_time*1000:.3f} ms, Comm time: {avg_comm_time*1000:.3f} ms")

    if check and local_batch_size > 0:
        ops.bmm(A, B, out=C)
        ops.synchronize()
        _report_check(rank, hostlib.parity_error(C, A, B), dtype, device)

    if overlap and comm_on and local_batch_size > 0 and ops.bmm_allreduce_overlap is not None:
        for _ in range(min(warmup_iterations, 3)):
            ops.bmm_allreduce_overlap(A, B, C, panels)
        ops.synchronize()
        dist.barrier()
        avg_overlap = _timed_loop(ops, lambda: ops.bmm_allreduce_overlap(A, B, C, panels), num_iterations)
        if rank == 0:
            print(f"    Overlapped end-to-end: {avg_overlap*1000:.3f} ms ({panels} panels; serial sum {avg_total_time*1000:.3f} ms)")
        avg_total_time = avg_overlap

    if fused and comm_on and local_batch_size > 0 and ops.bmm_allreduce_fused is not None:
        shared = (ops.reduce_buffer_like or ops.shared_like)(C)  # NVLS multicast buffer where available
        how = "reduced in the NVSwitch" if hasattr(shared, "mc_ptr") else "peer memory"
        for _ in range(min(warmup_iterations, 3)):
            ops.bmm_allreduce_fused(A, B, shared, panels)
        ops.synchronize()
        dist.barrier()
        avg_fused = _timed_loop(ops, lambda: ops.bmm_allreduce_fused(A, B, shared, panels), num_iterations)
        if rank == 0:
            print(f"    Fused GEMM+all_reduce ({how}) end-to-end: {avg_fused*1000:.3f} ms ({panels} panels; serial sum {(avg_compute_time + avg_comm_time)*1000:.3f} ms)")
        if check:
            # Against the NCCL all_reduce of the same rounded products (what the reference computes, :142-150).  Both are
            # correct roundings of one sum, but NCCL may round PARTIAL sums (ring / tree), which are bounded by the ranks'
            # terms, not by the final sum: the bound is per element, from every rank's product (hostlib.reduction_pair_error).
            ops.bmm(A, B, out=C)
            ops.all_reduce(C)
            ops.synchronize()

            def products():
                for r in range(world_size):
                    torch.manual_seed(r)  # the recipe above, for rank r
                    Ar = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
                    Br = torch.randn(local_batch_size, matrix_size, matrix_size, dtype=dtype, device=device)
                    yield ops.bmm(Ar, Br)

            err = hostlib.reduction_pair_error(shared.local, C, hostlib.term_magnitudes(products()), world_size)
            hostlib.report_check(rank, err, hostlib.CHECK_TOL[dtype], "of the fused all_reduce vs NCCL all_reduce of the same products",
                                 device=device)
        avg_total_time = avg_fused

    tflops_per_gpu = calculate_tflops(matrix_size, avg_total_time, num_ops=local_batch_size) if avg_total_time > 0 else 0.0
    return avg_total_time, tflops_per_gpu


def benchmark_matrix_parallel(matrix_size: int, dtype: torch.dtype, device: str, rank: int, world_size: int,
                              num_iterations: int = 50, warmup_iterations: int = 10,
                              ops: Optional[hostlib.Ops] = None, overlap: bool = False, panels: int = 4,
                              check: bool = False, fused: bool = False) -> Tuple[float, float]:
    """Mirror of matmul_scaling_benchmark.py:167-238: A replicated, B split by columns, C slabs gathered.

    The gathered result is ONE contiguous [world][N][N/world] buffer (the logical
    content of the reference's list of tensors, :219-221).  Inputs are seeded (0) so
    that A really is replicated and a global C exists to check against.
    """
    if world_size == 1:
        return benchmark_independent(matrix_size, dtype, device, rank, num_iterations, warmup_iterations, ops=ops,
                                     check=check)
    start_col, end_col = hostlib.split_columns(matrix_size, world_size, rank)
    width, pitch = end_col - start_col, hostlib.slab_pitch(matrix_size, world_size)
    if pitch != width and (overlap or fused):
        raise RuntimeError(f"--overlap/--fused need equal, 16-byte-pitched column slabs: matrix_size % {8 * world_size} == 0 "
                           f"(got {matrix_size})")
    # pitch == width whenever matrix_size % (8 * world) == 0: plain contiguous slabs.  Otherwise (the reference
    # breaks inside all_gather on its unequal slabs, :179-183,:219-221) every rank keeps its slab in storage of
    # the common pitch, so that the gather moves equal-size pieces and B_local / C_local rows stay 16-byte aligned.
    torch.manual_seed(0)
    A = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
    B_store = torch.zeros(matrix_size, pitch, dtype=dtype, device=device)
    if check:
        B_full = torch.randn(matrix_size, matrix_size, dtype=dtype, device=device)
        B_store[:, :width] = B_full[:, start_col:end_col]
    else:
        B_full = None
        torch.manual_seed(1000 + rank)
        B_store[:, :width] = torch.randn(matrix_size, width, dtype=dtype, device=device)
    C_store = torch.zeros(matrix_size, pitch, dtype=dtype, device=device)
    B_local, C_local = B_store[:, :width], C_store[:, :width]
    slabs = torch.zeros(world_size, matrix_size, pitch, dtype=dtype, device=device)

    for _ in range(warmup_iterations):
        ops.matmul(A, B_local, out=C_local)
        ops.all_gather_into(slabs, C_store)
    ops.synchronize()
    dist.barrier()

    compute_ms = comm_ms = 0.0
    for _ in range(num_iterations):
        compute_ms += _timed(ops, lambda: ops.matmul(A, B_local, out=C_local))
        comm_ms += _timed(ops, lambda: ops.all_gather_into(slabs, C_store))
    avg_compute_time = (compute_ms / 1000.0) / max(num_iterations, 1)
    avg_comm_time = (comm_ms / 1000.0) / max(num_iterations, 1)
    avg_total_time = avg_compute_time + avg_comm_time
    if rank == 0 and num_iterations > 0:
        print(f"    Compute time: {avg_compute_time*1000:.3f} ms, Comm time: {avg_comm_time*1000:.3f} ms")

    if overlap and ops.matmul_allgather_overlap is not None:
        slabs.zero_()
        for _ in range(min(warmup_iterations, 3)):
            ops.matmul_allgather_overlap(A, B_local, C_local, slabs, panels)
        ops.synchronize()
        dist.barrier()
        avg_overlap = _timed_loop(ops, lambda: ops.matmul_allgather_overlap(A, B_local, C_local, slabs, panels),
                                  num_iterations)
        if rank == 0:
            print(f"    Overlapped end-to-end: {avg_overlap*1000:.3f} ms ({panels} panels; serial sum {avg_total_time*1000:.3f} ms)")
        avg_total_time = avg_overlap

    if fused and ops.matmul_allgather_fused is not None:
        shared = (ops.gather_buffer or ops.shared_slabs)(matrix_size, end_col - start_col, dtype)
        how = "one multicast TMA store per chunk" if hasattr(shared, "mc_ptr") else "peer TMA stores"
        shared.local.zero_()
        ops.synchronize()
        dist.barrier()
        for _ in range(min(warmup_iterations, 3)):
            ops.matmul_allgather_fused(A, B_local, shared)
        ops.synchronize()
        dist.barrier()
        avg_fused = _timed_loop(ops, lambda: ops.matmul_allgather_fused(A, B_local, shared), num_iterations)
        if rank == 0:
            print(f"    Fused GEMM+all_gather end-to-end: {avg_fused*1000:.3f} ms (one kernel, {how}, + barrier; serial sum {(avg_compute_time + avg_comm_time)*1000:.3f} ms)")
        avg_total_time = avg_fused
        slabs = shared.local

    if check and B_full is not None:
        ops.synchronize()
        C_global = hostlib.assemble_slabs(slabs, matrix_size, world_size)  # [N][N]: slab r holds columns of rank r
        _report_check(rank, hostlib.parity_error(C_global, A, B_full), dtype, device)

    tflops_per_gpu = calculate_tflops(matrix_size, avg_total_time, num_ops=1) / world_size
    return avg_total_time, tflops_per_gpu


def _report_check(rank: int, err: float, dtype: torch.dtype, device: Optional[str] = None) -> None:
    hostlib.report_check(rank, err, hostlib.CHECK_TOL[dtype], "vs float64 reference", device=device)


def run_benchmarks(rank: int, world_size: int, matrix_sizes: List[int], dtype: torch.dtype, mode: ScalingMode,
                   num_iterations: int, warmup_iterations: int, ops: Optional[hostlib.Ops] = None,
                   device: Optional[str] = None, overlap: bool = False, panels: int = 4,
                   batch_size: int = hostlib.TOTAL_BATCH, per_gpu_batch: Optional[int] = None, check: bool = False,
                   bench_fns: Optional[dict] = None, results: Optional[list] = None, fused: bool = False) -> None:
    """Size loop, mode dispatch, aggregation and report of matmul_scaling_benchmark.py:251-347.

    `bench_fns` maps a ScalingMode to a stub `(size) -> (avg_s, tflops_per_gpu)`; tests use it to pin the report.
    """
    device = device or hostlib.rank_device(rank)
    total_batch = batch_size if per_gpu_batch is None else per_gpu_batch * world_size

    if rank == 0:
        bar = "=" * 70
        print(f"\n{bar}\nMatrix Multiplication Scaling Benchmark\n{bar}")
        print("Configuration:")
        print(f"  - Mode: {mode.value}")
        print(f"  - Number of GPUs: {world_size}")
        print(f"  - Data type: {dtype}")
        print(f"  - Iterations per test: {num_iterations}")
        print(f"  - Warmup iterations: {warmup_iterations}")
        print(f"{bar}\n")

    for size in matrix_sizes:
        if rank == 0:
            gib = size * size * hostlib.bytes_per_element(dtype) / (1024 ** 3)
            print(f"\nBenchmarking {size}x{size} matrix multiplication:")
            print(f"  - Memory per matrix: {gib:.2f} GB ({hostlib.dtype_name(dtype)})")
            print(f"  - Mode: {mode.value}")
        try:
            if bench_fns is not None:
                avg_time, tflops_per_gpu = bench_fns[mode](size)
            elif mode == ScalingMode.INDEPENDENT:
                avg_time, tflops_per_gpu = benchmark_independent(size, dtype, device, rank, num_iterations,
                                                                 warmup_iterations, ops=ops, check=check)
            elif mode == ScalingMode.BATCH_PARALLEL:
                avg_time, tflops_per_gpu = benchmark_batch_parallel(
                    size, total_batch, dtype, device, rank, world_size, num_iterations, warmup_iterations, ops=ops,
                    overlap=overlap, panels=panels, local_batch_size=per_gpu_batch, check=check, fused=fused)
            else:
                avg_time, tflops_per_gpu = benchmark_matrix_parallel(size, dtype, device, rank, world_size,
                                                                     num_iterations, warmup_iterations, ops=ops,
                                                                     overlap=overlap, panels=panels, check=check,
                                                                     fused=fused)

            avg_time_all = hostlib.metric_all_reduce(avg_time, device, "avg")
            tflops_all = hostlib.metric_all_reduce(tflops_per_gpu, device,
                                                   "sum" if mode == ScalingMode.INDEPENDENT else "avg")
            if rank == 0:
                print(f"\nResults for {size}x{size}:")
                print(f"  - Average time per operation: {avg_time_all*1000:.3f} ms")
                if mode == ScalingMode.INDEPENDENT:
                    print(f"  - TFLOPS per GPU: {tflops_per_gpu:.2f}")
                    print(f"  - Total system TFLOPS: {tflops_all:.2f}")
                    print(f"  - Scaling efficiency: {(tflops_all / (tflops_per_gpu * world_size)) * 100:.1f}%")
                    total_flops = 2.0 * size ** 3 * world_size
                elif mode == ScalingMode.BATCH_PARALLEL:
                    print(f"  - TFLOPS per GPU: {tflops_per_gpu:.2f}")
                    print(f"  - Total system TFLOPS: {tflops_per_gpu * world_size:.2f}")
                    print(f"  - Processing {total_batch} total batches across {world_size} GPU(s)")
                    total_flops = 2.0 * size ** 3 * total_batch
                else:
                    print(f"  - TFLOPS per GPU (portion): {tflops_per_gpu:.2f}")
                    print(f"  - Effective system TFLOPS: {tflops_all:.2f}")
                    print(f"  - Each GPU processes 1/{world_size} of the matrix")
                    total_flops = 2.0 * size ** 3
                actual = (total_flops / avg_time_all) / 1e12 if avg_time_all > 0 else 0.0
                print(f"  - Actual TFLOPS (total FLOPs / time): {actual:.2f}")
                if results is not None:
                    results.append({"size": size, "mode": mode.value, "n_gpus": world_size,
                                    "avg_time_ms": avg_time_all * 1000, "tflops_per_gpu": tflops_per_gpu,
                                    "aggregate_tflops": tflops_all, "actual_tflops": actual, "overlap": overlap,
                                    "fused": fused})
        except torch.cuda.OutOfMemoryError:
            if rank == 0:
                print(f"\n  ERROR: Out of memory for {size}x{size} matrices")
        except Exception as e:  # the reference's per-size net (:337-342)
            if rank == 0:
                print(f"\n  ERROR: {str(e)}")
        if ops is not None:
            ops.release_shared()  # peers unmap this size's shared buffers before empty_cache() frees them
        if torch.cuda.is_available():
            from . import ops as b2
            b2.release_workspaces()
            torch.cuda.empty_cache()
        if dist.is_initialized():
            dist.barrier()


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Matrix Multiplication Scaling Benchmark")
    hostlib.add_common_args(parser)
    parser.add_argument("--mode", type=str, default="independent",
                        choices=["independent", "batch_parallel", "matrix_parallel"],
                        help="Scaling mode (default: independent)")
    parser.add_argument("--overlap", action="store_true", help="Pipeline the collective behind compute (two streams)")
    parser.add_argument("--fused", action="store_true",
                        help="matrix_parallel: all_gather done by the GEMM epilogue over NVLink (one kernel)")
    parser.add_argument("--panels", type=int, default=4, help="Row panels per GEMM with --overlap (default: 4)")
    parser.add_argument("--batch-size", type=int, default=hostlib.TOTAL_BATCH,
                        help="Total batch for batch_parallel (default: 4, the reference's constant)")
    parser.add_argument("--per-gpu-batch", type=int, default=None,
                        help="Fixed local batch per GPU for batch_parallel (weak scaling)")
    return parser


def main(argv: Optional[List[str]] = None) -> None:
    args = build_parser().parse_args(argv)
    dtype = hostlib.DTYPE_MAP[args.dtype]
    mode = ScalingMode(args.mode)
    rank, world_size = setup_distributed()
    if rank == 0:
        hostlib.print_banner()
    device = hostlib.rank_device(rank)
    ops = hostlib.device_ops(rank, world_size, device)
    if world_size > 1 and not verify_collectives(rank, world_size, device, ops):
        if rank == 0:
            print("ERROR: Collective operations verification failed!")
        ops.destroy()
        cleanup_distributed()
        return
    results: list = []
    try:
        run_benchmarks(rank, world_size, args.sizes, dtype, mode, args.iterations, args.warmup, ops=ops, device=device,
                       overlap=args.overlap, panels=args.panels, batch_size=args.batch_size,
                       per_gpu_batch=args.per_gpu_batch, check=args.check, results=results, fused=args.fused)
    finally:
        ops.destroy()
        cleanup_distributed()
    if rank == 0:
        if args.json:
            with open(args.json, "w") as f:
                json.dump({"script": "matmul_scaling_benchmark", "mode": args.mode, "dtype": args.dtype,
                           "results": results}, f, indent=1)
        bar = "=" * 70
        print(f"\n{bar}\nBenchmark completed!\n{bar}\n")


if __name__ == "__main__":
    main()
