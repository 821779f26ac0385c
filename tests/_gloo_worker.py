"""Worker for test_multirank_gloo.py: the N>1 host logic on CPU tensors under gloo (world_size 2).

The device library is replaced by CPU stand-ins bound through hostlib.Ops — the oracle as the GEMM,
torch.distributed(gloo) as the collectives — so what runs here is the product's own mode bodies,
partitioning, gather layout, aggregation and report code."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["B2MM_BACKEND"] = "gloo"

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle import oracle, reference_port  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import hostlib  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import matmul_scaling_benchmark as ms  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import matmul_distributed_benchmark as md  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import matmul_overlap_benchmark as ov  # noqa: E402


class FakeEvent:
    def record(self):
        self.t = time.perf_counter()

    def elapsed_time(self, other):
        return (other.t - self.t) * 1000.0


def cpu_ops() -> hostlib.Ops:
    def matmul(A, B, out=None):
        C = oracle.gemm(A, B, threads=2)
        return C if out is None else out.copy_(C)

    def all_reduce(t):
        if t.numel():
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_gather_into(slabs, t):
        dist.all_gather(list(slabs.unbind(0)), t)  # slab r <- rank r, the layout b2mm_all_gather produces
        return slabs

    return hostlib.Ops(matmul=matmul, bmm=matmul, all_reduce=all_reduce, all_gather_into=all_gather_into,
                       synchronize=lambda: None, event=FakeEvent)


def main(out_dir: str):
    rank, ws = hostlib.setup_distributed()
    device = hostlib.rank_device(rank)
    assert device == "cpu" and ws == 2
    ops = cpu_ops()
    result = {"rank": rank, "verify": ms.verify_collectives(rank, ws, device, ops)}

    # metric aggregation (matmul_scaling_benchmark.py:297-306)
    result["sum"] = hostlib.metric_all_reduce(float(rank + 1), device, "sum")
    result["avg"] = hostlib.metric_all_reduce(float(rank + 1), device, "avg")

    # a parity verdict is collective: one rank over the bar makes EVERY rank raise (nobody walks on into a collective
    # the failing rank never reaches), and a pass returns the worst error over ranks everywhere
    real_stdout = sys.stdout
    sys.stdout = open(os.devnull, "w")
    try:
        try:
            hostlib.report_check(rank, 1.0 if rank == 1 else 0.0, 0.5, "collective verdict", device=device)
            result["check_raised"] = False
        except RuntimeError as e:
            result["check_raised"] = "1.000e+00 > 5.0e-01" in str(e)
        result["check_pass_value"] = hostlib.report_check(rank, 0.125 * (rank + 1), 0.5, "collective verdict", device=device)
    finally:
        sys.stdout.close()
        sys.stdout = real_stdout

    # batch_parallel: after bmm + all_reduce every rank holds the sum over ranks of its peers' products
    N, dt = 48, torch.float32
    A, B = reference_port.batch_parallel_inputs(rank, ws, N, dt)
    C = ops.bmm(A, B)
    ops.all_reduce(C)
    expect = sum(oracle.gemm(*reference_port.batch_parallel_inputs(r, ws, N, dt)).double() for r in range(ws))
    result["batch_parallel_err"] = oracle.rel_err(C, expect)

    # matrix_parallel: slabs [ws][N][N/ws] assemble to the global product
    A, Bfull, B_local, (s, e) = reference_port.matrix_parallel_inputs(rank, ws, N, dt)
    C_local = ops.matmul(A, B_local)
    slabs = torch.zeros(ws, N, e - s, dtype=dt)
    ops.all_gather_into(slabs, C_local)
    C_global = torch.cat(list(slabs.unbind(0)), dim=1)
    result["matrix_parallel_err"] = oracle.rel_err(C_global, A.double() @ Bfull.double())

    # the three modes end to end through run_benchmarks (rank 0 prints the report)
    log = open(os.path.join(out_dir, f"report_rank{rank}.txt"), "w")
    real_stdout = sys.stdout
    sys.stdout = log
    try:
        for mode in ms.ScalingMode:
            ms.run_benchmarks(rank, ws, [32, 64], torch.float32, mode, 2, 1, ops=ops, device=device, check=True)
        ms.run_benchmarks(rank, ws, [30], torch.bfloat16, ms.ScalingMode.MATRIX_PARALLEL, 1, 1, ops=ops, device=device)
        ms.run_benchmarks(rank, ws, [33], torch.float32, ms.ScalingMode.MATRIX_PARALLEL, 1, 1, ops=ops, device=device, check=True)
        ms.run_benchmarks(rank, ws, [33], torch.float32, ms.ScalingMode.MATRIX_PARALLEL, 1, 1, ops=ops, device=device, fused=True)
        ms.run_benchmarks(rank, 8, [32], torch.float32, ms.ScalingMode.BATCH_PARALLEL, 1, 1, ops=ops, device=device)
        # the earlier draft's modes (backup/matmul_distributed_benchmark.py); 33 % 2 != 0 exercises the unequal slabs
        for mode in md.BenchmarkMode:
            md.run_benchmarks(rank, ws, [32, 33], torch.float32, mode, 2, 1, ops=ops, device=device, check=True)
        # the serialised mode of backup/matmul_overlap_benchmark.py (the overlapped ones need CUDA streams)
        ov.run_benchmarks(rank, ws, [32], torch.float32, ov.BenchmarkMode.NO_OVERLAP, 2, 1, ops=ops, device=device, check=True)
    finally:
        sys.stdout = real_stdout
        log.close()
    with open(os.path.join(out_dir, f"result_rank{rank}.json"), "w") as f:
        json.dump(result, f)
    hostlib.cleanup_distributed()


if __name__ == "__main__":
    main(sys.argv[1])
