"""One torchrun launch that measures every multi-GPU mode at one size (default 16384 bf16).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=N --master-addr=127.0.0.1 scripts/multi_gpu_suite.py [N_size] [out.json]

independent        per-rank GEMM time (the weak-scaling reference T1)
batch_parallel     weak scaling (local batch 1): reference-style serial split, then the stream-pipelined
                   variant over a small sweep of (SMs left to NCCL) x (row panels); efficiency = T1 / T
matrix_parallel    serial split, stream-pipelined, and the fused GEMM+all_gather kernel; parity of the
                   assembled C against float64 on sampled rows
collectives        bus bandwidth of the all_reduce / all_gather payloads as a fraction of 900 GB/s
All times are device times (CUDA events), max over ranks.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from pytorch_distributed_matmul_benchmark_b200 import hostlib
import pytorch_distributed_matmul_benchmark_b200 as b2

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
OUT = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/multi_gpu_suite.json"
ITERS, WARM = 10, 3
dt = torch.bfloat16

rank, ws = hostlib.setup_distributed()
dev = hostlib.rank_device(rank)
comm = b2.Communicator(rank, ws, torch.device(dev))


def max_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def timed(fn, iters=ITERS, warm=WARM):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return max_ranks(s.elapsed_time(e) / iters)


def timed_each(fn, iters=12):
    """Per-iteration device times (ms) on this rank, no warm-up: shows ramp effects."""
    torch.cuda.synchronize()
    dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(iters)]


res = {"size": N, "world": ws, "dtype": "bfloat16", "iters": ITERS}
flop = 2.0 * N ** 3

# ---- independent ----
torch.manual_seed(rank)
A = torch.randn(N, N, dtype=dt, device=dev)
B = torch.randn(N, N, dtype=dt, device=dev)
C = torch.empty(N, N, dtype=dt, device=dev)
t1 = timed(lambda: b2.matmul(A, B, out=C))
res["independent"] = {"ms": t1, "tflops_per_gpu": flop / t1 / 1e9, "total_tflops": ws * flop / t1 / 1e9}

# ---- batch_parallel, weak scaling: local batch 1 ----
A3, B3, C3 = A.view(1, N, N), B.view(1, N, N), C.view(1, N, N)
t_gemm = timed(lambda: b2.bmm(A3, B3, out=C3))
t_ar = timed(lambda: comm.all_reduce(C3))
bytes_ar = C3.numel() * 2
bp = {"compute_ms": t_gemm, "comm_ms": t_ar, "serial_ms": t_gemm + t_ar, "serial_eff": t1 / (t_gemm + t_ar),
      "allreduce_busbw_GBs": bytes_ar * 2 * (ws - 1) / ws / (t_ar * 1e-3) / 1e9, "overlap": []}
# stream-pipelined with NCCL (16 SMs left to NCCL, 8 row panels) and the fused path (one GEMM + co-resident
# peer-memory reductions per 2048-row band), each measured in rounds interleaved with the GEMM alone so that
# the weak-scaling efficiency T(1 GPU) / T(N GPUs) compares times taken in the same thermal state
os.environ["B2MM_COMM_SMS"] = "16"
sharedC = comm.shared_like(C3)
t_p2p_alone = timed(lambda: comm.all_reduce_p2p(sharedC))
bp["p2p_allreduce_alone_ms"] = t_p2p_alone
bp["p2p_allreduce_busbw_GBs"] = bytes_ar * 2 * (ws - 1) / ws / (t_p2p_alone * 1e-3) / 1e9
mcC = comm.multicast_like(C3) if comm.nvls_available() else None  # reduced inside the NVSwitch
if mcC is not None:
    t_nvls_alone = timed(lambda: comm.all_reduce_nvls(mcC))
    bp["nvls_allreduce_alone_ms"] = t_nvls_alone
    bp["nvls_allreduce_busbw_GBs"] = bytes_ar * 2 * (ws - 1) / ws / (t_nvls_alone * 1e-3) / 1e9
rounds = []
for _ in range(5):
    tg = timed(lambda: b2.bmm(A3, B3, out=C3), iters=6, warm=2)
    tn = timed(lambda: comm.bmm_allreduce_overlap(A3, B3, C3, 8), iters=6, warm=2)
    tg2 = timed(lambda: b2.bmm(A3, B3, out=C3), iters=6, warm=2)
    tf = timed(lambda: comm.bmm_allreduce_fused(A3, B3, sharedC, 8), iters=6, warm=2)
    r = {"gemm_only_ms": tg, "nccl_pipelined_ms": tn, "nccl_eff": tg / tn,
         "gemm_only_again_ms": tg2, "fused_ms": tf, "fused_eff": tg2 / tf}
    if mcC is not None:
        tg3 = timed(lambda: b2.bmm(A3, B3, out=C3), iters=6, warm=2)
        tv = timed(lambda: comm.bmm_allreduce_fused(A3, B3, mcC, 8), iters=6, warm=2)
        r.update({"gemm_only_third_ms": tg3, "fused_nvls_ms": tv, "fused_nvls_eff": tg3 / tv})
    rounds.append(r)
bp["interleaved_rounds"] = rounds
bp["nccl_pipelined_weak_eff"] = sum(r["nccl_eff"] for r in rounds) / len(rounds)
bp["fused_weak_eff"] = sum(r["fused_eff"] for r in rounds) / len(rounds)
if mcC is not None:
    effs = sorted(r["fused_nvls_eff"] for r in rounds)
    bp["fused_nvls_weak_eff"] = sum(effs) / len(effs)
    bp["fused_nvls_weak_eff_min_median_max"] = [effs[0], effs[len(effs) // 2], effs[-1]]
    comm.bmm_allreduce_fused(A3, B3, mcC, 8)
    b2.bmm(A3, B3, out=C3)
    comm.all_reduce(C3)
    torch.cuda.synchronize()
    refn = C3[0, :256].double()
    bp["fused_nvls_vs_nccl_err"] = max_ranks(((mcC.local[0, :256].double() - refn).abs() / (refn.abs() + refn.pow(2).mean().sqrt())).max().item())
# parity of the fused path against NCCL's all_reduce of the same product
comm.bmm_allreduce_fused(A3, B3, sharedC, 8)
b2.bmm(A3, B3, out=C3)
comm.all_reduce(C3)
torch.cuda.synchronize()
ref = C3[0, :256].double()
bp["fused_vs_nccl_err"] = max_ranks(((sharedC.local[0, :256].double() - ref).abs() / (ref.abs() + ref.pow(2).mean().sqrt())).max().item())
res["batch_parallel_weak_b1"] = bp

# ---- matrix_parallel ----
if N % ws == 0:
    n_local = N // ws
    torch.manual_seed(0)
    A.copy_(torch.randn(N, N, dtype=dt, device=dev))
    Bfull = torch.randn(N, N, dtype=dt, device=dev)
    B_local = Bfull[:, rank * n_local:(rank + 1) * n_local].contiguous()
    C_local = torch.empty(N, n_local, dtype=dt, device=dev)
    slabs = torch.zeros(ws, N, n_local, dtype=dt, device=dev)
    t_c = timed(lambda: b2.matmul(A, B_local, out=C_local))
    t_g = timed(lambda: comm.all_gather_into(slabs, C_local))
    mp = {"compute_ms": t_c, "comm_ms": t_g, "serial_ms": t_c + t_g, "serial_actual_tflops": flop / (t_c + t_g) / 1e9,
          "allgather_busbw_GBs": slabs.numel() * 2 * (ws - 1) / ws / (t_g * 1e-3) / 1e9}
    mp["overlap_ms"] = timed(lambda: comm.matmul_allgather_overlap(A, B_local, C_local, slabs, 4))
    shared = comm.shared_slabs(N, n_local, dt)
    shared.local.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    mp["fused_ms"] = timed(lambda: comm.matmul_allgather_fused(A, B_local, shared))
    mp["fused_actual_tflops"] = flop / mp["fused_ms"] / 1e9
    os.environ["B2MM_M_BLOCKS"] = "2"  # tall single-buffered tiles, for comparison (default in fused mode: 1)
    mp["fused_tall_tiles_ms"] = timed(lambda: comm.matmul_allgather_fused(A, B_local, shared))
    os.environ.pop("B2MM_M_BLOCKS")
    t_bar = timed(lambda: comm.barrier(), iters=50)
    mp["barrier_ms"] = t_bar
    torch.cuda.synchronize()
    Cg = torch.cat(list(shared.local.unbind(0)), dim=1)
    mp["fused_parity_err"] = max_ranks(hostlib.parity_error(Cg, A, Bfull, rows=128))
    mp["fused_parity_tol"] = hostlib.CHECK_TOL[dt]
    # fraction of NVLink the fused kernel's remote stores need: (ws-1) slabs out per GPU
    mp["fused_nvlink_out_GBs"] = (ws - 1) * N * n_local * 2 / (mp["fused_ms"] * 1e-3) / 1e9
    res["matrix_parallel"] = mp

if rank == 0:
    os.makedirs(os.path.dirname(OUT) or ".", exist_ok=True)
    with open(OUT, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))
comm.destroy()
hostlib.cleanup_distributed()
