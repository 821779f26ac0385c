"""NVLS (multicast) all_reduce: set-up, correctness against NCCL and float64 sums, and a first timing.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=N --master-addr=127.0.0.1 scripts/nvls_check.py [size=16384]
"""
import json
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from pytorch_distributed_matmul_benchmark_b200 import hostlib, multicast
import pytorch_distributed_matmul_benchmark_b200 as b2

BIG = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
rank, ws = hostlib.setup_distributed()
dev = hostlib.rank_device(rank)
comm = b2.Communicator(rank, ws, torch.device(dev))
res = {"world": ws}


def log(*a):
    if rank == 0:
        print(*a, flush=True)


def max_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def timed(fn, iters=8, warm=3):
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


def rel(C, ref):
    ref = ref.double()
    return ((C.double() - ref).abs() / (ref.abs() + ref.pow(2).mean().sqrt())).max().item()


try:
    res["multicast_supported"] = multicast.multicast_supported(torch.device(dev).index)
    log("multicast supported:", res["multicast_supported"])
    # 1. plain all_reduce through the switch: known values, then random data against NCCL
    for dt in (torch.bfloat16, torch.float16, torch.float32):
        buf = comm.multicast_like(torch.empty(1 << 22, dtype=dt, device=dev))
        log(f"{dt}: buffer {buf.nbytes} bytes in {buf.size} mapped; unicast {buf.ptr:#x} multicast {buf.mc_ptr:#x}")
        buf.local.fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        comm.all_reduce_nvls(buf)
        torch.cuda.synchronize()
        ok_const = bool(torch.all(buf.local == float(ws * (ws + 1) // 2)))
        torch.manual_seed(50 + rank)
        x = torch.randn(1 << 22, dtype=dt, device=dev)
        buf.local.copy_(x)
        torch.cuda.synchronize()
        dist.barrier()
        comm.all_reduce_nvls(buf)
        comm.all_reduce(x)  # NCCL
        torch.cuda.synchronize()
        err = max_ranks(rel(buf.local, x))
        res[f"allreduce_{dt}"] = {"const_ok": ok_const, "vs_nccl": err}
        log(f"{dt}: constants {'OK' if ok_const else 'WRONG'}, vs NCCL {err:.3e}")
    # how does the switch round?  rank 0 holds 1.0, rank 1 holds x * ulp(1.0) (bf16 ulp = 2^-7), the others 0
    probe = comm.multicast_like(torch.empty(8, dtype=torch.bfloat16, device=dev))
    fracs = [0.25, 0.5, 0.75, 1.25, 1.5, 1.75, 2.5, 3.5]
    probe.local.zero_()
    if rank == 0:
        probe.local.fill_(1.0)
    elif rank == 1:
        probe.local.copy_(torch.tensor([f * 2.0 ** -7 for f in fracs], dtype=torch.bfloat16))
    torch.cuda.synchronize()
    dist.barrier()
    comm.all_reduce_nvls(probe)
    torch.cuda.synchronize()
    got = [round((v - 1.0) / 2.0 ** -7, 3) for v in probe.local.float().tolist()]
    res["rounding_probe"] = {"added_ulps": fracs, "result_ulps_above_one": got}
    log("rounding probe: 1.0 + x ulp ->", dict(zip(fracs, got)))
    comm.release_shared()

    # 2. fused GEMM + NVLS reduction, small, against the float64 sum of the per-rank (rounded) products
    N = 2048
    torch.manual_seed(rank)
    A = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
    B = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
    mcC = comm.multicast_like(torch.empty(1, N, N, dtype=torch.bfloat16, device=dev))
    for _ in range(3):
        comm.bmm_allreduce_fused(A, B, mcC, 2)
    torch.cuda.synchronize()
    want = 0
    for r in range(ws):
        torch.manual_seed(r)
        Ar = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
        Br = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
        want = want + b2.bmm(Ar, Br).double()
    err = max_ranks(rel(mcC.local, want))
    res["fused_nvls_err_2048"] = err
    log(f"fused GEMM + NVLS all_reduce at {N}: err vs float64 sum {err:.3e} (bf16 tol 4e-3)")
    comm.release_shared()

    # 3. timing at the benchmark size: GEMM alone, fused with peer pull/push, fused with NVLS; all_reduce alone both ways
    N = BIG
    torch.manual_seed(rank)
    A = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
    B = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
    C = torch.empty(1, N, N, dtype=torch.bfloat16, device=dev)
    p2pC = comm.shared_like(C)
    mcC = comm.multicast_like(C)
    bands = 8 if N % (8 * 512) == 0 else 2
    t = {}
    t["nccl_allreduce_ms"] = timed(lambda: comm.all_reduce(C))
    t["p2p_sync_allreduce_ms"] = timed(lambda: comm.all_reduce_sync(p2pC))
    t["nvls_allreduce_ms"] = timed(lambda: comm.all_reduce_nvls(mcC))
    for ctas in ("32", "64"):
        os.environ["B2MM_P2P_CTAS"] = ctas
        t[f"nvls_allreduce_{ctas}ctas_ms"] = timed(lambda: comm.all_reduce_nvls(mcC))
    os.environ.pop("B2MM_P2P_CTAS")
    rounds = []
    for _ in range(5):
        # every variant is preceded by the GEMM alone, so that each efficiency is a ratio of neighbouring times
        r = {}
        for name, fn in (("fused_p2p", lambda: comm.bmm_allreduce_fused(A, B, p2pC, bands)),
                         ("fused_nvls", lambda: comm.bmm_allreduce_fused(A, B, mcC, bands)),
                         ("nccl_pipelined", lambda: comm.bmm_allreduce_overlap(A, B, C, bands))):
            g = timed(lambda: b2.bmm(A, B, out=C), iters=5, warm=2)
            f = timed(fn, iters=5, warm=2)
            r[name] = {"gemm_ms": round(g, 3), "ms": round(f, 3), "eff": round(g / f, 4)}
        rounds.append(r)
        log(json.dumps(r))
    t["rounds"] = rounds
    for name in ("fused_p2p", "fused_nvls", "nccl_pipelined"):
        effs = sorted(r[name]["eff"] for r in rounds)
        t[f"{name}_eff_median"] = effs[len(effs) // 2]
        t[f"{name}_eff_mean"] = sum(effs) / len(effs)
    bytes_ar = C.numel() * 2
    for k in ("nccl", "p2p_sync", "nvls"):
        t[f"{k}_busbw_GBs"] = bytes_ar * 2 * (ws - 1) / ws / (t[f"{k}_allreduce_ms"] * 1e-3) / 1e9
    # parity of the two fused paths at full size
    comm.bmm_allreduce_fused(A, B, p2pC, bands)
    comm.bmm_allreduce_fused(A, B, mcC, bands)
    torch.cuda.synchronize()
    t["nvls_vs_p2p_err"] = max_ranks(rel(mcC.local[0, :512], p2pC.local[0, :512]))
    res["timing"] = t
    log(json.dumps({k: v for k, v in t.items() if k != "rounds"}))
except Exception:
    res["error"] = traceback.format_exc()
    print(f"[rank {rank}] FAILED\n{res['error']}", flush=True)

if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/nvls_check_{ws}gpu.json", "w"), indent=1)
    print("NVLS CHECK", "FAILED" if "error" in res else "DONE", flush=True)
try:
    comm.destroy()
    hostlib.cleanup_distributed()
except Exception:
    traceback.print_exc()
