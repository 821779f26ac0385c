"""Why does the pipelined batch_parallel time jump between ~6.9 and ~8.5 ms?  Per-iteration traces of the
GEMM + all_reduce pipeline for several NCCL CTA caps and SM-sharing schemes (8 GPUs, 16384 bf16, local batch 1)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from pytorch_distributed_matmul_benchmark_b200 import hostlib
import pytorch_distributed_matmul_benchmark_b200 as b2

N = 16384
rank, ws = hostlib.setup_distributed()
dev = hostlib.rank_device(rank)
torch.manual_seed(rank)
A = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
B = torch.randn(1, N, N, dtype=torch.bfloat16, device=dev)
C = torch.empty(1, N, N, dtype=torch.bfloat16, device=dev)


def per_iter(fn, iters):
    torch.cuda.synchronize(); dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn(); ev[i + 1].record()
    torch.cuda.synchronize()
    t = torch.tensor([ev[i].elapsed_time(ev[i + 1]) for i in range(iters)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [round(x, 2) for x in t.tolist()]


out = {}
comm0 = b2.Communicator(rank, ws, torch.device(dev))
sharedC = comm0.shared_like(C)
fused = {}
for panels in (2, 4, 8):
    fused[f"gemm_then_fused_p{panels}"] = {"gemm": per_iter(lambda: b2.bmm(A, B, out=C), 6),
                                           "pipe": per_iter(lambda: comm0.bmm_allreduce_fused(A, B, sharedC, panels), 16)}
# parity against NCCL's all_reduce of the same product
comm0.bmm_allreduce_fused(A, B, sharedC, 8)
b2.bmm(A, B, out=C); comm0.all_reduce(C); torch.cuda.synchronize()
ref = C[0, :512].double()
err = ((sharedC.local[0, :512].double() - ref).abs() / (ref.abs() + ref.pow(2).mean().sqrt())).max()
dist.all_reduce(err, op=dist.ReduceOp.MAX)
fused["max_err_vs_nccl_allreduce"] = err.item()
out["fused_coresident_p2p"] = fused
comm0.destroy()
for max_ctas in (0,):
    if max_ctas:
        os.environ["B2MM_NCCL_MAX_CTAS"] = str(max_ctas)
    else:
        os.environ.pop("B2MM_NCCL_MAX_CTAS", None)
    comm = b2.Communicator(rank, ws, torch.device(dev))
    key = f"nccl_max_ctas={max_ctas or 'default'}"
    res = {"allreduce_alone": per_iter(lambda: comm.all_reduce(C), 6)}
    schemes = [("reserve", max_ctas or 16)]
    for kind, val in schemes:
        os.environ.pop("B2MM_COMM_SMS", None); os.environ.pop("B2MM_TILES_PER_CLUSTER", None)
        os.environ["B2MM_COMM_SMS" if kind == "reserve" else "B2MM_TILES_PER_CLUSTER"] = str(val)
        res[f"gemm_then_{kind}{val}_p8"] = {"gemm": per_iter(lambda: b2.bmm(A, B, out=C), 6),
                                            "pipe": per_iter(lambda: comm.bmm_allreduce_overlap(A, B, C, 8), 16)}
    out[key] = res
    comm.destroy()
if rank == 0:
    json.dump(out, open("gpurun_out/batch_parallel_study.json", "w"), indent=1)
    for k, v in out.items():
        print(k)
        for kk, vv in v.items():
            print("   ", kk, vv)
hostlib.cleanup_distributed()
