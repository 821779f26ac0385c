"""Where does the fused matrix_parallel step spend its time?  (run on N GPUs, one process per GPU)

For the 16384 problem split over the world, every row times one thing, max over ranks, CUDA events:
  gemm_alone[tiles]          the slab GEMM into local memory, both tile shapes
  fused[multicast|peer][tiles]   GEMM whose epilogue stores into every rank's gather buffer, + barrier
  store_only[...]            the same kernels with K = 64: almost no math, so the time is the epilogue's store path
                             (TMA stores through the multicast address / to 8 peers) moving the same 512 MiB
  nccl_all_gather            ncclAllGather of the slabs alone
  barrier                    the trailing cross-rank barrier alone
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
rank, ws = hostlib.setup_distributed()
dev = hostlib.rank_device(rank)
comm = b2.Communicator(rank, ws, torch.device(dev))
dt = torch.bfloat16
n_local = N // ws
res = {"world": ws, "size": N}


def max_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return round(max_ranks(s.elapsed_time(e) / iters), 4)


torch.manual_seed(0)
A = torch.randn(N, N, dtype=dt, device=dev)
B_local = torch.randn(N, n_local, dtype=dt, device=dev)
C_local = torch.empty(N, n_local, dtype=dt, device=dev)
slabs = torch.empty(ws, N, n_local, dtype=dt, device=dev)
mc = comm.gather_buffer(N, n_local, dt)
peer = comm.shared_slabs(N, n_local, dt)
res["gather_buffer_is_multicast"] = hasattr(mc, "mc_ptr")
A64, B64 = A[:, :64].contiguous(), B_local[:64].contiguous()  # K = 64: the store path with (almost) no math

for rep in range(2):
    r = {}
    for tiles in ("1", "2"):
        os.environ["B2MM_M_BLOCKS"] = tiles
        tag = "256x256" if tiles == "1" else "512x256"
        r[f"gemm_alone[{tag}]"] = timed(lambda: b2.matmul(A, B_local, out=C_local))
        r[f"fused[multicast][{tag}]"] = timed(lambda: comm.matmul_allgather_fused(A, B_local, mc))
        r[f"fused[peer][{tag}]"] = timed(lambda: comm.matmul_allgather_fused(A, B_local, peer))
        r[f"store_only[local][{tag}]"] = timed(lambda: b2.matmul(A64, B64, out=C_local))
        r[f"store_only[multicast][{tag}]"] = timed(lambda: comm.matmul_allgather_fused(A64, B64, mc))
        r[f"store_only[peer][{tag}]"] = timed(lambda: comm.matmul_allgather_fused(A64, B64, peer))
    os.environ.pop("B2MM_M_BLOCKS")
    r["fused[multicast][default]"] = timed(lambda: comm.matmul_allgather_fused(A, B_local, mc))
    r["nccl_all_gather"] = timed(lambda: comm.all_gather_into(slabs, C_local))
    r["barrier"] = timed(lambda: comm.barrier())
    res[f"rep{rep}"] = r
    if rank == 0:
        print(json.dumps(r), flush=True)
gb = N * N * 2 / 1e9
res["bytes_into_each_gpu_GB"] = {"multicast (own slab comes back through the switch)": gb, "peer / nccl": gb * (ws - 1) / ws}
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/r2_gather_probe_{ws}gpu.json", "w"), indent=1)
comm.destroy()
hostlib.cleanup_distributed()
