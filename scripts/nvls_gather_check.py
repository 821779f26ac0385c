"""matrix_parallel with the gather done by multicast TMA stores: does a TMA store to a multicast address reach every
rank's copy?  Correctness vs float64 and timing against the peer-store (fan-out) kernel."""
import json, os, sys, traceback
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
    return max_ranks(s.elapsed_time(e) / iters)


try:
    for size in (2048, N):
        n_local = size // ws
        torch.manual_seed(0)
        A = torch.randn(size, size, dtype=dt, device=dev)
        Bfull = torch.randn(size, size, dtype=dt, device=dev)
        B_local = Bfull[:, rank * n_local:(rank + 1) * n_local].contiguous()
        mc = comm.multicast_like(torch.empty(ws, size, n_local, dtype=dt, device=dev))
        peer = comm.shared_slabs(size, n_local, dt)
        peer.local.zero_()
        torch.cuda.synchronize(); dist.barrier()
        comm.matmul_allgather_fused(A, B_local, mc)
        comm.matmul_allgather_fused(A, B_local, peer)
        torch.cuda.synchronize(); dist.barrier()
        Cg = torch.cat(list(mc.local.unbind(0)), dim=1)
        err = max_ranks(hostlib.parity_error(Cg, A, Bfull, rows=128))
        same = bool(torch.equal(mc.local, peer.local))
        r = {"parity_err": err, "identical_to_peer_store_path": same}
        if size == N:
            r["gemm_alone_ms"] = timed(lambda: b2.matmul(A, B_local))
            for _ in range(3):
                r.setdefault("peer_store_ms", []).append(round(timed(lambda: comm.matmul_allgather_fused(A, B_local, peer)), 4))
                r.setdefault("multicast_store_ms", []).append(round(timed(lambda: comm.matmul_allgather_fused(A, B_local, mc)), 4))
            os.environ["B2MM_M_BLOCKS"] = "2"
            r["multicast_store_tall_tiles_ms"] = round(timed(lambda: comm.matmul_allgather_fused(A, B_local, mc)), 4)
            os.environ.pop("B2MM_M_BLOCKS")
        res[f"size_{size}"] = r
        if rank == 0:
            print(size, json.dumps(r), flush=True)
        comm.release_shared()
except Exception:
    res["error"] = traceback.format_exc()
    print(f"[rank {rank}] FAILED\n{res['error']}", flush=True)
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/nvls_gather_check_{ws}gpu.json", "w"), indent=1)
    print("NVLS GATHER CHECK", "FAILED" if "error" in res else "DONE", flush=True)
try:
    comm.destroy(); hostlib.cleanup_distributed()
except Exception:
    traceback.print_exc()
