"""Does write-combined pinned host memory lift the host->device rate of the end-to-end path?  (all ranks at once)

Per rank: 1 GiB up (A and B of the 16384 problem) and 0.5 GiB down at the same time, as bench.py's copy_peak does, with
the upload source in (a) ordinary pinned memory and (b) cudaHostAllocWriteCombined memory (no CPU cache snooping on
the DMA reads).  One process per GPU; run with and without torch.distributed.run.
"""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
dev = torch.device("cuda", torch.cuda.current_device())
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
N = 16384
cudart = ctypes.CDLL("libcudart.so.12")


def host_alloc(nbytes: int, flags: int) -> torch.Tensor:
    p = ctypes.c_void_p()
    rc = cudart.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    assert rc == 0, f"cudaHostAlloc({flags}) -> {rc}"
    return torch.frombuffer((ctypes.c_uint8 * nbytes).from_address(p.value), dtype=torch.uint8)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def mx(x):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


dA = torch.empty(2 * N * N * 2, dtype=torch.uint8, device=dev)
dC = torch.empty(N * N * 2, dtype=torch.uint8, device=dev)
hC = torch.empty(N * N * 2, dtype=torch.uint8, pin_memory=True)
up, down, main = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.current_stream()
res = {"world": world}
for name, flags in (("pinned", 0), ("write_combined", 4), ("pinned_again", 0)):
    h = host_alloc(2 * N * N * 2, flags)
    res[name + "_is_pinned"] = bool(h.is_pinned())

    def once(both):
        up.wait_stream(main)
        down.wait_stream(main)
        with torch.cuda.stream(up):
            dA.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(down):
                hC.copy_(dC, non_blocking=True)
        main.wait_stream(up)
        main.wait_stream(down)

    for both in (False, True):
        once(both)
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(main)
        for _ in range(4):
            once(both)
        e.record(main)
        barrier()
        ms = mx(s.elapsed_time(e)) / 4
        res[f"{name}_{'up+down' if both else 'up_only'}_ms"] = round(ms, 3)
        res[f"{name}_{'up+down' if both else 'up_only'}_h2d_gbs_per_gpu"] = round(2 * N * N * 2 / ms / 1e6, 2)
    del h
if rank == 0:
    print(json.dumps(res), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/r2_wc_probe_{world}gpu.json", "w"), indent=1)
if world > 1:
    dist.destroy_process_group()
