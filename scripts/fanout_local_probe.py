"""Where does the fused gather's time go?  ONE GPU, the fan-out epilogue writing to LOCAL buffers (no NVLink):

    python scripts/fanout_local_probe.py [N=16384] [world=8]

Times C_local = A @ B_local (n_local = N / world) as: the plain GEMM (512x256 tiles), the plain GEMM with the
256x256 double-buffered tiles, and b2mm_gemm_allgather_fused with 1, 2, 4, 8 destinations that all live in this
GPU's memory.  If the 8-destination local run already costs what the 8-GPU fused run costs (1.02-1.07 ms at 16k),
the gap to the GEMM (0.73 ms) is the epilogue / TMA store engine, not the links.  ncu-friendly (one process).
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n_local = N // world
dev = "cuda:0"
dt = torch.bfloat16
torch.manual_seed(0)
A = torch.randn(N, N, dtype=dt, device=dev)
B_local = torch.randn(N, n_local, dtype=dt, device=dev)
C_local = torch.empty(N, n_local, dtype=dt, device=dev)
lib = _lib.load()
stream = torch.cuda.current_stream().cuda_stream


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def fused(n_dest):
    bufs = [torch.empty(world, N, n_local, dtype=dt, device=dev) for _ in range(n_dest)]
    ptrs = (ctypes.c_void_p * n_dest)(*[b.data_ptr() for b in bufs])

    def run():
        _lib.check(lib.b2mm_gemm_allgather_fused(A.data_ptr(), B_local.data_ptr(), ptrs, 0, n_dest, N, n_local, N, _lib.BF16,
                                                 None, 0, stream))
    ms = timed(run)
    ok = torch.equal(bufs[-1][0], b2.matmul(A, B_local))
    return ms, ok


rows = []
rows.append(("plain GEMM, default tiles", timed(lambda: b2.matmul(A, B_local, out=C_local)), True))
os.environ["B2MM_M_BLOCKS"] = "1"
rows.append(("plain GEMM, 256x256 double-buffered tiles", timed(lambda: b2.matmul(A, B_local, out=C_local)), True))
os.environ.pop("B2MM_M_BLOCKS")
for n_dest in (1, 2, 4, 8):
    if n_dest > world:
        break
    ms, ok = fused(n_dest)
    rows.append((f"fan-out epilogue, {n_dest} local destination(s)", ms, ok))
os.environ["B2MM_M_BLOCKS"] = "2"
ms, ok = fused(min(8, world))
rows.append((f"fan-out epilogue, {min(8, world)} local destinations, 512x256 tiles", ms, ok))
os.environ.pop("B2MM_M_BLOCKS")
flop = 2.0 * N * N * n_local
for name, ms, ok in rows:
    print(f"{name:62s} {ms * 1000:9.1f} us  {flop / ms / 1e9:7.0f} TFLOP/s  {'ok' if ok else 'WRONG'}", flush=True)
