"""matrix_parallel's slab GEMMs (A N x N, B_local N x N/ws) against cuBLAS on the same tensors, interleaved.

The narrow slabs are where wave quantisation bites: ws = 8 at N = 4096 is 32 tall tiles (or 64 tiles of 256 x 128) for
74 CTA pairs.  Rows: shape, the tile variant the library picked (or was forced to), microseconds per GEMM back to
back (median of rounds), and cuBLAS (torch.matmul, out=) in the same rounds.
"""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib

dt = torch.bfloat16
SHAPES = [(4096, 512), (4096, 1024), (4096, 2048), (8192, 1024), (8192, 2048), (8192, 4096), (16384, 2048), (16384, 4096)]
VARIANTS = [("auto", {}), ("256x128", {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "128"}), ("256x256", {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "256"}),
            ("512x256", {"B2MM_M_BLOCKS": "2"})]
rows = []
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, launches):
    for _ in range(5):
        fn()
    s.record()
    for _ in range(launches):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / launches * 1e3  # us


for N, n in SHAPES:
    torch.manual_seed(0)
    A = torch.randn(N, N, dtype=dt, device="cuda")
    B = torch.randn(N, n, dtype=dt, device="cuda")
    C = torch.empty(N, n, dtype=dt, device="cuda")
    launches = max(20, int(40000 / (2.0 * N * N * n / 1.4e15 * 1e6)))  # ~40 ms per measurement
    res = {name: [] for name, _ in VARIANTS}
    res["cublas"] = []
    kern = {}
    for r in range(3):
        for name, env in VARIANTS:
            for k in ("B2MM_M_BLOCKS", "B2MM_BLOCK_N"):
                os.environ.pop(k, None)
            os.environ.update(env)
            res[name].append(timed(lambda: b2.matmul(A, B, out=C), launches))
            kern[name] = _lib.load().b2mm_last_kernel().decode().split(" tile")[0].split(", ")[-1]
        res["cublas"].append(timed(lambda: torch.matmul(A, B, out=C), launches))
    for k in ("B2MM_M_BLOCKS", "B2MM_BLOCK_N"):
        os.environ.pop(k, None)
    row = {"N": N, "n_local": n, "world": N // n, "launches": launches, "auto_picks": kern["auto"],
           "us": {k: round(statistics.median(v), 2) for k, v in res.items()}}
    row["auto_vs_cublas"] = round(row["us"]["cublas"] / row["us"]["auto"], 4)
    row["tflops_auto"] = round(2.0 * N * N * n / row["us"]["auto"] / 1e6, 1)
    rows.append(row)
    print(json.dumps(row), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/r2_narrow_slab_ab.json", "w"), indent=1)
