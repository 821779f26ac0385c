"""Clock, power and throughput of the 16384^3 bf16 GEMM as a function of how many bytes it pulls from L2 per flop.

Each variant is the same arithmetic (same tcgen05 instruction shape, same flops) with a different tile, i.e. a
different L2->SM byte count per launch; every variant runs back to back for >= 2 s while NVML is sampled every 2 ms.
Together with the ncu byte counts of the same variants (scripts/gpu_r2_b.sh) this is the "does the clock move when
L2->SM traffic falls" table of DESIGN.md.  cuBLAS (torch.matmul) runs last as the kernel to beat.
"""
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib
from bench import NvmlSampler

N = 16384
SECONDS = float(sys.argv[1]) if len(sys.argv) > 1 else 2.5
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
nv = NvmlSampler(0)
VARIANTS = [("512x256 pair tile (default), 48 B/clk/SM", {}),
            ("256x256 pair tile, 64 B/clk/SM", {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "256"}),
            ("256x128 pair tile, 96 B/clk/SM", {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "128"}),
            ("128x256 single-CTA tile (no pairs), 96 B/clk/SM", {"B2MM_CTA_GROUP": "1"}),
            ("cuBLAS (torch.matmul)", None)]
rows = []
for rep in range(2):
    for name, env in VARIANTS:
        for k in ("B2MM_M_BLOCKS", "B2MM_BLOCK_N", "B2MM_CTA_GROUP"):
            os.environ.pop(k, None)
        if env:
            os.environ.update(env)
        fn = (lambda: torch.matmul(A, B, out=C)) if env is None else (lambda: b2.matmul(A, B, out=C))
        fn()
        torch.cuda.synchronize()
        time.sleep(2.0)  # cool down to the same starting point
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = int(SECONDS / 6.0e-3)
        t0 = time.time()
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        t1 = time.time()
        ms = s.elapsed_time(e) / n
        w = nv.window(t0 + 0.8, t1 - 0.05) if nv.ok else {}
        rows.append({"variant": name, "rep": rep, "ms": round(ms, 4), "tflops": round(2.0 * N ** 3 / ms / 1e9, 1),
                     "kernel": _lib.load().b2mm_last_kernel().decode() if env is not None else "cuBLAS", **w})
        print(json.dumps(rows[-1]), flush=True)
nv.stop()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/r2_power_table.json", "w"), indent=1)
