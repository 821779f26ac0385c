"""Does a power-capped GEMM lose throughput when it gets fewer SMs?  (decides whether 4-CTA clusters, which can
reach only 132 of the 148 SMs, could pay for themselves through TMA multicast)

Runs the 16384^3 bf16 GEMM back to back for ~2.5 s on all 148 SMs, then with 4 / 8 / 16 CTA pairs parked
(140 / 132 / 116 SMs), NVML sampled every 2 ms.  If the part sat on its minimum voltage, power would be linear in
the clock and throughput at the cap independent of the SM count; the numbers say how far from that it is.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib
from bench import NvmlSampler

N = 16384
SECONDS = 2.5
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
lib = _lib.load()
nv = NvmlSampler(0)
side = torch.cuda.Stream()
rows = []
for rep in range(2):
    for pairs in (0, 4, 8, 16):
        b2.matmul(A, B, out=C)
        torch.cuda.synchronize()
        time.sleep(2.0)
        if pairs:
            _lib.check(lib.b2mm_debug_occupy_sms(pairs, SECONDS + 1.5, side.cuda_stream))
            time.sleep(0.2)  # the parked CTAs are resident before the first GEMM asks for SMs
        n = int(SECONDS / 6.0e-3)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        t0 = time.time()
        ev[0].record()
        for i in range(n):
            b2.matmul(A, B, out=C)
            ev[i + 1].record()
        torch.cuda.current_stream().synchronize()
        t1 = time.time()
        side.synchronize()
        per = [ev[i].elapsed_time(ev[i + 1]) for i in range(n)]
        tail = per[len(per) // 2:]
        ms = sum(tail) / len(tail)
        w = nv.window(t0 + (t1 - t0) / 2, t1 - 0.05) if nv.ok else {}
        rows.append({"sms": 148 - 2 * pairs, "rep": rep, "first_launch_ms": round(per[0], 3), "sustained_ms": round(ms, 4),
                     "sustained_tflops": round(2.0 * N ** 3 / ms / 1e9, 1), **w})
        print(json.dumps(rows[-1]), flush=True)
nv.stop()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/r2_fewer_sms_probe.json", "w"), indent=1)
