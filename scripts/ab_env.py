"""Interleaved A/B of one environment knob of the GEMM, in one process (the knobs are read at every call).

    python scripts/ab_env.py B2MM_SPLIT_TAIL 0 1 [size=16384] [launches=60] [rounds=4]
"""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2

knob, va, vb = sys.argv[1], sys.argv[2], sys.argv[3]
N = int(sys.argv[4]) if len(sys.argv) > 4 else 16384
L = int(sys.argv[5]) if len(sys.argv) > 5 else 60
R = int(sys.argv[6]) if len(sys.argv) > 6 else 4
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
for _ in range(150):  # into the sustained power state first: both arms are then measured on the same footing
    b2.matmul(A, B, out=C)
torch.cuda.synchronize()
res = {va: [], vb: []}
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for r in range(R):
    for v in ((va, vb) if r % 2 == 0 else (vb, va)):
        os.environ[knob] = v
        for _ in range(5):
            b2.matmul(A, B, out=C)
        s.record()
        for _ in range(L):
            b2.matmul(A, B, out=C)
        e.record()
        torch.cuda.synchronize()
        res[v].append(round(2.0 * N ** 3 / (s.elapsed_time(e) / L) / 1e9, 1))
out = {"knob": knob, "size": N, "launches_per_round": L, "tflops": res, "median": {k: statistics.median(v) for k, v in res.items()}}
print(json.dumps(out), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/r2_ab_env.jsonl", "a") as f:
    f.write(json.dumps(out) + "\n")
