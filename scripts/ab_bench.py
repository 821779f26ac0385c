"""Interleaved A/B timing of b2mm vs cuBLAS in one process (same thermal state)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
N = int(sys.argv[1]); rounds = int(sys.argv[2]); iters = int(sys.argv[3])
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda"); B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
fs = {"b2mm": lambda: b2.matmul(A, B, out=C), "cublas": lambda: torch.matmul(A, B, out=C)}
for f in fs.values():
    for _ in range(3): f()
torch.cuda.synchronize()
for r in range(rounds):
    for name, f in fs.items():
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters): f()
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / iters
        print(f"round {r} {name} N={N}: {ms:.3f} ms {2*N**3/ms/1e9:.1f} TFLOPS", flush=True)
