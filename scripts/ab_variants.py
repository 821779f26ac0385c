"""Interleaved timing of kernel variants (env overrides) and cuBLAS at one size, several rounds, one process."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
N = int(sys.argv[1]); rounds = int(sys.argv[2]); iters = int(sys.argv[3])
variants = [dict(kv.split("=") for kv in v.split(",") if kv) for v in sys.argv[4:]] or [{}]
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda"); B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
def run(env):
    for k, v in env.items(): os.environ[k] = v
    for _ in range(3): b2.matmul(A, B, out=C) if env is not None else None
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(iters): b2.matmul(A, B, out=C)
    e.record(); torch.cuda.synchronize()
    for k in env: os.environ.pop(k)
    return s.elapsed_time(e) / iters
def run_cublas():
    for _ in range(3): torch.matmul(A, B, out=C)
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(iters): torch.matmul(A, B, out=C)
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters
for r in range(rounds):
    line = [f"N={N} round {r}:"]
    for env in variants:
        ms = run(env); line.append(f"{env or 'default'} {ms*1000:.1f}us {2*N**3/ms/1e9:.0f}TF")
    ms = run_cublas(); line.append(f"cublas {ms*1000:.1f}us {2*N**3/ms/1e9:.0f}TF")
    print("  ".join(line), flush=True)
