"""Run the b2mm GEMM a few times at one size (target for ncu / quick timing)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dtype = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}[sys.argv[2] if len(sys.argv) > 2 else "bfloat16"]
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
warmup = int(sys.argv[4]) if len(sys.argv) > 4 else 1
torch.manual_seed(0)
A = torch.randn(N, N, dtype=dtype, device="cuda")
B = torch.randn(N, N, dtype=dtype, device="cuda")
C = torch.empty(N, N, dtype=dtype, device="cuda")
for _ in range(warmup):
    b2.matmul(A, B, out=C)
torch.cuda.synchronize()
s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(iters):
    b2.matmul(A, B, out=C)
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / iters
print(f"N={N} {dtype} iters={iters}: {ms:.3f} ms {2*N**3/ms/1e9:.1f} TFLOPS "
      f"[GROUP_M={os.environ.get('B2MM_GROUP_M','8')} BLOCK_N={os.environ.get('B2MM_BLOCK_N','auto')}]", flush=True)
