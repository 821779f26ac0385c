"""One GEMM shape, b2mm or cuBLAS, optional padded leading dimensions (ncu target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
impl, M, N, K, pad, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
torch.manual_seed(0)
A = torch.randn(M, K + pad, dtype=torch.bfloat16, device="cuda")[:, :K]
B = torch.randn(K, N + pad, dtype=torch.bfloat16, device="cuda")[:, :N]
C = torch.empty(M, N + pad, dtype=torch.bfloat16, device="cuda")[:, :N]
f = (lambda: b2.matmul(A, B, out=C)) if impl == "b2mm" else (lambda: torch.matmul(A, B, out=C))
f(); torch.cuda.synchronize()
s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(iters): f()
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / iters
print(f"{impl} M={M} N={N} K={K} pad={pad}: {ms:.3f} ms {2*M*N*K/ms/1e9:.1f} TFLOPS env={ {k:v for k,v in os.environ.items() if k.startswith('B2MM')} }", flush=True)
