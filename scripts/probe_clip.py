"""Which columns past N does a GEMM touch?  (diagnostic for the TMA store clipping of ragged N)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2

dev = "cuda:0"
for dt in (torch.bfloat16, torch.float16, torch.float32):
    for N in (1, 2, 3, 5, 7, 9, 21, 22, 23, 24, 25, 33, 63, 65, 77, 127, 129, 255, 257):
        for K in (16, 10):
            M = 40
            ld = (N + 31) // 8 * 8
            A = torch.randn(M, 32, dtype=dt, device=dev)[:, :K]
            B = torch.randn(K, ld, dtype=dt, device=dev)[:, :N]
            out = torch.full((M + 8, ld), 7.0, dtype=dt, device=dev)
            b2.matmul(A, B, out=out[:M, :N])
            torch.cuda.synchronize()
            touched_cols = (out[:M, N:] != 7.0).any(0).nonzero().flatten().tolist()
            touched_rows = (out[M:] != 7.0).any(1).nonzero().flatten().tolist()
            ref = A.double() @ B.double()
            err = ((out[:M, :N].double() - ref).abs() / (ref.abs() + ref.pow(2).mean().sqrt())).max().item()
            if touched_cols or touched_rows or err > 5e-3:
                print(dt, "N", N, "K", K, "cols past N touched:", [N + c for c in touched_cols], "rows past M:", touched_rows, "err", err)
print("done")
