"""fp32 (3xTF32) with 64-byte k-blocks (SWIZZLE_64B, B2MM_FP32_K16=1) against the default 128-byte ones: parity and time."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib

res = {}
torch.manual_seed(0)
for M, N, K in ((640, 520, 1096), (300, 264, 76), (2048, 2048, 2048)):
    A = torch.randn(M, K, dtype=torch.float32, device="cuda")
    B = torch.randn(K, N, dtype=torch.float32, device="cuda")
    ref = A.double() @ B.double()
    rms = ref.pow(2).mean().sqrt()
    for v in ("0", "1"):
        os.environ["B2MM_FP32_K16"] = v
        C = b2.matmul(A, B)
        torch.cuda.synchronize()
        res[f"err_{M}x{N}x{K}_k16={v}"] = ((C.double() - ref).abs() / (ref.abs() + rms)).max().item()
    print(json.dumps(res), flush=True)
print(_lib.load().b2mm_last_kernel().decode())
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for n in (8192, 16384):
    A = torch.randn(n, n, dtype=torch.float32, device="cuda")
    B = torch.randn(n, n, dtype=torch.float32, device="cuda")
    C = torch.empty(n, n, dtype=torch.float32, device="cuda")
    out = {"0": [], "1": []}
    for r in range(3):
        for v in ("0", "1"):
            os.environ["B2MM_FP32_K16"] = v
            b2.matmul(A, B, out=C)
            s.record()
            for _ in range(5):
                b2.matmul(A, B, out=C)
            e.record()
            torch.cuda.synchronize()
            out[v].append(round(2.0 * n ** 3 / (s.elapsed_time(e) / 5) / 1e9, 1))
    res[f"tflops_{n}"] = out
    print(n, json.dumps(out), flush=True)
    del A, B, C
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/r2_fp32_k16_probe.json", "w"), indent=1)
