"""Programmatic dependent launch (B2MM_PDL=1): results must not change.  Chains of dependent GEMMs (each reads the
previous one's output) and write-after-write on one C, compared bit for bit with the serialised launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2

dev = "cuda:0"
ok = True
for N in (256, 1024, 2048, 4096):
    for dt in (torch.bfloat16, torch.float32):
        torch.manual_seed(N)
        X = torch.randn(N, N, dtype=dt, device=dev)
        W = (torch.randn(N, N, dtype=dt, device=dev) / N ** 0.5)
        outs = {}
        for pdl in ("0", "1"):
            os.environ["B2MM_PDL"] = pdl
            bufs = [torch.empty(N, N, dtype=dt, device=dev) for _ in range(2)]
            cur = X
            for i in range(24):  # a dependency chain: every GEMM reads what the previous one wrote
                b2.matmul(cur, W, out=bufs[i % 2])
                cur = bufs[i % 2]
            C = torch.empty(N, N, dtype=dt, device=dev)
            for i in range(24):  # write-after-write on one buffer, last writer wins
                b2.matmul(X if i % 2 else W, W, out=C)
            torch.cuda.synchronize()
            outs[pdl] = (cur.clone(), C.clone())
        same = torch.equal(outs["0"][0], outs["1"][0]) and torch.equal(outs["0"][1], outs["1"][1])
        finite = bool(torch.isfinite(outs["1"][0]).all())
        print(N, dt, "identical" if same else "DIFFERENT", "finite" if finite else "non-finite", flush=True)
        ok = ok and same and finite
os.environ.pop("B2MM_PDL", None)
print("PDL CHECK", "OK" if ok else "FAILED")
