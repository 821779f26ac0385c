"""Per-tile start-time trace of the persistent GEMM: how far do clusters drift apart, and which SMs are slow?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
from pytorch_distributed_matmul_benchmark_b200 import _lib
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
torch.manual_seed(0)
A = torch.randn(N, N, dtype=torch.bfloat16, device="cuda"); B = torch.randn(N, N, dtype=torch.bfloat16, device="cuda")
C = torch.empty(N, N, dtype=torch.bfloat16, device="cuda")
b2.matmul(A, B, out=C); torch.cuda.synchronize()
W, rows = 256, 400  # one column per SM id
buf = torch.zeros(rows * W, dtype=torch.int64, device="cuda")
_lib.load().b2mm_debug_set_trace(buf.data_ptr())
s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
s.record(); b2.matmul(A, B, out=C); e.record(); torch.cuda.synchronize()
_lib.load().b2mm_debug_set_trace(None)
print(f"traced launch: {s.elapsed_time(e):.3f} ms")
v = buf.view(rows, W).cpu()
sms = [i for i in range(W) if v[0][i] > 0]
t = v[1:, sms].double()
cnt = (t > 0).sum(dim=0)
print("SMs traced:", len(sms), " tiles per SM: min", int(cnt.min()), "max", int(cnt.max()), " total CTA-tiles", int(cnt.sum()))
t0 = t[0].min()
iters = int((t > 0).any(dim=1).sum())
for i in list(range(0, min(iters, 6))) + list(range(8, iters, 8)):
    row = t[i][t[i] > 0]
    print(f"iter {i:3d}: n={row.numel():3d} start min {1e-3*(row.min()-t0):9.1f} us  max {1e-3*(row.max()-t0):9.1f} us  spread {1e-3*(row.max()-row.min()):7.1f} us")
last = torch.stack([t[int(cnt[j]) - 1, j] for j in range(len(sms))])
print(f"last tile start: min {1e-3*(last.min()-t0):.1f} us max {1e-3*(last.max()-t0):.1f} us")
d1 = (t[1] - t[0]) * 1e-3
print("smid:tile0 duration us,tiles")
print(" ".join(f"{sms[j]}:{d1[j]:.1f},{int(cnt[j])}" for j in range(0, len(sms), 2)))
