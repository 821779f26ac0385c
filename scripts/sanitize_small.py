"""A handful of tiny GEMMs for compute-sanitizer (memcheck): every kernel variant, ragged edges, bmm, fp32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2
torch.manual_seed(0)
for dt, shape in ((torch.bfloat16, (520, 264, 136)), (torch.float16, (256, 512, 64)), (torch.float32, (200, 136, 264))):
    A = torch.randn(shape[0], shape[2], dtype=dt, device="cuda"); B = torch.randn(shape[2], shape[1], dtype=dt, device="cuda")
    C = b2.matmul(A, B); torch.cuda.synchronize()
    ref = A.double() @ B.double()
    print(dt, shape, "max err", ((C.double() - ref).abs() / (ref.abs() + ref.pow(2).mean().sqrt())).max().item(), flush=True)
A = torch.randn(2, 264, 72, dtype=torch.bfloat16, device="cuda"); B = torch.randn(2, 72, 136, dtype=torch.bfloat16, device="cuda")
print("bmm", (b2.bmm(A, B).double() - A.double() @ B.double()).abs().max().item())
for env in ({"B2MM_M_BLOCKS": "1"}, {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "128"}, {"B2MM_SCHED": "static"}, {"B2MM_CTA_GROUP": "1"}):
    os.environ.update(env)
    A = torch.randn(600, 136, dtype=torch.bfloat16, device="cuda"); B = torch.randn(136, 520, dtype=torch.bfloat16, device="cuda")
    C = b2.matmul(A, B); torch.cuda.synchronize()
    print(env, (C.double() - A.double() @ B.double()).abs().max().item(), flush=True)
    for k in env: os.environ.pop(k)
print("SANITIZE SCRIPT DONE")
