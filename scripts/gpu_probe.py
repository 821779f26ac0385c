"""First-contact GPU probe: run the GEMM on a ladder of shapes and print error statistics.

Not a test: a diagnostic that prints as much as possible per gpurun round trip.
"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pytorch_distributed_matmul_benchmark_b200 as b2

torch.manual_seed(0)
dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0), torch.cuda.get_device_properties(0).multi_processor_count, flush=True)

def check(M, N, K, dtype, batch=None, tag=""):
    shape_a = (M, K) if batch is None else (batch, M, K)
    shape_b = (K, N) if batch is None else (batch, K, N)
    A = torch.randn(shape_a, dtype=dtype, device=dev)
    B = torch.randn(shape_b, dtype=dtype, device=dev)
    C = b2.matmul(A, B)
    torch.cuda.synchronize()
    ref = torch.matmul(A.double(), B.double())
    rms = ref.pow(2).mean().sqrt().item()
    err = ((C.double() - ref).abs() / (ref.abs() + rms)).max().item()
    Ct = torch.matmul(A, B)
    err_t = ((Ct.double() - ref).abs() / (ref.abs() + rms)).max().item()
    mism = (C != Ct).float().mean().item()
    bad = ((C.double() - ref).abs() > 0.05 * rms)
    print(f"{tag} M={M} N={N} K={K} b={batch} {dtype}: e={err:.3e} (cublas e={err_t:.3e}) frac!=cublas={mism:.4f} bad={bad.float().mean().item():.4f}", flush=True)
    if bad.any():
        idx = bad.nonzero()[:8].tolist()
        print("   first bad idx:", idx, flush=True)
        if batch is None:
            rows = bad.any(dim=1).nonzero().flatten()
            cols = bad.any(dim=0).nonzero().flatten()
            print(f"   bad rows: n={rows.numel()} min={rows.min().item()} max={rows.max().item()}; bad cols: n={cols.numel()} min={cols.min().item()} max={cols.max().item()}", flush=True)
    return err

def bench(N, dtype, iters=20, warmup=5):
    A = torch.randn(N, N, dtype=dtype, device=dev)
    B = torch.randn(N, N, dtype=dtype, device=dev)
    C = torch.empty(N, N, dtype=dtype, device=dev)
    for f, name in ((lambda: b2.matmul(A, B, out=C), "b2mm"), (lambda: torch.matmul(A, B, out=C), "cublas")):
        for _ in range(warmup): f()
        torch.cuda.synchronize()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters): f()
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / iters
        print(f"bench {name} N={N} {dtype}: {ms:.3f} ms  {2*N**3/ms/1e9:.1f} TFLOPS", flush=True)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
for env in ("B2MM_CTA_GROUP", "B2MM_BLOCK_N", "B2MM_GROUP_M"):
    if os.environ.get(env): print(env, "=", os.environ[env])
if which in ("all", "small"):
    check(128, 256, 64, torch.bfloat16, tag="1tile1k")
    check(256, 256, 64, torch.bfloat16, tag="1tile")
    check(256, 256, 512, torch.bfloat16, tag="kloop")
    check(512, 512, 1024, torch.bfloat16)
    check(1024, 2048, 4096, torch.bfloat16)
    check(1000, 1096, 520, torch.bfloat16, tag="ragged")
    check(2048, 2048, 2048, torch.float16)
    check(512, 768, 1024, torch.bfloat16, batch=3, tag="bmm")
if which in ("all", "big"):
    check(4096, 4096, 4096, torch.bfloat16)
    check(8192, 8192, 8192, torch.bfloat16)
if which in ("all", "fp32"):
    check(256, 256, 256, torch.float32, tag="fp32")
    check(1024, 1024, 4096, torch.float32, tag="fp32")
    check(2048, 2048, 16384, torch.float32, tag="fp32-k16k")
if which in ("all", "bench"):
    for n in (4096, 8192, 16384):
        bench(n, torch.bfloat16)
    bench(8192, torch.float32, iters=5, warmup=2)
print("PROBE DONE", flush=True)
