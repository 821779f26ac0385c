"""Interleaved A/B timing of two builds of libb2mm.so in one process (same box, same thermal state).

    python scripts/ab_lib.py <old.so> <new.so> [N=16384] [rounds=6] [iters=10] [dtype=bf16]
"""
import ctypes
import sys

import torch

old_path, new_path = sys.argv[1], sys.argv[2]
N = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
rounds = int(sys.argv[4]) if len(sys.argv) > 4 else 6
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 10
dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[sys.argv[6] if len(sys.argv) > 6 else "bf16"]
code = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}[dt]
i64, vp = ctypes.c_int64, ctypes.c_void_p


def load(path):
    lib = ctypes.CDLL(path)
    lib.b2mm_gemm.restype = ctypes.c_int
    lib.b2mm_gemm.argtypes = [vp, vp, vp] + [i64] * 10 + [ctypes.c_int, vp, ctypes.c_size_t, vp]
    lib.b2mm_gemm_workspace_bytes.restype = ctypes.c_size_t
    lib.b2mm_gemm_workspace_bytes.argtypes = [i64] * 4 + [ctypes.c_int]
    lib.b2mm_last_error.restype = ctypes.c_char_p
    return lib


libs = {"old": load(old_path), "new": load(new_path)}
torch.manual_seed(0)
A = torch.randn(N, N, dtype=dt, device="cuda")
B = torch.randn(N, N, dtype=dt, device="cuda")
C = torch.empty(N, N, dtype=dt, device="cuda")
need = max(l.b2mm_gemm_workspace_bytes(N, N, N, 1, code) for l in libs.values())
ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
stream = torch.cuda.current_stream().cuda_stream


def run(lib):
    rc = lib.b2mm_gemm(A.data_ptr(), B.data_ptr(), C.data_ptr(), N, N, N, N, N, N, 1, 0, 0, 0, code, ws.data_ptr(), need, stream)
    assert rc == 0, lib.b2mm_last_error()


outs = {}
for name, lib in libs.items():
    for _ in range(3):
        run(lib)
    torch.cuda.synchronize()
    outs[name] = C.clone()
print("bit-identical results:", torch.equal(outs["old"], outs["new"]))
tot = {k: [] for k in libs}
for r in range(rounds):
    for name, lib in libs.items():
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            run(lib)
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / iters
        tot[name].append(ms)
        print(f"round {r} {name}: {ms:.4f} ms  {2 * N**3 / ms / 1e9:.1f} TFLOP/s", flush=True)
for name, v in tot.items():
    v = sorted(v)
    print(f"{name}: median {v[len(v)//2]:.4f} ms  best {v[0]:.4f} ms")
