"""ctypes front end of oracle/gemm_oracle.c (built by oracle/Makefile).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.  Every function names the
reference lines its C counterpart restates.  Inputs/outputs are torch CPU
tensors (torch is here only as an array container with bf16/fp16 storage).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor
from ctypes import c_double, c_int, c_int64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
DTYPE_CODE = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}
_lib = None


def build() -> str:
    """Compile the C restatement (gcc only; seconds)."""
    subprocess.check_call(["make", "-s", "-C", _HERE])
    return LIB_PATH


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = ctypes.CDLL(LIB_PATH)
        lib.oracle_gemm_rows.restype = c_int
        lib.oracle_gemm_rows.argtypes = [c_void_p, c_void_p, c_void_p] + [c_int64] * 10 + [c_int, c_int64, c_int64]
        lib.oracle_gemm.restype = c_int
        lib.oracle_gemm.argtypes = [c_void_p, c_void_p, c_void_p] + [c_int64] * 10 + [c_int]
        lib.oracle_gemm_block_f64.restype = c_int
        lib.oracle_gemm_block_f64.argtypes = [c_void_p, c_void_p, c_void_p] + [c_int64] * 7 + [c_int]
        lib.oracle_calculate_tflops.restype = c_double
        lib.oracle_calculate_tflops.argtypes = [c_int64, c_double, c_int64]
        lib.oracle_local_batch.restype = c_int64
        lib.oracle_local_batch.argtypes = [c_int64, c_int64]
        lib.oracle_col_split.restype = None
        lib.oracle_col_split.argtypes = [c_int64, c_int64, c_int64, ctypes.POINTER(c_int64), ctypes.POINTER(c_int64)]
        lib.oracle_f32_to_bf16.restype = ctypes.c_uint16
        lib.oracle_f32_to_bf16.argtypes = [ctypes.c_float]
        lib.oracle_f32_to_fp16.restype = ctypes.c_uint16
        lib.oracle_f32_to_fp16.argtypes = [ctypes.c_float]
        _lib = lib
    return _lib


def _cpu_rowmajor(t: torch.Tensor) -> torch.Tensor:
    t = t.detach().cpu()
    return t if (t.stride(-1) == 1 and t.is_contiguous()) else t.contiguous()


def gemm(A: torch.Tensor, B: torch.Tensor, threads: int | None = None) -> torch.Tensor:
    """C = A @ B (2-D) or bmm (3-D): exact products, double sum, one rounding.

    Restates torch.matmul at matmul_benchmark.py:62 / matmul_scaling_benchmark.py:92,211
    and torch.bmm at matmul_scaling_benchmark.py:142.
    """
    A, B = _cpu_rowmajor(A), _cpu_rowmajor(B)
    assert A.dtype == B.dtype and A.dim() == B.dim() and A.dim() in (2, 3)
    batched = A.dim() == 3
    if not batched:
        A3, B3 = A.unsqueeze(0), B.unsqueeze(0)
    else:
        A3, B3 = A, B
    b, M, K = A3.shape
    _, K2, N = B3.shape
    assert K == K2 and B3.shape[0] == b
    C = torch.empty((b, M, N), dtype=A.dtype)
    if b and M and N:
        lib = load()
        code = DTYPE_CODE[A.dtype]
        threads = threads or min(os.cpu_count() or 1, 32)
        step = max(1, (M + threads - 1) // threads)

        def run(r0: int) -> int:
            return lib.oracle_gemm_rows(A3.data_ptr(), B3.data_ptr(), C.data_ptr(), M, N, K, K, N, N, b, M * K, K * N,
                                        M * N, code, r0, min(M, r0 + step))

        with ThreadPoolExecutor(max_workers=threads) as pool:
            for rc in pool.map(run, range(0, M, step)):
                if rc != 0:
                    raise RuntimeError(f"oracle_gemm_rows failed with {rc}")
    return C if batched else C[0]


def gemm_block_f64(A: torch.Tensor, B: torch.Tensor, row0: int, rows: int, col0: int, cols: int) -> torch.Tensor:
    """Unrounded double-precision C[row0:row0+rows, col0:col0+cols] of a 2-D product (error denominator)."""
    A, B = _cpu_rowmajor(A), _cpu_rowmajor(B)
    assert A.dim() == 2 and B.dim() == 2 and A.dtype == B.dtype
    K = A.shape[1]
    out = torch.empty((rows, cols), dtype=torch.float64)
    rc = load().oracle_gemm_block_f64(A.data_ptr(), B.data_ptr(), out.data_ptr(), K, A.stride(0), B.stride(0), row0,
                                      rows, col0, cols, DTYPE_CODE[A.dtype])
    if rc != 0:
        raise RuntimeError(f"oracle_gemm_block_f64 failed with {rc}")
    return out


def rel_err(C: torch.Tensor, C64: torch.Tensor) -> float:
    """SURVEY.md 8(c) metric: max |C - C64| / (|C64| + rms(C64)).

    Plain element-wise relative error is meaningless for randn GEMMs (outputs
    near zero have rel-err > 1 even when perfectly rounded).
    """
    C64 = C64.double()
    rms = C64.pow(2).mean().sqrt()
    return ((C.detach().cpu().double() - C64).abs() / (C64.abs() + rms)).max().item()


def reduced_sum_err(got: torch.Tensor, terms64, final_roundings: float = 1.0, ring: bool = False) -> float:
    """Error of an all_reduce(SUM) of per-rank GEMM outputs, in units of the 8(c) tolerance of the dtype.

    What the reference computes (matmul_scaling_benchmark.py:142-150) is: each rank's product ROUNDED to the storage
    type, then a sum that is rounded again.  A correct implementation may therefore differ from the exact sum
    S = sum_r T_r of the unrounded products `terms64` by one 8(c)-grade error per TERM — measured against that term's
    own magnitude, |T_r| + rms(T_r), not against a sum that may have cancelled — plus `final_roundings` such errors
    on the sum itself (1: fp32 adds and one rounding to nearest; 2: one rounding by truncation, as the NVSwitch does;
    with `ring`, world - 1 roundings of partial sums, bounded by the terms' magnitudes).  Returns
        max |got - S| / (sum_r (|T_r| + rms(T_r)) + final_roundings * (|S| + rms(S)))
    which a correct result keeps <= the dtype's tolerance (4e-3 bf16, 5e-4 fp16, 1e-5 fp32)."""
    terms = [t.double() for t in terms64]
    S = sum(terms)
    per_term = sum(t.abs() + t.pow(2).mean().sqrt() for t in terms)
    last = per_term if ring else S.abs() + S.pow(2).mean().sqrt()
    return ((got.detach().cpu().double() - S).abs() / (per_term + final_roundings * last)).max().item()


def gemm_f64(A: torch.Tensor, B: torch.Tensor, threads: int | None = None) -> torch.Tensor:
    """Unrounded double-precision A @ B (2-D or batched 3-D), row blocks in parallel: the exact-sum side of reduced_sum_err."""
    A, B = _cpu_rowmajor(A), _cpu_rowmajor(B)
    if A.dim() == 3:
        return torch.stack([gemm_f64(A[i], B[i], threads) for i in range(A.shape[0])])
    M, N = A.shape[0], B.shape[1]
    threads = threads or min(os.cpu_count() or 1, 32)
    step = max(1, (M + threads - 1) // threads)
    with ThreadPoolExecutor(max_workers=threads) as pool:
        parts = list(pool.map(lambda r0: gemm_block_f64(A, B, r0, min(step, M - r0), 0, N), range(0, M, step)))
    return torch.cat(parts, dim=0)


def calculate_tflops(matrix_size: int, time_seconds: float, num_ops: int = 1) -> float:
    """matmul_benchmark.py:34-37 / matmul_scaling_benchmark.py:63-67."""
    return float(load().oracle_calculate_tflops(matrix_size, time_seconds, num_ops))


def local_batch(batch_size: int, world_size: int) -> int:
    """matmul_scaling_benchmark.py:111."""
    return int(load().oracle_local_batch(batch_size, world_size))


def col_split(matrix_size: int, world_size: int, rank: int) -> tuple[int, int]:
    """matmul_scaling_benchmark.py:179-181 (last rank takes the remainder)."""
    s, e = c_int64(), c_int64()
    load().oracle_col_split(matrix_size, world_size, rank, ctypes.byref(s), ctypes.byref(e))
    return s.value, e.value
