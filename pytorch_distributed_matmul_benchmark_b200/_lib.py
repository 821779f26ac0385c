"""ctypes binding of libb2mm.so — the C-ABI library declared in include/b2mm.h.

The product path has no fallback: if the shared library is missing or a call
fails, an exception is raised; nothing here ever routes to torch.matmul or to
the CPU oracle.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_size_t, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2mm.so")

OK = 0
ERR_BAD_ARG = 1
ERR_UNSUPPORTED = 2
ERR_CUDA = 3
ERR_NCCL = 4
ERR_WORKSPACE = 5

BF16, FP16, FP32 = 0, 1, 2
UNIQUE_ID_BYTES = 128
NOALLOC_SCRATCH_BYTES = 32 << 20

# name -> (restype, argtypes); must list every symbol include/b2mm.h declares
# (tests/test_abi.py checks the two against each other).
SIGNATURES = {
    "b2mm_version": (c_int, []),
    "b2mm_last_error": (c_char_p, []),
    "b2mm_launch_count": (c_uint64, []),
    "b2mm_last_kernel": (c_char_p, []),
    "b2mm_debug_set_trace": (None, [c_void_p]),
    "b2mm_debug_max_active_clusters": (c_int, [c_int]),
    "b2mm_debug_occupy_sms": (c_int, [c_int, ctypes.c_double, c_void_p]),
    "b2mm_gemm_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int64, c_int]),
    "b2mm_gemm": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                          c_int64, c_int64, c_int64, c_int64, c_int, c_void_p, c_size_t, c_void_p]),
    "b2mm_gemm_fp32_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                                       c_void_p]),
    "b2mm_comm_unique_id": (c_int, [c_void_p]),
    "b2mm_comm_init": (c_int, [c_void_p, c_int, c_int, ctypes.POINTER(c_void_p)]),
    "b2mm_comm_destroy": (c_int, [c_void_p]),
    "b2mm_all_reduce_sum": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b2mm_all_gather": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b2mm_ipc_export": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_int64)]),
    "b2mm_ipc_import": (c_int, [c_void_p, c_int64, ctypes.POINTER(c_void_p)]),
    "b2mm_ipc_release": (c_int, [c_void_p]),
    "b2mm_comm_barrier": (c_int, [c_void_p, c_void_p]),
    "b2mm_comm_barrier_signals": (c_int, [c_void_p, ctypes.POINTER(c_void_p), c_void_p]),
    "b2mm_gemm_allgather_fused": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_int, c_int64, c_int64,
                                          c_int64, c_int, c_void_p, c_size_t, c_void_p]),
    "b2mm_gemm_allgather_nvls": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int64, c_int64, c_int64, c_int,
                                         c_void_p, c_size_t, c_void_p]),
    "b2mm_all_reduce_p2p": (c_int, [c_void_p, ctypes.POINTER(c_void_p), c_int64, c_int, c_void_p]),
    "b2mm_all_reduce_p2p_sync": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p), c_int64, c_int, c_void_p]),
    "b2mm_gemm_allreduce_p2p": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int64, c_int64, c_int,
                                        c_int, c_void_p, c_size_t, c_void_p, c_void_p]),
    "b2mm_gemm_allreduce_fused": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p),
                                          c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p]),
    "b2mm_all_reduce_nvls_sync": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int64, c_int, c_void_p]),
    "b2mm_gemm_allreduce_nvls": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_void_p,
                                         c_int64, c_int64, c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p]),
    "b2mm_gemm_allgather_overlap": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                            c_int64, c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p]),
    "b2mm_gemm_allreduce_overlap": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int, c_int,
                                            c_void_p, c_size_t, c_void_p, c_void_p]),
}

_lib = None


class B2mmError(RuntimeError):
    """A libb2mm entry point returned a non-zero status."""

    def __init__(self, status: int, message: str):
        super().__init__(f"b2mm status {status}: {message}")
        self.status = status


def load() -> ctypes.CDLL:
    """Load libb2mm.so (once) and attach prototypes. Raises if it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.join(_HERE, 'csrc')}` "
                "(or __graft_entry__.build()). There is no fallback path.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(status: int) -> None:
    if status != OK:
        msg = load().b2mm_last_error()
        raise B2mmError(status, msg.decode() if msg else "")


def launch_count() -> int:
    return int(load().b2mm_launch_count())
