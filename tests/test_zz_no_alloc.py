"""b2mm_gemm_fp32_noalloc: the fp32 GEMM for hosts that cannot allocate device memory.

Some users of this library have no way to hand b2mm_gemm the fp32 workspace it asks for: the process
gave all GPU memory to its own pool at start-up, or it sits in a loop where cudaMalloc's device-wide
synchronisation is not acceptable, or it is a bare C/ctypes caller with nothing but the buffers it was
given.  For them the split scratch is a fixed 32 MiB array inside the library.  This test drives that
entry point through the C ABI exactly as such a caller would (inputs and outputs only), twice in a row
with different shapes and data — the array is shared between calls and keeps the previous call's
contents, so the second call is the one that would expose stale scratch — and compares each result
with the ordinary entry point (b2mm_gemm with caller-provided workspace).

Collected last (file name) so that a run stopping at the first failure has already run everything else;
it depends on no other test and passes alone in a fresh process.
"""
import pytest
import torch

from conftest import TOL, rel_err

pytestmark = pytest.mark.gpu

from pytorch_distributed_matmul_benchmark_b200 import _lib  # noqa: E402

DEV = "cuda:0"


def _reference_entry_point(A, B):
    """C = A @ B through the existing entry point, with workspace allocated by the caller (torch)."""
    lib = _lib.load()
    M, K = A.shape
    N = B.shape[1]
    C = torch.empty(M, N, dtype=torch.float32, device=A.device)
    need = lib.b2mm_gemm_workspace_bytes(M, N, K, 1, _lib.FP32)
    ws = torch.empty(need, dtype=torch.uint8, device=A.device)
    _lib.check(lib.b2mm_gemm(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, A.stride(0), B.stride(0), C.stride(0),
                             1, 0, 0, 0, _lib.FP32, ws.data_ptr(), need, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return C


def _noalloc(A, B):
    lib = _lib.load()
    M, K = A.shape
    N = B.shape[1]
    C = torch.full((M, N), float("nan"), dtype=torch.float32, device=A.device)
    _lib.check(lib.b2mm_gemm_fp32_noalloc(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, A.stride(0), B.stride(0),
                                          C.stride(0), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return C


def test_two_calls_in_a_row_match_the_workspace_entry_point():
    torch.cuda.set_device(0)
    # call 1 leaves a 1000-deep split of other data in the scratch; call 2 is smaller in M, larger in N,
    # so its operands land on top of call 1's at different offsets
    for seed, (M, N, K) in ((1, (384, 520, 1000)), (2, (200, 1032, 776))):
        torch.manual_seed(seed)
        A = torch.randn(M, K, dtype=torch.float32, device=DEV)
        B = torch.randn(K, N, dtype=torch.float32, device=DEV)
        got = _noalloc(A, B)
        want = _reference_entry_point(A, B)
        assert torch.isfinite(got).all()
        assert rel_err(got, want.double()) <= TOL[torch.float32]
        assert rel_err(got, A.double().cpu() @ B.double().cpu()) <= TOL[torch.float32]


def test_the_call_itself_takes_no_device_memory():
    """The property in the name: with inputs and output already in hand, a call leaves the device's free memory where it was.

    Free memory is read from the driver (cudaMemGetInfo), so an allocation made behind torch's back would show.  The
    first call on a device is the set-up call (it resolves the array's address, and with lazy module loading that is
    also when the driver makes room for the library's code and its 32 MiB array), so it is made before the measurement;
    the calls measured are the ones a caller repeats."""
    torch.cuda.set_device(0)
    lib = _lib.load()
    torch.manual_seed(3)
    M, N, K = 256, 264, 520
    A = torch.randn(M, K, dtype=torch.float32, device=DEV)
    B = torch.randn(K, N, dtype=torch.float32, device=DEV)
    C = torch.empty(M, N, dtype=torch.float32, device=DEV)
    stream = torch.cuda.current_stream().cuda_stream

    def call():
        _lib.check(lib.b2mm_gemm_fp32_noalloc(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, K, N, N, stream))
        torch.cuda.synchronize()

    call()  # set-up call
    free_before, _ = torch.cuda.mem_get_info()
    reserved_before = torch.cuda.memory_reserved()
    for _ in range(3):
        call()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, f"free device memory moved by {free_before - free_after} bytes across three calls"
    assert torch.cuda.memory_reserved() == reserved_before  # nor did it go through torch's allocator
    assert rel_err(C, A.double().cpu() @ B.double().cpu()) <= TOL[torch.float32]


def test_sizes_beyond_the_built_in_scratch_are_refused():
    lib = _lib.load()
    M = N = K = 2048  # needs 64 MiB of split scratch, the array holds 32 MiB
    assert lib.b2mm_gemm_workspace_bytes(M, N, K, 1, _lib.FP32) > _lib.NOALLOC_SCRATCH_BYTES
    A = torch.zeros(M, K, dtype=torch.float32, device=DEV)
    B = torch.zeros(K, N, dtype=torch.float32, device=DEV)
    C = torch.zeros(M, N, dtype=torch.float32, device=DEV)
    rc = lib.b2mm_gemm_fp32_noalloc(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, K, N, N, None)
    assert rc == _lib.ERR_WORKSPACE
    msg = lib.b2mm_last_error().decode()
    assert "B2MM_NOALLOC_SCRATCH_BYTES" in msg and str(_lib.NOALLOC_SCRATCH_BYTES) in msg
