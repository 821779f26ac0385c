// Internals shared by the translation units of libb2mm.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace b2mm {

// Record a message for b2mm_last_error() and return `code`.
int fail(int code, const char* fmt, ...);
void count_launch(uint64_t n);
int dtype_bytes(int dtype);
int device_sm_count();

size_t gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int64_t batch, int dtype);

// b2mm_gemm with a cap on the SMs the persistent grid may occupy (0 = all; the overlap drivers
// leave SMs free for the concurrently running collective) and an optional fan-out of C: when
// fanout > 0 the result is stored to fan_ptrs[0..fanout) (local or peer-mapped) and `C` is ignored.
int gemm_impl(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
              int64_t ldc, int64_t batch, int64_t sA, int64_t sB, int64_t sC, int dtype, void* workspace,
              size_t workspace_bytes, int max_sms, cudaStream_t stream, void* const* fan_ptrs = nullptr,
              int fanout = 0);

}  // namespace b2mm
