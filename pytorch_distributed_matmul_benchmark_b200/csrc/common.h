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

// Options of the internal GEMM entry beyond what b2mm_gemm exposes.
struct GemmOpts {
  int max_sms = 0;                    // > 0: static scheduler on at most this many SMs
  void* const* fan_ptrs = nullptr;    // fanout > 0: C is stored to fan_ptrs[0..fanout) (local or peer-mapped)
  int fanout = 0;
  int tiles_per_cluster = 0;          // dynamic scheduler: clusters retire after this many tiles (0 = never)
  int* band_counter = nullptr;        // progress counters, one per band of `band_rows` rows (see GemmParams)
  int band_rows = 0;
  bool head_split = false;            // hand out the first wave as half-width items (fused gather: earlier first stores)
  // out: filled by gemm_impl when band_counter is set
  bool fine_tail_ok = false;          // the caller can handle a last band split into one-tile-row bands (see out fields)
  mutable int signals_per_band = 0;   // value a band's counter reaches when the band is complete
  mutable int bands_per_batch = 0;    // counters per batch element: coarse bands, with the last one replaced by fine_bands
  mutable int fine_bands = 0;         // 0: all bands are band_rows high; n: the last band_rows are n bands of fine_rows
  mutable int fine_rows = 0;
  mutable int signals_per_fine_band = 0;
};

int gemm_impl(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
              int64_t ldc, int64_t batch, int64_t sA, int64_t sB, int64_t sC, int dtype, void* workspace,
              size_t workspace_bytes, cudaStream_t stream, const GemmOpts& opts = GemmOpts());

}  // namespace b2mm
