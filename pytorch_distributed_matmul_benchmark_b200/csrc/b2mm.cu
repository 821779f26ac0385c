// C-ABI entry points for the GEMM path (see include/b2mm.h).
// Host side only: argument checks, TMA tensor-map encoding, kernel selection
// and launch.  Nothing here synchronises the host with the device.
#include "../../include/b2mm.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.h"
#include "gemm_sm100.cuh"

namespace b2mm {

thread_local char g_err[512] = "";
thread_local char g_last_kernel[256] = "";
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

unsigned long long* g_trace = nullptr;  // set by b2mm_debug_set_trace (tools only)

void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int dtype_bytes(int dtype) {
  switch (dtype) {
    case B2MM_BF16:
    case B2MM_FP16: return 2;
    case B2MM_FP32: return 4;
    default: return 0;
  }
}

// ------------------------------------------------------------ driver entry
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 3-D map over a (batch, rows, cols) row-major array: dims {cols, rows, batch}.
// Encoding costs ~1 us on the host and a benchmark calls with the same buffers over and over, so the
// last few maps are kept per thread, keyed by everything that goes into them.
struct MapKey {
  const void* ptr;
  int64_t cols, rows, batch, ld, batch_stride;
  int dtype;
  uint32_t box_cols, box_rows;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && cols == o.cols && rows == o.rows && batch == o.batch && ld == o.ld &&
           batch_stride == o.batch_stride && dtype == o.dtype && box_cols == o.box_cols && box_rows == o.box_rows;
  }
};
static constexpr int kMapCache = 32;
static thread_local MapKey t_map_keys[kMapCache];
static thread_local CUtensorMap t_map_vals[kMapCache];
static thread_local int t_map_used = 0, t_map_next = 0;

static int make_map(CUtensorMap* map, const void* ptr, int dtype, int64_t cols, int64_t rows, int64_t batch,
                    int64_t ld, int64_t batch_stride, uint32_t box_cols, uint32_t box_rows) {
  if (batch <= 1) batch_stride = rows * ld;
  const MapKey key{ptr, cols, rows, batch, ld, batch_stride, dtype, box_cols, box_rows};
  for (int i = 0; i < t_map_used; ++i)
    if (t_map_keys[i] == key) {
      *map = t_map_vals[i];
      return B2MM_OK;
    }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(B2MM_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  const int es = dtype_bytes(dtype);
  CUtensorMapDataType dt = dtype == B2MM_BF16   ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                           : dtype == B2MM_FP16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)(batch > 0 ? batch : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * es, (cuuint64_t)batch_stride * es};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  // the box's inner extent is one swizzle span: 128 bytes everywhere except the 64-byte k-blocks of the fp32 variant
  const CUtensorMapSwizzle swz = (size_t)box_cols * es == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(map, dt, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(B2MM_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d): cols=%lld rows=%lld batch=%lld ld=%lld", (int)r,
                (long long)cols, (long long)rows, (long long)batch, (long long)ld);
  const int slot = t_map_used < kMapCache ? t_map_used++ : (t_map_next++ % kMapCache);
  t_map_keys[slot] = key;
  t_map_vals[slot] = *map;
  return B2MM_OK;
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

int device_sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cached[dev] = n;
  }
  return cached[dev];
}

struct Operands {
  const void *A, *B, *A_lo, *B_lo;
  void* C;
  int64_t M, N, K, lda, ldb, ldc, batch, sA, sB, sC;
  int dtype;
  void* const* fan_ptrs;  // fanout destinations for C (null: just C)
  int fanout;
  int tiles_per_cluster;
  int* band_counter;
  int band_rows;
  const GemmOpts* opts;
};

template <class Cfg>
static int launch_cfg(const Operands& o, int max_sms, cudaStream_t stream) {
  CUtensorMap tmA, tmB, tmC, tmA_lo, tmB_lo;
  int rc;
  if ((rc = make_map(&tmA, o.A, o.dtype, o.K, o.M, o.batch, o.lda, o.sA, Cfg::BLOCK_K, Cfg::CTA_M))) return rc;
  if (Cfg::kBKMajor) {  // B^T: N rows of K
    if ((rc = make_map(&tmB, o.B, o.dtype, o.K, o.N, o.batch, o.ldb, o.sB, Cfg::BLOCK_K, Cfg::LOAD_N))) return rc;
  } else {
    if ((rc = make_map(&tmB, o.B, o.dtype, o.N, o.K, o.batch, o.ldb, o.sB, Cfg::N_ATOM, Cfg::BLOCK_K))) return rc;
  }
  static FanoutMaps fan_none;  // zero-filled; never dereferenced when fanout == 0
  FanoutMaps fan_local;
  const FanoutMaps* fan = &fan_none;
  if (o.fanout > 0) {
    for (int d = 0; d < o.fanout; ++d)
      if ((rc = make_map(&fan_local.m[d], o.fan_ptrs[d], o.dtype, o.N, o.M, o.batch, o.ldc, o.sC, Cfg::OUT_COLS, 32))) return rc;
    for (int d = o.fanout; d < kMaxFanout; ++d) fan_local.m[d] = fan_local.m[0];
    fan = &fan_local;
    tmC = fan_local.m[0];
  } else if ((rc = make_map(&tmC, o.C, o.dtype, o.N, o.M, o.batch, o.ldc, o.sC, Cfg::OUT_COLS, 32))) {
    return rc;
  }
  if (Cfg::kSplit == 2) {
    if ((rc = make_map(&tmA_lo, o.A_lo, o.dtype, o.K, o.M, o.batch, o.lda, o.sA, Cfg::BLOCK_K, Cfg::CTA_M))) return rc;
    if (Cfg::kBKMajor) {
      if ((rc = make_map(&tmB_lo, o.B_lo, o.dtype, o.K, o.N, o.batch, o.ldb, o.sB, Cfg::BLOCK_K, Cfg::LOAD_N))) return rc;
    } else {
      if ((rc = make_map(&tmB_lo, o.B_lo, o.dtype, o.N, o.K, o.batch, o.ldb, o.sB, Cfg::N_ATOM, Cfg::BLOCK_K))) return rc;
    }
  } else {
    tmA_lo = tmA;
    tmB_lo = tmB;
    // tall tiles: the spare A map carries a 128-row box, used by the half tiles of the last wave (see num_work)
    if (Cfg::kMBlocks == 2)
      if ((rc = make_map(&tmA_lo, o.A, o.dtype, o.K, o.M, o.batch, o.lda, o.sA, Cfg::BLOCK_K, Cfg::CTA_M / 2))) return rc;
  }

  GemmParams p;
  p.M = (int)o.M; p.N = (int)o.N; p.K = (int)o.K; p.batch = (int)o.batch;
  p.m_tiles = (int)((o.M + Cfg::TILE_M - 1) / Cfg::TILE_M);
  p.n_tiles = (int)((o.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N);
  p.num_tiles = p.m_tiles * p.n_tiles * p.batch;
  p.num_kb = (int)((o.K + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K);
  // Raster band: tiles are handed out band by band, m fastest inside a band.  With the dynamic
  // scheduler the tiles in flight are ~#clusters consecutive indices, so a band of g tile-rows keeps
  // g A-panels live for a whole tile period plus a thin slice of B; 2048 rows (4 tall tiles) keeps
  // that under L2 at K = 16384 (measured: 7.0 GB of DRAM reads per 16k GEMM vs 8.0 GB at 4096 rows,
  // 10.2 GB at 1024 rows; cuBLAS 7.25 GB).
  p.group_m = env_int("B2MM_GROUP_M", 2048 / Cfg::TILE_M);
  if (p.group_m < 1) p.group_m = 1;
  const InKind kind = o.dtype == B2MM_BF16 ? InKind::BF16 : o.dtype == B2MM_FP16 ? InKind::FP16 : InKind::TF32;
  p.idesc = make_idesc(kind, 128 * Cfg::kCtaGroup, Cfg::BLOCK_N, Cfg::kBKMajor);
  static const uint64_t kHints[3] = {ptx::kEvictNormal, ptx::kEvictFirst, ptx::kEvictLast};
  p.hint_a = kHints[env_int("B2MM_HINT_A", 0) % 3];
  p.hint_b = kHints[env_int("B2MM_HINT_B", 0) % 3];
  p.trace = g_trace;
  p.debug_same_k = g_trace ? env_int("B2MM_DEBUG_SAMEK", 0) : 0;  // only together with the trace tool
  p.fanout = o.fanout;
  p.ragged_n = ((o.N * dtype_bytes(o.dtype)) % 16) != 0;
  p.c_ptr = o.C;
  p.ldc = o.ldc;
  p.c_stride = o.batch > 1 ? o.sC : 0;
  if (p.ragged_n && (o.fanout > 0 || o.band_counter))
    return fail(B2MM_ERR_UNSUPPORTED, "the fused collective paths need N * element size to be a multiple of 16 bytes (N=%lld)", (long long)o.N);
  p.max_tiles = o.tiles_per_cluster;
  p.band_counter = o.band_counter;
  p.band_tiles = 1;
  p.bands_per_batch = 1;
  p.fine_from = p.m_tiles;
  if (o.band_counter) {
    if (o.band_rows % Cfg::TILE_M) return fail(B2MM_ERR_UNSUPPORTED, "band of %d rows is not a whole number of %d-row tiles", o.band_rows, Cfg::TILE_M);
    p.band_tiles = o.band_rows / Cfg::TILE_M;
    p.bands_per_batch = (p.m_tiles + p.band_tiles - 1) / p.band_tiles;
    p.group_m = p.band_tiles;  // the raster band IS the signalling band: bands complete in order
    o.opts->signals_per_band = p.band_tiles * p.n_tiles * Cfg::kCtaGroup * Cfg::kEpiWarps;
    // The reduction of the LAST band is the one nothing hides: B2MM_FINE_TAIL=1 splits that band into bands of one
    // tile row (when the rows divide evenly and the caller allows it).  OFF by default: parity-tested on 2 GPUs, but
    // the 8-GPU A/B was within the noise (6.67 / 6.72 ms without, 6.70 / 6.47 with), so the proven geometry stays.
    o.opts->fine_bands = 0;
    if (o.opts->fine_tail_ok && p.band_tiles > 1 && p.m_tiles % p.band_tiles == 0 && p.m_tiles >= 2 * p.band_tiles &&
        o.M % Cfg::TILE_M == 0 && env_int("B2MM_FINE_TAIL", 0)) {
      p.fine_from = p.m_tiles - p.band_tiles;
      p.bands_per_batch += p.band_tiles - 1;
      o.opts->fine_bands = p.band_tiles;
      o.opts->fine_rows = Cfg::TILE_M;
      o.opts->signals_per_fine_band = p.n_tiles * Cfg::kCtaGroup * Cfg::kEpiWarps;
    }
    o.opts->bands_per_batch = p.bands_per_batch;
  }
  // Dynamic (hardware work-stealing) scheduling unless the caller capped the SMs.
  const char* sched = getenv("B2MM_SCHED");
  p.dynamic = (max_sms > 0 || (sched && !strcmp(sched, "static"))) ? 0 : 1;

  int sms = device_sm_count();
  if (sms <= 0) return fail(B2MM_ERR_CUDA, "cannot query SM count");
  if (max_sms > 0 && max_sms < sms) sms = max_sms;
  int clusters = sms / Cfg::kCtaGroup;
  // Last wave at half-tile granularity (tall tiles, dynamic scheduler; B2MM_SPLIT_TAIL=1, OFF by default): with one
  // tile per request the clusters finish spread over a whole tile period, so on average half a tile period of every
  // SM is idle at the end (1.7 % of a 16384^3 GEMM); handing out the last `clusters` tiles as 2 x `clusters` half
  // tiles halves that idle time.  Measured (round 2, interleaved A/B in the sustained state): 1365 vs 1370 TFLOP/s
  // at 16384^3, 1318 vs 1344 at 8192^3 — no gain, because under the 1 kW cap throughput is set by the energy per
  // flop, not by SM occupancy (idle SMs give their power to the busy ones), and half tiles pull a third more bytes
  // per flop from L2.  Kept for parts / runs that are not power limited.  Only when the problem is several waves
  // long, and not with progress counters (their targets count whole tiles).
  p.split_from = p.num_tiles;
  if (Cfg::kMBlocks == 2 && Cfg::kSplit == 1 && p.dynamic && !o.band_counter && o.tiles_per_cluster == 0 &&
      p.num_tiles >= 4 * clusters && env_int("B2MM_SPLIT_TAIL", 0))
    p.split_from = p.num_tiles - clusters;
  p.num_work = p.split_from + 2 * (p.num_tiles - p.split_from);
  // First wave in half-width items (256-wide double-buffered tiles, dynamic scheduler): asked for by the fused gather
  // (GemmOpts::head_split), whose remote stores are its bottleneck and cannot begin before the first tiles are through
  // all of K; B2MM_SPLIT_HEAD=1/0 forces it on / off (tools, tests).
  p.head_tiles = 0;
  p.idesc_half = make_idesc(kind, 128 * Cfg::kCtaGroup, Cfg::BLOCK_N / 2, Cfg::kBKMajor);
  const int want_head = env_int("B2MM_SPLIT_HEAD", o.opts->head_split ? 1 : 0);
  if (Cfg::kMBlocks == 1 && Cfg::kSplit == 1 && Cfg::BLOCK_N == 256 && Cfg::kCtaGroup == 2 && p.dynamic && !o.band_counter &&
      o.tiles_per_cluster == 0 && p.split_from == p.num_tiles && p.num_tiles >= 2 * clusters && want_head) {
    p.head_tiles = clusters < p.num_tiles ? clusters : p.num_tiles;
    p.num_work += p.head_tiles;
  }
  if (clusters > p.num_work || p.dynamic) clusters = p.num_work;
  if (clusters < 1) clusters = 1;

  auto kern = gemm_kernel<Cfg>;
  static thread_local int attr_dev_mask_lo = 0;  // devices 0..31 that already have the smem attribute
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 32 || !(attr_dev_mask_lo & (1 << dev))) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "cudaFuncSetAttribute(smem=%d): %s", Cfg::SMEM_BYTES, cudaGetErrorString(e));
    if (dev < 32) attr_dev_mask_lo |= (1 << dev);
  }

  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)(clusters * Cfg::kCtaGroup), 1, 1);
  cfg.blockDim = dim3(Cfg::kThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = Cfg::kCtaGroup;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  // Programmatic dependent launch: this grid's CTAs may become resident (and run their prologue) while the
  // previous kernel in the stream drains; they block in griddepcontrol.wait before touching global memory.
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = env_int("B2MM_PDL", 1) ? 2 : 1;  // measured: 2k 18.9 -> 17.4 us, 4k 96.1 -> 92.9 us per GEMM back to back
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, tmA_lo, tmB_lo, *fan, p);
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "gemm launch failed: %s", cudaGetErrorString(e));
  count_launch(1);
  snprintf(g_last_kernel, sizeof(g_last_kernel),
           "gemm_kernel<cta_group::%d, %dx%d tile per %s, %s, %d-stage TMA ring, %s TMEM%s> %s scheduler, grid %d CTAs",
           Cfg::kCtaGroup, Cfg::TILE_M, Cfg::BLOCK_N, Cfg::kCtaGroup == 2 ? "CTA pair" : "CTA",
           o.dtype == B2MM_BF16 ? "bf16" : o.dtype == B2MM_FP16 ? "fp16" : "fp32 as 3xTF32", Cfg::kStages,
           Cfg::kMBlocks == 2 ? "X/Y halves in" : "double-buffered", Cfg::kChunkKB > 0 ? ", fp32 register carry" : "",
           p.dynamic ? (p.head_tiles > 0 ? "dynamic (cluster launch control, first wave in half-width items)"
                                    : p.num_work > p.num_tiles ? "dynamic (cluster launch control, last wave in half tiles)" : "dynamic (cluster launch control)")
                     : "static",
           clusters * Cfg::kCtaGroup);
  return B2MM_OK;
}

// Units of "tile columns x waves": a cheap model for picking BLOCK_N.
static int64_t wave_cost(const Operands& o, int tile_m, int block_n, int clusters) {
  const int64_t tiles = ((o.M + tile_m - 1) / tile_m) * ((o.N + block_n - 1) / block_n) * o.batch;
  const int64_t waves = (tiles + clusters - 1) / clusters;
  return waves * block_n;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

size_t gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int64_t batch, int dtype) {
  if (dtype != B2MM_FP32 || batch <= 0) return 0;
  const size_t kp = align_up((size_t)K, 4);  // packed rows of the hi/lo splits keep a 16-byte pitch for TMA
  const size_t a = align_up((size_t)batch * M * kp * 4, 256);
  const size_t b = align_up((size_t)batch * N * kp * 4, 256);
  return 2 * a + 2 * b;
}

int gemm_impl(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
              int64_t ldc, int64_t batch, int64_t sA, int64_t sB, int64_t sC, int dtype, void* workspace,
              size_t workspace_bytes, cudaStream_t stream, const GemmOpts& opts) {
  const int es = dtype_bytes(dtype);
  const int max_sms = opts.max_sms;
  void* const* fan_ptrs = opts.fan_ptrs;
  const int fanout = opts.fanout;
  if (fanout < 0 || fanout > kMaxFanout) return fail(B2MM_ERR_UNSUPPORTED, "fan-out of %d destinations (max %d)", fanout, kMaxFanout);
  if (fanout > 0) {
    if (!fan_ptrs) return fail(B2MM_ERR_BAD_ARG, "null destination list");
    for (int d = 0; d < fanout; ++d)
      if (!fan_ptrs[d] || ((uintptr_t)fan_ptrs[d] & 15)) return fail(B2MM_ERR_BAD_ARG, "destination %d is null or not 16-byte aligned", d);
    C = fan_ptrs[0];
  }
  if (es == 0) return fail(B2MM_ERR_BAD_ARG, "unknown dtype %d", dtype);
  if (batch < 0 || M < 0 || N < 0 || K < 0) return fail(B2MM_ERR_BAD_ARG, "negative size");
  if (batch == 0 || M == 0 || N == 0) return B2MM_OK;  // empty output: nothing to do (torch.bmm on b=0)
  if (K == 0) {  // an empty sum: C = 0, as torch.matmul gives for (M x 0) @ (0 x N)
    if (fanout > 0) return fail(B2MM_ERR_UNSUPPORTED, "K == 0 with remote destinations");
    if (!C) return fail(B2MM_ERR_BAD_ARG, "null operand pointer");
    if (ldc < N) return fail(B2MM_ERR_BAD_ARG, "leading dimension smaller than row length");
    for (int64_t b = 0; b < batch; ++b) {
      cudaError_t e = cudaMemset2DAsync((char*)C + (size_t)b * sC * es, (size_t)ldc * es, 0, (size_t)N * es, (size_t)M, stream);
      if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "zero-fill of C failed: %s", cudaGetErrorString(e));
    }
    return B2MM_OK;
  }
  if (!A || !B || !C) return fail(B2MM_ERR_BAD_ARG, "null operand pointer");
  if (lda < K || ldb < N || ldc < N) return fail(B2MM_ERR_BAD_ARG, "leading dimension smaller than row length");
  if (((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) return fail(B2MM_ERR_UNSUPPORTED, "operands must be 16-byte aligned");
  if ((lda * es) % 16 || (ldb * es) % 16 || (ldc * es) % 16)
    return fail(B2MM_ERR_UNSUPPORTED, "leading dimensions must be multiples of 16 bytes (lda=%lld ldb=%lld ldc=%lld, %d-byte elements)",
                (long long)lda, (long long)ldb, (long long)ldc, es);
  if (batch > 1 && ((sA * es) % 16 || (sB * es) % 16 || (sC * es) % 16))
    return fail(B2MM_ERR_UNSUPPORTED, "batch strides must be multiples of 16 bytes");
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX || batch > 65535)
    return fail(B2MM_ERR_UNSUPPORTED, "dimension too large");

  Operands o{A, B, nullptr, nullptr, C, M, N, K, lda, ldb, ldc, batch, sA, sB, sC, dtype, fan_ptrs, fanout,
             opts.tiles_per_cluster, opts.band_counter, opts.band_rows, &opts};
  const int sms = device_sm_count();
  int clusters = (max_sms > 0 && max_sms < sms ? max_sms : sms) / 2;
  if (clusters < 1) clusters = 1;

  if (dtype == B2MM_FP32) {
    const size_t need = gemm_workspace_bytes(M, N, K, batch, dtype);
    if (!workspace || workspace_bytes < need)
      return fail(B2MM_ERR_WORKSPACE, "fp32 GEMM needs %zu bytes of workspace, got %zu", need, workspace_bytes);
    if ((uintptr_t)workspace & 255) return fail(B2MM_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    const int64_t kp = (int64_t)align_up((size_t)K, 4);
    const size_t a_bytes = align_up((size_t)batch * M * kp * 4, 256);
    const size_t b_bytes = align_up((size_t)batch * N * kp * 4, 256);
    char* w = (char*)workspace;
    float* a_hi = (float*)w;
    float* a_lo = (float*)(w + a_bytes);
    float* b_hi = (float*)(w + 2 * a_bytes);
    float* b_lo = (float*)(w + 2 * a_bytes + b_bytes);
    const int blocks = (sms > 0 ? sms : 148) * 8;
    split_tf32_kernel<<<blocks, 256, 0, stream>>>((const float*)A, a_hi, a_lo, M, K, lda, kp, batch, sA);
    split_tf32_transpose_kernel<<<blocks, 256, 0, stream>>>((const float*)B, b_hi, b_lo, K, N, ldb, kp, batch, sB);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "split launch failed: %s", cudaGetErrorString(e));
    count_launch(2);
    o.A = a_hi; o.A_lo = a_lo; o.lda = kp; o.sA = M * kp;
    o.B = b_hi; o.B_lo = b_lo; o.ldb = kp; o.sB = N * kp;  // B^T: N x K, packed
    // 256 x 128 tiles, tensor-core accumulation in chunks of 8 k-blocks (K = 256), fp32 register carry.
    // B2MM_FP32_K16=1 (experiment): 64-byte k-blocks (SWIZZLE_64B), 16 per register-carry chunk, 8 stages instead of 4
    if (env_int("B2MM_FP32_K16", 0))
      return launch_cfg<GemmConfig<2, 128, 4, 2, float, true, 1, /*kChunkKB=*/16, /*kSwz=*/64>>(o, max_sms, stream);
    return launch_cfg<GemmConfig<2, 128, 4, 2, float, /*kBKMajor=*/true, /*kMBlocks=*/1, /*kChunkKB=*/8>>(o, max_sms, stream);
  }

  int block_n = env_int("B2MM_BLOCK_N", 0);
  if (block_n != 128 && block_n != 256) {
    // 256-wide tiles move a third less data from L2 per flop (and the part is power
    // limited), so the narrow tile has to win clearly on wave quantisation to be picked.
    const int64_t c128 = wave_cost(o, 256, 128, clusters), c256 = wave_cost(o, 256, 256, clusters);
    block_n = (c128 * 100 < c256 * 85) ? 128 : 256;
  }
  const int cta_group = env_int("B2MM_CTA_GROUP", 2);

  // Tall tiles (512 x 256 per CTA pair) whenever the problem still fills the machine with them.
  int m_blocks = env_int("B2MM_M_BLOCKS", 0);
  if (m_blocks != 1 && m_blocks != 2) {
    const int64_t tall_tiles = ((M + 511) / 512) * ((N + 255) / 256) * batch;
    m_blocks = (tall_tiles >= clusters) ? 2 : 1;
    // With remote destinations (peer mappings, or one multicast address) the epilogue drains at NVLink
    // speed: keep TMEM double buffered so the next tile's main loop runs meanwhile.
    if (fanout > 0) m_blocks = 1;
  }

  if (dtype == B2MM_BF16) {
    if (cta_group == 1) return launch_cfg<GemmConfig<1, 256, 2, 1, __nv_bfloat16>>(o, max_sms, stream);
    if (m_blocks == 2) return launch_cfg<GemmConfig<2, 256, 2, 1, __nv_bfloat16, false, 2>>(o, max_sms, stream);
    if (block_n == 128) return launch_cfg<GemmConfig<2, 128, 2, 1, __nv_bfloat16>>(o, max_sms, stream);
    return launch_cfg<GemmConfig<2, 256, 2, 1, __nv_bfloat16>>(o, max_sms, stream);
  } else {
    if (cta_group == 1) return launch_cfg<GemmConfig<1, 256, 2, 1, __half>>(o, max_sms, stream);
    if (m_blocks == 2) return launch_cfg<GemmConfig<2, 256, 2, 1, __half, false, 2>>(o, max_sms, stream);
    if (block_n == 128) return launch_cfg<GemmConfig<2, 128, 2, 1, __half>>(o, max_sms, stream);
    return launch_cfg<GemmConfig<2, 256, 2, 1, __half>>(o, max_sms, stream);
  }
}

// ------------------------------------------------------- fixed scratch for b2mm_gemm_fp32_noalloc
// The hi/lo split operands of the fp32 path, for callers that cannot allocate (see include/b2mm.h).
// One array per device image; in host code its name is only a handle, the device address comes from
// cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B2MM_NOALLOC_SCRATCH_BYTES];

// Device address of g_noalloc_scratch on the CURRENT device; resolved once per device.
static int noalloc_scratch_address(void** out) {
  static std::mutex mu;
  static void* resolved[64] = {nullptr};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return fail(B2MM_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!resolved[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr)
      return fail(B2MM_ERR_CUDA, "cudaGetSymbolAddress(noalloc scratch) on device %d: %s", dev, cudaGetErrorString(e));
    resolved[dev] = ptr;
  }
  *out = resolved[dev];
  return B2MM_OK;
}

// Tools only: a CTA that owns a whole SM (all of its shared memory) and sleeps until `ns` have passed.
__global__ void __launch_bounds__(32, 1) occupy_kernel(unsigned long long ns) {
  extern __shared__ uint8_t occupy_smem[];
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  if (threadIdx.x == 0) occupy_smem[0] = 1;  // the allocation is what keeps other CTAs off this SM
  do {
    __nanosleep(20000);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (t - t0 < ns);
}

}  // namespace b2mm

extern "C" {

int b2mm_version(void) { return B2MM_VERSION; }

const char* b2mm_last_error(void) { return b2mm::g_err; }

uint64_t b2mm_launch_count(void) { return b2mm::g_launches.load(std::memory_order_relaxed); }

const char* b2mm_last_kernel(void) { return b2mm::g_last_kernel; }

size_t b2mm_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int64_t batch, int dtype) {
  return b2mm::gemm_workspace_bytes(M, N, K, batch, dtype);
}

int b2mm_debug_max_active_clusters(int cluster_size) {
  using Cfg = b2mm::GemmConfig<2, 256, 2, 1, __nv_bfloat16, false, 2>;
  if (cluster_size < 1 || cluster_size > 16) return -1;
  auto kern = b2mm::gemm_kernel<Cfg>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES) != cudaSuccess) return -1;
  if (cluster_size > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) return -1;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(148 * 16 / cluster_size * cluster_size, 1, 1);
  cfg.blockDim = dim3(Cfg::kThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cluster_size;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = -1;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return n;
}

int b2mm_debug_occupy_sms(int n_pairs, double seconds, b2mm_stream_t stream) {
  using namespace b2mm;
  if (n_pairs < 1 || n_pairs > 74 || seconds <= 0 || seconds > 30) return fail(B2MM_ERR_BAD_ARG, "occupy: 1..74 pairs, at most 30 s");
  constexpr int kSmem = 200 * 1024;
  cudaError_t e = cudaFuncSetAttribute(occupy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "occupy: %s", cudaGetErrorString(e));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * n_pairs, 1, 1);
  cfg.blockDim = dim3(32, 1, 1);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;  // pairs: whole TPCs are taken, as a GEMM cluster would take them
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  e = cudaLaunchKernelEx(&cfg, occupy_kernel, (unsigned long long)(seconds * 1e9));
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "occupy launch: %s", cudaGetErrorString(e));
  return B2MM_OK;
}

void b2mm_debug_set_trace(void* device_buffer) { b2mm::g_trace = (unsigned long long*)device_buffer; }

int b2mm_gemm(const void* A, const void* B, void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
              int64_t ldc, int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC, int dtype,
              void* workspace, size_t workspace_bytes, b2mm_stream_t stream) {
  return b2mm::gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, batch, strideA, strideB, strideC, dtype, workspace,
                         workspace_bytes, (cudaStream_t)stream);
}

int b2mm_gemm_fp32_noalloc(const float* A, const float* B, float* C, int64_t M, int64_t N, int64_t K, int64_t lda,
                           int64_t ldb, int64_t ldc, b2mm_stream_t stream) {
  using namespace b2mm;
  if (M < 0 || N < 0 || K < 0) return fail(B2MM_ERR_BAD_ARG, "negative size");
  const size_t need = gemm_workspace_bytes(M, N, K, 1, B2MM_FP32);
  if (need > (size_t)B2MM_NOALLOC_SCRATCH_BYTES)
    return fail(B2MM_ERR_WORKSPACE,
                "b2mm_gemm_fp32_noalloc: %lld x %lld x %lld needs %zu bytes of split scratch, the built-in array holds %zu "
                "(B2MM_NOALLOC_SCRATCH_BYTES); use b2mm_gemm with caller workspace for larger problems",
                (long long)M, (long long)N, (long long)K, need, (size_t)B2MM_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = noalloc_scratch_address(&scratch)) return rc;
  // Same kernels as b2mm_gemm(dtype = fp32); the split pre-pass rewrites every scratch byte the GEMM reads.
  return gemm_impl(A, B, C, M, N, K, lda, ldb, ldc, 1, 0, 0, 0, B2MM_FP32, scratch, (size_t)B2MM_NOALLOC_SCRATCH_BYTES,
                   (cudaStream_t)stream);
}

}  // extern "C"
