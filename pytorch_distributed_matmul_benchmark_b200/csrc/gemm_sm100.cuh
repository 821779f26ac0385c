// Persistent, warp-specialised tcgen05 GEMM for sm_100a (B200).
//
//   C[b] (M x N, row-major) = A[b] (M x K, row-major) * B[b] (K x N, row-major)
//
// This is what stands under torch.matmul / torch.bmm at the reference's timed
// call sites (matmul_benchmark.py:62, matmul_scaling_benchmark.py:92,142,211).
//
// Data movement and layout
//   A is the K-major UMMA operand: TMA box {BLOCK_K, CTA_M rows}, 128B swizzle.
//   B is row-major K x N, i.e. the MN-major UMMA operand: TMA boxes of
//   {N_ATOM columns (128 bytes), BLOCK_K rows}, 128B swizzle; the tensor core
//   transposes on read, so B is never transposed in memory.
//   Both land in a kStages-deep shared-memory ring guarded by full/empty
//   mbarriers.  With kCtaGroup == 2 a CTA pair (one cluster) works on one
//   (2 * CTA_M) x BLOCK_N tile: each CTA loads its own CTA_M rows of A and HALF
//   of the B columns; tcgen05.mma.cta_group::2 (M = 256) reads both halves.
//
// Tile shapes and TMEM.  Accumulators live in TMEM (512 fp32 columns per CTA).
//   kMBlocks == 1: CTA_M = 128, two BLOCK_N-column accumulators, double buffered: the epilogue
//   of tile i overlaps the main loop of tile i+1.
//   kMBlocks == 2 (the default for 16-bit inputs when the problem fills the machine): CTA_M = 256,
//   the two 128-row halves X and Y of the tile fill TMEM.  Per k-block a CTA loads 32 KB of A and
//   16 KB of B for 2 x 4 MMAs: 48 B/cycle/SM from L2 instead of 64, a quarter less L2->SM traffic
//   per flop on a part whose clock is set by its power cap.  TMEM is then single buffered, so the
//   MMA warp gets a kStagger-k-block head start on X of the next tile while the epilogue is still
//   draining Y.
//
// Warp roles (224 threads)
//   warp 0      TMA producer (one elected lane)
//   warp 1      TMEM allocator + MMA issuer (one elected lane, leader CTA only)
//   warps 2..5  epilogue: tcgen05.ld -> convert -> swizzled smem -> TMA store
//   warp 6      tile scheduler (one elected lane, leader CTA only; dynamic mode)
//
// Tile scheduling.  Static: grid = min(#SM, tiles) CTAs, cluster c takes tiles c, c + C, ...
// (used when the caller caps the SMs).  Dynamic (default): the grid has one cluster per tile;
// the clusters that get to run keep their TMEM/barriers/pipeline and ask the hardware for the
// next not-yet-launched cluster's index (clusterlaunchcontrol.try_cancel), so faster SMs simply
// process more tiles and the tiles in flight stay adjacent in raster order.  Measured on B200:
// per-SM tile times differ by up to 5 % with a fixed pattern over TPC position (whatever the
// L2 demand), so a static split finishes at the pace of the slowest SM and its clusters drift
// apart in k, which destroys their sharing of A/B panels in L2.
//
// Output.  Staged through 128B-swizzled shared memory and written with TMA stores, which clip rows exactly
// but columns only at 16-byte granularity; when N * sizeof(OutT) is not a multiple of 16 the chunk that
// straddles N is written with element stores instead (store_row_tail), so nothing outside M x N is touched.
// With GemmParams::fanout the same chunk goes to several destinations (peer mappings, or one multicast
// address that the NVSwitch replicates): the all_gather of matrix_parallel done by the epilogue.
//
// Launch.  Programmatic dependent launch: launch_dependents on entry, griddepcontrol.wait after the
// prologue and before the first global access, so the next GEMM's CTAs set up while this one drains.
//
// kSplit == 2 is the fp32 path (3xTF32): A and B^T each arrive as a (hi, lo)
// pair of TF32-exact fp32 arrays (written by the split kernels at the end of
// this file) and every k-step issues
//   acc += Alo*Bhi;  acc += Ahi*Blo;  acc += Ahi*Bhi
// with kind::tf32; kChunkKB bounds how long the tensor core accumulates before
// the sum is carried on in fp32 registers.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ptx.cuh"

namespace b2mm {

struct GemmParams {
  int M, N, K, batch;
  int m_tiles, n_tiles;  // tiles per batch element (cluster tiles)
  int num_tiles;         // m_tiles * n_tiles * batch
  // Work items handed out by the scheduler.  Items [0, split_from) are whole tiles; every later pair of items is the
  // upper and lower half (CTA_M / 2 rows per CTA, the X accumulator only) of one of the last tiles, so that the
  // final wave is balanced at half-tile granularity (kMBlocks == 2 only; split_from == num_tiles: no halves).
  int num_work;          // split_from + 2 * (num_tiles - split_from)   (+ head_tiles)
  int split_from;
  // kMBlocks == 1, BLOCK_N == 256 only: the FIRST head_tiles tiles are handed out as 2 x head_tiles half-width items
  // (128 columns each, MMA N = 128 through idesc_half), so that the first finished outputs exist after half a tile
  // period.  Used by the fused gather, whose stores are the bottleneck and cannot start before the first tiles have
  // run through all of K.  0: off.
  int head_tiles;
  uint32_t idesc_half;   // instruction descriptor with N = BLOCK_N / 2
  int num_kb;            // ceil(K / BLOCK_K)
  int group_m;           // tile-rasterisation group height (in tiles)
  uint32_t idesc;        // tcgen05 instruction descriptor
  uint64_t hint_a;       // L2 eviction-priority descriptors for the A / B loads
  uint64_t hint_b;
  // 0: static round-robin over a grid of <= #SM CTAs.  1: the grid has one cluster per tile and
  // running clusters obtain further tiles from the hardware (cluster launch control).
  int dynamic;
  int fanout;  // 0: store C through tmC; n > 0: store every chunk through fan.m[0..n)
  // Dynamic mode only: a cluster retires after this many tiles (0 = runs until no tile is left).
  // Retiring frees its two SMs, which lets pending blocks of a higher-priority stream (the
  // collective of the previous panel) in; the grid's remaining clusters are launched as usual.
  int max_tiles;
  // Optional progress counters (fused all_reduce): when non-null, every epilogue warp adds 1 to
  // band_counter[batch * bands_per_batch + m_blk / band_tiles] once its rows of a tile are in global
  // memory, so a co-resident consumer kernel can start on a row band as soon as the band is complete.
  int* band_counter;
  int band_tiles;       // tile rows per band
  int bands_per_batch;
  // Tile rows from fine_from on form bands (and raster groups) of ONE tile row each: the last coarse band of a fused
  // all_reduce is split so that the reduction left over when the GEMM ends is a quarter of a band, not a whole one.
  // fine_from == m_tiles: off.  A multiple of band_tiles (and of group_m).
  int fine_from;
  // Optional trace buffer (tools only; null in production), rows of 256 uint64 indexed by SM id:
  // row 0 = 1 + first tile index, row 1 + i = global-timer value at which the producer started
  // this CTA's i-th tile.
  unsigned long long* trace;
  int debug_same_k;  // tools only: every k-block loads k = 0 (all L2 hits; results are meaningless)
  // Ragged N (N * sizeof(OutT) not a multiple of 16): TMA stores clip at 16-byte granularity (measured: the
  // whole 16-byte piece holding column N-1 is written), which would clobber up to 15 bytes past column N-1 of
  // every row.  The one chunk of a tile row that straddles N is therefore written with element stores.
  int ragged_n;
  void* c_ptr;             // C as a plain pointer (only used by that element path)
  long long ldc, c_stride;  // row pitch and batch stride of C, in elements
};

// Element-exact tail of one staged row (see GemmParams::ragged_n): `n` leading elements of the 128-byte,
// 128B-swizzled staging row at `srow` (16-byte piece j sits at ((j ^ sw) << 4)) go to dst[0..n).
// Out of line on purpose: only edge tiles of ragged problems come here, and the main epilogue keeps its registers.
template <class OutT>
__device__ __noinline__ void store_row_tail(uint32_t srow, uint32_t sw, OutT* dst, int n) {
  constexpr int kPer = 16 / (int)sizeof(OutT);
  for (int e = 0; e < n; ++e) {
    const uint32_t addr = srow + ((((uint32_t)(e / kPer)) ^ sw) << 4) + (uint32_t)(e % kPer) * (uint32_t)sizeof(OutT);
    if constexpr (sizeof(OutT) == 4) {
      uint32_t v;
      asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
      *reinterpret_cast<uint32_t*>(dst + e) = v;
    } else {
      uint16_t v;
      asm volatile("ld.shared.b16 %0, [%1];" : "=h"(v) : "r"(addr));
      *reinterpret_cast<uint16_t*>(dst + e) = v;
    }
  }
}

enum class InKind : int { BF16 = 0, FP16 = 1, TF32 = 2 };

// Fan-out of the output: when GemmParams::fanout > 0 every finished chunk of C is stored to
// fanout destinations (tensor maps over local and peer-GPU buffers, NVLink-mapped) instead of
// to one — the all-gather of matrix_parallel done by the GEMM's own epilogue.
constexpr int kMaxFanout = 8;
struct FanoutMaps {
  CUtensorMap m[kMaxFanout];
};

template <int kCtaGroup_, int BLOCK_N_, int kElemBytes_, int kSplit_, typename OutT_, bool kBKMajor_ = false,
          int kMBlocks_ = 1, int kChunkKB_ = 0, int kSwz_ = 128>
struct GemmConfig {
  // Bytes of one operand row in shared memory = the TMA / UMMA swizzle span: 128 (SWIZZLE_128B, the default) or 64
  // (SWIZZLE_64B: k-blocks half as long, twice as many stages in the same shared memory).
  static constexpr int kSwz = kSwz_;
  static_assert(kSwz_ == 128 || kSwz_ == 64, "swizzle span");
  // > 0: the tensor core accumulates only kChunkKB_ k-blocks at a time in TMEM; the epilogue warps
  // add each finished chunk into fp32 REGISTERS (round-to-nearest) and store at the end of the
  // tile.  The tensor core's fp32 accumulator truncates on every add, a bias that grows linearly
  // with the accumulation length (measured: 3e-4 relative at K = 16384); short chunks keep it at
  // the level of K = chunk.  Used by the fp32 (3xTF32) path; needs BLOCK_N <= 128 and kMBlocks == 1.
  static constexpr int kChunkKB = kChunkKB_;
  // 128-row accumulator blocks per CTA.
  //   1: tile = (128 * kCtaGroup) x BLOCK_N, TMEM holds TWO accumulators (double buffered).
  //   2: tile = (256 * kCtaGroup) x BLOCK_N, TMEM holds the tile's two halves X and Y (single
  //      buffered); a third less L2->SM traffic per flop, and the MMA of the next tile gets a
  //      head start on X while the epilogue is still draining Y.
  static constexpr int kMBlocks = kMBlocks_;
  using OutT = OutT_;
  // false: B is K x N row-major (MN-major operand, transposed by the tensor core).
  // true:  B is given as B^T, N x K row-major (K-major operand, same layout as A).
  static constexpr bool kBKMajor = kBKMajor_;
  static constexpr int kCtaGroup = kCtaGroup_;
  static constexpr int BLOCK_N = BLOCK_N_;
  static constexpr int kElemBytes = kElemBytes_;
  static constexpr int kSplit = kSplit_;
  static constexpr bool kTf32 = (kElemBytes_ == 4);

  static constexpr int CTA_M = 128 * kMBlocks_;     // rows per CTA (kMBlocks_ x 128 TMEM lanes)
  static constexpr int TILE_M = CTA_M * kCtaGroup;  // rows per cluster tile
  static constexpr int BLOCK_K = kSwz_ / kElemBytes;  // one swizzle row of K
  static constexpr int UMMA_K = 32 / kElemBytes;    // K per tcgen05.mma
  static constexpr int K_STEPS = BLOCK_K / UMMA_K;  // 4
  static constexpr int LOAD_N = BLOCK_N / kCtaGroup;  // B columns this CTA loads
  static constexpr int N_ATOM = kSwz_ / kElemBytes;    // columns in one swizzle-row chunk
  static constexpr int B_CHUNKS = kBKMajor_ ? 1 : LOAD_N / N_ATOM;

  static constexpr int A_BYTES = CTA_M * kSwz_;        // 16 KB per 128 rows of 128 B
  static constexpr int B_CHUNK_BYTES = (kBKMajor_ ? LOAD_N : BLOCK_K) * kSwz_;  // rows of one swizzle span
  static constexpr int B_BYTES = B_CHUNKS * B_CHUNK_BYTES;
  static constexpr int STAGE_BYTES = kSplit * (A_BYTES + B_BYTES);

  static constexpr int kEpiWarps = 4;
  static constexpr int kThreads = 32 * (3 + kEpiWarps);  // + TMA, MMA and scheduler warps
  static constexpr int kClcSlots = 3;                    // ring of cluster-launch-control responses
  static constexpr int OUT_COLS = 128 / (int)sizeof(OutT);  // columns per 128-byte output chunk
  static constexpr int OUT_CHUNKS = BLOCK_N / OUT_COLS;
  static constexpr int OUT_BUF_BYTES = 32 * 128;  // 32 rows x 128 B, per warp
  static constexpr int OUT_BYTES = kEpiWarps * 2 * OUT_BUF_BYTES;

  static constexpr int BAR_BYTES = 1024;
  static constexpr int SMEM_LIMIT = 232448;  // 227 KB opt-in per CTA
  static constexpr int kStages = (SMEM_LIMIT - 1024 /*align slack*/ - BAR_BYTES - OUT_BYTES) / STAGE_BYTES;
  static constexpr int SMEM_BYTES = 1024 + BAR_BYTES + kStages * STAGE_BYTES + OUT_BYTES;

  static constexpr int TMEM_COLS = 2 * BLOCK_N;  // two accumulators (two buffers, or X and Y)
  static constexpr int kStagger = 2;             // k-blocks of head start for X when kMBlocks == 2
  static_assert(kMBlocks_ == 1 || kMBlocks_ == 2, "one or two accumulator blocks per CTA");
  static_assert(kChunkKB_ == 0 || (kMBlocks_ == 1 && BLOCK_N_ <= 128 && sizeof(OutT_) == 4),
                "register accumulation: one 128-row block, <= 128 fp32 columns per thread");
  static_assert(kMBlocks_ == 1 || kStagger < kStages, "head start must leave a free stage");
  static_assert(TMEM_COLS == 512 || TMEM_COLS == 256 || TMEM_COLS == 128 || TMEM_COLS == 64, "pow2 TMEM");
  static_assert(kStages >= 2, "pipeline too shallow");
  static_assert(LOAD_N % N_ATOM == 0, "B half must be whole 128-byte chunks");
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "UMMA N");
};

// ---- UMMA shared-memory descriptors (SWIZZLE_128B, sm_100 "version 1") ----
// bits [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout=2.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout = 2 /* 2: SWIZZLE_128B, 4: SWIZZLE_64B */) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// Host+device: instruction descriptor for kind::f16 / kind::tf32, fp32 accumulate,
// A K-major, B MN-major.
__host__ __device__ inline uint32_t make_idesc(InKind kind, int umma_m, int umma_n, bool b_k_major = false) {
  uint32_t fmt = (kind == InKind::FP16) ? 0u : (kind == InKind::BF16 ? 1u : 2u);
  uint32_t d = 0;
  d |= 1u << 4;                          // D format: F32
  d |= fmt << 7;                         // A format
  d |= fmt << 10;                        // B format
  d |= 0u << 15;                         // A major: K
  d |= (b_k_major ? 0u : 1u) << 16;      // B major: MN unless B^T was supplied
  d |= (uint32_t)(umma_n >> 3) << 17;    // N / 8
  d |= (uint32_t)(umma_m >> 4) << 24;    // M / 16
  return d;
}

__device__ __forceinline__ void decode_tile(const GemmParams& p, int tile, int& b, int& m_blk, int& n_blk) {
  const int per_batch = p.m_tiles * p.n_tiles;
  b = tile / per_batch;
  const int r = tile - b * per_batch;
  if (r >= p.fine_from * p.n_tiles) {  // the fine tail: one tile row after the other, left to right
    const int rr = r - p.fine_from * p.n_tiles;
    m_blk = p.fine_from + rr / p.n_tiles;
    n_blk = rr - (m_blk - p.fine_from) * p.n_tiles;
    return;
  }
  const int span = p.group_m * p.n_tiles;
  const int g = r / span;
  const int first_m = g * p.group_m;
  const int gm = min(p.group_m, p.m_tiles - first_m);
  const int in_g = r - g * span;
  n_blk = in_g / gm;
  m_blk = first_m + (in_g - n_blk * gm);
}

// Work item -> tile index and which part of it: -1 = the whole tile, 0 / 1 = its upper / lower half.
__device__ __forceinline__ int decode_work(const GemmParams& p, int work, int& half) {
  if (p.head_tiles > 0) {  // half = 2 / 3: the left / right half (in N) of one of the first tiles
    if (work < 2 * p.head_tiles) {
      half = 2 + (work & 1);
      return work >> 1;
    }
    half = -1;
    return work - p.head_tiles;
  }
  if (work < p.split_from) {
    half = -1;
    return work;
  }
  const int w = work - p.split_from;
  half = w & 1;
  return p.split_from + (w >> 1);
}

template <typename OutT>
__device__ __forceinline__ uint32_t pack2(uint32_t lo_bits, uint32_t hi_bits);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(uint32_t lo_bits, uint32_t hi_bits) {
  __nv_bfloat162 v = __floats2bfloat162_rn(__uint_as_float(lo_bits), __uint_as_float(hi_bits));
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(uint32_t lo_bits, uint32_t hi_bits) {
  __half2 v = __floats2half2_rn(__uint_as_float(lo_bits), __uint_as_float(hi_bits));
  return *reinterpret_cast<uint32_t*>(&v);
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmA_lo,
            const __grid_constant__ CUtensorMap tmB_lo, const __grid_constant__ FanoutMaps fan,
            const GemmParams p) {
  using namespace ptx;
  constexpr int G = Cfg::kCtaGroup;
  constexpr int kStages = Cfg::kStages;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  // barrier block
  const uint32_t full_bar = base;                                // kStages x 8 B
  const uint32_t empty_bar = base + 8u * kStages;                // kStages x 8 B
  const uint32_t tfull_bar = base + 16u * kStages;               // 2 x 8 B
  const uint32_t tempty_bar = base + 16u * kStages + 16u;        // 2 x 8 B
  const uint32_t tmem_slot = base + 16u * kStages + 32u;         // 4 B
  const uint32_t clc_full = base + 512u;                         // kClcSlots x 8 B
  const uint32_t clc_empty = base + 512u + 32u;                  // kClcSlots x 8 B
  const uint32_t clc_resp = base + 512u + 64u;                   // kClcSlots x 16 B
  const uint32_t tiles = base + Cfg::BAR_BYTES;                  // operand ring
  const uint32_t out_stage = tiles + kStages * Cfg::STAGE_BYTES;  // epilogue staging

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Let the next kernel in the stream be scheduled onto SMs as this grid drains (it waits below, after its
  // own prologue, until this grid has completed).  Only has an effect when that kernel asks for it at launch.
  if (threadIdx.x == 0) griddep_launch_dependents();
  const uint32_t cta_rank = (G == 1) ? 0u : cluster_ctarank();
  const int cluster_id = blockIdx.x / G;
  const int num_clusters = gridDim.x / G;
  constexpr uint16_t kPairMask = (G == 2) ? 0x3 : 0x1;

  if (warp == 0) {
    if (elect_one()) {
      for (int s = 0; s < kStages; ++s) {
        mbar_init(full_bar + 8u * s, 1);
        mbar_init(empty_bar + 8u * s, 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(tfull_bar + 8u * a, 1);
        mbar_init(tempty_bar + 8u * a, G * Cfg::kEpiWarps);
      }
      // Every role that walks the tile sequence consumes each response: in the leader CTA the
      // producer, the MMA issuer, the scheduler itself and the epilogue warps; in the peer the
      // producer and the epilogue warps.
      constexpr uint32_t kClcConsumers = (3 + Cfg::kEpiWarps) + (G == 2 ? 1 + Cfg::kEpiWarps : 0);
      for (int c = 0; c < Cfg::kClcSlots; ++c) {
        mbar_init(clc_full + 8u * c, 1);
        mbar_init(clc_empty + 8u * c, kClcConsumers);
      }
      fence_mbar_init();
      prefetch_tensormap(&tmA);
      prefetch_tensormap(&tmB);
      prefetch_tensormap(&tmC);
      if constexpr (Cfg::kSplit == 2) {
        prefetch_tensormap(&tmA_lo);
        prefetch_tensormap(&tmB_lo);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    tmem_alloc<G>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish<G>();
  }
  tc_fence_before();
  if constexpr (G == 2) cluster_sync(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw_addr));
  // Everything above touched only this CTA's own shared memory / TMEM; from here on global memory is read and
  // written, so wait for the preceding grid (a no-op unless launched with programmatic serialization).
  griddep_wait();

  // ---- tile sequence -------------------------------------------------------
  // Every role walks the same sequence of tiles: its cluster's own index first, then either
  // +num_clusters (static) or whatever the scheduler warp obtained from the hardware (dynamic;
  // responses are consumed in order from a small ring, one slot per tile).
  struct TileSeq {
    int tile;
    int slot;
    uint32_t phase;
    int done = 0;  // tiles already taken by this walker
  };
  const uint32_t clc_empty_leader = (G == 2) ? mapa(clc_empty, 0) : clc_empty;
  // kWarp: called by all 32 lanes of a warp (one lane releases the slot); else by a single thread.
  auto next_tile = [&](TileSeq& q, bool warp_collective) -> bool {
    if (!p.dynamic) {
      q.tile += num_clusters;
      return q.tile < p.num_work;
    }
    if (p.max_tiles > 0 && ++q.done >= p.max_tiles) return false;  // every walker counts the same sequence
    mbar_wait(clc_full + 8u * q.slot, q.phase);
    uint32_t ctaid_x;
    const bool ok = clc_query(clc_resp + 16u * q.slot, ctaid_x);
    fence_proxy_async_smem();  // our read is ordered before the slot's next asynchronous overwrite
    if (warp_collective) __syncwarp();
    if (!warp_collective || lane == 0) {
      if constexpr (G == 2) mbar_arrive_cluster(clc_empty_leader + 8u * q.slot);
      else mbar_arrive(clc_empty + 8u * q.slot);
    }
    if (++q.slot == Cfg::kClcSlots) { q.slot = 0; q.phase ^= 1u; }
    q.tile = (int)(ctaid_x / G);
    return ok;
  };

  if (warp == 0) {
    // ============================ TMA producer ============================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      TileSeq q{cluster_id, 0, 0u};
      int iter = 0;
      for (bool more = q.tile < p.num_work; more; more = next_tile(q, false), ++iter) {
        int half;
        const int tile = decode_work(p, q.tile, half);
        int b, m_blk, n_blk;
        decode_tile(p, tile, b, m_blk, n_blk);
        // a half tile: this CTA's 128 rows of the chosen half, through the 128-row box of tmA_lo
        const bool is_half = Cfg::kMBlocks == 2 && half >= 0;
        const int m_row = is_half ? m_blk * Cfg::TILE_M + half * (Cfg::TILE_M / 2) + (int)cta_rank * (Cfg::CTA_M / 2)
                                  : m_blk * Cfg::TILE_M + (int)cta_rank * Cfg::CTA_M;
        const bool n_half = Cfg::kMBlocks == 1 && half >= 2;  // half-width item: this CTA loads LOAD_N / 2 columns of B
        const uint32_t stage_tx = (uint32_t)(G * (is_half ? Cfg::STAGE_BYTES - Cfg::kSplit * Cfg::A_BYTES / 2
                                                 : n_half ? Cfg::STAGE_BYTES - Cfg::kSplit * Cfg::B_BYTES / 2 : Cfg::STAGE_BYTES));
        const CUtensorMap* mapA = is_half ? &tmA_lo : &tmA;
        const int n_col = n_half ? n_blk * Cfg::BLOCK_N + (half - 2) * (Cfg::BLOCK_N / 2) + (int)cta_rank * (Cfg::LOAD_N / 2)
                                 : n_blk * Cfg::BLOCK_N + (int)cta_rank * Cfg::LOAD_N;
        const int b_chunks = n_half ? Cfg::B_CHUNKS / 2 : Cfg::B_CHUNKS;
        if (p.trace) {  // row 0: SM id of each CTA; row 1 + i: start time of its i-th tile
          unsigned long long now;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
          uint32_t smid;
          asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
          // indexed by SM (not by block): with dynamic scheduling the grid is one cluster per tile
          p.trace[(size_t)(1 + iter) * 256 + smid] = now;
          if (iter == 0) p.trace[smid] = 1 + (unsigned long long)tile;
        }
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(empty_bar + 8u * stage, phase ^ 1u);
          const uint32_t fb_local = full_bar + 8u * stage;
          const uint32_t fb = (G == 2) ? mapa(fb_local, 0) : fb_local;
          if (cta_rank == 0) mbar_arrive_expect_tx(fb_local, stage_tx);
          const uint32_t sA = tiles + stage * Cfg::STAGE_BYTES;
          const uint32_t sB = sA + Cfg::kSplit * Cfg::A_BYTES;
          const int k0 = p.debug_same_k ? 0 : kb * Cfg::BLOCK_K;
          tma_load_3d<G>(sA, mapA, fb, k0, m_row, b, p.hint_a);
          if constexpr (Cfg::kSplit == 2) tma_load_3d<G>(sA + Cfg::A_BYTES, &tmA_lo, fb, k0, m_row, b, p.hint_a);
          if constexpr (Cfg::kBKMajor) {
            tma_load_3d<G>(sB, &tmB, fb, k0, n_col, b, p.hint_b);
            if constexpr (Cfg::kSplit == 2) tma_load_3d<G>(sB + Cfg::B_BYTES, &tmB_lo, fb, k0, n_col, b, p.hint_b);
          } else {
#pragma unroll
            for (int c = 0; c < Cfg::B_CHUNKS; ++c) {
              if (c >= b_chunks) break;
              tma_load_3d<G>(sB + c * Cfg::B_CHUNK_BYTES, &tmB, fb, n_col + c * Cfg::N_ATOM, k0, b, p.hint_b);
              if constexpr (Cfg::kSplit == 2)
                tma_load_3d<G>(sB + Cfg::B_BYTES + c * Cfg::B_CHUNK_BYTES, &tmB_lo, fb, n_col + c * Cfg::N_ATOM, k0, b,
                               p.hint_b);
            }
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
      // Tail: wait until every slot this CTA filled has been released, so no
      // tcgen05.commit arrival can target this CTA's smem after it exits.
      for (int s = 0; s < kStages; ++s) {
        mbar_wait(empty_bar + 8u * stage, phase ^ 1u);
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ============================= MMA issuer =============================
    if (cta_rank == 0 && elect_one()) {
      // K-major A: 8-row groups 1024 B apart (SBO); LBO unused under swizzle.
      constexpr uint32_t kSbo = 8u * Cfg::kSwz, kLayout = Cfg::kSwz == 128 ? 2u : 4u;  // 8-row group pitch; swizzle mode
      const uint64_t a_desc0 = make_smem_desc(tiles, 16, kSbo, kLayout);
      // MN-major B: 8-row K groups 1024 B apart (SBO); 128-byte column chunks
      // B_CHUNK_BYTES apart (LBO).
      // (K-major B^T: described exactly like A.)
      const uint64_t b_desc0 = Cfg::kBKMajor ? make_smem_desc(tiles + Cfg::kSplit * Cfg::A_BYTES, 16, kSbo, kLayout)
                                             : make_smem_desc(tiles + Cfg::kSplit * Cfg::A_BYTES, Cfg::B_CHUNK_BYTES, kSbo, kLayout);
      constexpr uint32_t A_KSTEP = (Cfg::UMMA_K * Cfg::kElemBytes) >> 4;  // 32 B along the swizzled row
      constexpr uint32_t B_KSTEP = Cfg::kBKMajor ? A_KSTEP : ((Cfg::UMMA_K * Cfg::kSwz) >> 4);  // or UMMA_K rows of one span
      constexpr uint32_t A_MBLOCK = (128 * Cfg::kSwz) >> 4;               // next 128 rows of A in the stage

      // All MMAs of one k-block for accumulator block `mb` (rows mb*128.. of this CTA's A slab).
      uint32_t idesc = p.idesc;  // switched to idesc_half for the half-width items of the first wave
      auto issue_kblock = [&](int stage, int mb, uint32_t d_tmem, bool first_kb) {
        const uint32_t soff = (uint32_t)(stage * Cfg::STAGE_BYTES) >> 4;
        const uint64_t a_hi = a_desc0 + soff + (uint32_t)mb * A_MBLOCK;
        const uint64_t b_hi = b_desc0 + soff;
#pragma unroll
        for (int k = 0; k < Cfg::K_STEPS; ++k) {
          const uint32_t acc_in = (first_kb && k == 0) ? 0u : 1u;
          if constexpr (Cfg::kSplit == 2) {
            const uint64_t a_lo = a_hi + (Cfg::A_BYTES >> 4);
            const uint64_t b_lo = b_hi + (Cfg::B_BYTES >> 4);
            umma<G, Cfg::kTf32>(d_tmem, a_lo + k * A_KSTEP, b_hi + k * B_KSTEP, idesc, acc_in);
            umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_lo + k * B_KSTEP, idesc, 1u);
            umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_hi + k * B_KSTEP, idesc, 1u);
          } else {
            umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_hi + k * B_KSTEP, idesc, acc_in);
          }
        }
      };

      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      TileSeq q{cluster_id, 0, 0u};
      for (bool more = q.tile < p.num_work; more; more = next_tile(q, false)) {
        if constexpr (Cfg::kMBlocks == 1) {
          idesc = (p.head_tiles > 0 && q.tile < 2 * p.head_tiles) ? p.idesc_half : p.idesc;
          // One accumulator per chunk of k-blocks (the whole K unless kChunkKB is set).
          const int chunk_kb = Cfg::kChunkKB > 0 ? Cfg::kChunkKB : p.num_kb;
          int in_chunk = 0;
          for (int kb = 0; kb < p.num_kb; ++kb) {
            if (in_chunk == 0) {
              mbar_wait<true>(tempty_bar + 8u * acc, acc_phase ^ 1u);
              tc_fence_after();
            }
            mbar_wait(full_bar + 8u * stage, phase);
            tc_fence_after();
            issue_kblock(stage, 0, tmem_base + (uint32_t)(acc * Cfg::BLOCK_N), in_chunk == 0);
            umma_commit<G>(empty_bar + 8u * stage, kPairMask);  // slot free once these MMAs retire
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
            if (++in_chunk == chunk_kb || kb == p.num_kb - 1) {
              umma_commit<G>(tfull_bar + 8u * acc, kPairMask);  // accumulator (chunk) complete
              if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
              in_chunk = 0;
            }
          }
        } else if (q.tile >= p.split_from) {
          // Half tile: the X accumulator only.  Y is waited for as well (and released untouched by the epilogue), which
          // keeps the two accumulators' phases in step with the whole tiles before and after.
          const uint32_t dX = tmem_base;
          mbar_wait<true>(tempty_bar, acc_phase ^ 1u);
          mbar_wait<true>(tempty_bar + 8u, acc_phase ^ 1u);
          tc_fence_after();
          for (int kb = 0; kb < p.num_kb; ++kb) {
            mbar_wait(full_bar + 8u * stage, phase);
            tc_fence_after();
            issue_kblock(stage, 0, dX, kb == 0);
            umma_commit<G>(empty_bar + 8u * stage, kPairMask);
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
          umma_commit<G>(tfull_bar, kPairMask);
          acc_phase ^= 1u;
        } else {
          const uint32_t dX = tmem_base, dY = tmem_base + (uint32_t)Cfg::BLOCK_N;
          const int head = p.num_kb < Cfg::kStagger ? p.num_kb : Cfg::kStagger;
          // X has been drained: start on it while the epilogue still reads Y.
          mbar_wait<true>(tempty_bar, acc_phase ^ 1u);
          tc_fence_after();
          {
            int s = stage;
            uint32_t ph = phase;
            for (int kb = 0; kb < head; ++kb) {
              mbar_wait(full_bar + 8u * s, ph);
              tc_fence_after();
              issue_kblock(s, 0, dX, kb == 0);
              if (++s == kStages) { s = 0; ph ^= 1u; }
            }
          }
          mbar_wait<true>(tempty_bar + 8u, acc_phase ^ 1u);
          tc_fence_after();
          for (int kb = 0; kb < head; ++kb) {  // Y catches up on the same stages, then they are released
            issue_kblock(stage, 1, dY, kb == 0);
            umma_commit<G>(empty_bar + 8u * stage, kPairMask);
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
          for (int kb = head; kb < p.num_kb; ++kb) {
            mbar_wait(full_bar + 8u * stage, phase);
            tc_fence_after();
            issue_kblock(stage, 0, dX, false);
            issue_kblock(stage, 1, dY, false);
            umma_commit<G>(empty_bar + 8u * stage, kPairMask);
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
          umma_commit<G>(tfull_bar, kPairMask);  // X and Y complete
          acc_phase ^= 1u;
        }
      }
    }
    __syncwarp();
  } else if (warp == 2 + Cfg::kEpiWarps) {
    // ============================== scheduler =============================
    // Leader CTA only: one request at a time (a request may only follow a successful one),
    // up to kClcSlots tiles ahead of the slowest consumer.
    if (p.dynamic && cta_rank == 0 && elect_one()) {
      TileSeq q{0, 0, 0u};
      for (;;) {
        if (p.max_tiles > 0 && q.done + 1 >= p.max_tiles) break;  // the cluster retires after this tile
        mbar_wait<true>(clc_empty + 8u * q.slot, q.phase ^ 1u);
        for (uint32_t c = 0; c < (uint32_t)G; ++c)
          mbar_arrive_expect_tx_cluster(G == 2 ? mapa(clc_full + 8u * q.slot, c) : clc_full + 8u * q.slot, 16u);
        clc_try_cancel(clc_resp + 16u * q.slot, clc_full + 8u * q.slot);
        if (!next_tile(q, false)) break;  // consumes its own copy of the response
      }
    }
    __syncwarp();
  } else {
    // ============================== epilogue ==============================
    using OutT = typename Cfg::OutT;
    const int quarter = warp & 3;  // TMEM lane quarter this warp may read
    const int ew = warp - 2;       // staging-buffer owner index
    const uint32_t my_stage = out_stage + (uint32_t)ew * 2u * Cfg::OUT_BUF_BYTES;
    const uint32_t tempty_leader = (G == 2) ? mapa(tempty_bar, 0) : tempty_bar;
    uint32_t buf = 0;

    // One 32-row x 128-byte chunk from staging to C — or to every destination of the fan-out.  Called by the
    // whole warp (lane i staged row i); lane 0 issues the TMA store.
    auto store_chunk = [&](uint32_t sbuf, int col, int row, int b) {
      if (p.ragged_n && col < p.N && col + Cfg::OUT_COLS > p.N) {  // warp-uniform: the chunk that straddles N
        if (row + (int)lane < p.M)
          store_row_tail<OutT>(sbuf + lane * 128u, lane & 7u,
                               static_cast<OutT*>(p.c_ptr) + (size_t)b * p.c_stride + (size_t)(row + (int)lane) * p.ldc + col,
                               p.N - col);
        if (lane == 0) tma_store_commit();  // an empty group keeps the wait_read<1> accounting uniform
        return;
      }
      if (lane == 0) {
        if (p.fanout == 0) {
          tma_store_3d(&tmC, sbuf, col, row, b);
        } else {
          for (int d = 0; d < p.fanout; ++d) tma_store_3d(&fan.m[d], sbuf, col, row, b);
        }
        tma_store_commit();
      }
    };

    // This warp's part of tile (b, m_blk) is complete in global memory: publish it.
    auto signal_band = [&](int b, int m_blk) {
      if (p.band_counter == nullptr) return;
      if (lane == 0) {
        tma_store_wait<0>();  // the bulk stores have been performed, not merely read out of smem
        __threadfence();
        const int band = m_blk < p.fine_from ? m_blk / p.band_tiles : p.fine_from / p.band_tiles + (m_blk - p.fine_from);
        int* ctr = p.band_counter + (size_t)b * p.bands_per_batch + band;
        asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(ctr) : "memory");
      }
      __syncwarp();
    };

    // Move this warp's 32 rows of one accumulator (BLOCK_N columns at t_acc) to C, 128 bytes
    // per row at a time; hand the accumulator back (slot) as soon as it has been read.
    auto drain = [&](uint32_t t_acc, int row0, int col0, int b, int slot, int n_chunks) {
#pragma unroll 1
      for (int c = 0; c < n_chunks; ++c) {
        const uint32_t sbuf = my_stage + buf * Cfg::OUT_BUF_BYTES;
        // the TMA store issued two chunks ago must have finished reading this buffer
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        const uint32_t srow = sbuf + (uint32_t)lane * 128u;
        const uint32_t sw = (uint32_t)(lane & 7);
        if constexpr (sizeof(OutT) == 2) {
          uint32_t v0[32], v1[32];
          tmem_ld_32x32(t_acc + (uint32_t)(c * 64), v0);
          tmem_ld_32x32(t_acc + (uint32_t)(c * 64 + 32), v1);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            st_shared_v4(srow + (((uint32_t)j ^ sw) << 4), pack2<OutT>(v0[8 * j + 0], v0[8 * j + 1]),
                         pack2<OutT>(v0[8 * j + 2], v0[8 * j + 3]), pack2<OutT>(v0[8 * j + 4], v0[8 * j + 5]),
                         pack2<OutT>(v0[8 * j + 6], v0[8 * j + 7]));
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            st_shared_v4(srow + (((uint32_t)(j + 4) ^ sw) << 4), pack2<OutT>(v1[8 * j + 0], v1[8 * j + 1]),
                         pack2<OutT>(v1[8 * j + 2], v1[8 * j + 3]), pack2<OutT>(v1[8 * j + 4], v1[8 * j + 5]),
                         pack2<OutT>(v1[8 * j + 6], v1[8 * j + 7]));
          }
        } else {
          uint32_t v0[32];
          tmem_ld_32x32(t_acc + (uint32_t)(c * 32), v0);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(srow + (((uint32_t)j ^ sw) << 4), v0[4 * j + 0], v0[4 * j + 1], v0[4 * j + 2], v0[4 * j + 3]);
        }
        if (c == n_chunks - 1) {
          // all of this warp's TMEM reads of the accumulator are done: hand it back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (G == 2) mbar_arrive_cluster(tempty_leader + 8u * slot);
            else mbar_arrive(tempty_bar + 8u * slot);
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        store_chunk(sbuf, col0 + c * Cfg::OUT_COLS, row0, b);
        buf ^= 1u;
      }
    };

    int acc = 0;
    uint32_t acc_phase = 0;
    TileSeq q{cluster_id, 0, 0u};
    for (bool more = q.tile < p.num_work; more; more = next_tile(q, true)) {
      int b, m_blk, n_blk, half;
      decode_tile(p, decode_work(p, q.tile, half), b, m_blk, n_blk);
      const int row0 = (Cfg::kMBlocks == 2 && half >= 0)
                           ? m_blk * Cfg::TILE_M + half * (Cfg::TILE_M / 2) + (int)cta_rank * (Cfg::CTA_M / 2) + quarter * 32
                           : m_blk * Cfg::TILE_M + (int)cta_rank * Cfg::CTA_M + quarter * 32;
      const int col0 = n_blk * Cfg::BLOCK_N;
      const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
      if constexpr (Cfg::kChunkKB > 0) {
        // Register accumulation: this thread owns one row of the tile, BLOCK_N fp32 sums.
        float sum[Cfg::BLOCK_N];
#pragma unroll
        for (int j = 0; j < Cfg::BLOCK_N; ++j) sum[j] = 0.f;
        const int n_chunks = (p.num_kb + Cfg::kChunkKB - 1) / Cfg::kChunkKB;
#pragma unroll 1
        for (int ch = 0; ch < n_chunks; ++ch) {
          mbar_wait(tfull_bar + 8u * acc, acc_phase);
          tc_fence_after();
          const uint32_t t_acc = lane_base + (uint32_t)(acc * Cfg::BLOCK_N);
#pragma unroll
          for (int c = 0; c < Cfg::BLOCK_N / 32; ++c) {
            uint32_t v[32];
            tmem_ld_32x32(t_acc + (uint32_t)(c * 32), v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              sum[c * 32 + j] = __fadd_rn(sum[c * 32 + j], __uint_as_float(v[j]));
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (G == 2) mbar_arrive_cluster(tempty_leader + 8u * acc);
            else mbar_arrive(tempty_bar + 8u * acc);
          }
          if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
        }
#pragma unroll
        for (int c = 0; c < Cfg::OUT_CHUNKS; ++c) {
          const uint32_t sbuf = my_stage + buf * Cfg::OUT_BUF_BYTES;
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          const uint32_t srow = sbuf + (uint32_t)lane * 128u;
          const uint32_t sw = (uint32_t)(lane & 7);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(srow + (((uint32_t)j ^ sw) << 4), __float_as_uint(sum[c * 32 + 4 * j + 0]),
                         __float_as_uint(sum[c * 32 + 4 * j + 1]), __float_as_uint(sum[c * 32 + 4 * j + 2]),
                         __float_as_uint(sum[c * 32 + 4 * j + 3]));
          fence_proxy_async_smem();
          __syncwarp();
          store_chunk(sbuf, col0 + c * Cfg::OUT_COLS, row0, b);
          buf ^= 1u;
        }
        signal_band(b, m_blk);
      } else if constexpr (Cfg::kMBlocks == 1) {
        mbar_wait(tfull_bar + 8u * acc, acc_phase);
        tc_fence_after();
        if (half >= 2)  // half-width item of the first wave: 128 columns, at the left or right half of the tile
          drain(lane_base + (uint32_t)(acc * Cfg::BLOCK_N), row0, col0 + (half - 2) * (Cfg::BLOCK_N / 2), b, acc, Cfg::OUT_CHUNKS / 2);
        else
          drain(lane_base + (uint32_t)(acc * Cfg::BLOCK_N), row0, col0, b, acc, Cfg::OUT_CHUNKS);
        signal_band(b, m_blk);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      } else {
        mbar_wait(tfull_bar, acc_phase);
        tc_fence_after();
        drain(lane_base, row0, col0, b, 0, Cfg::OUT_CHUNKS);                   // X: rows 0..127 of the slab
        if (half < 0) {
          drain(lane_base + (uint32_t)Cfg::BLOCK_N, row0 + 128, col0, b, 1, Cfg::OUT_CHUNKS);  // Y: rows 128..255
        } else if (lane == 0) {  // half tile: Y was not used, hand it straight back
          if constexpr (G == 2) mbar_arrive_cluster(tempty_leader + 8u);
          else mbar_arrive(tempty_bar + 8u);
        }
        signal_band(b, m_blk);
        acc_phase ^= 1u;
      }
    }
    if (lane == 0) tma_store_wait<0>();
    __syncwarp();
  }

  // ------------------------------ teardown ------------------------------
  tc_fence_before();
  if constexpr (G == 2) cluster_sync(); else __syncthreads();
  if (warp == 1) tmem_dealloc<G>(tmem_base, Cfg::TMEM_COLS);
}

// fp32 -> (hi, lo) with hi = tf32(x), lo = tf32(x - hi); both stored as fp32
// words whose low 13 mantissa bits are zero, so the tensor core's fp32->tf32
// read is exact whatever its rounding.
__global__ void split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo,
                                  int64_t rows, int64_t cols, int64_t ld, int64_t out_ld, int64_t batch,
                                  int64_t stride) {
  const int64_t cols4 = (cols + 3) >> 2;  // ld % 4 == 0 (16-byte rule), cols itself is free; out_ld = 4 * cols4
  const int64_t total = batch * rows * cols4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c4 = i % cols4;
    const int64_t r = (i / cols4) % rows;
    const int64_t b = i / (cols4 * rows);
    const float* src = x + b * stride + r * ld + c4 * 4;
    float in[4] = {0.f, 0.f, 0.f, 0.f};
    if (c4 * 4 + 4 <= cols) {
      const float4 v = *reinterpret_cast<const float4*>(src);
      in[0] = v.x; in[1] = v.y; in[2] = v.z; in[3] = v.w;
    } else {  // last, partial group of a row whose length is not a multiple of 4: never read past column cols-1
      for (int j = 0; j < 4; ++j)
        if (c4 * 4 + j < cols) in[j] = src[j];
    }
    float h[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t hb, lb;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(in[j]));
      h[j] = __uint_as_float(hb);
      const float rem = in[j] - h[j];
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(rem));
      l[j] = __uint_as_float(lb);
    }
    const int64_t o = (b * rows + r) * out_ld + c4 * 4;  // packed output, 16-byte-aligned pitch
    *reinterpret_cast<float4*>(hi + o) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + o) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

// Same split for B, written TRANSPOSED: x is rows x cols (K x N, leading
// dimension ld); hi/lo are cols x rows (N x K, row pitch out_ld >= rows).  32 x 32 tiles go
// through shared memory so that both the read and the write are coalesced.
__global__ void split_tf32_transpose_kernel(const float* __restrict__ x, float* __restrict__ hi,
                                            float* __restrict__ lo, int64_t rows, int64_t cols, int64_t ld,
                                            int64_t out_ld, int64_t batch, int64_t stride) {
  __shared__ float th[32][33];
  __shared__ float tl[32][33];
  const int64_t tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  const int64_t total = batch * tiles_r * tiles_c;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8 threads
  for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
    const int64_t tc = t % tiles_c;
    const int64_t tr = (t / tiles_c) % tiles_r;
    const int64_t b = t / (tiles_c * tiles_r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = tr * 32 + ty + 8 * i, c = tc * 32 + tx;
      float v = (r < rows && c < cols) ? x[b * stride + r * ld + c] : 0.f;
      uint32_t hb, lb;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(v));
      const float h = __uint_as_float(hb);
      const float rem = v - h;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(rem));
      th[ty + 8 * i][tx] = h;
      tl[ty + 8 * i][tx] = __uint_as_float(lb);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t c = tc * 32 + ty + 8 * i, r = tr * 32 + tx;  // output row = input column
      if (c < cols && r < rows) {
        const int64_t o = (b * cols + c) * out_ld + r;
        hi[o] = th[tx][ty + 8 * i];
        lo[o] = tl[tx][ty + 8 * i];
      }
    }
    __syncthreads();
  }
}

}  // namespace b2mm
