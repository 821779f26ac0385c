// Persistent, warp-specialised tcgen05 GEMM for sm_100a (B200).
//
//   C[b] (M x N, row-major) = A[b] (M x K, row-major) * B[b] (K x N, row-major)
//
// This is what stands under torch.matmul / torch.bmm at the reference's timed
// call sites (matmul_benchmark.py:62, matmul_scaling_benchmark.py:92,142,211).
//
// Data movement and layout
//   A is the K-major UMMA operand: TMA box {BLOCK_K, 128 rows}, 128B swizzle.
//   B is row-major K x N, i.e. the MN-major UMMA operand: TMA boxes of
//   {N_ATOM columns (128 bytes), BLOCK_K rows}, 128B swizzle; the tensor core
//   transposes on read, so B is never transposed in memory.
//   Both land in a kStages-deep shared-memory ring guarded by full/empty
//   mbarriers.  With kCtaGroup == 2 a CTA pair (one cluster) works on one
//   (256 x BLOCK_N) tile: each CTA loads its own 128 rows of A and HALF of the
//   B columns; tcgen05.mma.cta_group::2 reads both halves.
//
// Warp roles (192 threads)
//   warp 0      TMA producer (one elected lane)
//   warp 1      TMEM allocator + MMA issuer (one elected lane, leader CTA only)
//   warps 2..5  epilogue: tcgen05.ld -> convert -> swizzled smem -> TMA store
//
// Accumulators live in TMEM, double buffered (2 x BLOCK_N fp32 columns), so
// the epilogue of tile i overlaps the main loop of tile i+1.
//
// kSplit == 2 is the fp32 path (3xTF32): A and B each arrive as a (hi, lo)
// pair of TF32-exact fp32 arrays and every k-step issues
//   acc += Alo*Bhi;  acc += Ahi*Blo;  acc += Ahi*Bhi
// with kind::tf32.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ptx.cuh"

namespace b2mm {

struct GemmParams {
  int M, N, K, batch;
  int m_tiles, n_tiles;  // tiles per batch element (cluster tiles)
  int num_tiles;         // m_tiles * n_tiles * batch
  int num_kb;            // ceil(K / BLOCK_K)
  int group_m;           // tile-rasterisation group height (in tiles)
  uint32_t idesc;        // tcgen05 instruction descriptor
};

enum class InKind : int { BF16 = 0, FP16 = 1, TF32 = 2 };

template <int kCtaGroup_, int BLOCK_N_, int kElemBytes_, int kSplit_, typename OutT_>
struct GemmConfig {
  using OutT = OutT_;
  static constexpr int kCtaGroup = kCtaGroup_;
  static constexpr int BLOCK_N = BLOCK_N_;
  static constexpr int kElemBytes = kElemBytes_;
  static constexpr int kSplit = kSplit_;
  static constexpr bool kTf32 = (kElemBytes_ == 4);

  static constexpr int CTA_M = 128;                 // accumulator rows per CTA (TMEM lanes)
  static constexpr int TILE_M = CTA_M * kCtaGroup;  // rows per cluster tile
  static constexpr int BLOCK_K = 128 / kElemBytes;  // one 128-byte swizzle row of K
  static constexpr int UMMA_K = 32 / kElemBytes;    // K per tcgen05.mma
  static constexpr int K_STEPS = BLOCK_K / UMMA_K;  // 4
  static constexpr int LOAD_N = BLOCK_N / kCtaGroup;  // B columns this CTA loads
  static constexpr int N_ATOM = 128 / kElemBytes;     // columns in one 128-byte chunk
  static constexpr int B_CHUNKS = LOAD_N / N_ATOM;

  static constexpr int A_BYTES = CTA_M * 128;          // 16 KB
  static constexpr int B_CHUNK_BYTES = BLOCK_K * 128;  // BLOCK_K rows of 128 B
  static constexpr int B_BYTES = B_CHUNKS * B_CHUNK_BYTES;
  static constexpr int STAGE_BYTES = kSplit * (A_BYTES + B_BYTES);

  static constexpr int kEpiWarps = 4;
  static constexpr int kThreads = 32 * (2 + kEpiWarps);
  static constexpr int OUT_COLS = 128 / (int)sizeof(OutT);  // columns per 128-byte output chunk
  static constexpr int OUT_CHUNKS = BLOCK_N / OUT_COLS;
  static constexpr int OUT_BUF_BYTES = 32 * 128;  // 32 rows x 128 B, per warp
  static constexpr int OUT_BYTES = kEpiWarps * 2 * OUT_BUF_BYTES;

  static constexpr int BAR_BYTES = 1024;
  static constexpr int SMEM_LIMIT = 232448;  // 227 KB opt-in per CTA
  static constexpr int kStages = (SMEM_LIMIT - 1024 /*align slack*/ - BAR_BYTES - OUT_BYTES) / STAGE_BYTES;
  static constexpr int SMEM_BYTES = 1024 + BAR_BYTES + kStages * STAGE_BYTES + OUT_BYTES;

  static constexpr int TMEM_COLS = 2 * BLOCK_N;  // two accumulators
  static_assert(TMEM_COLS == 512 || TMEM_COLS == 256 || TMEM_COLS == 128 || TMEM_COLS == 64, "pow2 TMEM");
  static_assert(kStages >= 2, "pipeline too shallow");
  static_assert(LOAD_N % N_ATOM == 0, "B half must be whole 128-byte chunks");
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "UMMA N");
};

// ---- UMMA shared-memory descriptors (SWIZZLE_128B, sm_100 "version 1") ----
// bits [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout=2.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Host+device: instruction descriptor for kind::f16 / kind::tf32, fp32 accumulate,
// A K-major, B MN-major.
__host__ __device__ inline uint32_t make_idesc(InKind kind, int umma_m, int umma_n) {
  uint32_t fmt = (kind == InKind::FP16) ? 0u : (kind == InKind::BF16 ? 1u : 2u);
  uint32_t d = 0;
  d |= 1u << 4;                          // D format: F32
  d |= fmt << 7;                         // A format
  d |= fmt << 10;                        // B format
  d |= 0u << 15;                         // A major: K
  d |= 1u << 16;                         // B major: MN
  d |= (uint32_t)(umma_n >> 3) << 17;    // N / 8
  d |= (uint32_t)(umma_m >> 4) << 24;    // M / 16
  return d;
}

__device__ __forceinline__ void decode_tile(const GemmParams& p, int tile, int& b, int& m_blk, int& n_blk) {
  const int per_batch = p.m_tiles * p.n_tiles;
  b = tile / per_batch;
  const int r = tile - b * per_batch;
  const int span = p.group_m * p.n_tiles;
  const int g = r / span;
  const int first_m = g * p.group_m;
  const int gm = min(p.group_m, p.m_tiles - first_m);
  const int in_g = r - g * span;
  n_blk = in_g / gm;
  m_blk = first_m + (in_g - n_blk * gm);
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
            const __grid_constant__ CUtensorMap tmB_lo, const GemmParams p) {
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
  const uint32_t tiles = base + Cfg::BAR_BYTES;                  // operand ring
  const uint32_t out_stage = tiles + kStages * Cfg::STAGE_BYTES;  // epilogue staging

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
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

  if (warp == 0) {
    // ============================ TMA producer ============================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      // Barrier that collects this stage's bytes: the leader CTA's.
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        int b, m_blk, n_blk;
        decode_tile(p, tile, b, m_blk, n_blk);
        const int m_row = m_blk * Cfg::TILE_M + (int)cta_rank * Cfg::CTA_M;
        const int n_col = n_blk * Cfg::BLOCK_N + (int)cta_rank * Cfg::LOAD_N;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(empty_bar + 8u * stage, phase ^ 1u);
          const uint32_t fb_local = full_bar + 8u * stage;
          const uint32_t fb = (G == 2) ? mapa(fb_local, 0) : fb_local;
          if (cta_rank == 0) mbar_arrive_expect_tx(fb_local, (uint32_t)(G * Cfg::STAGE_BYTES));
          const uint32_t sA = tiles + stage * Cfg::STAGE_BYTES;
          const uint32_t sB = sA + Cfg::kSplit * Cfg::A_BYTES;
          const int k0 = kb * Cfg::BLOCK_K;
          tma_load_3d<G>(sA, &tmA, fb, k0, m_row, b, kEvictNormal);
          if constexpr (Cfg::kSplit == 2) tma_load_3d<G>(sA + Cfg::A_BYTES, &tmA_lo, fb, k0, m_row, b, kEvictNormal);
#pragma unroll
          for (int c = 0; c < Cfg::B_CHUNKS; ++c) {
            tma_load_3d<G>(sB + c * Cfg::B_CHUNK_BYTES, &tmB, fb, n_col + c * Cfg::N_ATOM, k0, b, kEvictNormal);
            if constexpr (Cfg::kSplit == 2)
              tma_load_3d<G>(sB + Cfg::B_BYTES + c * Cfg::B_CHUNK_BYTES, &tmB_lo, fb, n_col + c * Cfg::N_ATOM, k0, b,
                             kEvictNormal);
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
      const uint64_t a_desc0 = make_smem_desc(tiles, 16, 1024);
      // MN-major B: 8-row K groups 1024 B apart (SBO); 128-byte column chunks
      // B_CHUNK_BYTES apart (LBO).
      const uint64_t b_desc0 = make_smem_desc(tiles + Cfg::kSplit * Cfg::A_BYTES, Cfg::B_CHUNK_BYTES, 1024);
      constexpr uint32_t A_KSTEP = (Cfg::UMMA_K * Cfg::kElemBytes) >> 4;  // 32 B along the swizzled row
      constexpr uint32_t B_KSTEP = (Cfg::UMMA_K * 128) >> 4;              // UMMA_K rows of 128 B
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        mbar_wait<true>(tempty_bar + 8u * acc, acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * Cfg::BLOCK_N);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(full_bar + 8u * stage, phase);
          tc_fence_after();
          const uint32_t soff = (uint32_t)(stage * Cfg::STAGE_BYTES) >> 4;
          const uint64_t a_hi = a_desc0 + soff;
          const uint64_t b_hi = b_desc0 + soff;
#pragma unroll
          for (int k = 0; k < Cfg::K_STEPS; ++k) {
            const uint32_t first = (kb | k) != 0;
            if constexpr (Cfg::kSplit == 2) {
              const uint64_t a_lo = a_hi + (Cfg::A_BYTES >> 4);
              const uint64_t b_lo = b_hi + (Cfg::B_BYTES >> 4);
              umma<G, Cfg::kTf32>(d_tmem, a_lo + k * A_KSTEP, b_hi + k * B_KSTEP, p.idesc, first);
              umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_lo + k * B_KSTEP, p.idesc, 1u);
              umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_hi + k * B_KSTEP, p.idesc, 1u);
            } else {
              umma<G, Cfg::kTf32>(d_tmem, a_hi + k * A_KSTEP, b_hi + k * B_KSTEP, p.idesc, first);
            }
          }
          umma_commit<G>(empty_bar + 8u * stage, kPairMask);  // slot free once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        umma_commit<G>(tfull_bar + 8u * acc, kPairMask);  // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
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
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t buf = 0;
    for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
      int b, m_blk, n_blk;
      decode_tile(p, tile, b, m_blk, n_blk);
      const int row0 = m_blk * Cfg::TILE_M + (int)cta_rank * Cfg::CTA_M + quarter * 32;
      const int col0 = n_blk * Cfg::BLOCK_N;
      mbar_wait(tfull_bar + 8u * acc, acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * Cfg::BLOCK_N);
#pragma unroll 1
      for (int c = 0; c < Cfg::OUT_CHUNKS; ++c) {
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
        if (c == Cfg::OUT_CHUNKS - 1) {
          // all of this warp's TMEM reads for the tile are done: hand the accumulator back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (G == 2) mbar_arrive_cluster(tempty_leader + 8u * acc);
            else mbar_arrive(tempty_bar + 8u * acc);
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(&tmC, sbuf, col0 + c * Cfg::OUT_COLS, row0, b);
          tma_store_commit();
        }
        buf ^= 1u;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
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
                                  int64_t rows, int64_t cols, int64_t ld, int64_t batch, int64_t stride) {
  const int64_t cols4 = cols >> 2;  // cols % 4 == 0 is enforced by the 16-byte ld rule
  const int64_t total = batch * rows * cols4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c4 = i % cols4;
    const int64_t r = (i / cols4) % rows;
    const int64_t b = i / (cols4 * rows);
    const float4 v = *reinterpret_cast<const float4*>(x + b * stride + r * ld + c4 * 4);
    float in[4] = {v.x, v.y, v.z, v.w};
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
    const int64_t o = (b * rows + r) * cols + c4 * 4;  // packed output, ld == cols
    *reinterpret_cast<float4*>(hi + o) = make_float4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<float4*>(lo + o) = make_float4(l[0], l[1], l[2], l[3]);
  }
}

}  // namespace b2mm
