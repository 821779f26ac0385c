// All-reduce (SUM) over peer-mapped buffers, one launch per rank: every rank owns 1/world of the
// range; it PULLS its slice from every rank over NVLink, adds the `world` values in fp32 (one
// rounding to the storage type at the end — tighter than a chain of 16-bit adds), and PUSHES the
// result into every rank's buffer.  In place: a slice is read only by its owner, so nothing is
// overwritten before it has been read.
//
// HBM / NVLink bound: per element each rank reads `world` values (world-1 remote) and writes `world`
// (world-1 remote); 2 * (world-1)/world * bytes cross each direction of the link, the same as a
// ring.  The grid is small on purpose (it runs beside the GEMM of the next panel).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include <cstdio>

namespace b2mm {

struct PeerPtrs {
  void* p[8];
};

// Element `i` of a pointer table that arrives as a kernel parameter.  Indexing such a table with a run-time value
// makes the compiler copy it to local memory (LDL on every use); eight compare-selects on constant indices keep it
// in the constant bank / registers instead.
template <class Ptr>
__device__ __forceinline__ Ptr pick(const Ptr (&p)[8], int i) {
  Ptr v = p[0];
#pragma unroll
  for (int d = 1; d < 8; ++d)
    if (i == d) v = p[d];
  return v;
}

template <typename T>
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]);
template <>
__device__ __forceinline__ void unpack8<__nv_bfloat16>(const uint4& v, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <>
__device__ __forceinline__ void unpack8<__half>(const uint4& v, float (&f)[8]) {
  const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <typename T>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]);
template <>
__device__ __forceinline__ uint4 pack8<__nv_bfloat16>(const float (&f)[8]) {
  uint4 v;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return v;
}
template <>
__device__ __forceinline__ uint4 pack8<__half>(const float (&f)[8]) {
  uint4 v;
  __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return v;
}

// `vec_begin`/`vec_count`: this launch's range in 16-byte vectors (already restricted to the slice
// this rank owns).  kUnroll independent vectors per thread keep enough loads in flight for NVLink.
// The pull / add / push loop shared by both kernels below.  kWorld is the world size as a compile-time constant
// (2, 4, 8: every index into `ptrs` is then a constant and the table never leaves the parameter bank), or 0 for
// any other world <= 8 (the same 8-way unroll, each step predicated on the run-time `world`).
template <typename T, int kUnroll, int kWorld>
__device__ __forceinline__ void pull_add_push(const PeerPtrs& ptrs, int world, int64_t vec_begin, int64_t vec_count) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int kLanes = sizeof(T) == 4 ? 4 : 8;
  constexpr int kMax = kWorld > 0 ? kWorld : 8;  // kWorld == 0: any world <= 8, every step predicated on r < world
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < vec_count; base += stride * kUnroll) {
    float acc[kUnroll][kLanes];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
#pragma unroll
      for (int j = 0; j < kLanes; ++j) acc[u][j] = 0.f;
#pragma unroll
    for (int r = 0; r < kMax; ++r) {
      if (kWorld == 0 && r >= world) break;
      const uint4* src = reinterpret_cast<const uint4*>(ptrs.p[r]) + vec_begin;
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t i = base + u * stride;
        if (i < vec_count) asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i));
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (base + u * stride < vec_count) {
          if constexpr (sizeof(T) == 4) {
            acc[u][0] += __uint_as_float(v[u].x);
            acc[u][1] += __uint_as_float(v[u].y);
            acc[u][2] += __uint_as_float(v[u].z);
            acc[u][3] += __uint_as_float(v[u].w);
          } else {
            float f[8];
            unpack8<T>(v[u], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[u][j] += f[j];
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + u * stride;
      if (i < vec_count) {
        uint4 out;
        if constexpr (sizeof(T) == 4) {
          out = make_uint4(__float_as_uint(acc[u][0]), __float_as_uint(acc[u][1]), __float_as_uint(acc[u][2]), __float_as_uint(acc[u][3]));
        } else {
          out = pack8<T>(acc[u]);
        }
#pragma unroll
        for (int d = 0; d < kMax; ++d) {
          if (kWorld == 0 && d >= world) break;
          uint4* dst = reinterpret_cast<uint4*>(ptrs.p[d]) + vec_begin + i;
          asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(out.x), "r"(out.y), "r"(out.z), "r"(out.w) : "memory");
        }
      }
    }
  }
}

template <typename T, int kUnroll, int kWorld>
__global__ void __launch_bounds__(512) allreduce_p2p_kernel(const __grid_constant__ PeerPtrs ptrs, int world, int64_t vec_begin,
                                                            int64_t vec_count) {
  pull_add_push<T, kUnroll, kWorld>(ptrs, world, vec_begin, vec_count);
}

// ---- the same reduction with its cross-GPU synchronisation inside the kernel -----------------
// Signal pads: every rank owns int32[16] in peer-mapped memory: [0, 8) "panel ready" epochs and
// [8, 16) "pushes done" epochs, one slot per source rank, written by that rank over NVLink.
// Counters only grow, so nothing is ever reset.
struct SignalPtrs {
  int* p[8];
};

__device__ __forceinline__ void st_release_sys(int* addr, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* addr) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
// Spin until *addr >= epoch; a peer that never shows up traps instead of hanging the GPU.
__device__ __forceinline__ void wait_epoch(const int* addr, int epoch) {
  const long long t0 = clock64();
  while (ld_acquire_sys(addr) - epoch < 0) {
    if (clock64() - t0 > 20000000000ll) {
      printf("b2mm: peer signal watchdog (block %d thread %d, want %d, have %d)\n", (int)blockIdx.x, (int)threadIdx.x, epoch,
             ld_acquire_sys(addr));
      __trap();
    }
  }
}

// Common start of the self-synchronising reductions: wait (on the device) for the local GEMM to have
// published this row band, tell every rank, wait for every rank.
__device__ __forceinline__ void sync_reduce_prologue(const SignalPtrs& sig, int rank, int world, int epoch,
                                                     const int* band_counter, int band_target) {
  if (band_counter != nullptr && threadIdx.x == 0) {
    // Launched next to the GEMM that produces the data: wait until the GEMM's epilogues have published
    // this row band (GemmParams::band_counter) before telling the other ranks that it is ready.
    const long long t0 = clock64();
    int seen;
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(band_counter) : "memory");
      if (clock64() - t0 > 20000000000ll) {
        printf("b2mm: band watchdog (block %d, want %d, have %d)\n", (int)blockIdx.x, band_target, seen);
        __trap();
      }
    } while (seen < band_target);
  }
  if (band_counter != nullptr) __syncthreads();
  if ((int)threadIdx.x < world) {
    if (blockIdx.x == 0) st_release_sys(pick(sig.p, (int)threadIdx.x) + rank, epoch);  // my panel is complete (kernel order)
    wait_epoch(pick(sig.p, rank) + threadIdx.x, epoch);                                 // ... and so is everybody else's
  }
  __syncthreads();
}

// Needs no shared memory and few registers on purpose: its CTAs are meant to be co-resident with the
// GEMM's (which own all of an SM's shared memory but few of its threads), so the reduction of panel i
// runs on the same SMs as the GEMM of panel i+1 and nothing has to be reserved for it.
template <typename T, int kUnroll, int kWorld>
__global__ void __launch_bounds__(256) allreduce_p2p_sync_kernel(const __grid_constant__ PeerPtrs ptrs,
                                                                  const __grid_constant__ SignalPtrs sig, int rank, int world,
                                                                  int epoch, int64_t vec_begin, int64_t vec_count,
                                                                  const int* band_counter, int band_target) {
  sync_reduce_prologue(sig, rank, world, epoch, band_counter, band_target);
  pull_add_push<T, kUnroll, kWorld>(ptrs, world, vec_begin, vec_count);
}

// ---- the same, reduced INSIDE the NVSwitch (NVLS) ---------------------------------------------
// `mc` is a multicast address bound to every rank's buffer.  multimem.ld_reduce returns the sum over all
// ranks of one 16-byte vector (added in the switch, fp32 accumulation, one rounding); multimem.st writes the
// result into every rank's buffer.  Per GPU and direction 1 x the buffer crosses NVLink instead of
// 2 x (world-1)/world x world/... of the pull/push kernel above (half the traffic at world = 8), and the
// SMs do no arithmetic at all: they only issue the requests.
template <typename T>
__device__ __forceinline__ uint4 multimem_ld_reduce_add(const uint4* mc);
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce_add<__nv_bfloat16>(const uint4* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce_add<__half>(const uint4* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
template <>
__device__ __forceinline__ uint4 multimem_ld_reduce_add<float>(const uint4* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(uint4* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T, int kUnroll>
__global__ void __launch_bounds__(256) allreduce_nvls_sync_kernel(void* mc, const __grid_constant__ SignalPtrs sig, int rank, int world, int epoch,
                                                                   int64_t vec_begin, int64_t vec_count,
                                                                   const int* band_counter, int band_target) {
  sync_reduce_prologue(sig, rank, world, epoch, band_counter, band_target);
  uint4* slice = reinterpret_cast<uint4*>(mc) + vec_begin;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < vec_count; base += stride * kUnroll) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + u * stride;
      if (i < vec_count) v[u] = multimem_ld_reduce_add<T>(slice + i);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t i = base + u * stride;
      if (i < vec_count) multimem_st(slice + i, v[u]);
    }
  }
}

// After the last reduction of a call: tell every rank that this rank's pushes have landed, then wait for
// the same from every rank.  One warp.
__global__ void p2p_done_kernel(const __grid_constant__ SignalPtrs sig, int rank, int world, int epoch) {
  if ((int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(pick(sig.p, (int)threadIdx.x) + 8 + rank, epoch);
    wait_epoch(pick(sig.p, rank) + 8 + threadIdx.x, epoch);
  }
}

}  // namespace b2mm
