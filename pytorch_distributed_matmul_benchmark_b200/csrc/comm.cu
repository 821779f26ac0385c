// Collective entry points (see include/b2mm.h) and the stream-pipelined
// GEMM + collective drivers.  NCCL is loaded with dlopen so that libb2mm.so
// has no link-time dependency on it: in a torch process the already-loaded
// libnccl.so.2 (torch's) is the one that answers.
#include "../../include/b2mm.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "allreduce_p2p.cuh"
#include "common.h"

namespace b2mm {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitRankConfig)(ncclComm_t*, int, ncclUniqueId, int, ncclConfig_t*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

static NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) return;
#define B2MM_SYM(field, sym)                                              \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, sym)); \
  if (!api.field) return;
    B2MM_SYM(GetUniqueId, "ncclGetUniqueId")
    B2MM_SYM(CommInitRank, "ncclCommInitRank")
    B2MM_SYM(CommDestroy, "ncclCommDestroy")
    B2MM_SYM(AllReduce, "ncclAllReduce")
    B2MM_SYM(AllGather, "ncclAllGather")
    B2MM_SYM(Broadcast, "ncclBroadcast")
    B2MM_SYM(GroupStart, "ncclGroupStart")
    B2MM_SYM(GroupEnd, "ncclGroupEnd")
    B2MM_SYM(GetErrorString, "ncclGetErrorString")
#undef B2MM_SYM
    api.CommInitRankConfig = reinterpret_cast<decltype(api.CommInitRankConfig)>(dlsym(api.handle, "ncclCommInitRankConfig"));
    api.ok = true;
  });
  return api;
}

static constexpr int kMaxPanels = 64;

struct Comm {
  ncclComm_t nccl = nullptr;
  int rank = 0, world = 1;
  cudaEvent_t panel_done[kMaxPanels] = {};
  cudaEvent_t comm_done = nullptr;
  float* barrier_word = nullptr;  // 1 float on the device: payload of b2mm_comm_barrier
  int ready_epoch = 0;            // fused all_reduce: panels announced so far (same sequence on every rank)
  int done_epoch = 0;             // fused all_reduce: calls completed so far
};

static int nccl_fail(const char* what, ncclResult_t r) {
  return fail(B2MM_ERR_NCCL, "%s: %s", what, nccl().GetErrorString ? nccl().GetErrorString(r) : "nccl error");
}

static bool nccl_type(int dtype, ncclDataType_t* t) {
  switch (dtype) {
    case B2MM_BF16: *t = ncclBfloat16; return true;
    case B2MM_FP16: *t = ncclFloat16; return true;
    case B2MM_FP32: *t = ncclFloat32; return true;
    default: return false;
  }
}

// How the panel GEMMs of the stream-pipelined drivers make room for the collective kernels.
// Default: dynamic scheduler whose clusters retire every B2MM_TILES_PER_CLUSTER tiles (2), so that the
// pending blocks of the higher-priority comm stream get SMs within a tile or two and give them back
// when done.  B2MM_COMM_SMS=n > 0 selects the older scheme instead: static scheduler on (#SM - n) SMs.
static GemmOpts overlap_opts() {
  GemmOpts o;
  const char* v = getenv("B2MM_COMM_SMS");
  const int reserve = (v && *v) ? atoi(v) : 0;
  if (reserve > 0) {
    const int sms = device_sm_count();
    o.max_sms = sms - reserve > 2 ? sms - reserve : sms;
  } else {
    const char* t = getenv("B2MM_TILES_PER_CLUSTER");
    o.tiles_per_cluster = (t && *t) ? atoi(t) : 2;
    if (o.tiles_per_cluster < 0) o.tiles_per_cluster = 0;
  }
  return o;
}

static constexpr int kSyncUnroll = 4;

template <typename T>
static void carveout_for_type() {
  const int co = cudaSharedmemCarveoutMaxShared;
  cudaFuncSetAttribute(allreduce_p2p_sync_kernel<T, kSyncUnroll, 2>, cudaFuncAttributePreferredSharedMemoryCarveout, co);
  cudaFuncSetAttribute(allreduce_p2p_sync_kernel<T, kSyncUnroll, 4>, cudaFuncAttributePreferredSharedMemoryCarveout, co);
  cudaFuncSetAttribute(allreduce_p2p_sync_kernel<T, kSyncUnroll, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, co);
  cudaFuncSetAttribute(allreduce_p2p_sync_kernel<T, kSyncUnroll, 0>, cudaFuncAttributePreferredSharedMemoryCarveout, co);
  cudaFuncSetAttribute(allreduce_nvls_sync_kernel<T, kSyncUnroll>, cudaFuncAttributePreferredSharedMemoryCarveout, co);
}

}  // namespace b2mm

using namespace b2mm;

extern "C" {

int b2mm_comm_unique_id(void* id_out) {
  if (!id_out) return fail(B2MM_ERR_BAD_ARG, "null id buffer");
  static_assert(sizeof(ncclUniqueId) == B2MM_UNIQUE_ID_BYTES, "ncclUniqueId size");
  if (!nccl().ok) return fail(B2MM_ERR_NCCL, "libnccl.so.2 could not be loaded: %s", dlerror());
  ncclUniqueId id;
  ncclResult_t r = nccl().GetUniqueId(&id);
  if (r != ncclSuccess) return nccl_fail("ncclGetUniqueId", r);
  memcpy(id_out, &id, sizeof(id));
  return B2MM_OK;
}

int b2mm_comm_init(const void* id_bytes, int rank, int world, void** comm_out) {
  if (!id_bytes || !comm_out || world < 1 || rank < 0 || rank >= world) return fail(B2MM_ERR_BAD_ARG, "bad comm_init argument");
  if (!nccl().ok) return fail(B2MM_ERR_NCCL, "libnccl.so.2 could not be loaded");
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  Comm* c = new Comm();
  c->rank = rank;
  c->world = world;
  // B2MM_NCCL_MAX_CTAS=n caps the CTAs NCCL's kernels use on this communicator: in the pipelined modes
  // they run beside a GEMM that leaves only a few SMs free, and a collective that cannot get all of
  // its CTAs resident at once waits for the GEMM panel to finish.
  const char* mc = getenv("B2MM_NCCL_MAX_CTAS");
  const int max_ctas = (mc && *mc) ? atoi(mc) : 0;
  ncclResult_t r;
  if (max_ctas > 0 && nccl().CommInitRankConfig) {
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    cfg.maxCTAs = max_ctas;
    cfg.minCTAs = 1;
    r = nccl().CommInitRankConfig(&c->nccl, world, id, rank, &cfg);
  } else {
    r = nccl().CommInitRank(&c->nccl, world, id, rank);
  }
  if (r != ncclSuccess) {
    c->nccl = nullptr;
    b2mm_comm_destroy(c);
    return nccl_fail("ncclCommInitRank", r);
  }
  bool ok = cudaEventCreateWithFlags(&c->comm_done, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; ok && i < kMaxPanels; ++i)
    ok = cudaEventCreateWithFlags(&c->panel_done[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    b2mm_comm_destroy(c);  // releases whatever was created so far
    return fail(B2MM_ERR_CUDA, "cudaEventCreate failed");
  }
  if (cudaMalloc(&c->barrier_word, 256) != cudaSuccess || cudaMemset(c->barrier_word, 0, 256) != cudaSuccess) {
    b2mm_comm_destroy(c);
    return fail(B2MM_ERR_CUDA, "cudaMalloc(barrier word) failed");
  }
  *comm_out = c;
  return B2MM_OK;
}

int b2mm_comm_destroy(void* comm) {
  if (!comm) return B2MM_OK;
  Comm* c = (Comm*)comm;
  for (int i = 0; i < kMaxPanels; ++i)
    if (c->panel_done[i]) cudaEventDestroy(c->panel_done[i]);
  if (c->comm_done) cudaEventDestroy(c->comm_done);
  if (c->barrier_word) cudaFree(c->barrier_word);
  if (c->nccl) nccl().CommDestroy(c->nccl);
  delete c;
  return B2MM_OK;
}

int b2mm_all_reduce_sum(void* comm, void* buf, int64_t count, int dtype, b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  ncclDataType_t t;
  if (!c || count < 0 || !nccl_type(dtype, &t)) return fail(B2MM_ERR_BAD_ARG, "bad all_reduce argument");
  if (count == 0) return B2MM_OK;  // batch_parallel at ws=8: local batch 0 (matmul_scaling_benchmark.py:111)
  if (!buf) return fail(B2MM_ERR_BAD_ARG, "null buffer");
  ncclResult_t r = nccl().AllReduce(buf, buf, (size_t)count, t, ncclSum, c->nccl, (cudaStream_t)stream);
  if (r != ncclSuccess) return nccl_fail("ncclAllReduce", r);
  return B2MM_OK;
}

int b2mm_all_gather(void* comm, const void* send, void* recv_slabs, int64_t count_per_rank, int dtype,
                    b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  ncclDataType_t t;
  if (!c || count_per_rank < 0 || !nccl_type(dtype, &t)) return fail(B2MM_ERR_BAD_ARG, "bad all_gather argument");
  if (count_per_rank == 0) return B2MM_OK;
  if (!send || !recv_slabs) return fail(B2MM_ERR_BAD_ARG, "null buffer");
  ncclResult_t r = nccl().AllGather(send, recv_slabs, (size_t)count_per_rank, t, c->nccl, (cudaStream_t)stream);
  if (r != ncclSuccess) return nccl_fail("ncclAllGather", r);
  return B2MM_OK;
}

// ---- CUDA IPC: make a peer rank's buffer addressable from this rank's device ----------------
namespace {
struct IpcEntry {
  cudaIpcMemHandle_t h;
  void* base;
  int refs;
};
std::mutex g_ipc_mu;
constexpr int kMaxIpcOpen = 256;
IpcEntry g_ipc_open[kMaxIpcOpen];
int g_ipc_n = 0;

using GetRangeFn = int (*)(unsigned long long*, size_t*, unsigned long long);
GetRangeFn get_range_fn() {
  static GetRangeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<GetRangeFn>(p);
  });
  return fn;
}
}  // namespace

int b2mm_ipc_export(const void* ptr, void* handle_out_64_bytes, int64_t* offset_out) {
  static_assert(sizeof(cudaIpcMemHandle_t) == B2MM_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  if (!ptr || !handle_out_64_bytes || !offset_out) return fail(B2MM_ERR_BAD_ARG, "null argument");
  GetRangeFn range = get_range_fn();
  if (!range) return fail(B2MM_ERR_CUDA, "cuMemGetAddressRange not available from the driver");
  unsigned long long base = 0;
  size_t size = 0;
  if (range(&base, &size, (unsigned long long)(uintptr_t)ptr) != 0) return fail(B2MM_ERR_CUDA, "cuMemGetAddressRange failed");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base);
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  memcpy(handle_out_64_bytes, &h, sizeof(h));
  *offset_out = (int64_t)((unsigned long long)(uintptr_t)ptr - base);
  return B2MM_OK;
}

int b2mm_ipc_import(const void* handle_64_bytes, int64_t offset, void** ptr_out) {
  if (!handle_64_bytes || !ptr_out || offset < 0) return fail(B2MM_ERR_BAD_ARG, "bad ipc_import argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle_64_bytes, sizeof(h));
  std::lock_guard<std::mutex> lock(g_ipc_mu);
  for (int i = 0; i < g_ipc_n; ++i)  // an allocation can be opened only once per process
    if (!memcmp(&g_ipc_open[i].h, &h, sizeof(h))) {
      ++g_ipc_open[i].refs;
      *ptr_out = (char*)g_ipc_open[i].base + offset;
      return B2MM_OK;
    }
  if (g_ipc_n == kMaxIpcOpen) return fail(B2MM_ERR_UNSUPPORTED, "too many imported allocations (%d open); release some", kMaxIpcOpen);
  void* base = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
  g_ipc_open[g_ipc_n].h = h;
  g_ipc_open[g_ipc_n].base = base;
  g_ipc_open[g_ipc_n].refs = 1;
  ++g_ipc_n;
  *ptr_out = (char*)base + offset;
  return B2MM_OK;
}

int b2mm_ipc_release(const void* handle_64_bytes) {
  if (!handle_64_bytes) return fail(B2MM_ERR_BAD_ARG, "null handle");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle_64_bytes, sizeof(h));
  std::lock_guard<std::mutex> lock(g_ipc_mu);
  for (int i = 0; i < g_ipc_n; ++i)
    if (!memcmp(&g_ipc_open[i].h, &h, sizeof(h))) {
      if (--g_ipc_open[i].refs > 0) return B2MM_OK;
      cudaError_t e = cudaIpcCloseMemHandle(g_ipc_open[i].base);
      g_ipc_open[i] = g_ipc_open[--g_ipc_n];
      if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "cudaIpcCloseMemHandle: %s", cudaGetErrorString(e));
      return B2MM_OK;
    }
  return fail(B2MM_ERR_BAD_ARG, "handle was not imported (or already released)");
}

int b2mm_comm_barrier(void* comm, b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  if (!c) return fail(B2MM_ERR_BAD_ARG, "null communicator");
  ncclResult_t r = nccl().AllReduce(c->barrier_word, c->barrier_word, 1, ncclFloat32, ncclSum, c->nccl, (cudaStream_t)stream);
  if (r != ncclSuccess) return nccl_fail("ncclAllReduce(barrier)", r);
  return B2MM_OK;
}

int b2mm_gemm_allgather_fused(const void* A, const void* B_local, void* const* peer_slabs, int rank, int world,
                              int64_t M, int64_t n_local, int64_t K, int dtype, void* workspace,
                              size_t workspace_bytes, b2mm_stream_t stream) {
  const int es = dtype_bytes(dtype);
  if (es == 0) return fail(B2MM_ERR_BAD_ARG, "unknown dtype %d", dtype);
  if (!peer_slabs || world < 1 || world > 8 || rank < 0 || rank >= world)
    return fail(B2MM_ERR_BAD_ARG, "bad rank/world (%d/%d; at most 8 destinations) or null slab list", rank, world);
  // Destination d: slab `rank` of rank d's [world][M][n_local] buffer, as mapped into this process.
  void* dst[8];
  for (int d = 0; d < world; ++d) {
    if (!peer_slabs[d]) return fail(B2MM_ERR_BAD_ARG, "null slab pointer for rank %d", d);
    dst[d] = (char*)peer_slabs[d] + (size_t)rank * M * n_local * es;
  }
  GemmOpts opts;
  opts.fan_ptrs = dst;
  opts.fanout = world;
  opts.head_split = world > 4;  // only where the stores, not the math, bound the step (measured: -2.7 % at 8 GPUs, +2.5 % at 2)
  return gemm_impl(A, B_local, nullptr, M, n_local, K, K, n_local, n_local, 1, 0, 0, 0, dtype, workspace, workspace_bytes,
                   (cudaStream_t)stream, opts);
}

int b2mm_gemm_allgather_nvls(const void* A, const void* B_local, void* slabs_multicast, int rank, int world, int64_t M,
                             int64_t n_local, int64_t K, int dtype, void* workspace, size_t workspace_bytes,
                             b2mm_stream_t stream) {
  const int es = dtype_bytes(dtype);
  if (es == 0) return fail(B2MM_ERR_BAD_ARG, "unknown dtype %d", dtype);
  if (!slabs_multicast || world < 1 || rank < 0 || rank >= world) return fail(B2MM_ERR_BAD_ARG, "bad rank/world (%d/%d) or null buffer", rank, world);
  // ONE destination: slab `rank` of the [world][M][n_local] buffer at its multicast address; the switch
  // replicates every TMA store into all ranks' copies.
  void* dst[1] = {(char*)slabs_multicast + (size_t)rank * M * n_local * es};
  GemmOpts opts;
  opts.fan_ptrs = dst;
  opts.fanout = 1;
  opts.head_split = world > 4;  // only where the stores, not the math, bound the step (measured: -2.7 % at 8 GPUs, +2.5 % at 2)
  return gemm_impl(A, B_local, nullptr, M, n_local, K, K, n_local, n_local, 1, 0, 0, 0, dtype, workspace, workspace_bytes,
                   (cudaStream_t)stream, opts);
}

// One rank's part of the peer-memory all-reduce over elements [elem_off, elem_off + count) of the
// buffers in `peers` (pull, add in fp32, push); the cross-rank ordering is the caller's.
static int launch_allreduce_p2p(void* const* peers, int rank, int world, int64_t elem_off, int64_t count, int dtype,
                                cudaStream_t stream) {
  const int es = dtype_bytes(dtype);
  if ((elem_off * es) % 16 || (count * es) % 16) return fail(B2MM_ERR_UNSUPPORTED, "p2p all_reduce range must be 16-byte aligned");
  const int64_t vecs = count * es / 16;
  const int64_t per = vecs / world;
  const int64_t begin = (elem_off * es / 16) + per * rank;
  const int64_t mine = (rank == world - 1) ? vecs - per * (world - 1) : per;  // the last rank takes the remainder
  if (mine <= 0) return B2MM_OK;
  PeerPtrs pp;
  for (int d = 0; d < 8; ++d) pp.p[d] = peers[d < world ? d : 0];
  const char* v = getenv("B2MM_P2P_CTAS");
  // A small grid on purpose: it runs beside the next panel's GEMM (32 CTAs x 512 threads x 4 x 16 B = 1 MB of
  // loads in flight).  Measured round 1: correct but ~2x slower than NCCL's all_reduce at 2 GPUs; kept as the
  // fp32-accumulating alternative, NCCL stays the default of --overlap.
  int ctas = (v && *v) ? atoi(v) : 32;
  if (ctas < 1) ctas = 1;
  constexpr int kUnroll = 4;
  // world as a template argument (2 / 4 / 8; 0 = any other size, predicated): the peer table is then only ever
  // indexed by constants and stays in the parameter bank
#define B2MM_P2P(T, W) allreduce_p2p_kernel<T, kUnroll, W><<<ctas, 512, 0, stream>>>(pp, world, begin, mine)
#define B2MM_P2P_WORLD(T)                                                                      \
  switch (world) {                                                                             \
    case 2: B2MM_P2P(T, 2); break;                                                             \
    case 4: B2MM_P2P(T, 4); break;                                                             \
    case 8: B2MM_P2P(T, 8); break;                                                             \
    default: B2MM_P2P(T, 0); break;                                                            \
  }
  switch (dtype) {
    case B2MM_BF16: B2MM_P2P_WORLD(__nv_bfloat16) break;
    case B2MM_FP16: B2MM_P2P_WORLD(__half) break;
    default: B2MM_P2P_WORLD(float) break;
  }
#undef B2MM_P2P_WORLD
#undef B2MM_P2P
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "p2p all_reduce launch failed: %s", cudaGetErrorString(e));
  count_launch(1);
  return B2MM_OK;
}

int b2mm_all_reduce_p2p(void* comm, void* const* peer_bufs, int64_t count, int dtype, b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  if (!c || !peer_bufs || count < 0 || !dtype_bytes(dtype) || c->world > 8) return fail(B2MM_ERR_BAD_ARG, "bad p2p all_reduce argument");
  if (count == 0) return B2MM_OK;
  int rc;
  if ((rc = b2mm_comm_barrier(comm, stream))) return rc;  // every rank's input is complete
  if ((rc = launch_allreduce_p2p(peer_bufs, c->rank, c->world, 0, count, dtype, (cudaStream_t)stream))) return rc;
  return b2mm_comm_barrier(comm, stream);                 // every rank's pushes have landed
}

int b2mm_gemm_allreduce_p2p(void* comm, const void* A, const void* B, void* const* peer_C, int64_t N,
                            int64_t local_batch, int dtype, int n_panels, void* workspace, size_t workspace_bytes,
                            b2mm_stream_t compute_, b2mm_stream_t comm_stream_) {
  Comm* c = (Comm*)comm;
  if (!c || !peer_C || !dtype_bytes(dtype) || c->world > 8) return fail(B2MM_ERR_BAD_ARG, "bad comm, dtype or buffer list");
  if (local_batch == 0) return B2MM_OK;
  if (n_panels < 1 || n_panels * local_batch > kMaxPanels)
    return fail(B2MM_ERR_BAD_ARG, "n_panels * local_batch must be in [1, %d]", kMaxPanels);
  if (N % n_panels) return fail(B2MM_ERR_UNSUPPORTED, "N=%lld is not divisible by n_panels=%d", (long long)N, n_panels);
  cudaStream_t compute = (cudaStream_t)compute_, cs = (cudaStream_t)comm_stream_;
  const int es = dtype_bytes(dtype);
  const int64_t mp = N / n_panels;
  const GemmOpts opts = overlap_opts();
  char* C = (char*)peer_C[c->rank];
  int ev = 0, rc;
  for (int64_t b = 0; b < local_batch; ++b) {
    for (int i = 0; i < n_panels; ++i, ++ev) {
      const int64_t elem_off = b * N * N + (int64_t)i * mp * N;
      const char* a = (const char*)A + (size_t)elem_off * es;
      const char* bb = (const char*)B + (size_t)b * N * N * es;
      if ((rc = gemm_impl(a, bb, C + (size_t)elem_off * es, mp, N, N, N, N, N, 1, 0, 0, 0, dtype, workspace, workspace_bytes,
                          compute, opts)))
        return rc;
      if (cudaEventRecord(c->panel_done[ev], compute) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
      if (cudaStreamWaitEvent(cs, c->panel_done[ev], 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
      if ((rc = b2mm_comm_barrier(comm, cs))) return rc;  // panel finished on every rank
      if ((rc = launch_allreduce_p2p(peer_C, c->rank, c->world, elem_off, mp * N, dtype, cs))) return rc;
    }
  }
  if ((rc = b2mm_comm_barrier(comm, cs))) return rc;  // all pushes have landed everywhere
  if (cudaEventRecord(c->comm_done, cs) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
  if (cudaStreamWaitEvent(compute, c->comm_done, 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
  return B2MM_OK;
}

// CTAs of two kernels share an SM only if both ask for the same L1/shared-memory split: the GEMM runs with
// the maximum shared-memory carve-out, so the (shared-memory-free) reduction kernels ask for it too.
static void prefer_gemm_carveout() {
  static std::once_flag once[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return;
  std::call_once(once[dev], [] {
    carveout_for_type<__nv_bfloat16>();
    carveout_for_type<__half>();
    carveout_for_type<float>();
    cudaFuncSetAttribute(p2p_done_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  });
}

// One self-synchronising reduction launch (see allreduce_p2p.cuh) over this rank's slice of
// [elem_off, elem_off + count); band_counter may be null.
static int launch_sync_reduce(Comm* c, const PeerPtrs& pp, const SignalPtrs& sp, int64_t elem_off, int64_t count, int dtype,
                              const int* band_counter, int band_target, cudaStream_t cs, void* mc = nullptr) {
  const int es = dtype_bytes(dtype);
  const int64_t vecs = count * es / 16, per = vecs / c->world;
  const int64_t begin = elem_off * es / 16 + per * c->rank;
  const int64_t mine = (c->rank == c->world - 1) ? vecs - per * (c->world - 1) : per;
  // Fewer CTAs than SMs on purpose: should these CTAs ever fail to share SMs with a GEMM's, the GEMM still
  // finds free SMs and the device-side waits end (slowly) instead of deadlocking.
  const char* v = getenv("B2MM_P2P_CTAS");
  // NVLS: the SMs only issue requests, 32 CTAs already saturate the links (measured at 8 GPUs: 1.13 ms for
  // 512 MiB with 32, 64 or 140 CTAs), so a small grid leaves the co-resident GEMM alone.
  int ctas = (v && *v) ? atoi(v) : (mc ? 64 : device_sm_count() - 8);
  if (ctas < 1) ctas = 1;
  const int epoch = ++c->ready_epoch;
  if (mc) {  // reduced inside the NVSwitch through the multicast mapping of the buffers
    switch (dtype) {
      case B2MM_BF16: allreduce_nvls_sync_kernel<__nv_bfloat16, kSyncUnroll><<<ctas, 256, 0, cs>>>(mc, sp, c->rank, c->world, epoch, begin, mine, band_counter, band_target); break;
      case B2MM_FP16: allreduce_nvls_sync_kernel<__half, kSyncUnroll><<<ctas, 256, 0, cs>>>(mc, sp, c->rank, c->world, epoch, begin, mine, band_counter, band_target); break;
      default: allreduce_nvls_sync_kernel<float, kSyncUnroll><<<ctas, 256, 0, cs>>>(mc, sp, c->rank, c->world, epoch, begin, mine, band_counter, band_target); break;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "NVLS all_reduce launch failed: %s", cudaGetErrorString(e));
    count_launch(1);
    return B2MM_OK;
  }
#define B2MM_SYNC(T, W) \
  allreduce_p2p_sync_kernel<T, kSyncUnroll, W><<<ctas, 256, 0, cs>>>(pp, sp, c->rank, c->world, epoch, begin, mine, band_counter, band_target)
#define B2MM_SYNC_WORLD(T)                                                                     \
  switch (c->world) {                                                                          \
    case 2: B2MM_SYNC(T, 2); break;                                                            \
    case 4: B2MM_SYNC(T, 4); break;                                                            \
    case 8: B2MM_SYNC(T, 8); break;                                                            \
    default: B2MM_SYNC(T, 0); break;                                                           \
  }
  switch (dtype) {
    case B2MM_BF16: B2MM_SYNC_WORLD(__nv_bfloat16) break;
    case B2MM_FP16: B2MM_SYNC_WORLD(__half) break;
    default: B2MM_SYNC_WORLD(float) break;
  }
#undef B2MM_SYNC_WORLD
#undef B2MM_SYNC
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(B2MM_ERR_CUDA, "peer all_reduce launch failed: %s", cudaGetErrorString(e));
  count_launch(1);
  return B2MM_OK;
}

static int launch_done(Comm* c, const SignalPtrs& sp, cudaStream_t cs) {
  p2p_done_kernel<<<1, 32, 0, cs>>>(sp, c->rank, c->world, ++c->done_epoch);
  if (cudaGetLastError() != cudaSuccess) return fail(B2MM_ERR_CUDA, "done-signal launch failed");
  count_launch(1);
  return B2MM_OK;
}

int b2mm_all_reduce_p2p_sync(void* comm, void* const* peer_bufs, void* const* peer_signals, int64_t count, int dtype,
                             b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  if (!c || !peer_bufs || !peer_signals || count < 0 || !dtype_bytes(dtype) || c->world > 8)
    return fail(B2MM_ERR_BAD_ARG, "bad p2p all_reduce argument");
  if (count == 0) return B2MM_OK;
  if ((count * dtype_bytes(dtype)) % 16) return fail(B2MM_ERR_UNSUPPORTED, "p2p all_reduce needs a multiple of 16 bytes");
  PeerPtrs pp;
  SignalPtrs sp;
  for (int d = 0; d < 8; ++d) {
    pp.p[d] = peer_bufs[d < c->world ? d : 0];
    sp.p[d] = (int*)peer_signals[d < c->world ? d : 0];
  }
  prefer_gemm_carveout();
  int rc;
  if ((rc = launch_sync_reduce(c, pp, sp, 0, count, dtype, nullptr, 0, (cudaStream_t)stream))) return rc;
  return launch_done(c, sp, (cudaStream_t)stream);
}

int b2mm_all_reduce_nvls_sync(void* comm, void* multicast_buf, void* const* peer_signals, int64_t count, int dtype,
                              b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  if (!c || !multicast_buf || !peer_signals || count < 0 || !dtype_bytes(dtype) || c->world > 8)
    return fail(B2MM_ERR_BAD_ARG, "bad NVLS all_reduce argument");
  if (count == 0) return B2MM_OK;
  if ((count * dtype_bytes(dtype)) % 16 || ((uintptr_t)multicast_buf & 15))
    return fail(B2MM_ERR_UNSUPPORTED, "NVLS all_reduce needs a 16-byte aligned buffer of a multiple of 16 bytes");
  PeerPtrs pp;
  SignalPtrs sp;
  for (int d = 0; d < 8; ++d) {
    pp.p[d] = nullptr;
    sp.p[d] = (int*)peer_signals[d < c->world ? d : 0];
  }
  prefer_gemm_carveout();
  int rc;
  if ((rc = launch_sync_reduce(c, pp, sp, 0, count, dtype, nullptr, 0, (cudaStream_t)stream, multicast_buf))) return rc;
  return launch_done(c, sp, (cudaStream_t)stream);
}

int b2mm_comm_barrier_signals(void* comm, void* const* peer_signals, b2mm_stream_t stream) {
  Comm* c = (Comm*)comm;
  if (!c || !peer_signals || c->world > 8) return fail(B2MM_ERR_BAD_ARG, "bad signal-barrier argument");
  SignalPtrs sp;
  for (int d = 0; d < 8; ++d) {
    if (d < c->world && !peer_signals[d]) return fail(B2MM_ERR_BAD_ARG, "null signal pad for rank %d", d);
    sp.p[d] = (int*)peer_signals[d < c->world ? d : 0];
  }
  prefer_gemm_carveout();
  return launch_done(c, sp, (cudaStream_t)stream);
}

static int gemm_allreduce_fused_impl(Comm* c, const void* A, const void* B, void* C_local, void* const* peer_C, void* mc,
                                     void* const* peer_signals, int* band_counters, int64_t N, int64_t local_batch,
                                     int dtype, int n_bands, void* workspace, size_t workspace_bytes,
                                     b2mm_stream_t compute_, b2mm_stream_t comm_stream_);

int b2mm_gemm_allreduce_fused(void* comm, const void* A, const void* B, void* const* peer_C, void* const* peer_signals,
                              int* band_counters, int64_t N, int64_t local_batch, int dtype, int n_bands, void* workspace,
                              size_t workspace_bytes, b2mm_stream_t compute_, b2mm_stream_t comm_stream_) {
  Comm* c = (Comm*)comm;
  if (!c || !peer_C) return fail(B2MM_ERR_BAD_ARG, "bad comm, dtype or buffer list");
  return gemm_allreduce_fused_impl(c, A, B, peer_C[c->rank], peer_C, nullptr, peer_signals, band_counters, N, local_batch, dtype,
                                   n_bands, workspace, workspace_bytes, compute_, comm_stream_);
}

int b2mm_gemm_allreduce_nvls(void* comm, const void* A, const void* B, void* C_local, void* C_multicast,
                             void* const* peer_signals, int* band_counters, int64_t N, int64_t local_batch, int dtype,
                             int n_bands, void* workspace, size_t workspace_bytes, b2mm_stream_t compute_,
                             b2mm_stream_t comm_stream_) {
  Comm* c = (Comm*)comm;
  if (!c || !C_local || !C_multicast) return fail(B2MM_ERR_BAD_ARG, "bad comm or buffer");
  return gemm_allreduce_fused_impl(c, A, B, C_local, nullptr, C_multicast, peer_signals, band_counters, N, local_batch, dtype,
                                   n_bands, workspace, workspace_bytes, compute_, comm_stream_);
}

// One of peer_C (pull / add / push over peer mappings) and mc (reduced in the switch) is non-null.
static int gemm_allreduce_fused_impl(Comm* c, const void* A, const void* B, void* C_local, void* const* peer_C, void* mc,
                                     void* const* peer_signals, int* band_counters, int64_t N, int64_t local_batch,
                                     int dtype, int n_bands, void* workspace, size_t workspace_bytes,
                                     b2mm_stream_t compute_, b2mm_stream_t comm_stream_) {
  if (!c || (!peer_C && !mc) || !C_local || !peer_signals || !band_counters || !dtype_bytes(dtype) || c->world > 8)
    return fail(B2MM_ERR_BAD_ARG, "bad comm, dtype or buffer list");
  if (local_batch == 0) return B2MM_OK;
  if (n_bands < 1 || n_bands * local_batch > kMaxPanels)
    return fail(B2MM_ERR_BAD_ARG, "n_bands * local_batch must be in [1, %d]", kMaxPanels);
  if (N % n_bands) return fail(B2MM_ERR_UNSUPPORTED, "N=%lld is not divisible by n_bands=%d", (long long)N, n_bands);
  cudaStream_t compute = (cudaStream_t)compute_, cs = (cudaStream_t)comm_stream_;
  const int es = dtype_bytes(dtype);
  const int64_t band_rows = N / n_bands;
  if ((band_rows * N * es) % (16 * c->world)) return fail(B2MM_ERR_UNSUPPORTED, "band size is not a multiple of 16 bytes x world");
  PeerPtrs pp;
  SignalPtrs sp;
  for (int d = 0; d < 8; ++d) {
    pp.p[d] = peer_C ? peer_C[d < c->world ? d : 0] : nullptr;
    sp.p[d] = (int*)peer_signals[d < c->world ? d : 0];
  }
  prefer_gemm_carveout();
  // compute: clear the progress counters, then ONE GEMM over the whole batch on the whole device; its
  // epilogues count finished tiles per row band.
  // (the caller's counter array holds at least 256 ints: room for kMaxPanels bands however the last one is split)
  if (cudaMemsetAsync(band_counters, 0, sizeof(int) * 4 * kMaxPanels, compute) != cudaSuccess) return fail(B2MM_ERR_CUDA, "memset failed");
  if (cudaEventRecord(c->panel_done[0], compute) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
  if (cudaStreamWaitEvent(cs, c->panel_done[0], 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
  GemmOpts opts;
  opts.band_counter = band_counters;
  opts.band_rows = (int)band_rows;
  opts.fine_tail_ok = true;
  int rc = gemm_impl(A, B, C_local, N, N, N, N, N, N, local_batch, N * N, N * N, N * N, dtype, workspace, workspace_bytes,
                     compute, opts);
  if (rc) return rc;
  // The GEMM may have split the last band into `fine_bands` bands of one tile row: the reduction that nothing hides
  // (the one of the rows finished last) is then a fraction of a band.
  const int coarse = opts.fine_bands > 0 ? n_bands - 1 : n_bands;
  const int per_batch = coarse + opts.fine_bands;
  if (opts.bands_per_batch != per_batch || per_batch * local_batch > 4 * kMaxPanels)
    return fail(B2MM_ERR_UNSUPPORTED, "band geometry mismatch (%d vs %d)", opts.bands_per_batch, per_batch);
  if (opts.fine_bands > 0 && ((int64_t)opts.fine_rows * N * es) % (16 * c->world))
    return fail(B2MM_ERR_UNSUPPORTED, "fine band size is not a multiple of 16 bytes x world");
  // comm: one reduction per band, queued now, each waiting (on the device) for its band, then for its peers.
  for (int64_t b = 0; b < local_batch; ++b) {
    for (int i = 0; i < per_batch; ++i) {
      const bool fine = i >= coarse;
      const int64_t row0 = fine ? (int64_t)coarse * band_rows + (int64_t)(i - coarse) * opts.fine_rows : (int64_t)i * band_rows;
      const int64_t rows = fine ? opts.fine_rows : band_rows;
      if ((rc = launch_sync_reduce(c, pp, sp, b * N * N + row0 * N, rows * N, dtype, band_counters + b * per_batch + i,
                                   fine ? opts.signals_per_fine_band : opts.signals_per_band, cs, mc)))
        return rc;
    }
  }
  if ((rc = launch_done(c, sp, cs))) return rc;
  if (cudaEventRecord(c->comm_done, cs) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
  if (cudaStreamWaitEvent(compute, c->comm_done, 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
  return B2MM_OK;
}

int b2mm_gemm_allgather_overlap(void* comm, const void* A, const void* B_local, void* C_local, void* recv_slabs,
                                int64_t M, int64_t n_local, int64_t K, int dtype, int n_panels, void* workspace,
                                size_t workspace_bytes, b2mm_stream_t compute_, b2mm_stream_t comm_stream_) {
  Comm* c = (Comm*)comm;
  ncclDataType_t t;
  if (!c || !nccl_type(dtype, &t)) return fail(B2MM_ERR_BAD_ARG, "bad comm or dtype");
  if (n_panels < 1 || n_panels > kMaxPanels) return fail(B2MM_ERR_BAD_ARG, "n_panels must be in [1, %d]", kMaxPanels);
  if (M % n_panels) return fail(B2MM_ERR_UNSUPPORTED, "M=%lld is not divisible by n_panels=%d", (long long)M, n_panels);
  cudaStream_t compute = (cudaStream_t)compute_, cs = (cudaStream_t)comm_stream_;
  const int es = dtype_bytes(dtype);
  const int64_t mp = M / n_panels;
  const int64_t slab = M * n_local;  // elements per rank in recv_slabs
  const GemmOpts opts = overlap_opts();
  for (int i = 0; i < n_panels; ++i) {
    const char* a = (const char*)A + (size_t)i * mp * K * es;
    char* cl = (char*)C_local + (size_t)i * mp * n_local * es;
    int rc = gemm_impl(a, B_local, cl, mp, n_local, K, K, n_local, n_local, 1, 0, 0, 0, dtype, workspace,
                       workspace_bytes, compute, opts);
    if (rc) return rc;
    if (cudaEventRecord(c->panel_done[i], compute) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
    if (cudaStreamWaitEvent(cs, c->panel_done[i], 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
    // all_gather of row panel i: slab r receives rank r's panel at the same row offset.
    ncclResult_t r = nccl().GroupStart();
    if (r != ncclSuccess) return nccl_fail("ncclGroupStart", r);
    for (int root = 0; root < c->world; ++root) {
      char* dst = (char*)recv_slabs + ((size_t)root * slab + (size_t)i * mp * n_local) * es;
      r = nccl().Broadcast(cl, dst, (size_t)(mp * n_local), t, root, c->nccl, cs);
      if (r != ncclSuccess) { nccl().GroupEnd(); return nccl_fail("ncclBroadcast", r); }
    }
    r = nccl().GroupEnd();
    if (r != ncclSuccess) return nccl_fail("ncclGroupEnd", r);
  }
  if (cudaEventRecord(c->comm_done, cs) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
  if (cudaStreamWaitEvent(compute, c->comm_done, 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
  return B2MM_OK;
}

int b2mm_gemm_allreduce_overlap(void* comm, const void* A, const void* B, void* C, int64_t N, int64_t local_batch,
                                int dtype, int n_panels, void* workspace, size_t workspace_bytes,
                                b2mm_stream_t compute_, b2mm_stream_t comm_stream_) {
  Comm* c = (Comm*)comm;
  ncclDataType_t t;
  if (!c || !nccl_type(dtype, &t)) return fail(B2MM_ERR_BAD_ARG, "bad comm or dtype");
  if (local_batch == 0) return B2MM_OK;
  if (n_panels < 1 || n_panels * local_batch > kMaxPanels)
    return fail(B2MM_ERR_BAD_ARG, "n_panels * local_batch must be in [1, %d]", kMaxPanels);
  if (N % n_panels) return fail(B2MM_ERR_UNSUPPORTED, "N=%lld is not divisible by n_panels=%d", (long long)N, n_panels);
  cudaStream_t compute = (cudaStream_t)compute_, cs = (cudaStream_t)comm_stream_;
  const int es = dtype_bytes(dtype);
  const int64_t mp = N / n_panels;
  const GemmOpts opts = overlap_opts();
  int ev = 0;
  for (int64_t b = 0; b < local_batch; ++b) {
    for (int i = 0; i < n_panels; ++i, ++ev) {
      const char* a = (const char*)A + ((size_t)b * N * N + (size_t)i * mp * N) * es;
      const char* bb = (const char*)B + (size_t)b * N * N * es;
      char* cc = (char*)C + ((size_t)b * N * N + (size_t)i * mp * N) * es;
      int rc = gemm_impl(a, bb, cc, mp, N, N, N, N, N, 1, 0, 0, 0, dtype, workspace, workspace_bytes, compute, opts);
      if (rc) return rc;
      if (cudaEventRecord(c->panel_done[ev], compute) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
      if (cudaStreamWaitEvent(cs, c->panel_done[ev], 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
      ncclResult_t r = nccl().AllReduce(cc, cc, (size_t)(mp * N), t, ncclSum, c->nccl, cs);
      if (r != ncclSuccess) return nccl_fail("ncclAllReduce", r);
    }
  }
  if (cudaEventRecord(c->comm_done, cs) != cudaSuccess) return fail(B2MM_ERR_CUDA, "event record failed");
  if (cudaStreamWaitEvent(compute, c->comm_done, 0) != cudaSuccess) return fail(B2MM_ERR_CUDA, "stream wait failed");
  return B2MM_OK;
}

}  // extern "C"
