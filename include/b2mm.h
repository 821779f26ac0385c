/*
 * b2mm.h — C ABI of the B200-native dense-GEMM hot path.
 *
 * This is the drop-in boundary for the operators the reference benchmark
 * scripts reach on their timed paths.  The reference has no FFI of its own
 * (it is pure Python over torch); each entry point below therefore names the
 * torch call site it stands in for (paths relative to the reference repo).
 *
 * Conventions (all entry points)
 *   - plain pointers and sizes only; device pointers unless stated otherwise;
 *   - row-major, C = A * B, no alpha/beta, no transposes (what torch.matmul
 *     does on contiguous 2-D/3-D inputs);
 *   - asynchronous and ordered on `stream`; no host synchronisation inside,
 *     so CUDA-event timing around a call measures the kernel;
 *   - the library borrows pointers for the duration of the enqueued work and
 *     never frees or reallocates caller memory;
 *   - every function returns a b2mm_status; on failure a human-readable
 *     message is available from b2mm_last_error() (thread local);
 *   - there is no cuBLAS / CPU fallback: an unsupported shape is an error.
 */
#ifndef B2MM_H_
#define B2MM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2MM_VERSION 100 /* major*10000 + minor*100 + patch */

#if defined(__GNUC__)
#define B2MM_API __attribute__((visibility("default")))
#else
#define B2MM_API
#endif

typedef enum b2mm_status {
  B2MM_OK = 0,
  B2MM_ERR_BAD_ARG = 1,     /* null pointer, negative size, unknown dtype        */
  B2MM_ERR_UNSUPPORTED = 2, /* legal GEMM, but a shape/stride this build rejects  */
  B2MM_ERR_CUDA = 3,        /* a CUDA runtime/driver call failed                  */
  B2MM_ERR_NCCL = 4,        /* a NCCL call failed / libnccl could not be loaded   */
  B2MM_ERR_WORKSPACE = 5    /* caller-supplied workspace too small                */
} b2mm_status;

typedef enum b2mm_dtype {
  B2MM_BF16 = 0, /* torch.bfloat16: bf16 in, fp32 accumulate, bf16 out */
  B2MM_FP16 = 1, /* torch.float16:  fp16 in, fp32 accumulate, fp16 out */
  B2MM_FP32 = 2  /* torch.float32:  fp32 in/out, 3xTF32 split on tensor cores */
} b2mm_dtype;

/* CUDA stream handle, passed as an opaque pointer (cudaStream_t). */
typedef void* b2mm_stream_t;

/* Library version (B2MM_VERSION of the build). */
B2MM_API int b2mm_version(void);

/* Message for the last non-OK status returned on this thread ("" if none). */
B2MM_API const char* b2mm_last_error(void);

/* Number of kernels of THIS library launched by the calling process so far
 * (all threads).  bench.py reports the delta over its timed region. */
B2MM_API uint64_t b2mm_launch_count(void);

/* Which GEMM kernel configuration the last successful b2mm_gemm-family call ON THIS THREAD launched, as text
 * (tile shape, CTA pairing, scheduler, grid), e.g. for a benchmark record that must name the kernel it timed
 * rather than assume it ("" before the first launch).  No counterpart in the reference. */
B2MM_API const char* b2mm_last_kernel(void);

/* Tools only (scripts/cluster_occupancy.py): how many clusters of `cluster_size` CTAs of the main bf16 GEMM kernel
 * (one CTA per SM, 227 KB of shared memory) the current device can hold at once, as cudaOccupancyMaxActiveClusters
 * reports it; -1 on error.  Answers "would 4-CTA clusters (TMA multicast across two CTA pairs) still fill all
 * 148 SMs?" without launching anything.  No counterpart in the reference. */
B2MM_API int b2mm_debug_max_active_clusters(int cluster_size);

/* Tools only (scripts/fewer_sms_probe.py): park `n_pairs` idle CTA pairs (one whole SM each, sleeping) on the device for
 * `seconds`, so that what follows in other streams runs on 148 - 2 * n_pairs SMs.  Asynchronous.  Used to measure
 * how a power-capped GEMM trades SM count for clock.  No counterpart in the reference. */
B2MM_API int b2mm_debug_occupy_sms(int n_pairs, double seconds, b2mm_stream_t stream);

/* Tools only (scripts/trace_tiles.py): when non-null, every GEMM launched afterwards writes into this
 * device buffer, laid out as rows of 256 uint64 indexed by SM id: row 0 = 1 + the first tile index of the
 * CTA on that SM, row 1 + i = the global-timer value at which its producer started its i-th tile
 * (at least 2 + tiles / 74 rows).  Pass NULL to switch it off.  No counterpart in the reference. */
B2MM_API void b2mm_debug_set_trace(void* device_buffer);

/*
 * Bytes of scratch b2mm_gemm needs for this problem (0 for bf16/fp16).
 * fp32 needs room for the TF32 hi/lo split of both operands: 8 * (M + N) * K' bytes per batch entry,
 * K' = K rounded up to 4 (16-byte rows), each of the four arrays rounded up to 256 bytes.
 */
B2MM_API size_t b2mm_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int64_t batch, int dtype);

/*
 * C[b] = A[b] * B[b]  for b in [0, batch).
 *
 * Replaces torch.matmul(A, B) at matmul_benchmark.py:46,62 and
 * matmul_scaling_benchmark.py:79,92,188,211, and torch.bmm(A, B) at
 * matmul_scaling_benchmark.py:120,142 (batch > 1).
 *
 *   A: M x K, leading dimension lda (elements), batch stride strideA
 *   B: K x N, leading dimension ldb,            batch stride strideB
 *   C: M x N, leading dimension ldc,            batch stride strideC
 *   dtype: b2mm_dtype, same for A, B and C.
 *
 * Requirements: pointers 16-byte aligned; lda/ldb/ldc and batch strides
 * multiples of 16 bytes.  M, N, K themselves are free.  batch == 0, M == 0 or N == 0 is a no-op
 * (torch.bmm on empty tensors — matmul_scaling_benchmark.py:111 at ws=8); K == 0 zero-fills C
 * (an empty sum, as torch.matmul returns).
 * Rows/columns/K that do not fill a tile are handled in-kernel (TMA
 * zero-fill / clipped stores); nothing is padded by the caller, and nothing outside the M x N
 * elements of C is written, so C may be a view into a wider matrix.
 *
 * workspace: only read when b2mm_gemm_workspace_bytes(...) > 0.
 */
B2MM_API int b2mm_gemm(const void* A, const void* B, void* C,
              int64_t M, int64_t N, int64_t K,
              int64_t lda, int64_t ldb, int64_t ldc,
              int64_t batch, int64_t strideA, int64_t strideB, int64_t strideC,
              int dtype, void* workspace, size_t workspace_bytes,
              b2mm_stream_t stream);

/*
 * fp32 C = A * B (2-D, row-major, same layout rules as b2mm_gemm) for callers that must not
 * allocate device memory.  It has no counterpart in the reference, where torch allocates freely.
 *
 * Who this is for: b2mm_gemm's fp32 path wants 8*(M*K + K*N) bytes of workspace from its caller.
 * Some hosts of this library cannot provide it: a serving process that handed the whole GPU to its
 * own memory pool at start-up (a later cudaMalloc fails, or punches holes into that pool); a
 * latency-critical loop that cannot afford cudaMalloc's device-wide synchronisation; a thin C or
 * ctypes caller that owns only the buffers it was handed and has no allocator to ask.  For them the
 * scratch is a fixed array compiled into the library's device image.
 *
 * Limit: B2MM_NOALLOC_SCRATCH_BYTES (32 MiB) on every GPU that loads the library, held until the
 * process exits.  A problem whose split operands need more (8*(M*K + K*N) bytes, each operand's pair
 * rounded up to 256 bytes) is refused with B2MM_ERR_WORKSPACE and a message that names the limit;
 * e.g. 1024 x 1024 x 1024 needs 16 MiB and fits, 2048^3 does not.
 *
 * Set-up: the first call on a device resolves the array's device address (once per device, with
 * that device current).  Make that first call before any stream capture begins.
 *
 * Sharing: every call on a device uses the same array, and the array keeps what the previous call
 * left in it.  Nothing has to be cleared — a call overwrites every byte it later reads — but two
 * calls whose GPU work may overlap (different streams without an event between them) race on the
 * scratch.  Keep them on one stream, or order them with events.
 *
 * Performs no cudaMalloc/cudaFree and calls no allocating library routine.
 */
#define B2MM_NOALLOC_SCRATCH_BYTES (32u << 20)
B2MM_API int b2mm_gemm_fp32_noalloc(const float* A, const float* B, float* C,
                                    int64_t M, int64_t N, int64_t K,
                                    int64_t lda, int64_t ldb, int64_t ldc,
                                    b2mm_stream_t stream);

/* ------------------------------------------------------------------------
 * Collectives.  The communicator wraps an ncclComm_t owned by the library
 * (libnccl is dlopen'ed on first use; the same libnccl.so.2 torch loaded).
 * ---------------------------------------------------------------------- */

#define B2MM_UNIQUE_ID_BYTES 128

/* Fill a 128-byte NCCL unique id (host memory).  Rank 0 calls this and ships
 * the bytes to the other ranks through the existing torch process group —
 * the rendezvous matmul_scaling_benchmark.py:15-24 already performs. */
B2MM_API int b2mm_comm_unique_id(void* id_out_128_bytes);

/* Collective over all `world` ranks.  The current CUDA device is the rank's.
 * Stands where dist.init_process_group('nccl') creates the payload communicator
 * (matmul_scaling_benchmark.py:19). */
B2MM_API int b2mm_comm_init(const void* id_128_bytes, int rank, int world, void** comm_out);

/* Replaces dist.destroy_process_group() for this communicator
 * (matmul_scaling_benchmark.py:59-61). */
B2MM_API int b2mm_comm_destroy(void* comm);

/* In-place sum over ranks.  Replaces dist.all_reduce(C, op=SUM) at
 * matmul_scaling_benchmark.py:125,150. */
B2MM_API int b2mm_all_reduce_sum(void* comm, void* buf, int64_t count, int dtype, b2mm_stream_t stream);

/* recv is ONE contiguous buffer of world slabs, slab r = rank r's `send`
 * (count_per_rank elements each).  Replaces dist.all_gather(list, C_local) at
 * matmul_scaling_benchmark.py:194,221 without torch's flatten-and-copy. */
B2MM_API int b2mm_all_gather(void* comm, const void* send, void* recv_slabs,
                    int64_t count_per_rank, int dtype, b2mm_stream_t stream);

/* Peer mapping for the fused collectives (CUDA IPC, one process per GPU on one node).  The owner of
 * a device buffer exports a 64-byte handle plus the buffer's offset inside its allocation; the bytes
 * travel over the existing process group; every other rank imports them with ITS device current and
 * gets a pointer its kernels can load/store through NVLink.  No counterpart in the reference (torch's
 * NCCL process group does the equivalent internally for dist.all_gather, matmul_scaling_benchmark.py:221).
 * An allocation is mapped once per process and reference counted: every b2mm_ipc_import must be paired
 * with one b2mm_ipc_release of the same handle, and the last release unmaps it.  All importers must have
 * released (and their kernels finished) BEFORE the owner frees the allocation — CUDA leaves the other
 * order undefined — so callers put a cross-rank barrier between the releases and the free. */
#define B2MM_IPC_HANDLE_BYTES 64
B2MM_API int b2mm_ipc_export(const void* ptr, void* handle_out_64_bytes, int64_t* offset_out);
B2MM_API int b2mm_ipc_import(const void* handle_64_bytes, int64_t offset, void** ptr_out);
B2MM_API int b2mm_ipc_release(const void* handle_64_bytes);

/* Cross-rank barrier on `stream` (a one-word ncclAllReduce): everything enqueued on `stream` before
 * it, on every rank, has completed before anything after it starts, on any rank.  Stands where the
 * reference calls dist.barrier() (matmul_scaling_benchmark.py:82,129,198); required after
 * b2mm_gemm_allgather_fused before any rank reads its slabs. */
B2MM_API int b2mm_comm_barrier(void* comm, b2mm_stream_t stream);

/* The same barrier without a NCCL kernel: one warp writes this rank's next epoch into every rank's signal pad over
 * NVLink (st.release.sys after a system fence) and waits until every rank's epoch has arrived in its own pad.
 * `peer_signals[d]`: rank d's pad (int32[16], zero at start, peer-mapped as for b2mm_all_reduce_p2p_sync; the barrier
 * uses words 8..15, shared with the completion signal of the fused all_reduce, so all ranks must issue both kinds of
 * call in the same order).  About a third of the NCCL barrier's latency (7 vs 20 us on 8 GPUs), which is what closes a
 * fused GEMM + all_gather step: the all_gather of matmul_scaling_benchmark.py:221 is complete on every rank when
 * this returns on the stream.  world <= 8. */
B2MM_API int b2mm_comm_barrier_signals(void* comm, void* const* peer_signals, b2mm_stream_t stream);

/*
 * matrix_parallel as ONE kernel: C_local = A * B_local (matmul_scaling_benchmark.py:211) whose epilogue
 * writes every finished tile straight into slab `rank` of EVERY rank's gather buffer over NVLink
 * (TMA stores to peer-mapped memory), i.e. the all_gather of :221 overlapped with the math tile by
 * tile, with no collective kernel and no SMs taken from the GEMM.
 *
 *   peer_slabs[d]: base of rank d's [world][M][n_local] buffer as mapped into THIS process (d == rank:
 *   the local buffer; others: CUDA-IPC / peer mappings).  world <= 8.
 *
 * The stores of other ranks into the caller's buffer are complete only after a b2mm_comm_barrier
 * (or any cross-rank synchronisation) that follows this call on every rank; likewise nobody may still
 * be reading the buffers when the call starts.
 */
B2MM_API int b2mm_gemm_allgather_fused(const void* A, const void* B_local, void* const* peer_slabs,
                                       int rank, int world, int64_t M, int64_t n_local, int64_t K,
                                       int dtype, void* workspace, size_t workspace_bytes,
                                       b2mm_stream_t stream);

/*
 * The same with ONE store per chunk: `slabs_multicast` is the multicast address (see b2mm_all_reduce_nvls_sync) of the
 * [world][M][n_local] gather buffer every rank allocated; the epilogue's TMA store goes to that address and the
 * NVSwitch replicates it into all ranks' copies, so a GPU sends each byte of its slab once instead of world - 1 times.
 * Replaces C_local = torch.matmul(A, B_local) + dist.all_gather (matmul_scaling_benchmark.py:211 + :221), like the call above.
 */
B2MM_API int b2mm_gemm_allgather_nvls(const void* A, const void* B_local, void* slabs_multicast, int rank, int world,
                                      int64_t M, int64_t n_local, int64_t K, int dtype, void* workspace,
                                      size_t workspace_bytes, b2mm_stream_t stream);

/*
 * In-place SUM over ranks of `count` elements, done by this library's own kernel over peer-mapped
 * buffers instead of NCCL (replaces dist.all_reduce(C, op=SUM), matmul_scaling_benchmark.py:125,150):
 * each rank pulls its 1/world slice from every rank over NVLink, adds in fp32, rounds once, and pushes
 * the result into every rank's buffer.  peer_bufs[r] = rank r's buffer as mapped here (b2mm_ipc_import;
 * own rank: the local pointer).  Includes the cross-rank barriers before and after.  world <= 8.
 */
B2MM_API int b2mm_all_reduce_p2p(void* comm, void* const* peer_bufs, int64_t count, int dtype, b2mm_stream_t stream);

/*
 * The same reduction with its cross-GPU synchronisation inside the kernels (epoch counters in the
 * peer-mapped signal pads, see b2mm_gemm_allreduce_fused) instead of NCCL barriers, and built to be
 * co-resident with a GEMM running on another stream: this is what overlaps dist.all_reduce(C) with the
 * next matmul in the double-buffered and pipelined modes of backup/matmul_overlap_benchmark.py:93-278.
 * Complete, on `stream`, when every rank's result has landed in every buffer.
 */
B2MM_API int b2mm_all_reduce_p2p_sync(void* comm, void* const* peer_bufs, void* const* peer_signals, int64_t count,
                                      int dtype, b2mm_stream_t stream);

/*
 * batch_parallel as a pipeline of this library's own kernels (matmul_scaling_benchmark.py:142 + :150):
 * C = bmm(A, B) in row panels on `compute`, written into peer_C[rank]; behind each panel, on
 * `comm_stream`: barrier, then the peer-memory all-reduce of that panel (as b2mm_all_reduce_p2p).
 * The panel GEMMs use the dynamic scheduler with retiring clusters so that the small reduce kernels
 * find SMs.  On return `compute` waits for the last reduction.  A, B, C: (local_batch, N, N).
 */
B2MM_API int b2mm_gemm_allreduce_p2p(void* comm, const void* A, const void* B, void* const* peer_C,
                                     int64_t N, int64_t local_batch, int dtype, int n_panels,
                                     void* workspace, size_t workspace_bytes,
                                     b2mm_stream_t compute, b2mm_stream_t comm_stream);

/*
 * batch_parallel with the reduction fused next to the GEMM (matmul_scaling_benchmark.py:142 + :150).
 * ONE GEMM over the whole local batch runs on the whole device; its epilogues count finished tiles per
 * row band (N / n_bands rows).  The reductions, one kernel per band, are queued at the same time on
 * `comm_stream`; their CTAs use no shared memory and few registers so that they are CO-RESIDENT with the
 * GEMM's CTAs: each waits on the device for its band, then for the same band on every rank (epoch counters
 * in peer-mapped signal pads), then pulls / adds in fp32 / pushes as b2mm_all_reduce_p2p.  No NCCL kernel,
 * nothing reserved, no panel launches.
 *   peer_signals[r]: rank r's signal pad, int32[16], zero-initialised once, mapped like peer_C.
 *   band_counters:   local scratch, int32[n_bands * local_batch] (cleared by the call).
 *   band rows must be a multiple of the GEMM's tile height (512).
 */
B2MM_API int b2mm_gemm_allreduce_fused(void* comm, const void* A, const void* B, void* const* peer_C,
                                       void* const* peer_signals, int* band_counters, int64_t N,
                                       int64_t local_batch, int dtype, int n_bands, void* workspace,
                                       size_t workspace_bytes, b2mm_stream_t compute, b2mm_stream_t comm_stream);

/*
 * The two calls above with the reduction done INSIDE the NVSwitch (NVLS): `multicast_buf` / `C_multicast` is a
 * multicast address (cuMulticastCreate / cuMulticastBindMem, mapped by the caller) bound to the same buffer
 * on every rank; each rank issues multimem.ld_reduce (sum over all ranks, fp32 accumulation in the switch, one
 * rounding) and multimem.st (result into every rank's buffer) for its 1/world slice.  Half the NVLink traffic
 * of pull/add/push at world = 8 and no arithmetic on the SMs.  Synchronisation (signal pads, band counters) and
 * completion semantics are those of b2mm_all_reduce_p2p_sync / b2mm_gemm_allreduce_fused; C_local is this rank's
 * ordinary (unicast) mapping of the same memory, which the GEMM writes.  Stands where dist.all_reduce(C, op=SUM)
 * (matmul_scaling_benchmark.py:125,150) lets NCCL pick its NVLS algorithm.
 */
B2MM_API int b2mm_all_reduce_nvls_sync(void* comm, void* multicast_buf, void* const* peer_signals, int64_t count, int dtype,
                                       b2mm_stream_t stream);
B2MM_API int b2mm_gemm_allreduce_nvls(void* comm, const void* A, const void* B, void* C_local, void* C_multicast,
                                      void* const* peer_signals, int* band_counters, int64_t N, int64_t local_batch,
                                      int dtype, int n_bands, void* workspace, size_t workspace_bytes,
                                      b2mm_stream_t compute, b2mm_stream_t comm_stream);

/*
 * matrix_parallel with the gather pipelined behind compute
 * (after backup/matmul_overlap_benchmark.py:93-180, made race-free):
 * C_local = A * B_local is computed in `n_panels` row panels on `compute`;
 * as each panel finishes, its rows are all-gathered on `comm_stream` into
 * recv_slabs ([world][M][n_local], the same layout b2mm_all_gather produces).
 * On return `compute` has been made to wait for the last gather.
 */
B2MM_API int b2mm_gemm_allgather_overlap(void* comm,
                                const void* A, const void* B_local, void* C_local, void* recv_slabs,
                                int64_t M, int64_t n_local, int64_t K,
                                int dtype, int n_panels,
                                void* workspace, size_t workspace_bytes,
                                b2mm_stream_t compute, b2mm_stream_t comm_stream);

/*
 * batch_parallel (matmul_scaling_benchmark.py:142 bmm + :150 all_reduce) with the
 * reduction pipelined behind compute, the double-buffered idea of
 * backup/matmul_overlap_benchmark.py:93-180: each local batch element is cut into
 * `n_panels` row panels; each panel is all-reduced on `comm_stream` as soon as
 * its GEMM has finished on `compute`.  A, B, C are (local_batch, N, N) contiguous.
 */
B2MM_API int b2mm_gemm_allreduce_overlap(void* comm,
                                const void* A, const void* B, void* C,
                                int64_t N, int64_t local_batch,
                                int dtype, int n_panels,
                                void* workspace, size_t workspace_bytes,
                                b2mm_stream_t compute, b2mm_stream_t comm_stream);

#ifdef __cplusplus
}
#endif
#endif /* B2MM_H_ */
