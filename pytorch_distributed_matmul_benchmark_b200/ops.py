"""Host-side operators with the signatures the reference scripts use.

`matmul` / `bmm` stand where the reference calls `torch.matmul` / `torch.bmm`
(matmul_benchmark.py:62, matmul_scaling_benchmark.py:92,142,211); `Communicator`
stands where it calls `dist.all_reduce` / `dist.all_gather`
(matmul_scaling_benchmark.py:150,221).  torch is used for device memory and
streams only: every operator hands raw device pointers to libb2mm.so.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import _lib

_DTYPES = {torch.bfloat16: _lib.BF16, torch.float16: _lib.FP16, torch.float32: _lib.FP32}


def _dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"b2mm supports bfloat16/float16/float32, got {t.dtype}") from None


def _stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("b2mm operators take CUDA tensors only (there is no CPU fallback)")


_workspaces = {}


def workspace_for(M: int, N: int, K: int, batch: int, dtype_code: int, device: torch.device):
    """Scratch for the fp32 (3xTF32) path; one cached buffer per device, grown on demand."""
    if dtype_code != _lib.FP32:  # bf16 / fp16 need none: skip the library call on the hot path
        return None, 0
    need = int(_lib.load().b2mm_gemm_workspace_bytes(M, N, K, batch, dtype_code))
    if need == 0:
        return None, 0
    key = (device.type, device.index)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < need:
        _workspaces[key] = buf = torch.empty(need, dtype=torch.uint8, device=device)
    return buf, need


def release_workspaces() -> None:
    """Drop cached scratch (the reference calls torch.cuda.empty_cache() after each size)."""
    _workspaces.clear()


def _rows_ok(t: torch.Tensor) -> bool:
    return t.stride(-1) == 1 and t.stride(-2) >= t.shape[-1]


def _call_gemm(A, B, out, M, N, K, lda, ldb, ldc, batch, sA, sB, sC, code) -> None:
    """One b2mm_gemm on the current stream of A's device.  Kept lean: at 2k and below the kernel is shorter
    than a careless host path (the reference's loop is launch-bound there too)."""
    ws, ws_bytes = workspace_for(M, N, K, batch, code, A.device)
    gemm = _lib.load().b2mm_gemm
    dev = A.device.index
    if dev == torch.cuda.current_device():
        rc = gemm(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, lda, ldb, ldc, batch, sA, sB, sC, code,
                  ws.data_ptr() if ws is not None else None, ws_bytes, torch.cuda.current_stream().cuda_stream)
    else:
        with torch.cuda.device(dev):
            rc = gemm(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, lda, ldb, ldc, batch, sA, sB, sC, code,
                      ws.data_ptr() if ws is not None else None, ws_bytes, torch.cuda.current_stream().cuda_stream)
    if rc:
        _lib.check(rc)


def matmul(A: torch.Tensor, B: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C = A @ B for 2-D row-major tensors; allocates C like torch.matmul unless `out` is given."""
    if A.dim() == 3 and B.dim() == 3:
        return bmm(A, B, out=out)
    if A.dim() != 2 or B.dim() != 2:
        raise RuntimeError(f"matmul expects 2-D (or 3-D batched) operands, got {A.dim()}-D and {B.dim()}-D")
    if not (A.is_cuda and B.is_cuda):
        _require_cuda(A, B)
    dtype = A.dtype
    if dtype != B.dtype:
        raise RuntimeError(f"expected both operands to have the same dtype, got {A.dtype} and {B.dtype}")
    M, K = A.shape
    K2, N = B.shape
    if K != K2:
        raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({M}x{K} and {K2}x{N})")
    code = _DTYPES.get(dtype)
    if code is None:
        _dtype_code(A)
    lda, a1 = A.stride()
    if a1 != 1 or lda < K:
        A = A.contiguous()
        lda = K
    ldb, b1 = B.stride()
    if b1 != 1 or ldb < N:
        B = B.contiguous()
        ldb = N
    if out is None:
        out = torch.empty((M, N), dtype=dtype, device=A.device)
    elif out.shape != (M, N) or out.dtype != dtype or not _rows_ok(out):
        raise RuntimeError("out has the wrong shape, dtype or layout")
    if M == 0 or N == 0:
        return out
    _call_gemm(A, B, out, M, N, K, lda, ldb, out.stride(0), 1, 0, 0, 0, code)
    return out


def bmm(A: torch.Tensor, B: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """C[b] = A[b] @ B[b]; mirrors torch.bmm, including empty batches."""
    if A.dim() != 3 or B.dim() != 3:
        raise RuntimeError("batch1 and batch2 must be 3D tensors")
    _require_cuda(A, B)
    if A.dtype != B.dtype:
        raise RuntimeError(f"expected both operands to have the same dtype, got {A.dtype} and {B.dtype}")
    b, M, K = A.shape
    b2, K2, N = B.shape
    if b != b2 or K != K2:
        raise RuntimeError(f"batch1 and batch2 shapes cannot be multiplied ({tuple(A.shape)} and {tuple(B.shape)})")
    if not _rows_ok(A):
        A = A.contiguous()
    if not _rows_ok(B):
        B = B.contiguous()
    if out is None:
        out = torch.empty((b, M, N), dtype=A.dtype, device=A.device)
    elif out.shape != (b, M, N) or out.dtype != A.dtype or not _rows_ok(out):
        raise RuntimeError("out has the wrong shape, dtype or layout")
    code = _dtype_code(A)
    if b == 0 or M == 0 or N == 0:
        return out
    _call_gemm(A, B, out, M, N, K, A.stride(1), B.stride(1), out.stride(1), b, A.stride(0), B.stride(0), out.stride(0), code)
    return out


class HostGemm:
    """C = A @ B for operands that live in (pinned) HOST memory: the end-to-end call.

    The transfers are pipelined with the math inside one call: B goes up first, then A in row
    panels; the GEMM of panel i starts as soon as panel i has landed, and its rows of C go back
    down while panel i+1 is still arriving (PCIe is full duplex).  Device staging buffers, three
    streams and the events are owned by this object and reused from call to call.
    """

    def __init__(self, M: int, N: int, K: int, dtype: torch.dtype, device="cuda", panels: int = 8):
        if M % panels:
            raise RuntimeError(f"M={M} must be divisible by panels={panels}")
        self.M, self.N, self.K, self.dtype, self.panels = M, N, K, dtype, panels
        self.device = torch.device(device)
        self.dA = torch.empty(M, K, dtype=dtype, device=self.device)
        # B is read by every panel's GEMM, so it is double buffered: the next call's B goes up while this call's
        # GEMMs are still running.  A and C are reused panel by panel (per-panel events below).
        self.dB = [torch.empty(K, N, dtype=dtype, device=self.device) for _ in range(2)]
        self.dC = torch.empty(M, N, dtype=dtype, device=self.device)
        self.up = torch.cuda.Stream(device=self.device)
        self.down = torch.cuda.Stream(device=self.device)
        self.ev_b = torch.cuda.Event()
        self.ev_a = [torch.cuda.Event() for _ in range(panels)]
        self.ev_c = [torch.cuda.Event() for _ in range(panels)]  # GEMM of panel i finished (dA[i], dB read; dC[i] written)
        self.ev_d = [torch.cuda.Event() for _ in range(panels)]  # C panel i has left the device
        self.ev_b_free = [torch.cuda.Event(), torch.cuda.Event()]  # last GEMM that read dB[j] finished
        self.ev_done = torch.cuda.Event()
        self.calls = 0

    def __call__(self, hA: torch.Tensor, hB: torch.Tensor, hC: torch.Tensor) -> torch.Tensor:
        for t in (hA, hB, hC):
            if t.is_cuda or not t.is_pinned():
                raise RuntimeError("HostGemm takes pinned host tensors")
        main = torch.cuda.current_stream(self.device)
        rows = self.M // self.panels
        first = self.calls == 0
        dB = self.dB[self.calls % 2]
        self.down.wait_stream(main)
        # Back-to-back calls overlap: the upload stream never waits for the previous call as a whole, only for the
        # GEMM that last read the buffer it is about to overwrite (recorded two calls ago for dB, one call ago per
        # A panel), so in steady state the host->device link stays busy and a call costs one upload of A and B.
        with torch.cuda.stream(self.up):
            if self.calls >= 2:
                self.up.wait_event(self.ev_b_free[self.calls % 2])
            dB.copy_(hB, non_blocking=True)
            self.ev_b.record()
            for i in range(self.panels):
                if not first:
                    self.up.wait_event(self.ev_c[i])
                self.dA[i * rows:(i + 1) * rows].copy_(hA[i * rows:(i + 1) * rows], non_blocking=True)
                self.ev_a[i].record()
        main.wait_event(self.ev_b)
        for i in range(self.panels):
            main.wait_event(self.ev_a[i])
            if not first:
                main.wait_event(self.ev_d[i])  # the previous call's C panel has been copied out of dC[i]
            matmul(self.dA[i * rows:(i + 1) * rows], dB, out=self.dC[i * rows:(i + 1) * rows])
            self.ev_c[i].record(main)
            with torch.cuda.stream(self.down):
                self.down.wait_event(self.ev_c[i])
                hC[i * rows:(i + 1) * rows].copy_(self.dC[i * rows:(i + 1) * rows], non_blocking=True)
                self.ev_d[i].record()
        self.ev_b_free[self.calls % 2].record(main)
        self.ev_done.record(self.down)
        main.wait_event(self.ev_done)  # on the caller's stream, the call is complete when hC is
        self.calls += 1
        return hC


class SharedSlabs:
    """A rank's buffer plus the pointers under which every rank's buffer is visible here.

    The peer mappings are reference-counted imports held by libb2mm.so; `close()` gives them back.  Peers must
    close before the owner's memory is freed, so Communicator keeps every SharedSlabs it made alive until
    `release_shared()` has closed them on all ranks and passed a barrier.
    """

    def __init__(self, local: torch.Tensor, peer_ptrs, imported_handles: List[bytes]):
        self.local, self.peer_ptrs, self._imported = local, peer_ptrs, imported_handles

    def close(self) -> None:
        lib = _lib.load()
        for h in self._imported:
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            _lib.check(lib.b2mm_ipc_release(ctypes.cast(buf, ctypes.c_void_p)))
        self._imported = []
        self.peer_ptrs = None


class Communicator:
    """An ncclComm_t owned by libb2mm.so, bootstrapped over an existing torch process group.

    torch.distributed performs the rendezvous exactly as the reference does
    (matmul_scaling_benchmark.py:15-24); it is then only used to ship the NCCL
    unique id.  Payload collectives go through the C ABI.
    """

    def __init__(self, rank: int, world_size: int, device: torch.device):
        import torch.distributed as dist
        self.rank, self.world_size, self.device = rank, world_size, device
        self._comm = ctypes.c_void_p()
        lib = _lib.load()
        id_buf = (ctypes.c_ubyte * _lib.UNIQUE_ID_BYTES)()
        if rank == 0:
            _lib.check(lib.b2mm_comm_unique_id(ctypes.cast(id_buf, ctypes.c_void_p)))
        payload = [bytes(id_buf)]
        dist.broadcast_object_list(payload, src=0)
        id_buf = (ctypes.c_ubyte * _lib.UNIQUE_ID_BYTES).from_buffer_copy(payload[0])
        with torch.cuda.device(device):
            _lib.check(lib.b2mm_comm_init(ctypes.cast(id_buf, ctypes.c_void_p), rank, world_size,
                                          ctypes.byref(self._comm)))
        self.comm_stream = torch.cuda.Stream(device=device, priority=-1)
        self._shared: List[SharedSlabs] = []
        self._multicast: list = []
        self._signals: Optional[SharedSlabs] = None

    def release_shared(self) -> None:
        """Unmap every buffer handed out by share() / shared_slabs() / shared_like() and let the owners free them.

        Collective: every rank calls it at the same point (the scripts do after each matrix size, before
        torch.cuda.empty_cache()).  Order: this rank's kernels finish, its imports are closed, all ranks meet at a
        barrier, and only then are the local tensors dropped — an owner never frees memory a peer still maps.
        """
        import torch.distributed as dist
        torch.cuda.synchronize(self.device)
        for sh in self._shared:
            sh.close()
        if dist.is_initialized():  # unconditional: ranks may disagree on whether anything was shared after an error
            dist.barrier()
        self._shared = []
        for mc in self._multicast:  # collective as well (every rank made the same buffers in the same order)
            mc.close()
        self._multicast = []

    def destroy(self) -> None:
        if self._comm:
            if self._signals is not None:
                self._shared.append(self._signals)
                self._signals = None
            self.release_shared()
            _lib.check(_lib.load().b2mm_comm_destroy(self._comm))
            self._comm = ctypes.c_void_p()

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        """In-place SUM, like dist.all_reduce(t, op=ReduceOp.SUM)."""
        _require_cuda(t)
        if not t.is_contiguous():
            raise RuntimeError("all_reduce needs a contiguous tensor")
        _lib.check(_lib.load().b2mm_all_reduce_sum(self._comm, t.data_ptr() if t.numel() else None, t.numel(),
                                                   _dtype_code(t), _stream_ptr()))
        return t

    def all_gather_into(self, slabs: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        """slabs[r] = rank r's t; `slabs` is one contiguous (world, *t.shape) tensor."""
        _require_cuda(slabs, t)
        if not (slabs.is_contiguous() and t.is_contiguous()):
            raise RuntimeError("all_gather needs contiguous tensors")
        if slabs.numel() != t.numel() * self.world_size or slabs.dtype != t.dtype:
            raise RuntimeError("slabs must hold world_size copies of the input")
        _lib.check(_lib.load().b2mm_all_gather(self._comm, t.data_ptr(), slabs.data_ptr(), t.numel(), _dtype_code(t),
                                               _stream_ptr()))
        return slabs

    def all_gather(self, t: torch.Tensor) -> List[torch.Tensor]:
        """Returns the list dist.all_gather(list, t) would fill (views of one slab buffer)."""
        slabs = torch.empty((self.world_size,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        self.all_gather_into(slabs, t)
        return list(slabs.unbind(0))

    def barrier(self) -> None:
        """Cross-rank barrier on the current stream (dist.barrier() for enqueued work)."""
        _lib.check(_lib.load().b2mm_comm_barrier(self._comm, _stream_ptr()))

    def barrier_signals(self) -> None:
        """The same barrier through the ranks' peer-mapped signal pads (one warp, no NCCL kernel): a third of the latency.
        B2MM_SIGNAL_BARRIER=0 falls back to the NCCL barrier."""
        import os
        if os.environ.get("B2MM_SIGNAL_BARRIER", "1") == "0":
            return self.barrier()
        _lib.check(_lib.load().b2mm_comm_barrier_signals(self._comm, self._signal_pads().peer_ptrs, _stream_ptr()))

    def share(self, local: torch.Tensor, track: bool = True) -> "SharedSlabs":
        """Map every rank's `local` (same shape on all ranks) into every other rank (CUDA IPC over NVLink).

        torch allocates (device memory is torch's business); the 64-byte handles travel through the
        existing process group; the library only ever sees the resulting raw pointers.
        """
        import torch.distributed as dist
        lib = _lib.load()
        handle = (ctypes.c_ubyte * 64)()
        offset = ctypes.c_int64()
        _lib.check(lib.b2mm_ipc_export(local.data_ptr(), ctypes.cast(handle, ctypes.c_void_p), ctypes.byref(offset)))
        gathered = [None] * self.world_size
        dist.all_gather_object(gathered, (bytes(handle), offset.value))
        ptrs, imported = [], []
        with torch.cuda.device(self.device):
            for r, (h, off) in enumerate(gathered):
                if r == self.rank:
                    ptrs.append(local.data_ptr())
                else:
                    p = ctypes.c_void_p()
                    buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                    _lib.check(lib.b2mm_ipc_import(ctypes.cast(buf, ctypes.c_void_p), off, ctypes.byref(p)))
                    ptrs.append(p.value)
                    imported.append(h)
        shared = SharedSlabs(local, (ctypes.c_void_p * self.world_size)(*ptrs), imported)
        if track:
            self._shared.append(shared)
        return shared

    def shared_slabs(self, M: int, n_local: int, dtype: torch.dtype) -> "SharedSlabs":
        """A [world][M][n_local] gather buffer on every rank, each visible to every other rank."""
        return self.share(torch.empty((self.world_size, M, n_local), dtype=dtype, device=self.device))

    def gather_buffer(self, M: int, n_local: int, dtype: torch.dtype):
        """The [world][M][n_local] buffer of the fused matrix_parallel path: multicast (one TMA store per chunk, replicated
        by the NVSwitch) where the hardware has it, else peer-mapped (one store per destination).  Collective."""
        buf = self._try_multicast(torch.empty((self.world_size, M, n_local), dtype=dtype, device="meta"))
        return buf if buf is not None else self.shared_slabs(M, n_local, dtype)

    def shared_like(self, t: torch.Tensor) -> "SharedSlabs":
        """A buffer shaped like `t` on every rank, each visible to every other rank (for the p2p all_reduce)."""
        return self.share(torch.empty_like(t))

    def multicast_like(self, t: torch.Tensor):
        """A buffer shaped like `t` on every rank, all bound to one NVSwitch multicast object (NVLS): `.local` is this
        rank's tensor, `.mc_ptr` the address through which the switch reduces / broadcasts.  Collective."""
        from .multicast import MulticastBuffer
        buf = MulticastBuffer(self.rank, self.world_size, self.device, tuple(t.shape), t.dtype)
        self._multicast.append(buf)
        return buf

    def nvls_available(self) -> bool:
        """True when the reductions can be done inside the NVSwitch (multicast objects supported; B2MM_NVLS=0 opts out)."""
        import os
        if os.environ.get("B2MM_NVLS", "1") == "0" or self.world_size < 2:
            return False
        if getattr(self, "_nvls", None) is None:
            from .multicast import multicast_supported
            try:
                self._nvls = multicast_supported(self.device.index if self.device.index is not None else torch.cuda.current_device())
            except Exception:  # no cuda-python bindings, or a driver without the attribute
                self._nvls = False
        return self._nvls

    def reduce_buffer_like(self, t: torch.Tensor):
        """The buffer the fused / co-resident all_reduce paths work on: a multicast (NVLS) buffer where the hardware
        has it, otherwise a peer-mapped one (pull / add / push).  Collective; `.local` is this rank's tensor."""
        buf = self._try_multicast(t)
        return buf if buf is not None else self.shared_like(t)

    def _try_multicast(self, t: torch.Tensor):
        """A multicast buffer, or None when the box cannot make one; the outcome is the same on every rank (rank 0's
        failure to create the object is broadcast), so all ranks take the peer-mapped path together."""
        if not self.nvls_available():
            return None
        try:
            return self.multicast_like(t)
        except RuntimeError as e:
            if self.rank == 0:
                print(f"b2mm: NVLS multicast buffer unavailable ({e}); using peer-mapped buffers")
            self._nvls = False
            return None

    def all_reduce_nvls(self, buf, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """In-place SUM of buf.local over ranks, reduced inside the NVSwitch (multimem.ld_reduce / multimem.st)."""
        t = buf.local
        _lib.check(_lib.load().b2mm_all_reduce_nvls_sync(self._comm, buf.mc_ptr, self._signal_pads().peer_ptrs, t.numel(),
                                                         _dtype_code(t), _stream_ptr(stream)))
        return t

    def all_reduce_p2p(self, shared: "SharedSlabs") -> torch.Tensor:
        """In-place SUM of shared.local over ranks by the library's own NVLink kernel (fp32 adds, one rounding)."""
        t = shared.local
        _lib.check(_lib.load().b2mm_all_reduce_p2p(self._comm, shared.peer_ptrs, t.numel(), _dtype_code(t), _stream_ptr()))
        return t

    def bmm_allreduce_p2p(self, A: torch.Tensor, B: torch.Tensor, shared_C: "SharedSlabs", n_panels: int = 4) -> torch.Tensor:
        """C = bmm(A, B) into shared_C.local in row panels, each panel summed over ranks by the p2p kernel behind it."""
        _require_cuda(A, B)
        A, B = A.contiguous(), B.contiguous()  # the C entry points take densely packed batches (no-op when they are)
        b, N, _ = A.shape
        code = _dtype_code(A)
        ws, ws_bytes = workspace_for(N // n_panels, N, N, 1, code, A.device)
        self.comm_stream.wait_stream(torch.cuda.current_stream())
        _lib.check(_lib.load().b2mm_gemm_allreduce_p2p(
            self._comm, A.data_ptr(), B.data_ptr(), shared_C.peer_ptrs, N, b, code, n_panels,
            ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr(), self.comm_stream.cuda_stream))
        return shared_C.local

    def _signal_pads(self) -> "SharedSlabs":
        if self._signals is None:
            pad = torch.zeros(64, dtype=torch.int32, device=self.device)
            torch.cuda.synchronize(self.device)
            self._signals = self.share(pad, track=False)  # lives as long as the communicator
            self._band_counters = torch.zeros(256, dtype=torch.int32, device=self.device)
        return self._signals

    def all_reduce_sync(self, shared: "SharedSlabs", stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """In-place SUM over ranks by the self-synchronising peer-memory kernels (no NCCL); may run beside a GEMM."""
        if hasattr(shared, "mc_ptr"):
            return self.all_reduce_nvls(shared, stream)
        t = shared.local
        _lib.check(_lib.load().b2mm_all_reduce_p2p_sync(self._comm, shared.peer_ptrs, self._signal_pads().peer_ptrs, t.numel(),
                                                        _dtype_code(t), _stream_ptr(stream)))
        return t

    def bmm_allreduce_fused(self, A: torch.Tensor, B: torch.Tensor, shared_C: "SharedSlabs", n_panels: int = 4) -> torch.Tensor:
        """C = bmm(A, B) as one GEMM whose epilogue publishes finished row bands; each band is summed over ranks by a
        self-synchronising kernel (in the NVSwitch, or over peer mappings) that shares the SMs with the rest of the
        GEMM (no NCCL, nothing reserved).  `n_panels` = row bands per batch element."""
        _require_cuda(A, B)
        A, B = A.contiguous(), B.contiguous()  # the C entry points take densely packed batches (no-op when they are)
        self._signal_pads()
        b, N, _ = A.shape
        code = _dtype_code(A)
        # the library runs ONE batched GEMM over the whole local batch (the bands are only its signalling unit), so the
        # fp32 split scratch is sized for the whole batch, not for one row panel
        ws, ws_bytes = workspace_for(N, N, N, b, code, A.device)
        self.comm_stream.wait_stream(torch.cuda.current_stream())
        if hasattr(shared_C, "mc_ptr"):  # MulticastBuffer: the bands are reduced inside the NVSwitch
            _lib.check(_lib.load().b2mm_gemm_allreduce_nvls(
                self._comm, A.data_ptr(), B.data_ptr(), shared_C.local.data_ptr(), shared_C.mc_ptr, self._signals.peer_ptrs,
                self._band_counters.data_ptr(), N, b, code, n_panels,
                ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr(), self.comm_stream.cuda_stream))
            return shared_C.local
        _lib.check(_lib.load().b2mm_gemm_allreduce_fused(
            self._comm, A.data_ptr(), B.data_ptr(), shared_C.peer_ptrs, self._signals.peer_ptrs,
            self._band_counters.data_ptr(), N, b, code, n_panels,
            ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr(), self.comm_stream.cuda_stream))
        return shared_C.local

    def matmul_allgather_fused(self, A: torch.Tensor, B_local: torch.Tensor, shared: "SharedSlabs") -> torch.Tensor:
        """C_local = A @ B_local with the all_gather done by the GEMM epilogue (peer TMA stores), then a barrier."""
        _require_cuda(A, B_local)
        M, K = A.shape
        n_local = B_local.shape[1]
        if tuple(shared.local.shape) != (self.world_size, M, n_local) or shared.local.dtype != A.dtype:
            raise RuntimeError("shared slabs do not match the problem")
        if not (A.is_contiguous() and B_local.is_contiguous()):  # the C entry points take lda = K, ldb = n_local
            A, B_local = A.contiguous(), B_local.contiguous()
        code = _dtype_code(A)
        ws, ws_bytes = workspace_for(M, n_local, K, 1, code, A.device)
        if hasattr(shared, "mc_ptr"):  # MulticastBuffer: one store per chunk, replicated by the NVSwitch
            _lib.check(_lib.load().b2mm_gemm_allgather_nvls(
                A.data_ptr(), B_local.data_ptr(), shared.mc_ptr, self.rank, self.world_size, M, n_local, K, code,
                ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr()))
            self.barrier_signals()
            return shared.local
        _lib.check(_lib.load().b2mm_gemm_allgather_fused(
            A.data_ptr(), B_local.data_ptr(), shared.peer_ptrs, self.rank, self.world_size, M, n_local, K, code,
            ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr()))
        self.barrier_signals()
        return shared.local

    def matmul_allgather_overlap(self, A: torch.Tensor, B_local: torch.Tensor, C_local: torch.Tensor,
                                 slabs: torch.Tensor, n_panels: int = 4) -> torch.Tensor:
        """C_local = A @ B_local in row panels, each panel gathered on the comm stream while the next computes."""
        _require_cuda(A, B_local, C_local, slabs)
        A, B_local = A.contiguous(), B_local.contiguous()
        if not (C_local.is_contiguous() and slabs.is_contiguous()):
            raise RuntimeError("C_local and slabs must be contiguous")
        M, K = A.shape
        n_local = B_local.shape[1]
        code = _dtype_code(A)
        ws, ws_bytes = workspace_for(M // n_panels, n_local, K, 1, code, A.device)
        self.comm_stream.wait_stream(torch.cuda.current_stream())
        _lib.check(_lib.load().b2mm_gemm_allgather_overlap(
            self._comm, A.data_ptr(), B_local.data_ptr(), C_local.data_ptr(), slabs.data_ptr(), M, n_local, K, code,
            n_panels, ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr(),
            self.comm_stream.cuda_stream))
        return slabs

    def bmm_allreduce_overlap(self, A: torch.Tensor, B: torch.Tensor, C: torch.Tensor, n_panels: int = 4) -> torch.Tensor:
        """C = bmm(A, B) with each panel all-reduced on the comm stream while the next computes."""
        _require_cuda(A, B, C)
        A, B = A.contiguous(), B.contiguous()  # the C entry points take densely packed batches (no-op when they are)
        b, N, _ = A.shape
        code = _dtype_code(A)
        ws, ws_bytes = workspace_for(N // n_panels, N, N, 1, code, A.device)
        self.comm_stream.wait_stream(torch.cuda.current_stream())
        _lib.check(_lib.load().b2mm_gemm_allreduce_overlap(
            self._comm, A.data_ptr(), B.data_ptr(), C.data_ptr(), N, b, code, n_panels,
            ws.data_ptr() if ws is not None else None, ws_bytes, _stream_ptr(), self.comm_stream.cuda_stream))
        return C
