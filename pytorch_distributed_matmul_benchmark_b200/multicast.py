"""Multicast (NVLS) buffers: one allocation per rank, all bound to one NVSwitch multicast object.

Device memory is plumbing, so this lives in the Python host layer next to torch's allocator: the buffer is made
with the CUDA virtual-memory API (cuda-python bindings of cuMemCreate / cuMulticastCreate / cuMulticastBindMem),
wrapped as a torch tensor, and libb2mm.so only ever sees two raw pointers: the rank's own (unicast) mapping, which
the GEMM writes, and the multicast mapping, through which `multimem.ld_reduce` / `multimem.st` reach every
rank's copy (b2mm_all_reduce_nvls_sync, b2mm_gemm_allreduce_nvls in include/b2mm.h).

The multicast object is created by rank 0 and handed to the other ranks as a POSIX file descriptor over a Unix
socket (SCM_RIGHTS); the socket path travels over the existing process group.  No counterpart in the reference:
torch's NCCL process group does the equivalent internally when it picks its NVLS algorithm for
dist.all_reduce (matmul_scaling_benchmark.py:150).
"""
from __future__ import annotations

import os
import socket
import tempfile
from typing import Tuple

import torch
import torch.distributed as dist


def _drv():
    try:
        from cuda.bindings import driver
    except ImportError:  # older cuda-python layout
        from cuda import cuda as driver
    return driver


def _ok(result, what: str):
    """cuda-python returns (CUresult, value...) tuples; raise on failure, unwrap the value."""
    err = result[0]
    if int(err) != 0:
        name = getattr(err, "name", None) or str(err)  # the enum's own name; no driver call needed to report an error
        raise RuntimeError(f"{what} failed: {name} ({int(err)})")
    rest = result[1:]
    return rest[0] if len(rest) == 1 else rest


def multicast_supported(device_index: int) -> bool:
    drv = _drv()
    dev = _ok(drv.cuDeviceGet(device_index), "cuDeviceGet")
    return bool(_ok(drv.cuDeviceGetAttribute(drv.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev),
                    "cuDeviceGetAttribute(MULTICAST_SUPPORTED)"))


class _Blob:
    """Just enough of __cuda_array_interface__ for torch.as_tensor to wrap raw device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


SOCKET_TIMEOUT_S = 60.0  # a peer that died before connecting must not leave the others waiting for ever


def _send_fd_to_peers(fd: int, n_peers: int, path: str) -> None:
    try:
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as srv:
            srv.settimeout(SOCKET_TIMEOUT_S)
            srv.bind(path)
            srv.listen(n_peers)
            dist.barrier()  # peers connect only once the socket exists
            for _ in range(n_peers):
                conn, _ = srv.accept()
                with conn:
                    socket.send_fds(conn, [b"m"], [fd])
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass


def _recv_fd(path: str) -> int:
    dist.barrier()
    with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
        c.settimeout(SOCKET_TIMEOUT_S)
        c.connect(path)
        _, fds, _, _ = socket.recv_fds(c, 1, 1)
    if not fds:
        raise RuntimeError("no file descriptor received for the multicast object")
    return fds[0]


def _agree(phase: str, error) -> None:
    """Collective end of a set-up phase: every rank reports its own outcome, and if ANY rank failed, EVERY rank raises
    (the same message everywhere), so nobody is left waiting in the next barrier for a rank that has already given up."""
    outcomes = [None] * dist.get_world_size()
    dist.all_gather_object(outcomes, None if error is None else str(error))
    failed = [(r, e) for r, e in enumerate(outcomes) if e is not None]
    if failed:
        raise RuntimeError(f"multicast set-up failed in phase '{phase}': " + "; ".join(f"rank {r}: {e}" for r, e in failed))


class MulticastBuffer:
    """`local`: this rank's tensor (its own physical memory); `mc_ptr`: the multicast address of the same bytes.

    Collective constructor and `close()`: every rank of the process group calls them at the same point.
    """

    def __init__(self, rank: int, world_size: int, device: torch.device, shape: Tuple[int, ...], dtype: torch.dtype):
        drv = _drv()
        self.rank, self.world_size, self.device = rank, world_size, device
        self._closed = False
        torch.cuda.set_device(device)
        torch.zeros(1, device=device)  # makes sure torch's primary context is current on this thread
        index = device.index if device.index is not None else torch.cuda.current_device()
        dev = _ok(drv.cuDeviceGet(index), "cuDeviceGet")
        if not multicast_supported(index):
            raise RuntimeError("this GPU / driver does not support multicast objects (NVLS)")
        nbytes = int(torch.empty(0, dtype=dtype).element_size())
        for s in shape:
            nbytes *= int(s)
        fd_type = drv.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR

        alloc = drv.CUmemAllocationProp()
        alloc.type = drv.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        alloc.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        alloc.location.id = index
        alloc.requestedHandleTypes = fd_type
        gran_mem = _ok(drv.cuMemGetAllocationGranularity(
            alloc, drv.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity")

        prop = drv.CUmulticastObjectProp()
        prop.numDevices = world_size
        prop.size = max(nbytes, 1)
        prop.handleTypes = fd_type
        prop.flags = 0
        gran_mc = _ok(drv.cuMulticastGetGranularity(
            prop, drv.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity")
        gran = max(int(gran_mem), int(gran_mc))
        size = (max(nbytes, 1) + gran - 1) // gran * gran
        prop.size = size
        self.size, self.nbytes = size, nbytes

        # rank 0 creates the multicast object; the others import it from its file descriptor.  A failure on rank 0
        # (e.g. multicast disabled by the fabric manager) is broadcast, so that every rank raises instead of waiting
        payload = [None]
        fd = -1
        if rank == 0:
            try:
                self._mc = _ok(drv.cuMulticastCreate(prop), "cuMulticastCreate")
                fd = int(_ok(drv.cuMemExportToShareableHandle(self._mc, fd_type, 0), "cuMemExportToShareableHandle"))
                payload = [os.path.join(tempfile.gettempdir(), f"b2mm_mc_{os.getpid()}_{id(self):x}.sock")]
            except RuntimeError as e:
                payload = [f"ERROR: {e}"]
                if getattr(self, "_mc", None) is not None:  # created but not exportable: do not leak the object
                    drv.cuMemRelease(self._mc)
                    self._mc = None
        dist.broadcast_object_list(payload, src=0)
        if payload[0].startswith("ERROR: "):
            self._closed = True
            raise RuntimeError(payload[0][len("ERROR: "):])
        # From here on every phase ends in _agree(): a failure on one rank (import, out of memory, bind, a socket that
        # timed out) makes all ranks unwind what they built and raise together; Communicator._try_multicast then takes
        # the peer-mapped path on every rank.
        self._mem = self._va = self._mc_va = None
        self._bound = self._mapped = self._mc_mapped = False
        err = None
        try:
            if rank == 0:
                try:
                    _send_fd_to_peers(fd, world_size - 1, payload[0])
                finally:
                    os.close(fd)
            else:
                fd = _recv_fd(payload[0])
                try:
                    self._mc = _ok(drv.cuMemImportFromShareableHandle(fd, fd_type), "cuMemImportFromShareableHandle")
                finally:
                    os.close(fd)
            _ok(drv.cuMulticastAddDevice(self._mc, dev), "cuMulticastAddDevice")
        except (RuntimeError, OSError) as e:
            err = e
        self._phase_end("hand-over of the multicast object / add device", err)  # every device added before anybody binds

        # this rank's physical memory, its ordinary mapping, and its binding into the multicast object
        access = drv.CUmemAccessDesc()
        access.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        access.location.id = index
        access.flags = drv.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
        try:
            self._mem = _ok(drv.cuMemCreate(size, alloc, 0), "cuMemCreate")
            self._va = _ok(drv.cuMemAddressReserve(size, gran, 0, 0), "cuMemAddressReserve")
            _ok(drv.cuMemMap(self._va, size, 0, self._mem, 0), "cuMemMap")
            self._mapped = True
            _ok(drv.cuMemSetAccess(self._va, size, [access], 1), "cuMemSetAccess")
            _ok(drv.cuMulticastBindMem(self._mc, 0, self._mem, 0, size, 0), "cuMulticastBindMem")
            self._bound = True
        except RuntimeError as e:
            err = e
        self._phase_end("allocate / map / bind", err)  # every rank's memory is bound before the multicast address is used

        try:
            self._mc_va = _ok(drv.cuMemAddressReserve(size, gran, 0, 0), "cuMemAddressReserve(multicast)")
            _ok(drv.cuMemMap(self._mc_va, size, 0, self._mc, 0), "cuMemMap(multicast)")
            self._mc_mapped = True
            _ok(drv.cuMemSetAccess(self._mc_va, size, [access], 1), "cuMemSetAccess(multicast)")
        except RuntimeError as e:
            err = e
        self._phase_end("map the multicast address", err)

        self.ptr, self.mc_ptr = int(self._va), int(self._mc_va)
        flat = torch.as_tensor(_Blob(self.ptr, nbytes), device=device)
        self.local = flat.view(dtype).view(*shape)
        self.local.zero_()
        torch.cuda.synchronize(device)
        dist.barrier()

    def _phase_end(self, phase: str, error) -> None:
        try:
            _agree(phase, error)
        except RuntimeError:
            self._unwind()
            raise

    def _unwind(self) -> None:
        """Give back whatever exists, newest first; driver errors here are ignored (we are already failing)."""
        drv = _drv()
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        steps = []
        if self._mc_va is not None:
            if self._mc_mapped:
                steps.append(lambda: drv.cuMemUnmap(self._mc_va, self.size))
            steps.append(lambda: drv.cuMemAddressFree(self._mc_va, self.size))
        if self._bound:
            steps.append(lambda: drv.cuMulticastUnbind(self._mc, drv.cuDeviceGet(index)[1], 0, self.size))
        if self._va is not None:
            if self._mapped:
                steps.append(lambda: drv.cuMemUnmap(self._va, self.size))
            steps.append(lambda: drv.cuMemAddressFree(self._va, self.size))
        if self._mem is not None:
            steps.append(lambda: drv.cuMemRelease(self._mem))
        if getattr(self, "_mc", None) is not None:
            steps.append(lambda: drv.cuMemRelease(self._mc))
        for step in steps:
            try:
                step()
            except Exception:
                pass
        self._mem = self._va = self._mc_va = self._mc = None
        self._bound = self._mapped = self._mc_mapped = False
        self.local = None
        self._closed = True

    def close(self) -> None:
        if self._closed:
            return
        drv = _drv()
        torch.cuda.synchronize(self.device)
        dist.barrier()  # nobody is still reducing through the multicast address
        self.local = None
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        dev = _ok(drv.cuDeviceGet(index), "cuDeviceGet")
        _ok(drv.cuMemUnmap(self._mc_va, self.size), "cuMemUnmap(multicast)")
        _ok(drv.cuMemAddressFree(self._mc_va, self.size), "cuMemAddressFree(multicast)")
        _ok(drv.cuMulticastUnbind(self._mc, dev, 0, self.size), "cuMulticastUnbind")
        _ok(drv.cuMemUnmap(self._va, self.size), "cuMemUnmap")
        _ok(drv.cuMemAddressFree(self._va, self.size), "cuMemAddressFree")
        _ok(drv.cuMemRelease(self._mem), "cuMemRelease")
        _ok(drv.cuMemRelease(self._mc), "cuMemRelease(multicast)")
        self._closed = True
