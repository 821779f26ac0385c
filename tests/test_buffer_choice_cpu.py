"""Which buffer the fused paths get (multicast where the box has NVLS, peer-mapped otherwise) — the decision logic of
ops.Communicator, exercised without a device."""
import torch

from pytorch_distributed_matmul_benchmark_b200 import ops


class _Multicast:
    mc_ptr = 0x1000


def _bare_comm():
    c = ops.Communicator.__new__(ops.Communicator)  # no NCCL, no device: only the choice logic is under test
    c.rank, c.world_size, c.device = 0, 2, torch.device("cpu")
    c._multicast, c._shared, c._nvls = [], [], None
    c.shared_like = lambda t: "peer-mapped"
    c.shared_slabs = lambda M, n, dt: "peer-mapped slabs"
    return c


def test_multicast_buffers_are_used_where_available():
    c = _bare_comm()
    c.nvls_available = lambda: True
    c.multicast_like = lambda t: _Multicast()
    assert hasattr(c.reduce_buffer_like(torch.empty(4, 4)), "mc_ptr")
    assert hasattr(c.gather_buffer(8, 4, torch.bfloat16), "mc_ptr")


def test_peer_mapped_buffers_otherwise():
    c = _bare_comm()
    c.nvls_available = lambda: False
    assert c.reduce_buffer_like(torch.empty(4, 4)) == "peer-mapped"
    assert c.gather_buffer(8, 4, torch.bfloat16) == "peer-mapped slabs"


def test_a_failed_multicast_setup_falls_back_once_and_for_all(capsys):
    """Rank 0's failure is broadcast by MulticastBuffer, so every rank gets the same RuntimeError and takes this path."""
    c = _bare_comm()
    calls = []

    def failing(t):
        calls.append(1)
        raise RuntimeError("cuMulticastCreate failed: CUDA_ERROR_NOT_SUPPORTED (801)")

    c.multicast_like = failing
    c.nvls_available = ops.Communicator.nvls_available.__get__(c)
    c._nvls = True  # as if the attribute query had said yes
    assert c.reduce_buffer_like(torch.empty(4, 4)) == "peer-mapped"
    assert "NVLS multicast buffer unavailable" in capsys.readouterr().out
    assert c._nvls is False and c.gather_buffer(8, 4, torch.bfloat16) == "peer-mapped slabs" and len(calls) == 1


def test_nvls_opt_out(monkeypatch):
    c = _bare_comm()
    c._nvls = True
    monkeypatch.setenv("B2MM_NVLS", "0")
    assert ops.Communicator.nvls_available(c) is False
    monkeypatch.delenv("B2MM_NVLS")
    assert ops.Communicator.nvls_available(c) is True
    c.world_size = 1
    assert ops.Communicator.nvls_available(c) is False


def test_a_setup_phase_fails_on_every_rank_when_one_rank_failed(monkeypatch):
    """multicast._agree: all ranks report their outcome; one failure anywhere raises everywhere, naming the rank."""
    import pytest
    from pytorch_distributed_matmul_benchmark_b200 import multicast

    class FakeDist:
        def __init__(self, others):
            self.others = others

        def get_world_size(self):
            return len(self.others) + 1

        def all_gather_object(self, out, mine):
            out[:] = [mine] + self.others

    monkeypatch.setattr(multicast, "dist", FakeDist([None, None]))
    multicast._agree("bind", None)  # everybody fine: no exception
    monkeypatch.setattr(multicast, "dist", FakeDist([None, "cuMemCreate failed: CUDA_ERROR_OUT_OF_MEMORY (2)"]))
    with pytest.raises(RuntimeError, match=r"phase 'bind'.*rank 2: cuMemCreate failed"):
        multicast._agree("bind", None)  # THIS rank succeeded and still raises
    with pytest.raises(RuntimeError, match="rank 0: boom"):
        multicast._agree("bind", RuntimeError("boom"))


def test_unwind_releases_what_was_built_newest_first(monkeypatch):
    from pytorch_distributed_matmul_benchmark_b200 import multicast
    log = []

    class FakeDrv:
        def __getattr__(self, name):
            def call(*a):
                log.append(name)
                return (0, 0)
            return call

    monkeypatch.setattr(multicast, "_drv", lambda: FakeDrv())
    b = multicast.MulticastBuffer.__new__(multicast.MulticastBuffer)
    b.device, b.size = torch.device("cpu"), 4096
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    b._mc, b._mem, b._va, b._mc_va = 11, 22, 33, None  # failed while mapping the multicast address was never reached
    b._bound, b._mapped, b._mc_mapped = True, True, False
    b._unwind()
    assert log == ["cuDeviceGet", "cuMulticastUnbind", "cuMemUnmap", "cuMemAddressFree", "cuMemRelease", "cuMemRelease"]
    assert b._closed and b._mc is None and b.local is None
    b.close()  # a closed buffer's close() is a no-op
