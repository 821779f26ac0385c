"""B200-native dense-GEMM hot path of pytorch-distributed-matmul-benchmark.

Host-side mirror of the reference's operators over the C-ABI library
`libb2mm.so` (see include/b2mm.h).  Importing the package does not load the
library; the first operator call does, and fails loudly if it is not built.
"""
from . import _lib  # noqa: F401
from .ops import Communicator, HostGemm, bmm, matmul, release_workspaces  # noqa: F401

__all__ = ["matmul", "bmm", "HostGemm", "Communicator", "release_workspaces"]
