"""Shared test plumbing.

Markers
  gpu   needs a CUDA device (run on the B200 box: `pytest -m gpu`); everything else
        runs on a GPU-less host (`pytest -m "not gpu"`).

The oracle (oracle/) is test infrastructure: tests import it freely, the product
package never does (see test_no_oracle_in_product.py).
"""
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

DTYPES = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}
# max |C - C64| / (|C64| + rms(C64)); SURVEY.md 8(c).  One output ulp for the 16-bit types
# (2^-8, 2^-11 relative to |C64| + rms, plus fp32 accumulation noise), and 1e-5 for float32 at any K <= 16384
# (the survey's bar; the 3xTF32 kernel measures 4e-6 .. 5e-6 there, cuBLAS SGEMM 1.4e-5, torch CPU 2e-6).
TOL = {torch.bfloat16: 4e-3, torch.float16: 5e-4, torch.float32: 1e-5}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _from_np(a: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    if dtype in (torch.bfloat16, torch.float16):
        return torch.from_numpy(a.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(a.copy())


def load_golden(name: str, dtype_name: str) -> dict:
    """A golden npz as torch tensors of the recorded dtype (16-bit types are stored as bit patterns)."""
    dt = DTYPES[dtype_name]
    out = {}
    with np.load(os.path.join(GOLDEN, name)) as z:
        for k in z.files:
            a = z[k]
            out[k] = torch.from_numpy(a.copy()) if k.startswith("cols_") else _from_np(a, dt)
    return out


@pytest.fixture(scope="session")
def host_logic():
    with open(os.path.join(GOLDEN, "host_logic.json")) as f:
        return json.load(f)


def rel_err(C: torch.Tensor, C64: torch.Tensor) -> float:
    C64 = C64.double().cpu()
    rms = C64.pow(2).mean().sqrt()
    return ((C.detach().cpu().double() - C64).abs() / (C64.abs() + rms)).max().item()
