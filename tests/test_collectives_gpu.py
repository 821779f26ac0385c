"""The library's collectives and fused compute+collective paths on two GPUs (skipped on a single-GPU box)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_collectives_and_fused_paths_two_gpus(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = tmp_path / "res.json"
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr=127.0.0.1", f"--master-port={port}", os.path.join(ROOT, "tests", "_nccl_worker.py"),
                        str(out)], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    for path in (out, str(out) + ".rank1"):
        res = json.load(open(path))
        bad = [k for k, v in res.items() if v is False]
        assert not bad, (path, bad)
        assert res["launches"] > 0
