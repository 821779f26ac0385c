"""bench.py end to end on the GPU: the JSON contract at N = 1, and — where the box has two GPUs — the sharded modes,
which bench.py parity-checks itself (it exits non-zero if one of its checks fails)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _last_json(text: str) -> dict:
    lines = [l for l in text.splitlines() if l.startswith("{")]
    assert len(lines) == 1, f"bench.py must print ONE JSON line, got {len(lines)}"
    return json.loads(lines[0])


def test_single_gpu_line_keeps_the_contract():
    p = subprocess.run([sys.executable, "bench.py", "--steps", "3", "--warmup", "3", "--quick", "--no-cpu-baseline"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    d = _last_json(p.stdout)
    assert d["metric"] == "bf16_16k_gemm_tflops" and d["unit"] == "TFLOP/s" and d["n_gpus"] == 1 and d["steps"] == 3
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["dtype"] == "bf16" and d["vs_baseline"] is None
    assert d["config"] == {"workload": d["config"]["workload"], "sizes": [16384], "mode": "independent"}
    assert d["gpu_launches"] == 3, "one kernel of this library per timed step"
    r = d["roofline"]
    assert r["bound"] == "tensor" and r["frac"] == pytest.approx(r["achieved"] / r["peak"])
    assert "512x256 tile per CTA pair" in r["kernel"] and len(r["per_launch_ms"]) == 3
    assert r["achieved"] == pytest.approx(2 * 16384 ** 3 / (r["kernel_ms"] * 1e-3) / 1e12)
    assert 800 < d["value"] < 2250, "a B200 runs this GEMM between 0.8 and 2.25 PFLOP/s"
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 2 * 16384 * 16384 * 2 and e["d2h_bytes_per_step"] == 16384 * 16384 * 2
    assert e["value"] < d["value"] and 0.5 < e["frac_of_copy_peak"] < 1.3
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_two_gpu_line_carries_parity_checked_modes():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr=127.0.0.1",
                        f"--master-port={port}", "bench.py", "--gpus", "2", "--steps", "3", "--warmup", "3"],
                       cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    d = _last_json(p.stdout)
    assert d["n_gpus"] == 2 and d["modes_parity_ok"] is True and "modes_error" not in d
    modes = {(m["mode"], m["variant"].split(" ")[0]) for m in d["modes"]}
    assert {("batch_parallel", "serial"), ("batch_parallel", "fused"), ("matrix_parallel", "serial"), ("matrix_parallel", "fused"),
            ("overlap (iteration level)", "no_overlap:"), ("overlap (iteration level)", "overlap:")} <= modes
    for m in d["modes"]:
        assert m["parity_ok"] is True and m["parity_err"] <= m["parity_tol"] and m["ms"] > 0 and m["gemm_alone_ms"] > 0
    fused = [m for m in d["modes"] if m["mode"] == "batch_parallel" and m["variant"].startswith("fused")][0]
    assert fused["bit_identical_on_all_ranks"] is True and 0.5 < fused["weak_scaling_eff"] < 1.2
