"""bench.py helpers that can be checked without a GPU: the clock / throttle-reason parser and the peak table."""
import json
import os
import sys

from conftest import ROOT

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _sampler(lines):
    s = bench.ClockSampler.__new__(bench.ClockSampler)  # no nvidia-smi here: feed the parser directly
    s.lines, s.proc, s.thread = lines, None, None
    return s


def test_clock_window_takes_the_median_and_collects_reasons():
    rows = [(10.00, "1965, 1965, 310.5, Not Active, Not Active, Not Active, Not Active"),
            (10.05, "1410, 1965, 990.1, Not Active, Not Active, Not Active, Active"),
            (10.10, "1395, 1965, 998.7, Not Active, Not Active, Not Active, Active"),
            (10.15, "1380, 1965, 1001.2, Not Active, Not Active, Not Active, Active"),
            (11.00, "1965, 1965, 200.0, Not Active, Not Active, Not Active, Not Active")]
    w = _sampler(rows).window(10.04, 10.16)
    assert w["samples"] == 3 and w["sm_mhz"] == 1395.0 and w["sm_max_mhz"] == 1965.0
    assert w["reasons"] == ["sw_power_cap"] and w["power_w_max"] == 1001.2
    assert not set(w["reasons"]) & set(bench.BAD_REASONS)  # sw_power_cap is kept and noted, not a rejection


def test_clock_window_flags_thermal_slowdown_and_survives_garbage():
    rows = [(1.0, "1200, 1965, 700.0, Active, Active, Not Active, Active"), (1.05, "[N/A], 1965, [N/A], x"), (1.1, "")]
    w = _sampler(rows).window(0.9, 1.2)
    assert w["samples"] == 1 and set(w["reasons"]) >= {"hw_slowdown", "hw_thermal_slowdown"}
    assert set(w["reasons"]) & set(bench.BAD_REASONS)
    assert _sampler([]).window(0, 1) == {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}


def test_peaks_come_from_the_driver_file_or_the_stated_fallback(tmp_path, monkeypatch):
    p = bench.measured_peaks()
    assert p["source"] in ("measured", "fallback") and p["bf16_tflops"] > 1000 and p["hbm_gbs"] > 5000
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))  # no MEASURED_PEAKS.json there
    assert bench.measured_peaks()["source"] == "fallback"
    (tmp_path / "MEASURED_PEAKS.json").write_text(json.dumps({"hbm_gbs": 6500.0, "bf16_tflops": 1650.0}))
    assert bench.measured_peaks() == {"hbm_gbs": 6500.0, "bf16_tflops": 1650.0, "source": "measured"}


def test_both_arms_share_one_config_dict():
    """The driver compares the two arms' `config`: both take it from one function, extras live under `notes`."""
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": shared_config()') == 2 and src.count('"workload": WORKLOAD') == 1
    cfg = bench.shared_config()
    assert cfg == {"workload": bench.WORKLOAD, "sizes": [16384], "mode": "independent"}
    assert "16384" in bench.WORKLOAD and bench.METRIC == "bf16_16k_gemm_tflops"


def test_reference_arm_uses_every_core_it_may_and_reports_the_timed_step():
    """torch.distributed.run exports OMP_NUM_THREADS=1: the arm must undo that, and ms_per_step is what was timed."""
    import torch
    before = torch.get_num_threads()
    try:
        torch.set_num_threads(1)
        bench.N, keep = 512, bench.N  # a small stand-in problem: the logic is size independent
        try:
            r = bench.cpu_reference_arm(steps=2, warmup=1, budget_s=1.0)
        finally:
            bench.N = keep
        assert r["cores"] == bench.host_threads() == torch.get_num_threads()
        assert r["kind"] == "port" and 0 < r["sample_fraction_of_step"] <= 1.0
        assert abs(r["full_step_ms_extrapolated"] * r["sample_fraction_of_step"] - r["ms_per_step"]) < 1e-9
    finally:
        torch.set_num_threads(before)


def test_nvml_sampler_window_summarises_samples():
    s = bench.NvmlSampler.__new__(bench.NvmlSampler)
    s.samples = [(1.00, 1965, 300.0), (1.01, 1500, 900.0), (1.02, 1400, 1000.0), (2.0, 1965, 250.0)]
    w = s.window(1.005, 1.5)
    assert w == {"samples": 2, "sm_mhz_median": 1450.0, "sm_mhz_min": 1400, "power_w_max": 1000.0, "power_w_mean": 950.0}
    assert s.window(5, 6) == {"samples": 0}
