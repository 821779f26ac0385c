#!/bin/bash
# Round-2 GPU call G (4 GPUs): collectives / multicast tests after the set-up rewrite, bench at N=4.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_collectives_gpu.py "tests/test_scripts_gpu.py::test_scaling_script_two_gpus" -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2g_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2g_bench_n4.json 2> gpurun_out/r2g_bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2g_bench_n4.json") if l.startswith("{")][-1])
    print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["frac_of_copy_peak"], "modes_parity_ok", d.get("modes_parity_ok"))
    for m in d.get("modes", []):
        print(" ", m["mode"], "|", m["variant"][:40], "| ms", round(m["ms"], 3), "gemm", round(m["gemm_alone_ms"], 3), "eff", m.get("weak_scaling_eff"), "bus", m.get("bus_gbs"), "err", m["parity_err"], m["parity_ok"], m.get("store_path_alone_ms"), m.get("frac_of_bound"))
except Exception as e:
    print("no line:", e)
PY
tail -5 gpurun_out/r2g_bench_n4.err
