#!/bin/bash
# Round-2 GPU call E (8 GPUs): bench at N=8 with the sharded modes, the gather probe, and the drop-in scripts with --check.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2e_bench_n8.json 2> gpurun_out/r2e_bench_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2e_bench_n8.json").read().strip().splitlines()[-1])
    print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["frac_of_copy_peak"], "modes_parity_ok", d.get("modes_parity_ok"))
    for m in d.get("modes", []):
        print(" ", m["mode"], "|", m["variant"][:40], "| ms", round(m["ms"], 3), "gemm", round(m["gemm_alone_ms"], 3), "eff", m.get("weak_scaling_eff"), "bus", m.get("bus_gbs"), "err", m["parity_err"], m["parity_ok"])
except Exception as e:
    print("no line:", e)
PY
tail -5 gpurun_out/r2e_bench_n8.err
timeout 600 $TR --master-port 29612 scripts/gather_probe.py 16384 > gpurun_out/r2e_gather_probe.log 2>&1; echo "gather probe rc=$?"; grep "^{" gpurun_out/r2e_gather_probe.log | cut -c1-1500
timeout 600 $TR --master-port 29613 matmul_scaling_benchmark.py --mode matrix_parallel --fused --check --sizes 4096 8192 16384 --iterations 20 --warmup 5 > gpurun_out/r2e_sc8_matrix_fused.txt 2>&1; echo "matrix script rc=$?"; grep -E "Compute time|Fused|Parity|ERROR|Actual" gpurun_out/r2e_sc8_matrix_fused.txt
timeout 600 $TR --master-port 29614 matmul_scaling_benchmark.py --mode batch_parallel --per-gpu-batch 1 --fused --panels 8 --check --sizes 16384 --iterations 20 --warmup 5 > gpurun_out/r2e_sc8_batch_fused.txt 2>&1; echo "batch script rc=$?"; grep -E "Compute time|Fused|Parity|ERROR|Actual" gpurun_out/r2e_sc8_batch_fused.txt
