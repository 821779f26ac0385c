#!/bin/bash
# Round-2 GPU call H (4 GPUs): bench at N=4 and the scripts' --check at a world size where NCCL rounds partial sums.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2h_bench_n4.json 2> gpurun_out/r2h_bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
try:
    d = json.loads([l for l in open("gpurun_out/r2h_bench_n4.json") if l.startswith("{")][-1])
    print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["frac_of_copy_peak"], "modes_parity_ok", d.get("modes_parity_ok"))
    for m in d.get("modes", []):
        print(" ", m["mode"], "|", m["variant"][:40], "| ms", round(m["ms"], 3), "gemm", round(m["gemm_alone_ms"], 3), "eff", m.get("weak_scaling_eff"), "err", m["parity_err"], m.get("err_vs_nccl_path_full_tensor"), m["parity_ok"])
except Exception as e:
    print("no line:", e)
PY
tail -3 gpurun_out/r2h_bench_n4.err
B2MM_NVLS=0 timeout 600 $TR --master-port 29616 bench.py --gpus 4 --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][-1]); print(\"NVLS=0 bench:\", d.get(\"modes_parity_ok\"), [(m[\"variant\"][:14], round(m[\"ms\"],3), m[\"parity_ok\"]) for m in d[\"modes\"]])"
timeout 600 $TR --master-port 29612 matmul_scaling_benchmark.py --mode batch_parallel --fused --panels 2 --check --sizes 2048 4096 --iterations 5 --warmup 2 2>&1 | grep -E "Parity|ERROR|Fused" 
timeout 600 $TR --master-port 29613 matmul_overlap_benchmark.py --mode no_overlap --check --sizes 2048 4096 --iterations 5 --warmup 2 2>&1 | grep -E "Parity|ERROR"
timeout 600 $TR --master-port 29614 matmul_overlap_benchmark.py --mode overlap --check --sizes 2048 --iterations 5 --warmup 2 2>&1 | grep -E "Parity|ERROR"
timeout 600 $TR --master-port 29615 matmul_distributed_benchmark.py --mode data_parallel --check --sizes 2048 --dtype float32 --iterations 5 --warmup 2 2>&1 | grep -E "Parity|ERROR"
