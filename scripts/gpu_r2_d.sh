#!/bin/bash
# Round-2 GPU call D (2 GPUs): bench N=2 after the timing-protocol change; narrow-slab A/B vs cuBLAS on GPU 0.
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2d_bench_n2.json 2> gpurun_out/r2d_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2d_bench_n2.json").read().strip().splitlines()[-1])
    print("value", d["value"], "ms", d["ms_per_step"], "modes_parity_ok", d.get("modes_parity_ok"))
    for m in d.get("modes", []):
        print(" ", m["mode"], "|", m["variant"][:40], "| ms", round(m["ms"], 3), "gemm", round(m["gemm_alone_ms"], 3), "eff", m.get("weak_scaling_eff"), "err", m["parity_err"], m["parity_ok"])
except Exception as e:
    print("no line:", e)
PY
tail -5 gpurun_out/r2d_bench_n2.err
timeout 600 python scripts/narrow_slab_ab.py > gpurun_out/r2d_narrow.log 2>&1; echo "narrow rc=$?"; cut -c1-420 gpurun_out/r2d_narrow.log
