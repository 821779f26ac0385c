#!/bin/bash
# Round-2 GPU call F (1 GPU): smoke, the driver's bench command, ncu launch list of the same program, one full ncu capture of the top kernel.
set -u
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2f_bench_n1.json 2> gpurun_out/r2f_bench_n1.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r2f_bench_n1.json; echo
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv python bench.py --steps 2 --warmup 1 --quick --no-cpu-baseline > gpurun_out/r2f_ncu_bench.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -f -o gpurun_out/r2_gemm_bf16_16k python scripts/prof_gemm.py 16384 bfloat16 1 2 > gpurun_out/r2f_ncu_full.log 2>&1; echo "ncu full rc=$?"; ls -la gpurun_out/r2_gemm_bf16_16k.ncu-rep
