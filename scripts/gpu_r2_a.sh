#!/bin/bash
# Round-2 GPU call A (2 GPUs): the whole -m gpu suite (incl. the >= 2-GPU tests), bench at N=1 and N=2, reference arm under torchrun.
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_gpus.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2a_bench_n1.json 2> gpurun_out/r2a_bench_n1.err; echo "bench n1 rc=$?"
tail -c 3000 gpurun_out/r2a_bench_n1.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2a_bench_n2.json 2> gpurun_out/r2a_bench_n2.err; echo "bench n2 rc=$?"
tail -c 6000 gpurun_out/r2a_bench_n2.json; tail -20 gpurun_out/r2a_bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 --ref-budget-s 30 > gpurun_out/r2a_ref_n2.json 2> gpurun_out/r2a_ref_n2.err; echo "ref n2 rc=$?"
cat gpurun_out/r2a_ref_n2.json
