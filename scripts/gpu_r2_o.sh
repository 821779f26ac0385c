#!/bin/bash
# signal-pad barrier closing the fused gather: correctness tests, then A/B against the NCCL barrier through the drop-in script.
set -u
mkdir -p gpurun_out
NG=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
if [ "$NG" = "2" ]; then
timeout 900 python -m pytest tests/test_collectives_gpu.py "tests/test_scripts_gpu.py::test_scaling_script_two_gpus" tests/test_bench_gpu.py -m gpu -x -q 2>&1 | tail -3
fi
for sb in 0 1 0 1; do
  echo "B2MM_SIGNAL_BARRIER=$sb"
  B2MM_SIGNAL_BARRIER=$sb timeout 600 $TR --master-port 2961$sb matmul_scaling_benchmark.py --mode matrix_parallel --fused --check --sizes 4096 8192 16384 --iterations 30 --warmup 5 2>&1 | grep -E "Fused|Parity|ERROR" | cut -c1-110
done
