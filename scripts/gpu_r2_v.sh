#!/bin/bash
# A/B of the fine last band of the fused batch_parallel step (B2MM_FINE_TAIL) through the drop-in script, 8 GPUs.
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for ft in 0 1 0 1; do
  echo "B2MM_FINE_TAIL=$ft"
  B2MM_FINE_TAIL=$ft timeout 300 $TR --master-port 2963$ft matmul_scaling_benchmark.py --mode batch_parallel --per-gpu-batch 1 --fused --panels 8 --sizes 16384 --iterations 30 --warmup 5 2>&1 | grep -E "Compute time|Fused|ERROR" | cut -c1-120
done
