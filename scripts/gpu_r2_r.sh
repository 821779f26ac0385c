#!/bin/bash
# first wave in half-width items: parity on 1 GPU, the multi-GPU tests, A/B of the fused gather through the drop-in script.
set -u
NG=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
if [ "$NG" = "2" ]; then
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_collectives_gpu.py "tests/test_scripts_gpu.py::test_scaling_script_two_gpus" -m gpu -x -q 2>&1 | tail -3
fi
for sh in 0 1 0 1; do
  echo "B2MM_SPLIT_HEAD=$sh"
  B2MM_SPLIT_HEAD=$sh timeout 600 $TR --master-port 2962$sh matmul_scaling_benchmark.py --mode matrix_parallel --fused --check --sizes 8192 16384 --iterations 30 --warmup 5 2>&1 | grep -E "Fused|Parity|ERROR" | cut -c1-60
done
