#!/bin/bash
# Drop-in for the reference launcher of the same name:
#   run_scaling_benchmark.sh [NUM_GPUS=2] [MODE=independent] [DTYPE=bfloat16] [extra flags...]
# Extra flags (--overlap, --fused, --panels 8, --per-gpu-batch 1, --check, --json f) go to the script unchanged.
source "$(dirname "$0")/run_common.sh"
NUM_GPUS=${1:-2}
MODE=${2:-independent}
DTYPE=${3:-bfloat16}
shift $(( $# < 3 ? $# : 3 ))
printf 'Matrix Multiplication Scaling Benchmark\n  GPUs: %s\n  Mode: %s (independent, batch_parallel, matrix_parallel)\n  Data type: %s\n\n' \
    "$NUM_GPUS" "$MODE" "$DTYPE"
b2mm_launch "$NUM_GPUS" 29503 matmul_scaling_benchmark.py "${B2MM_DEFAULT_SWEEP[@]}" --mode "$MODE" --dtype "$DTYPE" "$@"
