#!/bin/bash
# Drop-in for the reference's backup/run_distributed_benchmark.sh:
#   run_distributed_benchmark.sh [NUM_GPUS=2] [MODE=data_parallel] [DTYPE=bfloat16] [extra flags...]
source "$(dirname "$0")/run_common.sh"
NUM_GPUS=${1:-2}
MODE=${2:-data_parallel}
DTYPE=${3:-bfloat16}
shift $(( $# < 3 ? $# : 3 ))
printf 'Starting distributed matrix multiplication benchmark\n  GPUs: %s\n  Mode: %s\n  Data type: %s\n\n' \
    "$NUM_GPUS" "$MODE" "$DTYPE"
b2mm_launch "$NUM_GPUS" 29501 matmul_distributed_benchmark.py "${B2MM_DEFAULT_SWEEP[@]}" --mode "$MODE" --dtype "$DTYPE" "$@"
