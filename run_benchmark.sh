#!/bin/bash
# Drop-in for the reference launcher of the same name: run_benchmark.sh [NUM_GPUS=1] [DTYPE=bfloat16] [extra flags...]
# (extra flags, e.g. --check or --json out.json, go to matmul_benchmark.py unchanged).
source "$(dirname "$0")/run_common.sh"
NUM_GPUS=${1:-1}
DTYPE=${2:-bfloat16}
shift $(( $# < 2 ? $# : 2 ))
printf 'Starting distributed matrix multiplication benchmark with %s GPU(s)\nData type: %s\n\n' "$NUM_GPUS" "$DTYPE"
b2mm_launch "$NUM_GPUS" 29500 matmul_benchmark.py "${B2MM_DEFAULT_SWEEP[@]}" --dtype "$DTYPE" "$@"
