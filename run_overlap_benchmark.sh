#!/bin/bash
# Drop-in for the reference's backup/run_overlap_benchmark.sh:
#   run_overlap_benchmark.sh [NUM_GPUS=2] [MODE=overlap] [DTYPE=bfloat16] [extra flags...]
source "$(dirname "$0")/run_common.sh"
NUM_GPUS=${1:-2}
MODE=${2:-overlap}
DTYPE=${3:-bfloat16}
shift $(( $# < 3 ? $# : 3 ))
printf 'Starting overlapped communication/computation benchmark\n  GPUs: %s\n  Mode: %s (no_overlap, overlap, pipeline)\n  Data type: %s\n\n' \
    "$NUM_GPUS" "$MODE" "$DTYPE"
b2mm_launch "$NUM_GPUS" 29502 matmul_overlap_benchmark.py "${B2MM_DEFAULT_SWEEP[@]}" --mode "$MODE" --dtype "$DTYPE" "$@"
