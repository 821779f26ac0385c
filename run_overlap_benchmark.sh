#!/bin/bash
# Same positional interface as the reference's backup/run_overlap_benchmark.sh: NUM_GPUS [MODE] [DTYPE].

NUM_GPUS=${1:-2}
MODE=${2:-overlap}
DTYPE=${3:-bfloat16}
shift $(( $# < 3 ? $# : 3 ))
cd "$(dirname "$0")"

echo "Starting overlapped communication/computation benchmark"
echo "  GPUs: $NUM_GPUS"
echo "  Mode: $MODE (no_overlap, overlap, pipeline)"
echo "  Data type: $DTYPE"
echo ""

ARGS="--sizes 4096 8192 16384 --iterations 50 --warmup 10 --mode $MODE --dtype $DTYPE"
if [ "$NUM_GPUS" -eq 1 ]; then
    echo "Running in single GPU mode..."
    python3 matmul_overlap_benchmark.py $ARGS "$@"
else
    echo "Running in distributed mode with $NUM_GPUS GPUs..."
    python3 -m torch.distributed.run --nnodes=1 --nproc_per_node=$NUM_GPUS \
        --master_addr=127.0.0.1 --master_port=29502 \
        matmul_overlap_benchmark.py $ARGS "$@"
fi
