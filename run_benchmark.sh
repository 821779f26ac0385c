#!/bin/bash
# Same positional interface as the reference launcher: NUM_GPUS [DTYPE].
# Anything after the two positionals is passed through (e.g. --check, --json out.json).

NUM_GPUS=${1:-1}
DTYPE=${2:-bfloat16}
shift $(( $# < 2 ? $# : 2 ))
cd "$(dirname "$0")"

echo "Starting distributed matrix multiplication benchmark with $NUM_GPUS GPU(s)"
echo "Data type: $DTYPE"
echo ""

ARGS="--sizes 4096 8192 16384 --iterations 50 --warmup 10 --dtype $DTYPE"
if [ "$NUM_GPUS" -eq 1 ]; then
    echo "Running in single GPU mode..."
    python3 matmul_benchmark.py $ARGS "$@"
else
    echo "Running in distributed mode with $NUM_GPUS GPUs..."
    python3 -m torch.distributed.run --nnodes=1 --nproc_per_node=$NUM_GPUS \
        --master_addr=127.0.0.1 --master_port=29500 \
        matmul_benchmark.py $ARGS "$@"
fi
