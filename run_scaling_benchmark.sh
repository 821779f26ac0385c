#!/bin/bash
# Same positional interface as the reference launcher: NUM_GPUS [MODE] [DTYPE].
# Anything after the three positionals is passed through (e.g. --overlap --panels 8 --check).

NUM_GPUS=${1:-2}
MODE=${2:-independent}
DTYPE=${3:-bfloat16}
shift $(( $# < 3 ? $# : 3 ))
cd "$(dirname "$0")"

echo "Matrix Multiplication Scaling Benchmark"
echo "  GPUs: $NUM_GPUS"
echo "  Mode: $MODE (independent, batch_parallel, matrix_parallel)"
echo "  Data type: $DTYPE"
echo ""

ARGS="--sizes 4096 8192 16384 --iterations 50 --warmup 10 --mode $MODE --dtype $DTYPE"
if [ "$NUM_GPUS" -eq 1 ]; then
    echo "Running in single GPU mode..."
    python3 matmul_scaling_benchmark.py $ARGS "$@"
else
    echo "Running in distributed mode with $NUM_GPUS GPUs..."
    python3 -m torch.distributed.run --nnodes=1 --nproc_per_node=$NUM_GPUS \
        --master_addr=127.0.0.1 --master_port=29503 \
        matmul_scaling_benchmark.py $ARGS "$@"
fi
