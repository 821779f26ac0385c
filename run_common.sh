#!/bin/bash
# Shared by run_benchmark.sh / run_scaling_benchmark.sh / run_overlap_benchmark.sh.
#   b2mm_launch <gpus> <master-port> <script.py> <script args...>
# One process when gpus == 1, otherwise one process per GPU under torch.distributed.run on this node
# (rendezvous on 127.0.0.1: container hostnames do not always resolve).
B2MM_DEFAULT_SWEEP=(--sizes 4096 8192 16384 --iterations 50 --warmup 10)

b2mm_launch() {
    local gpus=$1 port=$2 script=$3
    shift 3
    cd "$(dirname "${BASH_SOURCE[0]}")" || exit 1
    if [ "$gpus" -eq 1 ]; then
        echo "Running in single GPU mode..."
        exec python3 "$script" "$@"
    fi
    echo "Running in distributed mode with $gpus GPUs..."
    exec python3 -m torch.distributed.run --nnodes=1 --nproc_per_node="$gpus" \
        --master_addr=127.0.0.1 --master_port="$port" "$script" "$@"
}
