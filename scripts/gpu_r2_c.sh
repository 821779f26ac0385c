#!/bin/bash
# Round-2 GPU call C (1 GPU): half-tile tail parity + A/B, fewer-SMs probe.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2c_pytest.log
timeout 300 python scripts/ab_env.py B2MM_SPLIT_TAIL 0 1 16384 60 6 2>&1 | tail -3
timeout 300 python scripts/ab_env.py B2MM_SPLIT_TAIL 0 1 8192 200 6 2>&1 | tail -3
timeout 300 python scripts/fewer_sms_probe.py > gpurun_out/r2c_fewer_sms.log 2>&1; echo "fewer sms rc=$?"; cat gpurun_out/r2c_fewer_sms.log | cut -c1-300
