#!/usr/bin/env python3
"""Drop-in for the reference's backup/matmul_distributed_benchmark.py (same flags, modes and report); see the package module."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from pytorch_distributed_matmul_benchmark_b200.matmul_distributed_benchmark import main  # noqa: E402

if __name__ == "__main__":
    main()
