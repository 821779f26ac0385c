"""CPU oracle for the GEMM hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
legs may import this package, and only as the checker or the reported CPU
baseline.  The product package never does (tests/test_no_oracle_in_product.py
enforces it).
"""
