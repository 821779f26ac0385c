"""How many SMs can a cluster size reach?  cudaOccupancyMaxActiveClusters for the main GEMM kernel (1 CTA per SM).

Decides whether 4-CTA clusters (TMA multicast of A across two CTA pairs) are worth building: a cluster must sit
inside one GPC, and the B200's GPCs do not all hold a multiple of 4 SMs.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_distributed_matmul_benchmark_b200 import _lib

torch.cuda.init()
torch.zeros(1, device="cuda")
lib = _lib.load()
sms = torch.cuda.get_device_properties(0).multi_processor_count
rows = []
for cs in (1, 2, 4, 8, 16):
    n = lib.b2mm_debug_max_active_clusters(cs)
    rows.append({"cluster_size": cs, "max_active_clusters": n, "sms_reached": n * cs if n > 0 else None,
                 "sms_stranded": sms - n * cs if n > 0 else None})
    print(rows[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"device": torch.cuda.get_device_name(0), "sms": sms, "rows": rows}, open("gpurun_out/r2_cluster_occupancy.json", "w"), indent=1)
