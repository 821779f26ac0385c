"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (the share of the step each kernel takes).

    python scripts/summarise_launches.py profiles/r1_bench_launches.csv > profiles/r1_bench_launches_summary.txt
"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = defaultdict(lambda: [0, 0.0, 1e30, 0.0])
for r in rows[1:]:
    ns = float(r[vi].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ui], 1)
    a = agg[r[ki]]
    a[0] += 1
    a[1] += ns
    a[2] = min(a[2], ns)
    a[3] = max(a[3], ns)
total = sum(a[1] for a in agg.values())
print(f"source: {sys.argv[1]} ({len(rows) - 1} launches, {total / 1e6:.3f} ms of kernel time; per-launch times under ncu are")
print("cold-cache and serialised, so it is the SHARE of each kernel that is compared with the live run, not the absolute)\n")
print(f"{'launches':>8} {'total ms':>10} {'share':>7} {'min ms':>9} {'max ms':>9}  kernel")
for name, (n, t, lo, hi) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    short = name if len(name) <= 150 else name[:147] + "..."
    print(f"{n:8d} {t / 1e6:10.3f} {100 * t / total:6.1f}% {lo / 1e6:9.3f} {hi / 1e6:9.3f}  {short}")
