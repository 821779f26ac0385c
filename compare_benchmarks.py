#!/usr/bin/env python3
"""Drop-in for the reference's backup/compare_benchmarks.py: run the four launcher configurations it compares and
show their lines for one matrix size side by side.

    ./compare_benchmarks.py [--gpus 2] [--dtype bfloat16] [--size 16384] [--dry-run] [extra launcher flags ...]

The reference (backup/compare_benchmarks.py:30-63) shells out to run_benchmark.sh, run_distributed_benchmark.sh
(data_parallel) and run_overlap_benchmark.sh (no_overlap, overlap) at 2 GPUs and greps the 16384 block of each report
for "Results for" / "Average time" / "TFLOPS" / "overhead" (:22-26).  Same four runs and the same filter here, plus
one table at the end built from those lines.  `scripts/sweep.py` is the JSON-based sweep over more configurations.
"""
import argparse
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
KEEP = ("Results for", "Average time", "Total time", "TFLOPS", "overhead")

# (title, launcher, positional mode or None) — the reference's TEST 1..4
RUNS = [
    ("Original benchmark - Independent (no communication)", "run_benchmark.sh", None),
    ("Distributed - Data Parallel (with allreduce)", "run_distributed_benchmark.sh", "data_parallel"),
    ("Overlap Benchmark - No Overlap", "run_overlap_benchmark.sh", "no_overlap"),
    ("Overlap Benchmark - With Overlap", "run_overlap_benchmark.sh", "overlap"),
]


def command(launcher: str, gpus: int, mode, dtype: str, extra) -> list:
    """Launcher positionals as the reference passes them: NUM_GPUS [MODE] DTYPE, then pass-through flags."""
    return ["bash", os.path.join(ROOT, launcher), str(gpus)] + ([mode] if mode else []) + [dtype] + list(extra)


def block_for_size(report: str, size: int) -> list:
    """The lines of the `Results for SIZExSIZE` block that the reference's filter keeps."""
    lines = report.splitlines()
    tag = f"Results for {size}x{size}"
    for i, line in enumerate(lines):
        if tag in line:
            out = [line.strip()]
            for nxt in lines[i + 1:i + 15]:
                if nxt.startswith("Benchmarking ") or nxt.startswith("====") or nxt.startswith("Results for"):
                    break
                if any(k in nxt for k in KEEP):
                    out.append(nxt.strip())
            return out
    return []


def headline(block: list) -> tuple:
    """(milliseconds, TFLOPS) from a kept block: the first time line and the first TFLOPS figure."""
    ms = tf = None
    for line in block:
        m = re.search(r"time per \w+: ([0-9.]+) ms", line)
        if m and ms is None:
            ms = float(m.group(1))
        m = re.search(r"TFLOPS[^:]*: ([0-9.]+)", line)
        if m and tf is None:
            tf = float(m.group(1))
    return ms, tf


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Run the four benchmark configurations of the reference's comparison and tabulate them")
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--dtype", default="bfloat16", choices=["float32", "float16", "bfloat16"])
    ap.add_argument("--size", type=int, default=16384, help="matrix size whose block is shown (must be one the launchers sweep)")
    ap.add_argument("--dry-run", action="store_true", help="print the commands instead of running them")
    args, extra = ap.parse_known_args(argv)
    bar = "=" * 80
    print(f"\n{bar}\nCOMPREHENSIVE BENCHMARK COMPARISON\n{bar}")
    table = []
    for n, (title, launcher, mode) in enumerate(RUNS, 1):
        cmd = command(launcher, args.gpus, mode, args.dtype, extra)
        print(f"\n### TEST {n}: {title}")
        print(f"{'=' * 70}\nRunning: {' '.join(cmd[1:])}\n{'=' * 70}")
        if args.dry_run:
            continue
        p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
        block = block_for_size(p.stdout, args.size)
        if p.returncode != 0 or not block:
            print(f"  (no {args.size}x{args.size} results; exit code {p.returncode})")
            print("\n".join("  " + l for l in (p.stdout + p.stderr).splitlines()[-5:]))
            table.append((title, None, None))
            continue
        print("\n".join(block))
        table.append((title,) + headline(block))
    if not args.dry_run:
        print(f"\n{bar}\nSUMMARY ({args.size}x{args.size}, {args.gpus} GPU(s), {args.dtype})\n{bar}")
        for title, ms, tf in table:
            print(f"  {title:55s} {'-' if ms is None else f'{ms:9.3f} ms'}  {'-' if tf is None else f'{tf:9.2f} TFLOPS'}")
        print("\n  1 = no communication (the ceiling); 2 = GEMM then all_reduce, timed separately; 3 = the same serialised in one\n"
              "  loop; 4 = the all_reduce of step i overlapped with the GEMM of step i+1.  4 should beat 3 and approach 1.")
    return 0


if __name__ == "__main__":
    sys.exit(main())
