#!/usr/bin/env python3
"""Sweep driver: run the drop-in launchers over modes x GPU counts x dtypes and print ONE table.

The reference's backup/compare_benchmarks.py (:30-63) runs four launcher configurations through subprocess and
greps the 16k lines out of stdout; this does the same job from the scripts' --json output.

    python scripts/sweep.py --gpus 1 2 8 --dtypes bfloat16 --sizes 16384 --out gpurun_out/sweep.md
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(script, gpus, extra, sizes, iters, warmup, port):
    with tempfile.NamedTemporaryFile(suffix=".json", delete=False) as f:
        path = f.name
    base = [script, "--sizes", *map(str, sizes), "--iterations", str(iters), "--warmup", str(warmup), "--json", path] + extra
    cmd = ([sys.executable] + base) if gpus == 1 else (
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr=127.0.0.1",
         f"--master-port={port}"] + base)
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
    try:
        return json.load(open(path))["results"] if p.returncode == 0 else [{"error": p.stderr[-300:]}]
    except (OSError, ValueError):
        return [{"error": (p.stdout + p.stderr)[-300:]}]
    finally:
        if os.path.exists(path):
            os.unlink(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, nargs="+", default=[1])
    ap.add_argument("--dtypes", nargs="+", default=["bfloat16"])
    ap.add_argument("--sizes", type=int, nargs="+", default=[4096, 8192, 16384])
    ap.add_argument("--iterations", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rows = ["| GPUs | mode | dtype | size | ms | TFLOP/s per GPU | aggregate TFLOP/s |", "|---|---|---|---|---|---|---|"]
    port = 29600
    for g in a.gpus:
        for dt in a.dtypes:
            configs = [("matmul_benchmark.py", "basic", [])] if g == 1 else []
            configs += [("matmul_scaling_benchmark.py", "independent", ["--mode", "independent"])]
            if g > 1:
                configs += [("matmul_scaling_benchmark.py", "batch_parallel (weak, fused)",
                             ["--mode", "batch_parallel", "--per-gpu-batch", "1", "--fused", "--panels", "8"]),
                            ("matmul_scaling_benchmark.py", "matrix_parallel (fused)", ["--mode", "matrix_parallel", "--fused"])]
            for script, label, extra in configs:
                port += 1
                for r in run(script, g, extra + ["--dtype", dt], a.sizes, a.iterations, a.warmup, port):
                    if "error" in r:
                        rows.append(f"| {g} | {label} | {dt} | - | ERROR | {r['error'][-80:]!r} | |")
                        continue
                    agg = r.get("total_tflops", r.get("actual_tflops", r.get("aggregate_tflops")))
                    rows.append(f"| {g} | {label} | {dt} | {r['size']} | {r['avg_time_ms']:.3f} | {r['tflops_per_gpu']:.1f} | {agg:.1f} |")
    text = "\n".join(rows)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        open(a.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
