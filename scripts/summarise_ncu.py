"""Turn an .ncu-rep (gpurun_out/, scratch) into the small, committed summaries under profiles/.

    python scripts/summarise_ncu.py gpurun_out/r1_gemm_bf16_16k.ncu-rep profiles/r1_gemm_bf16_16k [--traffic-key KEY]
"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "sm__warps_active.avg.per_cycle_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_uniform.sum", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    key = sys.argv[sys.argv.index("--traffic-key") + 1] if "--traffic-key" in sys.argv else None
    raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    lines, traffic = [], []
    for r in data:
        lines.append(f"kernel: {r[name_i][:160]}")
        vals = {}
        for i, h in enumerate(hdr):
            if h in KEEP:
                lines.append(f"  {h:78s} {r[i]:>18s} {units[i]}")
                vals[h] = (r[i], units[i])
        def to_bytes(v, u):
            x = float(v.replace(",", ""))
            return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]
        if "dram__bytes_read.sum" in vals:
            traffic.append(to_bytes(*vals["dram__bytes_read.sum"]) + to_bytes(*vals["dram__bytes_write.sum"]))
        lines.append("")
    with open(out + "_summary.txt", "w") as f:
        f.write(f"source: {rep} (ncu --set full --clock-control none --import-source on)\n\n" + "\n".join(lines))
    if key and traffic:
        path = "profiles/roofline_traffic.json"
        try:
            d = json.load(open(path))
        except OSError:
            d = {}
        d[key] = sum(traffic) / len(traffic)
        d[key + "_source"] = out + "_summary.txt"
        json.dump(d, open(path, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
