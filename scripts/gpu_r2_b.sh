#!/bin/bash
# Round-2 GPU call B (2 GPUs): full -m gpu suite, bench N=2 with the sharded modes, cluster-size occupancy, host topology,
# power/clock table of the tile variants and their ncu byte counts.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2b_pytest.log
tail -4 gpurun_out/r2b_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2b_bench_n2.json 2> gpurun_out/r2b_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/r2b_bench_n2.json").read().strip().splitlines()[-1])
    print("value", d["value"], "modes_parity_ok", d.get("modes_parity_ok"))
    for m in d.get("modes", []):
        print(" ", m["mode"], "|", m["variant"][:40], "| ms", round(m["ms"], 3), "eff", m.get("weak_scaling_eff"), "err", m["parity_err"], m["parity_ok"])
except Exception as e:
    print("no line:", e)
PY
tail -5 gpurun_out/r2b_bench_n2.err
timeout 120 python scripts/cluster_occupancy.py 2>&1 | tail -8
{ lscpu | head -30; echo; ls /sys/devices/system/node/; for d in /sys/bus/pci/devices/*; do v=$(cat $d/vendor 2>/dev/null); c=$(cat $d/class 2>/dev/null); if [ "$v" = "0x10de" ]; then echo "$d class=$c numa=$(cat $d/numa_node) "; fi; done; nvidia-smi topo -m; free -g; cat /sys/fs/cgroup/cpuset.cpus.effective 2>/dev/null; cat /sys/fs/cgroup/cpuset.mems.effective 2>/dev/null; } > gpurun_out/r2b_host_topology.txt 2>&1
timeout 300 python scripts/power_table.py 2.5 > gpurun_out/r2b_power_table.log 2>&1; echo "power table rc=$?"; cat gpurun_out/r2b_power_table.log | cut -c1-400
M="l1tex__m_xbar2l1tex_read_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum,sm__cycles_elapsed.avg.per_second,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
i=0
for v in "" "B2MM_M_BLOCKS=1 B2MM_BLOCK_N=256" "B2MM_M_BLOCKS=1 B2MM_BLOCK_N=128" "B2MM_CTA_GROUP=1"; do
  i=$((i+1))
  env $v timeout 300 ncu --metrics $M --clock-control none -k regex:gemm_kernel -s 2 -c 1 --csv --log-file gpurun_out/r2b_ncu_variant$i.csv python scripts/prof_gemm.py 16384 bfloat16 1 2 > gpurun_out/r2b_ncu_variant$i.log 2>&1
  echo "variant $i [$v] ncu rc=$?"
done
