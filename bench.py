#!/usr/bin/env python3
"""bench.py — the headline measurement: bf16 16384^3 GEMM TFLOP/s on N B200s (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N = 1 is BASELINE.json's configs[1] (the matmul_benchmark.py shape on one B200); for N > 1 `value` is
configs[2], `independent` mode: one full GEMM per GPU, no data-path collective (weak scaling), so the 1 -> 8
curve compares like with like.  A "step" is one C = A @ B over one 16384 x 16384 pair of synthetic randn
matrices.  For N > 1 the same run ALSO parity-checks and times the two modes that really shard
(configs[3] batch_parallel + all_reduce, configs[4] matrix_parallel + all_gather, and the iteration-level
overlap of backup/matmul_overlap_benchmark.py) and reports them under `modes`; a parity failure there makes
the process exit non-zero after the line has been printed.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      tensor-core roofline of the GEMM kernel: achieved = 2*N^3 / mean launch duration (CUDA events
                on the launching stream over the timed region); peak = the measured cuBLAS bf16 BURST figure of
                MEASURED_PEAKS.json (the timed region is a fraction of a second); `sustained` = a second,
                self-contained >= 2 s back-to-back loop against the measured SUSTAINED figure, with its own
                clock record; `cublas_same_run` = torch.matmul on the same tensors, interleaved, off the
                measured path (the Blackwell kernel to beat); traffic = DRAM bytes per launch from the committed
                ncu capture (profiles/roofline_traffic.json) or null.
  cpu_baseline  the reference's CPU path (torch.matmul on CPU, timed as matmul_benchmark.py:69-74 does;
                restated in oracle/reference_port.py) on the box's host cores, bounded sample (N = 1 only).
  e2e           the same metric through the public host-buffer call (ops.HostGemm): every step copies A and B
                from pinned host memory, runs the GEMM and copies C back — all inside the timed region; with
                the host<->device rates it reached and the plain pinned-copy peak measured in the same run at
                the same N (all ranks at once), so that the bound (PCIe / host fabric) is visible.
  modes         N > 1: one record per sharded mode and variant (see sharded_modes()).
  notes         free-form context (parallelism, L2 policy, per-GPU figures); `config` is the same in both arms.

`--impl reference` times the reference's CPU implementation (the port above; the reference is pure Python over
torch and cannot travel to the GPU box) with all the host threads the process may use; each step is a bounded
row-slab sample of the same 16384^3 product.  Under torchrun rank 0 alone runs it.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N = 16384
METRIC = "bf16_16k_gemm_tflops"
UNIT = "TFLOP/s"
FLOP = 2.0 * N ** 3
SPEC_PEAK = 2250.0  # B200 dense bf16, TFLOP/s
NVLINK_GBS = 900.0  # NVLink 5, per direction per GPU
BAD_REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")
# One workload for both arms (the driver divides one by the other): the 16384^3 bf16 product, per GPU.
WORKLOAD = f"{N}x{N}x{N} bfloat16 GEMM per GPU (BASELINE.json configs[1]; independent mode for N>1)"
TOL_BF16 = 4e-3  # max |C - C64| / (|C64| + rms(C64)), SURVEY.md 8(c): one bf16 output ulp


def shared_config() -> dict:
    """`config` of BOTH arms — identical by construction, so the driver's same_config holds."""
    return {"workload": WORKLOAD, "sizes": [N], "mode": "independent"}


def measured_peaks() -> dict:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return dict(json.load(f), source="measured")
    except OSError:
        # fallback stated in the B200 profiling recipe
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the warm-up + timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.lines, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                          "50", "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def window(self, t0: float, t1: float) -> dict:
        rows = [l.split(", ") for t, l in self.lines if t0 <= t <= t1 + 0.15 and l]
        rows = [r for r in rows if len(r) == 7] or [l.split(", ") for _, l in self.lines[-3:] if len(l.split(", ")) == 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[3:]) if v.strip().lower().startswith("active")})

        def num(x):
            try:
                return float(x)
            except ValueError:
                return None
        sm = [v for v in (num(r[0]) for r in rows) if v is not None]
        pw = [v for v in (num(r[2]) for r in rows) if v is not None]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": num(rows[0][1]),
                "power_w_max": max(pw) if pw else None, "reasons": reasons, "samples": len(rows)}

    def stop(self):
        if self.proc:
            self.proc.terminate()


class NvmlSampler:
    """SM clock and board power straight from NVML every ~2 ms: nvidia-smi cannot resolve a 0.1 s timed region
    (2-4 samples), this can (tens of samples).  Optional: without the bindings the record just says so."""

    def __init__(self, index: int):
        self.samples, self.ok, self._stop = [], False, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
            self.ok = True
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()
        except Exception:
            self.ok = False

    def _run(self):
        nv, h = self.nv, self.h
        while not self._stop.is_set():
            try:
                self.samples.append((time.time(), nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetPowerUsage(h) / 1000.0))
            except Exception:
                pass
            time.sleep(0.002)

    def window(self, t0: float, t1: float) -> dict:
        rows = [(c, p) for t, c, p in self.samples if t0 <= t <= t1]
        if not rows:
            return {"samples": 0}
        return {"samples": len(rows), "sm_mhz_median": statistics.median(c for c, _ in rows),
                "sm_mhz_min": min(c for c, _ in rows), "power_w_max": round(max(p for _, p in rows), 1),
                "power_w_mean": round(sum(p for _, p in rows) / len(rows), 1)}

    def stop(self):
        self._stop.set()


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_arm(steps: int, warmup: int, budget_s: float = 150.0) -> dict:
    """The reference's CPU path on a bounded sample: rows [0, m) of the 16384^3 bf16 product per step.

    torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the reference run by hand has every core, so
    the thread count is set back to what the process may use and recorded."""
    import torch
    from oracle import reference_port  # the one other place bench.py may execute oracle/

    torch.set_num_threads(host_threads())
    cores = torch.get_num_threads()
    torch.manual_seed(0)
    A = torch.randn(N, N, dtype=torch.bfloat16)
    B = torch.randn(N, N, dtype=torch.bfloat16)
    # calibrate on a 512-row slab (after one throw-away call), then size the per-step slab to the budget
    torch.matmul(A[:256], B)
    t = time.perf_counter()
    torch.matmul(A[:512], B)
    rate = 512 / (time.perf_counter() - t)  # rows per second
    m = int(rate * budget_s / max(steps + warmup, 1)) // 128 * 128
    m = max(128, min(N, m))
    slab = A[:m]
    for _ in range(warmup):
        torch.matmul(slab, B)
    t0 = time.perf_counter()
    for _ in range(steps):
        torch.matmul(slab, B)  # timed exactly as matmul_benchmark.py:69-74 (perf_counter around the loop)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    tflops = 2.0 * m * N * N / dt / 1e12
    assert abs(reference_port.calculate_tflops(N, dt * N / m) - tflops) < 1e-9 * max(tflops, 1)
    del A, B, slab
    # BASELINE.json configs[0], the reference's own CPU-runnable case: 4096^3 float32, benchmark_matmul's CPU branch
    c1_time, c1_tflops = reference_port.benchmark_matmul_cpu(4096, torch.float32, num_iterations=5, warmup_iterations=1, seed=0)
    config1 = {"workload": "4096x4096x4096 float32, torch.matmul on CPU (BASELINE.json configs[0]), 1 warm-up + 5 timed",
               "value": c1_tflops, "unit": UNIT, "ms": c1_time * 1000, "cores": cores}
    return {"value": tflops, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"rows [0,{m}) of the {N}^3 bf16 product per step ({m / N:.4f} of a step), torch.matmul on CPU, "
                      f"{warmup} warm-up + {steps} timed", "ms_per_step": dt * 1000, "full_step_ms_extrapolated": dt * 1000 * N / m,
            "sample_fraction_of_step": m / N, "host_cpus": os.cpu_count(), "config1": config1}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU arm; the other ranks exit 0 without work
    base = cpu_reference_arm(args.steps, args.warmup, budget_s=args.ref_budget_s)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": base["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": shared_config(),
            "notes": {"arm": "reference CPU path: torch.matmul on the host cores, timed as matmul_benchmark.py:69-74, rank 0 only",
                      "ms_per_step": "the TIMED time of one step's sample; full_step_ms_extrapolated scales it to all 16384 rows",
                      "full_step_ms_extrapolated": base["full_step_ms_extrapolated"],
                      "sample_fraction_of_step": base["sample_fraction_of_step"]},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "config1")},
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# N > 1: the modes that shard (BASELINE.json configs[3], configs[4]) — parity first, then time
# ------------------------------------------------------------------------------------------------------------------
def rel_err_gpu(got, ref, chunk_rows: int = 2048) -> float:
    """max |got - ref| / (|ref| + rms(ref)) over whole 2-D tensors, on the device, in float64, chunk by chunk."""
    import torch
    g2, r2 = got.reshape(-1, got.shape[-1]), ref.reshape(-1, ref.shape[-1])
    sq = 0.0
    for i in range(0, r2.shape[0], chunk_rows):
        sq += r2[i:i + chunk_rows].double().pow(2).sum().item()
    rms = (sq / r2.numel()) ** 0.5
    worst = 0.0
    for i in range(0, r2.shape[0], chunk_rows):
        r = r2[i:i + chunk_rows].double()
        worst = max(worst, ((g2[i:i + chunk_rows].double() - r).abs() / (r.abs() + rms)).max().item())
    return worst


def sharded_modes(args, rank: int, world: int, dev, t_independent_ms: float) -> list:
    """Parity-check and time batch_parallel, matrix_parallel and the iteration-level overlap on these `world` GPUs.

    Every record: {"mode", "variant", "ms", "gemm_alone_ms" (timed interleaved, same process), "weak_scaling_eff"
    (= gemm_alone_ms / ms where the per-GPU work is that GEMM), "bus_gbs", "frac_of_900", "parity_err",
    "parity_tol", "parity_ok", ...}.  Device times, CUDA events, max over ranks.

    Parity is established BEFORE timing, on the full 16384 problem, three ways: (a) sub-blocks against the CPU
    oracle (oracle/ is the checker here, as in tests/: exact products, one rounding per rank's product as the
    reference reduces rounded tensors, float64 sum); (b) the whole tensor against the reference-style NCCL
    path on the same inputs; (c) the result is bit-identical on every rank."""
    import torch
    import torch.distributed as dist
    import pytorch_distributed_matmul_benchmark_b200 as b2
    from pytorch_distributed_matmul_benchmark_b200 import hostlib
    from pytorch_distributed_matmul_benchmark_b200 import matmul_overlap_benchmark as ov
    from pytorch_distributed_matmul_benchmark_b200.matmul_scaling_benchmark import _timed
    from oracle import oracle  # checker only: nothing timed below touches it

    dt = torch.bfloat16
    iters = max(3, min(args.steps, 10))
    ops = hostlib.device_ops(rank, world, str(dev))
    comm = ops.comm
    recs = []

    def max_ranks(x: float) -> float:
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def timed(fn, n=iters, warm=3) -> float:
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        return max_ranks(s.elapsed_time(e) / n)

    def same_on_all_ranks(t) -> bool:
        """Bit-identical on every rank: compare position-weighted 64-bit checksums of the raw bit patterns."""
        v = t.reshape(-1, t.shape[-1]).view(torch.int32)  # two 16-bit values per word
        rs, cs = v.sum(dim=1, dtype=torch.int64), v.sum(dim=0, dtype=torch.int64)
        sig = torch.stack([(rs * torch.arange(1, rs.numel() + 1, device=dev)).sum(),
                           (cs * torch.arange(1, cs.numel() + 1, device=dev)).sum(), rs.sum()])
        lo, hi = sig.clone(), sig.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        return bool(torch.equal(lo, hi))

    def blocks_for(rank_: int, k: int = 2):
        """k 64 x 64 sub-block origins, different on every rank, 16-byte aligned columns, inside the matrix."""
        out = []
        for j in range(k):
            r0 = (rank_ * 2029 + j * 7919 + 321) % (N - 64)
            c0 = ((rank_ * 4099 + j * 5003 + 777) % (N - 64)) // 8 * 8
            out.append((r0, c0))
        return out

    # ---------------- batch_parallel, per-GPU batch 1 (weak scaling; the reference's 4 // ws is 0 at ws = 8) -------
    torch.manual_seed(rank)  # matmul_scaling_benchmark.py:114-116
    A = torch.randn(1, N, N, dtype=dt, device=dev)
    B = torch.randn(1, N, N, dtype=dt, device=dev)
    C = torch.empty_like(A)
    red = comm.reduce_buffer_like(C)  # multicast (NVLS) where the box has it, else peer-mapped
    nvls = hasattr(red, "mc_ptr")
    fused_variant = "fused NVLS (GEMM + in-switch multimem reduction, one pass)" if nvls else "fused peer (GEMM + pull/add/push over peer memory)"
    bands = 8
    ar_bytes = C.numel() * 2

    comm.bmm_allreduce_fused(A, B, red, bands)
    b2.bmm(A, B, out=C)
    comm.all_reduce(C)  # the reference's serial path: bmm, then dist.all_reduce (:142-150)
    torch.cuda.synchronize()
    identical = same_on_all_ranks(red.local)
    # (a) oracle sub-blocks: the exact (float64) product of every rank on this block; oracle.reduced_sum_err applies
    # the reference's error budget (one rounding per rank's product, each against that product's own magnitude, then
    # the roundings of the sum) and returns the error in units of the bf16 tolerance
    blks = blocks_for(rank)
    terms = [[] for _ in blks]
    term_mag = torch.zeros(1, N, N, dtype=torch.float32, device=dev)  # sum_r |T_r| + rms(T_r): bounds every partial sum
    for r in range(world):
        torch.manual_seed(r)
        Ar = torch.randn(1, N, N, dtype=dt, device=dev)
        Br = torch.randn(1, N, N, dtype=dt, device=dev)
        for j, (r0, c0) in enumerate(blks):
            a = Ar[0, r0:r0 + 64, :].cpu()
            b = Br[0, :, c0:c0 + 64].contiguous().cpu()
            terms[j].append(oracle.gemm_block_f64(a, b, 0, 64, 0, 64))
        Tr = b2.bmm(Ar, Br).float()  # magnitudes only (the denominator of the full-tensor comparison below)
        term_mag += Tr.abs_() + Tr.pow(2).mean().sqrt()
        del Ar, Br, Tr
    # (b) the whole tensor against the NCCL path: two correct reductions of the same products; NCCL may round partial
    # sums (ring / tree), which the ranks' terms bound, not the final sum — hostlib.reduction_pair_error, in tolerance units
    err_vs_nccl = max_ranks(hostlib.reduction_pair_error(red.local, C, term_mag, world))
    del term_mag
    # roundings of the sum: one to nearest on the peer path; one by truncation in the switch (measured: within one
    # ulp, so it counts double); NCCL: one in the switch (NVLS) or one per ring step
    k_fused = 2.0 if nvls else 1.0
    k_nccl = float(max(world - 1, 2))

    def block_err(t, k, ring=False):
        return max_ranks(max(oracle.reduced_sum_err(t[0, r0:r0 + 64, c0:c0 + 64], terms[j], k, ring) for j, (r0, c0) in enumerate(blks)))

    err_oracle, err_nccl_oracle = block_err(red.local, k_fused), block_err(C, k_nccl, ring=True)
    vs = (f"CPU oracle, {len(blks)} 64x64 sub-blocks per rank: exact float64 products of all {world} ranks; error = |got - sum| / "
          "(sum_r(|T_r| + rms T_r) + k (|S| + rms S)), k = roundings of the sum")
    parity_fused = {"parity_err": err_oracle, "parity_tol": TOL_BF16, "sum_roundings_allowed": k_fused, "parity_vs": vs,
                    "err_vs_nccl_path_full_tensor": err_vs_nccl, "tol_vs_nccl_path": TOL_BF16,
                    "vs_nccl_path_bound": "|fused - nccl| / ((world-1) sum_r(|T_r| + rms T_r) + 2 (|S| + rms S)), every element",
                    "bit_identical_on_all_ranks": identical,
                    "parity_ok": bool(err_oracle <= TOL_BF16 and err_vs_nccl <= TOL_BF16 and identical)}
    parity_serial = {"parity_err": err_nccl_oracle, "parity_tol": TOL_BF16, "sum_roundings_allowed": k_nccl, "parity_vs": vs,
                     "parity_ok": bool(err_nccl_oracle <= TOL_BF16)}

    # timing.  Reference-style split: every phase between synchronisations, compute + comm summed (:135-157)
    for _ in range(2):
        b2.bmm(A, B, out=C)
        comm.all_reduce(C)
    torch.cuda.synchronize()
    dist.barrier()
    comp = comm_ms = 0.0
    for _ in range(iters):
        comp += _timed(ops, lambda: b2.bmm(A, B, out=C))
        comm_ms += _timed(ops, lambda: comm.all_reduce(C))
    comp, comm_ms = max_ranks(comp / iters), max_ranks(comm_ms / iters)
    bus = lambda ms: ar_bytes * 2 * (world - 1) / world / (ms * 1e-3) / 1e9  # noqa: E731  (all_reduce bus bandwidth)
    recs.append(dict(mode="batch_parallel", variant="serial NCCL (reference-style: bmm, sync, all_reduce, sync)", per_gpu_batch=1,
                     ms=comp + comm_ms, compute_ms=comp, comm_ms=comm_ms, gemm_alone_ms=comp, weak_scaling_eff=comp / (comp + comm_ms),
                     bus_gbs=bus(comm_ms), frac_of_900=bus(comm_ms) / NVLINK_GBS, total_tflops=world * FLOP / ((comp + comm_ms) * 1e-3) / 1e12,
                     **parity_serial))
    # fused.  `ms` and `gemm_alone_ms` follow the headline's own protocol (a pause, then args.warmup untimed and
    # args.steps timed calls), so that they are comparable with `ms_per_step` of this and of the N = 1 run; the
    # interleaved rounds after that give the ratio of neighbouring measurements in one power state.
    def like_headline(fn) -> float:
        torch.cuda.synchronize()
        time.sleep(1.5)
        return timed(fn, n=args.steps, warm=args.warmup)

    g_head = like_headline(lambda: b2.bmm(A, B, out=C))
    f_head = like_headline(lambda: comm.bmm_allreduce_fused(A, B, red, bands))
    rounds = []
    for _ in range(3):
        g = timed(lambda: b2.bmm(A, B, out=C), n=max(iters // 2, 3), warm=2)
        f = timed(lambda: comm.bmm_allreduce_fused(A, B, red, bands), n=max(iters // 2, 3), warm=2)
        rounds.append((g, f))
    g_med = statistics.median(g for g, _ in rounds)
    f_med = statistics.median(f for _, f in rounds)
    ar_alone = timed(lambda: comm.all_reduce_sync(red))
    recs.append(dict(mode="batch_parallel", variant=fused_variant, per_gpu_batch=1, bands=bands, ms=f_head, gemm_alone_ms=g_head,
                     weak_scaling_eff=g_head / f_head, weak_scaling_eff_vs_independent_of_this_run=t_independent_ms / f_head,
                     protocol=f"ms and gemm_alone_ms: 1.5 s pause, {args.warmup} warm-up + {args.steps} timed calls each (the headline's protocol)",
                     interleaved={"gemm_alone_ms": g_med, "fused_ms": f_med, "weak_scaling_eff": g_med / f_med,
                                  "rounds_ms": [{"gemm_alone": round(g, 4), "fused": round(f, 4)} for g, f in rounds]},
                     exposed_comm_ms=f_head - g_head, allreduce_alone_ms=ar_alone, bus_gbs=bus(ar_alone), frac_of_900=bus(ar_alone) / NVLINK_GBS,
                     bus_note="bus_gbs is the stand-alone reduction (same kernels, no GEMM beside it): fused, the transfer hides behind the GEMM",
                     total_tflops=world * FLOP / (f_head * 1e-3) / 1e12, **parity_fused))

    # ---------------- iteration-level overlap (backup/matmul_overlap_benchmark.py:36-180): all_reduce of C_i beside GEMM i+1 ----
    # parity of the reduction it uses (the self-synchronising all_reduce on its own): C_r into the buffer, reduce, compare
    b2.bmm(A, B, out=red.local)
    torch.cuda.synchronize()
    dist.barrier()
    comm.all_reduce_sync(red)
    torch.cuda.synchronize()
    err_sync = block_err(red.local, k_fused)
    sync_identical = same_on_all_ranks(red.local)
    parity_ov = {"parity_err": err_sync, "parity_tol": TOL_BF16, "sum_roundings_allowed": k_fused, "parity_vs": vs,
                 "bit_identical_on_all_ranks": sync_identical, "parity_ok": bool(err_sync <= TOL_BF16 and sync_identical)}
    del red
    comm.release_shared()
    torch.manual_seed(4000 + rank)
    t_no, _, _ = ov.benchmark_no_overlap(N, dt, str(dev), rank, iters, 3, ops=ops)
    t_no = max_ranks(t_no * 1000)
    comm.release_shared()
    torch.cuda.empty_cache()
    t_ov, tf_ov, _ = ov.benchmark_overlap(N, dt, str(dev), rank, iters, 3, ops=ops)
    t_ov = max_ranks(t_ov * 1000)
    g_ov = FLOP / (tf_ov * 1e12) * 1e3  # the compute-only GEMM time that function measures (:167-178)
    g_ov = max_ranks(g_ov)
    comm.release_shared()
    torch.cuda.empty_cache()
    recs.append(dict(mode="overlap (iteration level)", variant="no_overlap: GEMM, sync, NCCL all_reduce, sync", ms=t_no, gemm_alone_ms=g_ov,
                     weak_scaling_eff=g_ov / t_no, bus_gbs=None, frac_of_900=None, **parity_serial))
    recs.append(dict(mode="overlap (iteration level)", variant="overlap: all_reduce of C_i " + ("in the NVSwitch" if nvls else "over peer memory")
                     + " on a second stream beside GEMM i+1 (double-buffered, every edge an event)", ms=t_ov, gemm_alone_ms=g_ov,
                     weak_scaling_eff=g_ov / t_ov, bus_gbs=bus(ar_alone), frac_of_900=bus(ar_alone) / NVLINK_GBS, **parity_ov))
    del A, B, C
    torch.cuda.empty_cache()

    # ---------------- matrix_parallel: A replicated, B column-split, C slabs gathered (:167-238) ---------------------
    torch.manual_seed(0)  # seeded so that A really is replicated and a global C exists to check against
    A = torch.randn(N, N, dtype=dt, device=dev)
    B_full = torch.randn(N, N, dtype=dt, device=dev)
    s_col, e_col = hostlib.split_columns(N, world, rank)
    n_local = e_col - s_col
    B_local = B_full[:, s_col:e_col].contiguous()
    C_local = torch.empty(N, n_local, dtype=dt, device=dev)
    slabs = torch.zeros(world, N, n_local, dtype=dt, device=dev)
    gbuf = comm.gather_buffer(N, n_local, dt)
    mc_gather = hasattr(gbuf, "mc_ptr")
    gbuf.local.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    comm.matmul_allgather_fused(A, B_local, gbuf)
    b2.matmul(A, B_local, out=C_local)
    comm.all_gather_into(slabs, C_local)
    torch.cuda.synchronize()
    same_as_nccl = bool(torch.equal(gbuf.local, slabs))
    t = torch.tensor([1.0 if same_as_nccl else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    same_as_nccl = bool(t.item() == 1.0)
    err_mp = err_mp_serial = 0.0
    for slab in range(world):  # one sub-block in every rank's slab, rows differ per rank
        r0 = (rank * 2029 + slab * 997 + 55) % (N - 64)
        c0 = ((slab * 131 + rank * 17) % (n_local - 64)) // 8 * 8
        a = A[r0:r0 + 64, :].cpu()
        b = B_full[:, slab * n_local + c0: slab * n_local + c0 + 64].contiguous().cpu()
        ref = oracle.gemm_block_f64(a, b, 0, 64, 0, 64)
        err_mp = max(err_mp, oracle.rel_err(gbuf.local[slab, r0:r0 + 64, c0:c0 + 64], ref))
        err_mp_serial = max(err_mp_serial, oracle.rel_err(slabs[slab, r0:r0 + 64, c0:c0 + 64], ref))
    err_mp, err_mp_serial = max_ranks(err_mp), max_ranks(err_mp_serial)
    Cg = hostlib.assemble_slabs(gbuf.local, N, world)
    err_f64_rows = max_ranks(hostlib.parity_error(Cg, A, B_full, rows=128))  # float64 on the device, first + last 128 rows, all columns
    del Cg
    parity_mp = {"parity_err": err_mp, "parity_tol": TOL_BF16, "parity_vs": f"CPU oracle (float64), one 64x64 sub-block in each of the {world} slabs per rank",
                 "err_vs_float64_rows_on_device": err_f64_rows, "bit_identical_to_nccl_gather": same_as_nccl,
                 "parity_ok": bool(err_mp <= TOL_BF16 and err_f64_rows <= TOL_BF16 and same_as_nccl)}
    parity_mp_serial = {"parity_err": err_mp_serial, "parity_tol": TOL_BF16, "parity_vs": parity_mp["parity_vs"], "parity_ok": bool(err_mp_serial <= TOL_BF16)}

    for _ in range(2):
        b2.matmul(A, B_local, out=C_local)
        comm.all_gather_into(slabs, C_local)
    torch.cuda.synchronize()
    dist.barrier()
    comp = comm_ms = 0.0
    for _ in range(iters):
        comp += _timed(ops, lambda: b2.matmul(A, B_local, out=C_local))
        comm_ms += _timed(ops, lambda: comm.all_gather_into(slabs, C_local))
    comp, comm_ms = max_ranks(comp / iters), max_ranks(comm_ms / iters)
    ag_bytes = N * N * 2
    ag_bus = lambda ms: ag_bytes * (world - 1) / world / (ms * 1e-3) / 1e9  # noqa: E731  (bytes each GPU receives / time)
    link_floor_ms = ag_bytes * (world - 1) / world / (NVLINK_GBS * 1e9) * 1e3
    recs.append(dict(mode="matrix_parallel", variant="serial NCCL (reference-style: matmul, sync, all_gather, sync)", ms=comp + comm_ms,
                     compute_ms=comp, comm_ms=comm_ms, gemm_alone_ms=comp, weak_scaling_eff=None, speedup_vs_independent_of_this_run=t_independent_ms / (comp + comm_ms),
                     bus_gbs=ag_bus(comm_ms), frac_of_900=ag_bus(comm_ms) / NVLINK_GBS, total_tflops=FLOP / ((comp + comm_ms) * 1e-3) / 1e12,
                     **parity_mp_serial))
    rounds = []
    for _ in range(3):
        g = timed(lambda: b2.matmul(A, B_local, out=C_local), warm=2)
        f = timed(lambda: comm.matmul_allgather_fused(A, B_local, gbuf), warm=2)
        rounds.append((g, f))
    g_med = statistics.median(g for g, _ in rounds)
    f_med = statistics.median(f for _, f in rounds)
    # the store path on its own: the same kernel with K = 64 has (almost) no math left, so its time is what the epilogue's
    # stores need to put all N x N outputs into every rank's buffer over NVLink — the measured link floor of this mode
    A64, B64 = A[:, :64].contiguous(), B_local[:64].contiguous()  # the fused entry points take densely packed operands
    store_only = timed(lambda: comm.matmul_allgather_fused(A64, B64, gbuf), warm=2)
    del A64, B64
    comm.matmul_allgather_fused(A, B_local, gbuf)  # leave the real product in the buffer
    recs.append(dict(mode="matrix_parallel", variant="fused (GEMM epilogue stores each tile " + ("through ONE multicast TMA store, replicated by the NVSwitch"
                     if mc_gather else "into every rank's buffer by peer TMA stores") + "; + barrier)", ms=f_med, gemm_alone_ms=g_med,
                     weak_scaling_eff=None, store_path_alone_ms=store_only, bound_ms=max(g_med, store_only), frac_of_bound=max(g_med, store_only) / f_med,
                     bound_note="bound = max(GEMM alone, store path alone [same kernel, K = 64]); the stores cannot start before the first tiles "
                                "have run through all of K, so one tile period of the link's time is structurally lost",
                     link_floor_nominal_ms=link_floor_ms, speedup_vs_independent_of_this_run=t_independent_ms / f_med,
                     rounds_ms=[{"gemm_alone": round(g, 4), "fused": round(f, 4)} for g, f in rounds],
                     bus_gbs=ag_bus(f_med), frac_of_900=ag_bus(f_med) / NVLINK_GBS,
                     bus_note="bytes each GPU receives / whole fused time (the transfer runs under the math, so this is a floor on the link rate reached)",
                     total_tflops=FLOP / (f_med * 1e-3) / 1e12, **parity_mp))
    del gbuf
    ops.destroy()
    return recs


def copy_peak(torch, dev, hA, hB, hC, barrier, max_over_ranks, world: int) -> dict:
    """Plain pinned cudaMemcpyAsync of one step's bytes (A and B up, C down, both directions at once), all ranks
    at the same time: the most any end-to-end path over these host buffers could reach on this box at this N."""
    dA = torch.empty(N, N, dtype=torch.bfloat16, device=dev)
    dB = torch.empty(N, N, dtype=torch.bfloat16, device=dev)
    dC = torch.empty(N, N, dtype=torch.bfloat16, device=dev)
    up, down = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream()

    def once():
        up.wait_stream(main)
        down.wait_stream(main)
        with torch.cuda.stream(up):
            dA.copy_(hA, non_blocking=True)
            dB.copy_(hB, non_blocking=True)
        with torch.cuda.stream(down):
            hC.copy_(dC, non_blocking=True)
        main.wait_stream(up)
        main.wait_stream(down)

    once()
    barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    s.record(main)
    for _ in range(reps):
        once()
    e.record(main)
    barrier()
    ms = max_over_ranks(s.elapsed_time(e)) / reps
    h2d, d2h = 2 * N * N * 2, N * N * 2
    return {"ms_per_step_bytes": ms, "h2d_gbs_per_gpu": h2d / (ms * 1e-3) / 1e9, "d2h_gbs_per_gpu": d2h / (ms * 1e-3) / 1e9,
            "aggregate_gbs": world * (h2d + d2h) / (ms * 1e-3) / 1e9,
            "how": "pinned cudaMemcpyAsync, A+B up and C down concurrently on two streams, all ranks at once, 3 repetitions"}


def run_b2mm(args) -> None:
    import torch
    import torch.distributed as dist
    import pytorch_distributed_matmul_benchmark_b200 as b2
    from pytorch_distributed_matmul_benchmark_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local % torch.cuda.device_count())
    dev = torch.device("cuda", torch.cuda.current_device())
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()  # fail loudly here if the CUDA extension is missing

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # independent-mode recipe: seeded per rank (matmul_scaling_benchmark.py:73-75)
    torch.manual_seed(rank)
    A = torch.randn(N, N, dtype=torch.bfloat16, device=dev)
    B = torch.randn(N, N, dtype=torch.bfloat16, device=dev)
    C = torch.empty(N, N, dtype=torch.bfloat16, device=dev)
    stream = torch.cuda.current_stream()
    sampler = ClockSampler(dev.index) if rank == 0 else None
    nvml = NvmlSampler(dev.index) if rank == 0 else None

    def timed_device_region():
        w0 = time.time()  # clocks are sampled from the first warm-up launch on: same kernel, same load
        for _ in range(args.warmup):
            b2.matmul(A, B, out=C)
        barrier()
        t0 = time.time()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        launches0 = _lib.launch_count()
        ev[0].record(stream)
        for i in range(args.steps):
            b2.matmul(A, B, out=C)
            ev[i + 1].record(stream)
        barrier()
        w1 = time.time()
        per_launch = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
        total_ms = ev[0].elapsed_time(ev[-1])
        return total_ms, per_launch, _lib.launch_count() - launches0, (w0, w1), (t0, w1)

    total_ms, per_launch, launches, win, timed_win = timed_device_region()
    clocks = sampler.window(*win) if sampler else None
    if clocks and any(r in BAD_REASONS for r in clocks["reasons"]):  # rejected: take it again, once
        time.sleep(5)
        total_ms, per_launch, launches, win, timed_win = timed_device_region()
        clocks = sampler.window(*win)
        clocks["remeasured"] = True
    if clocks is not None and nvml is not None:
        clocks["nvml_timed_region"] = nvml.window(*timed_win) if nvml.ok else {"samples": 0, "note": "NVML bindings unavailable"}
    kernel_name = lib.b2mm_last_kernel().decode()
    total_ms = max_over_ranks(total_ms)
    ms_per_step = total_ms / args.steps
    value = world * FLOP / (ms_per_step * 1e-3) / 1e12
    kernel_ms = sum(per_launch) / len(per_launch)

    # ---- roofline extras (rank 0 of a single-GPU run): a >= 2 s sustained loop and cuBLAS on the same tensors --------
    sustained = cublas_same_run = None
    if world == 1 and not args.quick:
        time.sleep(1.0)
        t0 = time.time()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_sus = max(int(2200.0 / kernel_ms) + 1, 50)
        s.record(stream)
        for _ in range(n_sus):
            b2.matmul(A, B, out=C)
        e.record(stream)
        torch.cuda.synchronize()
        t1 = time.time()
        sus_ms = s.elapsed_time(e) / n_sus
        sus_clk = sampler.window(t0 + 0.5, t1 - 0.05) if sampler else {}
        sustained = {"launches": n_sus, "seconds": (t1 - t0), "ms_per_launch": sus_ms, "tflops": FLOP / (sus_ms * 1e-3) / 1e12,
                     "clocks": sus_clk, "nvml": nvml.window(t0 + 0.5, t1 - 0.05) if nvml and nvml.ok else None}
        # the Blackwell kernel to beat, on the same tensors, interleaved round by round (off the measured path)
        time.sleep(1.0)
        Cc = torch.empty_like(C)
        rounds = []
        for _ in range(3):
            pair = {}
            for name, fn in (("b2mm", lambda: b2.matmul(A, B, out=C)), ("cublas", lambda: torch.matmul(A, B, out=Cc))):
                fn()
                s.record(stream)
                for _ in range(10):
                    fn()
                e.record(stream)
                torch.cuda.synchronize()
                pair[name] = FLOP / (s.elapsed_time(e) / 10 * 1e-3) / 1e12
            rounds.append({k: round(v, 1) for k, v in pair.items()})
        cublas_same_run = {"rounds_tflops": rounds, "b2mm_median": statistics.median(r["b2mm"] for r in rounds),
                           "cublas_median": statistics.median(r["cublas"] for r in rounds),
                           "bit_identical_results": bool(torch.equal(C, Cc)),
                           "how": "3 rounds x (1 warm-up + 10 timed) of each, alternating, same A and B, after the sustained loop (warm part)"}
        del Cc

    # ---- end to end: host buffers, copies inside the timed region -------------------------------
    hA = torch.empty(N, N, dtype=torch.bfloat16, pin_memory=True)
    hB = torch.empty(N, N, dtype=torch.bfloat16, pin_memory=True)
    hC = torch.empty(N, N, dtype=torch.bfloat16, pin_memory=True)
    hA.copy_(A)
    hB.copy_(B)
    del C
    host_gemm = b2.HostGemm(N, N, N, torch.bfloat16, device=dev, panels=8)  # the public host-buffer call
    e2e_steps = max(3, min(args.steps, 10))

    def e2e_step():
        host_gemm(hA, hB, hC)

    for _ in range(2):
        e2e_step()
    barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(stream)
    for _ in range(e2e_steps):
        e2e_step()
    e.record(stream)
    barrier()
    e2e_ms = max_over_ranks(s.elapsed_time(e)) / e2e_steps
    e2e_value = world * FLOP / (e2e_ms * 1e-3) / 1e12
    bytes_in, bytes_out = hA.numel() * 2 + hB.numel() * 2, hC.numel() * 2
    del host_gemm
    torch.cuda.empty_cache()
    peak_copy = copy_peak(torch, dev, hA, hB, hC, barrier, max_over_ranks, world)
    del hA, hB, hC, A, B
    torch.cuda.empty_cache()
    if sampler:
        sampler.stop()
    if nvml:
        nvml.stop()

    # ---- N > 1: the sharded modes, parity first -------------------------------------------------
    modes = None
    modes_error = None
    if world > 1 and not args.no_modes:
        try:
            modes = sharded_modes(args, rank, world, dev, ms_per_step)
        except Exception as ex:  # reported in the line, and fatal below
            import traceback
            modes_error = f"rank {rank}: {type(ex).__name__}: {ex}"
            traceback.print_exc()

    if modes_error is not None and rank != 0:
        # the other ranks may be inside a collective this rank never reaches: leave at once, the launcher ends them
        sys.stderr.flush()
        os._exit(1)
    if rank != 0:
        bad = modes is not None and not all(m["parity_ok"] for m in modes)
        dist.destroy_process_group()
        if bad:
            sys.exit(1)
        return

    peaks = measured_peaks()
    achieved = FLOP / (kernel_ms * 1e-3) / 1e12
    timed_s = total_ms * 1e-3
    # a fraction-of-a-second region is a burst measurement: divide by the burst figure (long regions: sustained)
    burst = timed_s < 1.0 or not peaks.get("bf16_tflops_sustained")
    peak = float(peaks["bf16_tflops"] if burst else peaks["bf16_tflops_sustained"])
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get("gemm_bf16_16k_dram_bytes_per_launch")
    except OSError:
        pass
    roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": UNIT, "frac": achieved / peak,
                "traffic": traffic,
                "peak_source": f"{peaks['source']} cuBLAS bf16 {'burst' if burst else 'sustained'} "
                               f"(timed region {timed_s:.2f} s: {'shorter' if burst else 'longer'} than 1 s)",
                "peak_sustained": peaks.get("bf16_tflops_sustained"), "frac_of_spec_dense_peak": achieved / SPEC_PEAK,
                "spec_dense_peak": SPEC_PEAK, "kernel": kernel_name,
                "algorithmic_flop_per_launch": FLOP, "algorithmic_bytes_per_launch": 3 * N * N * 2,
                "kernel_ms": kernel_ms, "per_launch_ms": [round(x, 4) for x in per_launch]}
    if sustained is not None:
        ps = peaks.get("bf16_tflops_sustained")
        sustained["peak"] = ps
        sustained["frac"] = sustained["tflops"] / ps if ps else None
        sustained["frac_of_spec_dense_peak"] = sustained["tflops"] / SPEC_PEAK
        roofline["sustained"] = sustained
    if cublas_same_run is not None:
        roofline["cublas_same_run"] = cublas_same_run

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        c = cpu_reference_arm(steps=3, warmup=1, budget_s=20.0)
        cpu = {k: c[k] for k in ("value", "unit", "cores", "kind", "sample", "config1")}

    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": bytes_out,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "h2d_gbs_per_gpu": bytes_in / (e2e_ms * 1e-3) / 1e9,
           "d2h_gbs_per_gpu": bytes_out / (e2e_ms * 1e-3) / 1e9,
           "aggregate_copy_gbs": world * (bytes_in + bytes_out) / (e2e_ms * 1e-3) / 1e9,
           "copy_peak_same_run": peak_copy, "frac_of_copy_peak": peak_copy["ms_per_step_bytes"] / e2e_ms,
           "bound": "host<->device copies (PCIe Gen5 x16 per GPU, shared host fabric across GPUs): the step cannot be shorter than "
                    "moving its 1 GiB in and 0.5 GiB out; the GEMM (a quarter of the step) runs under them"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "config": shared_config(),
            "notes": {"parallelism": f"replicas x{world}, no data-path collective",
                      "l2": "inputs (2 x 512 MiB) and output (512 MiB) exceed the 126 MB L2; no flush needed",
                      "per_gpu_tflops": value / world, "pct_of_spec_dense_peak": 100.0 * value / world / SPEC_PEAK},
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches}
    ok = True
    if modes is not None:
        line["modes"] = modes
        ok = all(m["parity_ok"] for m in modes)
        line["modes_parity_ok"] = ok
    if modes_error is not None:
        line["modes_error"] = modes_error
        ok = False
    print(json.dumps(line), flush=True)
    if modes_error is not None:
        os._exit(1)
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["b2mm", "reference"], default="b2mm")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~20 s CPU baseline leg")
    ap.add_argument("--quick", action="store_true", help="skip the sustained loop and the cuBLAS comparison (N = 1)")
    ap.add_argument("--no-modes", action="store_true", help="N > 1: skip the sharded-mode records")
    ap.add_argument("--ref-budget-s", type=float, default=150.0, help="--impl reference: CPU seconds the whole run may take")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b2mm" else args.warmup
    if args.gpus > 1 and "RANK" not in os.environ:
        # run by hand without a launcher: start one process per GPU the way the driver does
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b2mm(args)


if __name__ == "__main__":
    main()
