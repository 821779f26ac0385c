"""Worker for test_collectives_gpu.py: the library's collectives and fused paths on >= 2 GPUs (one process per GPU).

Every result is checked on every rank against values computed independently: the oracle (exact products, double
sum) for GEMM content, and plain float64 sums of the per-rank inputs for the reductions."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle import oracle, reference_port  # noqa: E402
import pytorch_distributed_matmul_benchmark_b200 as b2  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import _lib, hostlib  # noqa: E402

TOL = {torch.bfloat16: 4e-3, torch.float16: 5e-4, torch.float32: 1e-5}


def rel(C, ref):
    ref = ref.double().cpu()
    rms = ref.pow(2).mean().sqrt()
    return ((C.double().cpu() - ref).abs() / (ref.abs() + rms)).max().item()


def main(out_path):
    rank, ws = hostlib.setup_distributed()
    dev = hostlib.rank_device(rank)
    comm = b2.Communicator(rank, ws, torch.device(dev))
    res = {}

    # all_reduce / all_gather through NCCL owned by the library
    for dt in (torch.bfloat16, torch.float32):
        torch.manual_seed(100 + rank)
        x = torch.randn(3, 1000, dtype=dt, device=dev)
        want = sum(torch.randn(3, 1000, dtype=dt, generator=torch.Generator().manual_seed(0)) * 0 for _ in range(1))  # placeholder shape
        parts = []
        for r in range(ws):
            torch.manual_seed(100 + r)
            parts.append(torch.randn(3, 1000, dtype=dt, device=dev).double())
        want = sum(parts)
        torch.manual_seed(100 + rank)
        x = torch.randn(3, 1000, dtype=dt, device=dev)
        comm.all_reduce(x)
        torch.cuda.synchronize()
        res[f"all_reduce_{dt}"] = rel(x, want) <= (ws - 1) * TOL[dt] + 1e-6
        slabs = torch.zeros(ws, 3, 1000, dtype=dt, device=dev)
        torch.manual_seed(100 + rank)
        mine = torch.randn(3, 1000, dtype=dt, device=dev)
        comm.all_gather_into(slabs, mine)
        torch.cuda.synchronize()
        res[f"all_gather_{dt}"] = all(torch.equal(slabs[r].double(), parts[r]) for r in range(ws))
        # own peer-memory all_reduce: fp32 adds, one rounding
        sh = comm.shared_like(mine)
        sh.local.copy_(mine)
        torch.cuda.synchronize()
        dist.barrier()
        comm.all_reduce_p2p(sh)
        torch.cuda.synchronize()
        res[f"all_reduce_p2p_{dt}"] = rel(sh.local, want) <= TOL[dt]
    comm.all_reduce(torch.empty(0, dtype=torch.bfloat16, device=dev))  # local batch 0 (ws = 8 in the reference): no-op
    res["empty_all_reduce"] = True

    # matrix_parallel, fused GEMM + all_gather: every rank ends up with the global product
    N = 1024
    A, Bfull, B_local, (s, e) = reference_port.matrix_parallel_inputs(rank, ws, N, torch.bfloat16, device=dev)
    shared = comm.shared_slabs(N, e - s, torch.bfloat16)
    shared.local.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    for _ in range(2):
        comm.matmul_allgather_fused(A, B_local, shared)
    torch.cuda.synchronize()
    Cg = torch.cat(list(shared.local.unbind(0)), dim=1)
    res["fused_allgather"] = rel(Cg, oracle.gemm_block_f64(A.cpu(), Bfull.cpu(), 0, N, 0, N)) <= TOL[torch.bfloat16]
    slabs = torch.zeros(ws, N, e - s, dtype=torch.bfloat16, device=dev)
    C_local = b2.matmul(A, B_local)
    comm.all_gather_into(slabs, C_local)
    torch.cuda.synchronize()
    res["fused_allgather_equals_nccl_path"] = torch.equal(slabs, shared.local)

    # batch_parallel, fused GEMM + all_reduce: sum over ranks of the per-rank products.  The expectation comes from
    # the ORACLE, never from the kernel under test: the exact (float64, unrounded) product of every rank, and the
    # reference's error budget on top — one rounding per rank's product, then the roundings of the sum
    # (matmul_scaling_benchmark.py:142-150 reduces rounded C tensors); oracle.reduced_sum_err states the bound.
    def oracle_terms(dt, n, local_b):
        terms = []
        for r in range(ws):
            Ar, Br = reference_port.batch_parallel_inputs(r, ws, n, dt, batch_size=local_b * ws, device="cpu")
            terms.append(oracle.gemm_f64(Ar, Br))
        return terms

    Ab, Bb = reference_port.batch_parallel_inputs(rank, ws, 1024, torch.bfloat16, batch_size=2 * ws, device="cpu")
    Ab, Bb = Ab.to(dev), Bb.to(dev)  # drawn on the CPU so that the oracle sees the very same inputs
    terms = oracle_terms(torch.bfloat16, 1024, 2)
    want = sum(terms)
    shC = comm.shared_like(torch.empty(2, 1024, 1024, dtype=torch.bfloat16, device=dev))
    for _ in range(2):
        comm.bmm_allreduce_fused(Ab, Bb, shC, 2)
    torch.cuda.synchronize()
    res["fused_allreduce"] = oracle.reduced_sum_err(shC.local, terms, 1.0) <= TOL[torch.bfloat16]  # fp32 adds, ONE rounding to nearest
    # the same with the last row band split into one-tile-row bands (B2MM_FINE_TAIL=1): other band geometry, same sums
    coarse = shC.local.clone()
    os.environ["B2MM_FINE_TAIL"] = "1"
    for _ in range(2):
        comm.bmm_allreduce_fused(Ab, Bb, shC, 2)
    torch.cuda.synchronize()
    os.environ.pop("B2MM_FINE_TAIL")
    res["fused_allreduce_fine_tail_same_bits"] = bool(torch.equal(shC.local, coarse))
    Cn = b2.bmm(Ab, Bb)
    comm.all_reduce(Cn)
    torch.cuda.synchronize()
    res["nccl_allreduce_of_products"] = oracle.reduced_sum_err(Cn, terms, max(ws - 1, 1), ring=True) <= TOL[torch.bfloat16]
    res["fused_allreduce_vs_nccl"] = rel(shC.local, Cn.double()) <= 2 * TOL[torch.bfloat16]  # two roundings of one sum
    # float32: the fused path runs ONE batched GEMM, its split scratch must cover the whole local batch
    Af, Bf = reference_port.batch_parallel_inputs(rank, ws, 512, torch.float32, batch_size=2 * ws, device="cpu")
    Af, Bf = Af.to(dev), Bf.to(dev)
    terms32 = oracle_terms(torch.float32, 512, 2)
    shF = comm.shared_like(torch.empty(2, 512, 512, dtype=torch.float32, device=dev))
    for _ in range(2):
        comm.bmm_allreduce_fused(Af, Bf, shF, 1)
    torch.cuda.synchronize()
    res["fused_allreduce_float32"] = oracle.reduced_sum_err(shF.local, terms32, 1.0) <= TOL[torch.float32]
    # the same reductions done inside the NVSwitch (multicast buffers), where the box has NVLS
    res["nvls_available"] = "yes" if comm.nvls_available() else "no"
    if comm.nvls_available():
        for dt in (torch.bfloat16, torch.float32):
            buf = comm.multicast_like(torch.empty(3, 4096, dtype=dt, device=dev))
            buf.local.fill_(float(rank + 1))
            torch.cuda.synchronize()
            dist.barrier()
            comm.all_reduce_nvls(buf)
            torch.cuda.synchronize()
            res[f"nvls_constants_{dt}"] = bool(torch.all(buf.local == float(ws * (ws + 1) // 2)))
            parts = []
            for r in range(ws):
                torch.manual_seed(300 + r)
                parts.append(torch.randn(3, 4096, dtype=dt, device=dev))
            dist.barrier()  # every rank has read the previous result before anybody overwrites its buffer
            buf.local.copy_(parts[rank])
            torch.cuda.synchronize()
            dist.barrier()
            comm.all_reduce_sync(buf)  # dispatches to the NVLS kernel for a multicast buffer
            torch.cuda.synchronize()
            # the switch adds in fp32 and rounds once, but (measured) not always to nearest: allow 1 ulp instead of 1/2
            res[f"nvls_all_reduce_{dt}"] = rel(buf.local, sum(p.double() for p in parts)) <= 2 * TOL[dt]
        mcS = comm.gather_buffer(N, e - s, torch.bfloat16)  # matrix_parallel: one multicast TMA store per chunk
        res["gather_buffer_is_multicast"] = hasattr(mcS, "mc_ptr")
        for _ in range(2):
            comm.matmul_allgather_fused(A, B_local, mcS)
        torch.cuda.synchronize()
        res["fused_allgather_nvls_equals_nccl_path"] = torch.equal(slabs, mcS.local)
        mcC = comm.multicast_like(torch.empty(2, 1024, 1024, dtype=torch.bfloat16, device=dev))
        for _ in range(2):
            comm.bmm_allreduce_fused(Ab, Bb, mcC, 2)
        torch.cuda.synchronize()
        res["fused_allreduce_nvls"] = oracle.reduced_sum_err(mcC.local, terms, 2.0) <= TOL[torch.bfloat16]  # one rounding, by truncation
        res["fused_allreduce_nvls_vs_peer_path"] = rel(mcC.local, shC.local.double()) <= 2 * TOL[torch.bfloat16]
        mcF = comm.multicast_like(torch.empty(2, 512, 512, dtype=torch.float32, device=dev))
        for _ in range(2):
            comm.bmm_allreduce_fused(Af, Bf, mcF, 1)
        torch.cuda.synchronize()
        res["fused_allreduce_nvls_float32"] = oracle.reduced_sum_err(mcF.local, terms32, 2.0) <= TOL[torch.float32]
        comm.release_shared()

    # share -> use -> release -> free, more often than the import table is long: mappings are given back (an owner
    # never frees memory a peer still maps) and a freed-and-reallocated block can be shared again
    ok = True
    for i in range(300):  # the table holds 256 open allocations
        sh = comm.shared_like(torch.empty(6 * 1024 * 1024, dtype=torch.bfloat16, device=dev))  # 12 MiB: its own block
        sh.local.fill_(float(rank + 1))
        torch.cuda.synchronize()
        dist.barrier()
        comm.all_reduce_p2p(sh)
        if i % 50 == 0:
            torch.cuda.synchronize()
            ok = ok and bool(torch.all(sh.local == float(ws * (ws + 1) // 2)))
        del sh
        comm.release_shared()
        torch.cuda.empty_cache()
    res["share_release_cycles"] = ok
    res["launches"] = _lib.launch_count()

    comm.destroy()
    hostlib.cleanup_distributed()
    if rank == 0:
        json.dump(res, open(out_path, "w"))
    else:
        json.dump(res, open(out_path + f".rank{rank}", "w"))


if __name__ == "__main__":
    main(sys.argv[1])
