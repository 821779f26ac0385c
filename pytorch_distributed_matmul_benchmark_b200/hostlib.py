"""Pieces shared by the two drop-in benchmark scripts.

Everything the reference re-defines in each of its scripts lives here once:
process-group setup (matmul_benchmark.py:9-32, matmul_scaling_benchmark.py:15-61),
the FLOP convention (matmul_benchmark.py:34-37, matmul_scaling_benchmark.py:63-67),
dtype parsing (:169-174 / :366-371) and the device banner (:178-190 / :376-386).

`Ops` is the seam between the host layer and the device library: the product
binds it to libb2mm.so; the CPU tests bind it to stand-ins so that the
multi-rank host logic can run under gloo without a GPU.
"""
from __future__ import annotations

import dataclasses
import os
from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist

DTYPE_MAP = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}
DEFAULT_SIZES = [4096, 8192, 16384]
DEFAULT_ITERATIONS = 50
DEFAULT_WARMUP = 10
TOTAL_BATCH = 4  # matmul_scaling_benchmark.py:283

# Dense tensor-core peaks the efficiency line divides by (the reference hard-codes
# RTX 6000 Ada / RX 7900 XTX numbers, matmul_benchmark.py:131-139).  TFLOP/s.
B200_PEAK_TFLOPS = {torch.bfloat16: 2250.0, torch.float16: 2250.0, torch.float32: 1125.0 / 3.0}
B200_PEAK_NOTE = {torch.bfloat16: "B200", torch.float16: "B200", torch.float32: "B200 (TF32 tensor peak / 3, 3xTF32 split)"}


def dtype_name(dtype: torch.dtype) -> str:
    return "float32" if dtype == torch.float32 else ("float16" if dtype == torch.float16 else "bfloat16")


def bytes_per_element(dtype: torch.dtype) -> int:
    return 4 if dtype == torch.float32 else 2


def calculate_tflops(matrix_size: int, time_seconds: float, num_ops: int = 1) -> float:
    """2 * N^3 * num_ops / t / 1e12 — the FLOP convention of every reported number."""
    return (2.0 * (matrix_size ** 3) * num_ops / time_seconds) / 1e12


def backend_name() -> str:
    """NCCL, as the reference; B2MM_BACKEND=gloo is a test hook for GPU-less hosts."""
    return os.environ.get("B2MM_BACKEND", "nccl")


def setup_distributed(announce_single: bool = True) -> Tuple[int, int]:
    """env:// rendezvous driven by torch.distributed.run, one process per GPU."""
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        rank, world_size = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        backend = backend_name()
        dist.init_process_group(backend=backend)
        if backend == "nccl":
            torch.cuda.set_device(rank % torch.cuda.device_count())
        return rank, world_size
    if announce_single:
        print("Running in single GPU mode")
    return 0, 1


def cleanup_distributed() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


def rank_device(rank: int) -> str:
    if backend_name() != "nccl" and not torch.cuda.is_available():
        return "cpu"
    return f"cuda:{rank % torch.cuda.device_count()}"


def print_banner() -> None:
    """Version / device banner of the reference mains (matmul_benchmark.py:178-190, matmul_scaling_benchmark.py:376-386)."""
    print(f"PyTorch version: {torch.__version__}")
    print(f"CUDA available: {torch.cuda.is_available()}")
    if torch.cuda.is_available():
        print(f"CUDA version: {torch.version.cuda}")
        print(f"Number of CUDA devices: {torch.cuda.device_count()}")
        for i in range(torch.cuda.device_count()):
            props = torch.cuda.get_device_properties(i)
            print(f"  GPU {i}: {torch.cuda.get_device_name(i)}")
            print(f"    Memory: {props.total_memory / (1024**3):.2f} GB")
            print(f"    SMs: {props.multi_processor_count}")


@dataclasses.dataclass
class Ops:
    """Operators the mode bodies call; field names follow the torch calls they replace."""
    matmul: Callable  # (A, B, out=None) -> C
    bmm: Callable  # (A, B, out=None) -> C
    all_reduce: Callable  # (t) -> t, in-place SUM            [dist.all_reduce]
    all_gather_into: Callable  # (slabs, t) -> slabs           [dist.all_gather]
    matmul_allgather_overlap: Optional[Callable] = None  # (A, B_local, C_local, slabs, n_panels)
    bmm_allreduce_overlap: Optional[Callable] = None  # (A, B, C, n_panels)
    shared_slabs: Optional[Callable] = None  # (M, n_local, dtype) -> SharedSlabs
    gather_buffer: Optional[Callable] = None  # (M, n_local, dtype) -> MulticastBuffer (NVLS) or SharedSlabs
    matmul_allgather_fused: Optional[Callable] = None  # (A, B_local, shared) -> slabs
    shared_like: Optional[Callable] = None  # (t) -> SharedSlabs
    reduce_buffer_like: Optional[Callable] = None  # (t) -> MulticastBuffer (NVLS) or SharedSlabs; has .local
    nvls_available: Callable = lambda: False
    bmm_allreduce_fused: Optional[Callable] = None  # (A, B, shared_C, n_panels) -> C
    all_reduce_sync: Optional[Callable] = None  # (shared, stream=None) -> t; peer-memory all_reduce, co-resident with a GEMM
    release_shared: Callable = lambda: None  # collective: unmap peer buffers before their owners free them
    synchronize: Callable = torch.cuda.synchronize
    event: Callable = lambda: torch.cuda.Event(enable_timing=True)
    destroy: Callable = lambda: None
    comm: object = None  # the ops.Communicator behind the collective fields (None on a single GPU / in CPU tests)


def device_ops(rank: int, world_size: int, device: str) -> Ops:
    """Bind Ops to libb2mm.so.  Raises if the library is missing: there is no fallback."""
    from . import ops as b2
    comm = b2.Communicator(rank, world_size, torch.device(device)) if world_size > 1 else None

    def no_comm(*_a, **_k):
        raise RuntimeError("collective called without a process group")

    return Ops(
        matmul=b2.matmul,
        bmm=b2.bmm,
        all_reduce=comm.all_reduce if comm else no_comm,
        all_gather_into=comm.all_gather_into if comm else no_comm,
        matmul_allgather_overlap=comm.matmul_allgather_overlap if comm else None,
        bmm_allreduce_overlap=comm.bmm_allreduce_overlap if comm else None,
        shared_slabs=comm.shared_slabs if comm else None,
        gather_buffer=comm.gather_buffer if comm else None,
        matmul_allgather_fused=comm.matmul_allgather_fused if comm else None,
        shared_like=comm.shared_like if comm else None,
        reduce_buffer_like=comm.reduce_buffer_like if comm else None,
        nvls_available=comm.nvls_available if comm else (lambda: False),
        bmm_allreduce_fused=comm.bmm_allreduce_fused if comm else None,
        all_reduce_sync=comm.all_reduce_sync if comm else None,
        release_shared=comm.release_shared if comm else (lambda: None),
        destroy=comm.destroy if comm else (lambda: None),
        comm=comm,
    )


def metric_all_reduce(value: float, device: str, op: str) -> float:
    """Cross-rank aggregation of a scalar metric (matmul_scaling_benchmark.py:297-306).

    AVG is computed as SUM / world so that the gloo test hook works too (the reference
    makes the same substitution in matmul_benchmark.py:114-118).
    """
    if not dist.is_initialized():
        return value
    t = torch.tensor([value], dtype=torch.float32, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    out = t.item()
    return out / dist.get_world_size() if op == "avg" else out


def split_columns(matrix_size: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Column slab of rank `rank`; the last rank takes the remainder (matmul_scaling_benchmark.py:179-181)."""
    per = matrix_size // world_size
    start = rank * per
    return start, (start + per if rank < world_size - 1 else matrix_size)


def slab_pitch(matrix_size: int, world_size: int) -> int:
    """Row pitch (elements) of every rank's column-slab storage: the widest slab, rounded up to 16 bytes of any dtype.

    Equal to the slab width when matrix_size % (8 * world_size) == 0 (then nothing is padded); otherwise the
    unequal slabs of split_columns travel through equal-size storage, which is what all_gather needs.
    """
    widest = matrix_size - (world_size - 1) * (matrix_size // world_size)
    return (widest + 7) // 8 * 8


def assemble_slabs(slabs: torch.Tensor, matrix_size: int, world_size: int) -> torch.Tensor:
    """[world][N][pitch] gathered storage -> the N x N product (slab r holds the columns of rank r)."""
    parts = []
    for r in range(world_size):
        s, e = split_columns(matrix_size, world_size, r)
        parts.append(slabs[r, :, :e - s])
    return torch.cat(parts, dim=1)


def add_common_args(parser) -> None:
    parser.add_argument("--sizes", type=int, nargs="+", default=list(DEFAULT_SIZES),
                        help="Matrix sizes to benchmark (default: 4096 8192 16384)")
    parser.add_argument("--iterations", type=int, default=DEFAULT_ITERATIONS,
                        help="Number of iterations per benchmark (default: 50)")
    parser.add_argument("--warmup", type=int, default=DEFAULT_WARMUP,
                        help="Number of warmup iterations (default: 10)")
    parser.add_argument("--dtype", type=str, default="bfloat16", choices=["float32", "float16", "bfloat16"],
                        help="Data type for matrices (default: bfloat16)")
    # additive flags (not in the reference)
    parser.add_argument("--check", action="store_true",
                        help="After timing each size, compare the result with torch.matmul on the same tensors")
    parser.add_argument("--json", type=str, default=None, help="Also write per-size results to this JSON file")


# max |C - C64| / (|C64| + rms(C64)), SURVEY.md 8(c): one output ulp for the 16-bit types, 1e-5 for float32
# (the 3xTF32 kernel measures 4e-6 .. 5e-6 at K = 16384; cuBLAS SGEMM 1.4e-5).
CHECK_TOL = {torch.bfloat16: 4e-3, torch.float16: 5e-4, torch.float32: 1e-5}


def report_check(rank: int, err: float, tol: float, what: str, indent: str = "    ", device: Optional[str] = None) -> float:
    """Print one parity line and raise on a miss — on EVERY rank together.

    The decision is collective (MAX of the error over ranks) so that no rank leaves the others inside the next
    collective; `err` and `tol` are printed as measured, never rescaled."""
    if dist.is_initialized():
        on = "cpu"
        if dist.get_backend() == "nccl":
            on = device if device is not None else f"cuda:{torch.cuda.current_device()}"
        t = torch.tensor([err if err == err else float("inf")], dtype=torch.float64, device=on)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        err = t.item()
    if rank == 0:
        print(f"{indent}Parity {what}: max err {err:.3e} (tolerance {tol:.1e}) -> {'OK' if err <= tol else 'FAIL'}")
    if not err <= tol:
        raise RuntimeError(f"parity check failed ({what}): {err:.3e} > {tol:.1e}")
    return err


def parity_error(C: torch.Tensor, A: torch.Tensor, B: torch.Tensor, rows: int = 256) -> float:
    """max |C - C64| / (|C64| + rms(C64)) over the first and last `rows` rows, C64 in float64 on the device."""
    M = A.shape[-2]
    idx = torch.cat([torch.arange(0, min(rows, M)), torch.arange(max(0, M - rows), M)]).unique().to(A.device)
    ref = torch.matmul(A.index_select(-2, idx).double(), B.double())
    got = C.index_select(-2, idx).double()
    rms = ref.pow(2).mean().sqrt()
    return ((got - ref).abs() / (ref.abs() + rms)).max().item()


def reduction_pair_error(got: torch.Tensor, ref: torch.Tensor, term_mag: torch.Tensor, world_size: int,
                         chunk_rows: int = 2048) -> float:
    """Two all_reduce(SUM) results of the same per-rank products, compared in units of the dtype's tolerance.

    `ref` is the NCCL result (one rounding in the switch, or one per ring / tree step — each of a PARTIAL sum, which
    can be far larger than the final, possibly cancelled, sum), `got` the library's own reduction (fp32 adds, one
    rounding).  `term_mag` = sum over ranks of |T_r| + rms(T_r) bounds every partial sum, so two correct results obey
        |got - ref| <= tol * ((world - 1) * term_mag + 2 * (|ref| + rms(ref)))      element by element;
    the returned value is the largest ratio of the left side to the bracket (float64, chunked, on the device).
    A flat bar against |ref| + rms(ref) alone is wrong for world > 2: measured at 4 GPUs, NCCL's tree result sits
    3.4 such "ulps" from the exactly rounded sum where a term is large and the sum is not."""
    g2, r2, m2 = (t.reshape(-1, t.shape[-1]) for t in (got, ref, term_mag))
    sq = 0.0
    for i in range(0, r2.shape[0], chunk_rows):
        sq += r2[i:i + chunk_rows].double().pow(2).sum().item()
    rms = (sq / max(r2.numel(), 1)) ** 0.5
    worst = 0.0
    for i in range(0, r2.shape[0], chunk_rows):
        r = r2[i:i + chunk_rows].double()
        bound = max(world_size - 1, 1) * m2[i:i + chunk_rows].double() + 2.0 * (r.abs() + rms)
        worst = max(worst, ((g2[i:i + chunk_rows].double() - r).abs() / bound).max().item())
    return worst


def term_magnitudes(products) -> torch.Tensor:
    """sum_r (|T_r| + rms(T_r)) in float32 for an iterable of per-rank products (consumed one at a time)."""
    mag = None
    for T in products:
        t = T.float()
        m = t.abs() + t.pow(2).mean().sqrt()
        mag = m if mag is None else mag.add_(m)
    return mag
