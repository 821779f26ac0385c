"""The oracle is pinned against the reference itself: tests/golden/ holds outputs of the
reference's own operator call (torch.matmul / torch.bmm on CPU) on inputs drawn with the
reference's own recipes, and values of its own host arithmetic (make_golden.py).  CPU only."""
import glob
import os

import pytest
import torch

from conftest import DTYPES, GOLDEN, TOL, load_golden, rel_err
from oracle import oracle, reference_port


@pytest.mark.parametrize("dname", list(DTYPES))
@pytest.mark.parametrize("stem", ["matmul_independent_{d}_r0", "matmul_independent_{d}_r1", "matmul_longk_{d}"])
def test_oracle_matches_reference_matmul(stem, dname):
    g = load_golden(stem.format(d=dname) + ".npz", dname)
    C = oracle.gemm(g["A"], g["B"])
    assert C.shape == g["C"].shape and C.dtype == g["C"].dtype
    C64 = g["A"].double() @ g["B"].double()
    dt = DTYPES[dname]
    # both the oracle and the reference's output are within tolerance of the exact product ...
    assert rel_err(C, C64) <= TOL[dt]
    assert rel_err(g["C"], C64) <= TOL[dt]
    # ... and of each other: identical except where the fp32-accumulated sum of the reference's
    # backend falls on the other side of a rounding boundary (one output ulp apart; for outputs
    # near zero an ulp is tiny, so the distance is measured against |C| + rms like every error here).
    if dt == torch.float32:
        assert rel_err(C, g["C"].double()) <= 5e-6
    else:
        assert rel_err(C, g["C"].double()) <= 2 * TOL[dt]  # two roundings of one value: at most 1 ulp apart
        assert (C != g["C"]).float().mean().item() < 0.01


@pytest.mark.parametrize("dname", list(DTYPES))
@pytest.mark.parametrize("rank", [0, 1])
def test_oracle_matches_reference_bmm(dname, rank):
    g = load_golden(f"bmm_batch_parallel_{dname}_ws2_r{rank}.npz", dname)
    C = oracle.gemm(g["A"], g["B"])
    assert C.shape == g["C"].shape
    for b in range(C.shape[0]):
        C64 = g["A"][b].double() @ g["B"][b].double()
        assert rel_err(C[b], C64) <= TOL[DTYPES[dname]]
        assert rel_err(g["C"][b], C64) <= TOL[DTYPES[dname]]
        assert rel_err(C[b], g["C"][b].double()) <= 2 * TOL[DTYPES[dname]]


@pytest.mark.parametrize("dname", list(DTYPES))
def test_matrix_parallel_slabs_assemble_to_global_product(dname):
    """Column slabs computed per rank (matmul_scaling_benchmark.py:179-183,:211) tile the global C."""
    g = load_golden(f"matmul_matrix_parallel_{dname}_ws3.npz", dname)
    N, ws = g["A"].shape[0], 3
    pieces = []
    for r in range(ws):
        s, e = oracle.col_split(N, ws, r)
        assert [s, e] == g[f"cols_r{r}"].tolist()
        C_local = oracle.gemm(g["A"], g["B"][:, s:e].contiguous())
        assert rel_err(C_local, g[f"C_local_r{r}"].double()) <= 2 * TOL[DTYPES[dname]]
        pieces.append(C_local)
    C = torch.cat(pieces, dim=1)
    assert C.shape == g["C"].shape
    assert rel_err(C, g["A"].double() @ g["B"].double()) <= TOL[DTYPES[dname]]
    assert e == N  # the last rank took the remainder (136 = 45 + 45 + 46)


def test_reference_recipes_reproduce_golden_inputs():
    """oracle/reference_port.py draws the same tensors the reference drew when the fixtures were made."""
    for dname, dt in DTYPES.items():
        for rank in (0, 1):
            g = load_golden(f"matmul_independent_{dname}_r{rank}.npz", dname)
            A, B = reference_port.independent_inputs(rank, g["A"].shape[0], dt)
            assert torch.equal(A, g["A"]) and torch.equal(B, g["B"])
            g = load_golden(f"bmm_batch_parallel_{dname}_ws2_r{rank}.npz", dname)
            A, B = reference_port.batch_parallel_inputs(rank, 2, g["A"].shape[1], dt)
            assert torch.equal(A, g["A"]) and torch.equal(B, g["B"])
        g = load_golden(f"matmul_matrix_parallel_{dname}_ws3.npz", dname)
        A, B, B_local, (s, e) = reference_port.matrix_parallel_inputs(2, 3, g["A"].shape[0], dt)
        assert torch.equal(A, g["A"]) and torch.equal(B, g["B"]) and [s, e] == g["cols_r2"].tolist()


def test_conversions_match_torch():
    lib = oracle.load()
    torch.manual_seed(3)
    x = torch.cat([torch.randn(4000) * 100, torch.randn(2000) * 1e-6, torch.tensor([0.0, -0.0, 65504.0, 65519.9, 65520.0,
                   1e-8, 6.1e-5, 5.96e-8, 2.98e-8, 2.99e-8, float("inf"), -float("inf"), 1.0039062, 1.00390625])])
    bf = x.to(torch.bfloat16).view(torch.int16).tolist()
    hf = x.to(torch.float16).view(torch.int16).tolist()
    for v, eb, eh in zip(x.tolist(), bf, hf):
        assert lib.oracle_f32_to_bf16(v) == (eb & 0xFFFF), v
        assert lib.oracle_f32_to_fp16(v) == (eh & 0xFFFF), v


def test_block_f64_equals_double_product():
    torch.manual_seed(5)
    A = torch.randn(96, 300, dtype=torch.bfloat16)
    B = torch.randn(300, 120, dtype=torch.bfloat16)
    blk = oracle.gemm_block_f64(A, B, 8, 40, 16, 64)
    ref = (A.double() @ B.double())[8:48, 16:80]
    assert torch.allclose(blk, ref, rtol=0, atol=1e-9)


def test_host_arithmetic_matches_reference(host_logic):
    for n, t, v in host_logic["calculate_tflops_mb"]:
        assert oracle.calculate_tflops(n, t) == pytest.approx(v, rel=1e-12)
        assert reference_port.calculate_tflops(n, t) == pytest.approx(v, rel=1e-12)
    for n, t, k, v in host_logic["calculate_tflops_ms"]:
        assert oracle.calculate_tflops(n, t, k) == pytest.approx(v, rel=1e-12)
        assert reference_port.calculate_tflops(n, t, k) == pytest.approx(v, rel=1e-12)
    for total, ws, local in host_logic["local_batch"]:
        assert oracle.local_batch(total, ws) == local
    for N, ws, rank, s, e in host_logic["col_split"]:
        assert oracle.col_split(N, ws, rank) == (s, e)


def test_reference_cpu_path_port_runs():
    """The CPU branch of benchmark_matmul (matmul_benchmark.py:69-74), restated: returns (avg_s, 2N^3/avg/1e12)."""
    avg, tf = reference_port.benchmark_matmul_cpu(256, torch.float32, num_iterations=2, warmup_iterations=1, seed=0)
    assert avg > 0 and tf == pytest.approx(2.0 * 256 ** 3 / avg / 1e12, rel=1e-12)


def test_empty_and_degenerate_shapes():
    assert oracle.gemm(torch.zeros(0, 8), torch.zeros(8, 4)).shape == (0, 4)
    assert oracle.gemm(torch.zeros(0, 4, 4), torch.zeros(0, 4, 4)).shape == (0, 4, 4)  # batch_parallel at ws=8
    A = torch.randn(3, 5)
    assert torch.equal(oracle.gemm(A, torch.zeros(5, 0)), torch.zeros(3, 0))


def test_all_fixtures_are_covered():
    names = {os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "*.npz"))}
    expect = set()
    for d in DTYPES:
        expect |= {f"matmul_independent_{d}_r0.npz", f"matmul_independent_{d}_r1.npz", f"matmul_longk_{d}.npz",
                   f"bmm_batch_parallel_{d}_ws2_r0.npz", f"bmm_batch_parallel_{d}_ws2_r1.npz",
                   f"matmul_matrix_parallel_{d}_ws3.npz"}
    assert names == expect
