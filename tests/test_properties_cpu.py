"""Property tests (hypothesis) of the host-side sharding arithmetic and of the oracle.  CPU only."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from oracle import oracle
from pytorch_distributed_matmul_benchmark_b200 import hostlib

sizes = st.integers(min_value=1, max_value=5000)
worlds = st.integers(min_value=1, max_value=8)


@settings(max_examples=200, deadline=None)
@given(N=sizes, ws=worlds)
def test_column_split_partitions_the_columns(N, ws):
    """matmul_scaling_benchmark.py:179-181: equal slabs of N // ws, the last rank takes the remainder."""
    edges = [hostlib.split_columns(N, ws, r) for r in range(ws)]
    assert edges[0][0] == 0 and edges[-1][1] == N
    for (s0, e0), (s1, _) in zip(edges, edges[1:]):
        assert e0 == s1 and e0 - s0 == N // ws
    assert edges[-1][1] - edges[-1][0] == N - (ws - 1) * (N // ws)
    # the C restatement in the oracle agrees
    for r in range(ws):
        assert oracle.col_split(N, ws, r) == edges[r]


@settings(max_examples=200, deadline=None)
@given(N=sizes, ws=worlds)
def test_slab_pitch_holds_every_slab_at_a_16_byte_row_pitch(N, ws):
    pitch = hostlib.slab_pitch(N, ws)
    widths = [e - s for s, e in (hostlib.split_columns(N, ws, r) for r in range(ws))]
    assert pitch % 8 == 0 and max(widths) <= pitch < max(widths) + 8
    if N % (8 * ws) == 0:
        assert pitch == N // ws  # nothing padded: the plain contiguous [ws][N][N/ws] layout


@settings(max_examples=40, deadline=None)
@given(N=st.integers(min_value=1, max_value=96), ws=worlds, seed=st.integers(0, 2 ** 16))
def test_assemble_slabs_inverts_the_column_split(N, ws, seed):
    g = torch.Generator().manual_seed(seed)
    full = torch.randn(N, N, generator=g)
    pitch = hostlib.slab_pitch(N, ws)
    slabs = torch.full((ws, N, pitch), float("nan"))
    for r in range(ws):
        s, e = hostlib.split_columns(N, ws, r)
        slabs[r, :, :e - s] = full[:, s:e]
    assert torch.equal(hostlib.assemble_slabs(slabs, N, ws), full)


@settings(max_examples=100, deadline=None)
@given(N=st.integers(1, 20000), t=st.floats(1e-6, 10.0), k=st.integers(0, 8))
def test_calculate_tflops_convention(N, t, k):
    """2 * N^3 * num_ops / t / 1e12 (matmul_scaling_benchmark.py:63-67), in the host layer and in the oracle."""
    want = 2.0 * N ** 3 * k / t / 1e12
    assert hostlib.calculate_tflops(N, t, k) == want
    assert abs(oracle.calculate_tflops(N, t, k) - want) <= 1e-12 * max(want, 1e-300)


@settings(max_examples=30, deadline=None)
@given(M=st.integers(1, 24), N=st.integers(1, 24), K=st.integers(1, 200), seed=st.integers(0, 2 ** 16))
def test_oracle_is_exact_products_double_sum_one_rounding(M, N, K, seed):
    """The oracle's contract, against numpy: float64 accumulation of exact products, rounded once to the output type.
    (Summation order in float64 may differ by ~1e-16 relative, so allow the last ulp of the output type.)"""
    g = torch.Generator().manual_seed(seed)
    for dt, eps in ((torch.float32, 2.0 ** -23), (torch.bfloat16, 2.0 ** -7), (torch.float16, 2.0 ** -10)):
        A = torch.randn(M, K, generator=g).to(dt)
        B = torch.randn(K, N, generator=g).to(dt)
        ref64 = A.double().numpy() @ B.double().numpy()
        got = oracle.gemm(A, B).double().numpy()
        want = torch.from_numpy(ref64).to(dt).double().numpy()
        assert np.all(np.abs(got - want) <= eps * np.maximum(np.abs(want), 2.0 ** -14))
        # and the float64 block the GPU tests compare against is that same product
        blk = oracle.gemm_block_f64(A, B, 0, M, 0, N).numpy()
        assert np.allclose(blk, ref64, rtol=1e-13, atol=1e-13)


def test_two_correct_reductions_pass_the_pair_bound_and_a_wrong_one_does_not():
    """hostlib.reduction_pair_error: a tree of bf16 adds (what NCCL may do) and an fp32 sum rounded once (what the
    library's own reductions do) are both correct and must agree within the tolerance, for every world size, although
    a flat bar against the final sum does not hold; a result that lost one rank's contribution must not pass."""
    import torch
    from pytorch_distributed_matmul_benchmark_b200 import hostlib
    tol = hostlib.CHECK_TOL[torch.bfloat16]
    torch.manual_seed(5)
    for world in (2, 4, 8):
        terms = [(torch.randn(256, 384) * 128).to(torch.bfloat16) for _ in range(world)]
        level = list(terms)
        while len(level) > 1:  # pairwise tree, every partial sum rounded to bf16
            level = [level[i] + level[i + 1] for i in range(0, len(level), 2)]
        tree = level[0]
        once = sum(t.float() for t in terms).to(torch.bfloat16)
        mag = hostlib.term_magnitudes(iter(terms))
        assert mag.dtype == torch.float32 and mag.shape == tree.shape
        assert hostlib.reduction_pair_error(once, tree, mag, world) <= tol
        lost = (sum(t.float() for t in terms[1:])).to(torch.bfloat16)  # rank 0's product never arrived
        assert hostlib.reduction_pair_error(lost, tree, mag, world) > tol
