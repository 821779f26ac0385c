"""Parity of the CUDA path with the oracle, through the C ABI, on a B200.

Small and medium shapes are compared element-wise with the oracle (exact products, double sum,
one rounding) and with the committed golden outputs of the reference's own call.  At BASELINE.json's
full size (16384^3) the oracle cannot produce the whole matrix in seconds, so parity is shown through
oracle sub-blocks (double precision, any rows/columns) and size-independent properties."""
import ctypes

import pytest
import torch

from conftest import DTYPES, TOL, load_golden, rel_err
from oracle import oracle, reference_port

pytestmark = pytest.mark.gpu

import pytorch_distributed_matmul_benchmark_b200 as b2  # noqa: E402
from pytorch_distributed_matmul_benchmark_b200 import _lib  # noqa: E402

DEV = "cuda:0"


def _native_loaded():
    assert _lib.load() is not None
    with open("/proc/self/maps") as f:
        assert "libb2mm.so" in f.read(), "the CUDA extension is not the thing that ran"


@pytest.mark.parametrize("dname", list(DTYPES))
@pytest.mark.parametrize("stem", ["matmul_independent_{d}_r0", "matmul_independent_{d}_r1", "matmul_longk_{d}"])
def test_golden_matmul(stem, dname):
    """Same inputs as the reference's recorded torch.matmul call; output within one rounding of it."""
    g = load_golden(stem.format(d=dname) + ".npz", dname)
    before = _lib.launch_count()
    C = b2.matmul(g["A"].to(DEV), g["B"].to(DEV))
    torch.cuda.synchronize()
    assert _lib.launch_count() > before
    dt = DTYPES[dname]
    C64 = g["A"].double() @ g["B"].double()
    assert rel_err(C, C64) <= TOL[dt]
    assert rel_err(C, g["C"].double()) <= 2 * TOL[dt]
    assert rel_err(C, oracle.gemm(g["A"], g["B"]).double()) <= 2 * TOL[dt]
    if dt != torch.float32:  # 16-bit outputs: bit-identical to the oracle except at rounding boundaries
        assert (C.cpu() != oracle.gemm(g["A"], g["B"])).float().mean().item() < 0.01
    _native_loaded()


@pytest.mark.parametrize("dname", list(DTYPES))
@pytest.mark.parametrize("rank", [0, 1])
def test_golden_bmm(dname, rank):
    g = load_golden(f"bmm_batch_parallel_{dname}_ws2_r{rank}.npz", dname)
    C = b2.bmm(g["A"].to(DEV), g["B"].to(DEV))
    for b in range(C.shape[0]):
        assert rel_err(C[b], g["A"][b].double() @ g["B"][b].double()) <= TOL[DTYPES[dname]]
        assert rel_err(C[b], g["C"][b].double()) <= 2 * TOL[DTYPES[dname]]


@pytest.mark.parametrize("dname", list(DTYPES))
def test_golden_matrix_parallel_slabs(dname):
    """Column slabs with the reference's split rule (last rank wider, 136 = 45+45+46) tile the recorded global C."""
    g = load_golden(f"matmul_matrix_parallel_{dname}_ws3.npz", dname)
    A = g["A"].to(DEV)
    pieces = []
    for r in range(3):
        s, e = g[f"cols_r{r}"].tolist()
        Bl = g["B"][:, s:e]
        if (Bl.shape[1] * Bl.element_size()) % 16:  # 45 bf16 columns = 90-byte rows: outside the 16-byte rule
            with pytest.raises(_lib.B2mmError) as ei:
                b2.matmul(A, Bl.contiguous().to(DEV))
            assert ei.value.status == _lib.ERR_UNSUPPORTED
            pad = torch.zeros(Bl.shape[0], (Bl.shape[1] + 7) // 8 * 8, dtype=Bl.dtype)  # padded row pitch, same columns
            pad[:, : Bl.shape[1]] = Bl
            out = torch.empty(A.shape[0], pad.shape[1], dtype=Bl.dtype, device=DEV)
            C_local = b2.matmul(A, pad.to(DEV)[:, : Bl.shape[1]], out=out[:, : Bl.shape[1]])
        else:
            C_local = b2.matmul(A, Bl.contiguous().to(DEV))
        assert rel_err(C_local, g[f"C_local_r{r}"].double()) <= 2 * TOL[DTYPES[dname]]
        pieces.append(C_local.cpu())
    assert rel_err(torch.cat(pieces, dim=1), g["A"].double() @ g["B"].double()) <= TOL[DTYPES[dname]]


SHAPES = [(128, 256, 64), (256, 256, 64), (512, 256, 128), (512, 512, 1024), (1000, 1096, 520), (8, 8, 8), (1, 16, 24),
          (136, 8, 4104), (520, 264, 72), (2048, 1536, 1024)]


@pytest.mark.parametrize("dname", list(DTYPES))
@pytest.mark.parametrize("M,N,K", SHAPES)
def test_oracle_parity_seeded(M, N, K, dname):
    dt = DTYPES[dname]
    torch.manual_seed(M * 131 + N * 17 + K)
    A = torch.randn(M, K, dtype=dt)
    B = torch.randn(K, N, dtype=dt)
    C = b2.matmul(A.to(DEV), B.to(DEV))
    ref = oracle.gemm(A, B)
    C64 = oracle.gemm_block_f64(A, B, 0, M, 0, N)
    assert rel_err(C, C64) <= TOL[dt]
    assert rel_err(C, ref.double()) <= 2 * TOL[dt]


@pytest.mark.parametrize("env", [{}, {"B2MM_M_BLOCKS": "1"}, {"B2MM_M_BLOCKS": "1", "B2MM_BLOCK_N": "128"},
                                 {"B2MM_SCHED": "static"}, {"B2MM_CTA_GROUP": "1"}])
def test_every_kernel_variant_agrees(env, monkeypatch):
    """Tile shape / scheduler variants are different instantiations of one kernel: same bits out."""
    torch.manual_seed(11)
    A = torch.randn(1536, 1160, dtype=torch.bfloat16, device=DEV)
    B = torch.randn(1160, 1288, dtype=torch.bfloat16, device=DEV)
    base = b2.matmul(A, B)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    assert torch.equal(b2.matmul(A, B), base)


@pytest.mark.parametrize("dname", ["bfloat16", "float16"])
def test_last_wave_in_half_tiles_gives_the_same_bits(dname, monkeypatch):
    """Problems several waves long hand out their last tiles as half tiles (GemmParams::num_work): same arithmetic,
    same bits as whole tiles — also with ragged edges, where a lower half may lie entirely outside the matrix —
    and right against the oracle on rows from the split region and from the unsplit one."""
    dt = DTYPES[dname]
    torch.manual_seed(17)
    for M, N, K in ((8192, 4864, 192), (8000, 4872, 328)):  # 16 x 19 = 304 tall tiles >= 4 x 74 clusters
        A = torch.randn(M, K, dtype=dt, device=DEV)
        B = torch.randn(K, N, dtype=dt, device=DEV)
        monkeypatch.setenv("B2MM_SPLIT_TAIL", "1")
        C = torch.full((M, N), 3.0, dtype=dt, device=DEV)
        b2.matmul(A, B, out=C)
        assert "half tiles" in _lib.load().b2mm_last_kernel().decode()
        monkeypatch.setenv("B2MM_SPLIT_TAIL", "0")
        C0 = b2.matmul(A, B)
        assert "half tiles" not in _lib.load().b2mm_last_kernel().decode()
        assert torch.equal(C, C0)
        for r0 in (0, M - 300):
            ref = oracle.gemm_block_f64(A, B, r0, 300, 0, N)
            assert rel_err(C[r0:r0 + 300], ref) <= TOL[dt]


@pytest.mark.parametrize("dname", ["bfloat16", "float16"])
def test_first_wave_in_half_width_items_gives_the_same_bits(dname, monkeypatch):
    """The fused gather hands out its first wave as half-width items (GemmParams::head_tiles) so that its stores start
    earlier; forced here on a plain GEMM (B2MM_SPLIT_HEAD=1 with the 256x256 tiles it applies to): same bits as whole
    tiles, also where the right half of an edge tile lies almost or entirely outside the matrix."""
    dt = DTYPES[dname]
    torch.manual_seed(19)
    monkeypatch.setenv("B2MM_M_BLOCKS", "1")
    monkeypatch.setenv("B2MM_BLOCK_N", "256")
    for M, N, K in ((4096, 2560, 320), (4000, 2440, 136), (4096, 2312, 72)):  # >= 148 tiles of 256 x 256
        A = torch.randn(M, K, dtype=dt, device=DEV)
        B = torch.randn(K, N, dtype=dt, device=DEV)
        monkeypatch.setenv("B2MM_SPLIT_HEAD", "0")
        C0 = b2.matmul(A, B)
        assert "half-width" not in _lib.load().b2mm_last_kernel().decode()
        monkeypatch.setenv("B2MM_SPLIT_HEAD", "1")
        C = torch.full((M, N), 5.0, dtype=dt, device=DEV)
        b2.matmul(A, B, out=C)
        assert "first wave in half-width items" in _lib.load().b2mm_last_kernel().decode()
        assert torch.equal(C, C0)
        assert rel_err(C[:300], oracle.gemm_block_f64(A, B, 0, 300, 0, N)) <= TOL[dt]


def test_fp32_with_64_byte_k_blocks_gives_the_same_bits(monkeypatch):
    """B2MM_FP32_K16=1 runs the 3xTF32 kernel on SWIZZLE_64B tiles (k-blocks of 16): the same sums in the same order."""
    torch.manual_seed(23)
    A = torch.randn(520, 1096, dtype=torch.float32, device=DEV)
    B = torch.randn(1096, 392, dtype=torch.float32, device=DEV)
    base = b2.matmul(A, B)
    monkeypatch.setenv("B2MM_FP32_K16", "1")
    C = b2.matmul(A, B)
    assert "8-stage" in _lib.load().b2mm_last_kernel().decode()
    assert torch.equal(C, base)
    assert rel_err(C, oracle.gemm_block_f64(A, B, 0, 520, 0, 392)) <= TOL[torch.float32]


def test_independent_recipe_on_device_matches_oracle():
    """benchmark_independent's recipe on the GPU generator (matmul_scaling_benchmark.py:73-75), rank 0 and 3."""
    for rank in (0, 3):
        A, B = reference_port.independent_inputs(rank, 1024, torch.bfloat16, device=DEV)
        C = b2.matmul(A, B)
        assert rel_err(C, oracle.gemm_block_f64(A, B, 0, 1024, 0, 1024)) <= TOL[torch.bfloat16]


def test_strided_views_and_out_argument():
    torch.manual_seed(2)
    A = torch.randn(300, 520 + 24, dtype=torch.float16, device=DEV)[:, :520]
    B = torch.randn(520, 256 + 8, dtype=torch.float16, device=DEV)[:, :256]
    out = torch.full((300, 256 + 16), 7.0, dtype=torch.float16, device=DEV)
    b2.matmul(A, B, out=out[:, :256])
    assert rel_err(out[:, :256], A.double().cpu() @ B.double().cpu()) <= TOL[torch.float16]
    assert torch.all(out[:, 256:] == 7.0), "stores must be clipped to the N columns of C"


def test_empty_and_error_cases():
    z = b2.bmm(torch.empty(0, 64, 64, dtype=torch.bfloat16, device=DEV), torch.empty(0, 64, 64, dtype=torch.bfloat16, device=DEV))
    assert z.shape == (0, 64, 64)  # batch_parallel at ws=8 (matmul_scaling_benchmark.py:111)
    assert b2.matmul(torch.empty(0, 16, dtype=torch.float32, device=DEV), torch.empty(16, 8, dtype=torch.float32, device=DEV)).shape == (0, 8)
    with pytest.raises(RuntimeError, match="cannot be multiplied"):
        b2.matmul(torch.zeros(8, 16, device=DEV), torch.zeros(8, 16, device=DEV))
    with pytest.raises(RuntimeError, match="same dtype"):
        b2.matmul(torch.zeros(8, 8, device=DEV), torch.zeros(8, 8, device=DEV, dtype=torch.bfloat16))
    with pytest.raises(TypeError):
        b2.matmul(torch.zeros(8, 8, device=DEV, dtype=torch.float64), torch.zeros(8, 8, device=DEV, dtype=torch.float64))
    with pytest.raises(_lib.B2mmError):  # 12 bf16 = 24-byte rows
        b2.matmul(torch.zeros(8, 12, device=DEV, dtype=torch.bfloat16), torch.zeros(12, 12, device=DEV, dtype=torch.bfloat16))
    lib = _lib.load()
    rc = lib.b2mm_gemm(ctypes.c_void_p(256), ctypes.c_void_p(256), ctypes.c_void_p(256), 64, 64, 64, 64, 64, 64, 1, 0, 0, 0,
                       _lib.FP32, None, 0, None)
    assert rc == _lib.ERR_WORKSPACE and b"workspace" in lib.b2mm_last_error()


@pytest.mark.parametrize("dname", list(DTYPES))
def test_k_that_is_not_a_multiple_of_the_vector_width(dname):
    """K, N, M are free as long as the row PITCH is 16-byte aligned: K = 10 / 1030 inside wider rows.  For fp32
    this goes through the split pre-pass, whose packed rows are padded to 4 floats (TMA zero-fills past K)."""
    dt = DTYPES[dname]
    for M, N, K in ((33, 21, 10), (130, 77, 1030), (5, 3, 1)):
        torch.manual_seed(K)
        Aw = torch.randn(M, K + 16 - K % 8, dtype=dt)  # wide rows: slice on the device to keep the padded pitch
        Bw = torch.randn(K, N + 16 - N % 8, dtype=dt)
        out = torch.full((M, N + 16 - N % 8), 7.0, dtype=dt, device=DEV)
        b2.matmul(Aw.to(DEV)[:, :K], Bw.to(DEV)[:, :N], out=out[:, :N])
        want = oracle.gemm_block_f64(Aw[:, :K].contiguous(), Bw[:, :N].contiguous(), 0, M, 0, N)
        assert rel_err(out[:, :N], want) <= TOL[dt]
        assert torch.all(out[:, N:] == 7.0)


def test_empty_inner_dimension_zero_fills():
    """(M x 0) @ (0 x N) is an empty sum: C = 0, written by the library (cudaMemset2DAsync), padding untouched."""
    for dt in DTYPES.values():
        out = torch.full((40, 24), 7.0, dtype=dt, device=DEV)
        C = b2.matmul(torch.empty(40, 0, dtype=dt, device=DEV), torch.empty(0, 16, dtype=dt, device=DEV), out=out[:, :16])
        assert torch.all(C == 0) and torch.all(out[:, 16:] == 7.0)
    C = b2.bmm(torch.empty(3, 8, 0, dtype=torch.bfloat16, device=DEV), torch.empty(3, 0, 8, dtype=torch.bfloat16, device=DEV))
    assert C.shape == (3, 8, 8) and torch.all(C == 0)


def test_identity_and_linearity_properties():
    """Size-independent properties: A @ I == A bit-exactly; (A1 + A2) @ B == A1 @ B + A2 @ B to rounding."""
    torch.manual_seed(4)
    n = 2048
    A = torch.randn(n, n, dtype=torch.bfloat16, device=DEV)
    eye = torch.eye(n, dtype=torch.bfloat16, device=DEV)
    assert torch.equal(b2.matmul(A, eye), A)
    assert torch.equal(b2.matmul(eye, A), A)
    # powers of two keep every bf16 sum exact, so linearity must hold bit for bit in fp32 accumulation
    A1 = torch.randint(-1, 2, (n, n), device=DEV).to(torch.bfloat16)
    A2 = torch.randint(-1, 2, (n, n), device=DEV).to(torch.bfloat16)
    Bm = torch.randint(-1, 2, (n, n), device=DEV).to(torch.bfloat16)
    P1, P2, S = b2.matmul(A1, Bm), b2.matmul(A2, Bm), b2.matmul(A1 + A2, Bm)
    e1, e2 = A1.double() @ Bm.double(), A2.double() @ Bm.double()
    # integer inputs: every product and partial sum is exact in fp32, so wherever the exact result fits
    # bf16's 8 significant bits the kernel must return it bit for bit, and the three results must add up
    ok = (e1.abs() <= 256) & (e2.abs() <= 256) & ((e1 + e2).abs() <= 256)
    assert ok.float().mean().item() > 0.5
    assert torch.equal(P1.double()[ok], e1[ok]) and torch.equal(P2.double()[ok], e2[ok])
    assert torch.equal(S.double()[ok], (P1.double() + P2.double())[ok])


@pytest.mark.parametrize("dname", ["bfloat16", "float16", "float32"])
def test_full_size_16k_sub_blocks(dname):
    """BASELINE.json config 2 (16384^3): oracle sub-blocks in double precision at the corners, the middle and
    across the 512-row / 256-column tile seams, plus bit-equality of two runs (determinism)."""
    dt = DTYPES[dname]
    N = 16384
    A, B = reference_port.independent_inputs(0, N, dt, device=DEV)
    C = b2.matmul(A, B)
    torch.cuda.synchronize()
    Ac, Bc = A.cpu(), B.cpu()
    worst = 0.0
    for r0, c0 in [(0, 0), (N - 24, N - 40), (500, 240), (8180, 8170), (1016, N - 264), (N - 520, 3)]:
        blk = oracle.gemm_block_f64(Ac, Bc, r0, 24, c0, 40)
        worst = max(worst, rel_err(C[r0:r0 + 24, c0:c0 + 40], blk))
    assert worst <= TOL[dt], worst
    assert torch.equal(b2.matmul(A, B), C)


def test_bmm_matches_per_matrix_matmul_and_oracle():
    torch.manual_seed(9)
    A = torch.randn(3, 520, 1032, dtype=torch.bfloat16, device=DEV)
    B = torch.randn(3, 1032, 776, dtype=torch.bfloat16, device=DEV)
    C = b2.bmm(A, B)
    for b in range(3):
        assert torch.equal(C[b], b2.matmul(A[b], B[b]))
    assert rel_err(C[1], oracle.gemm_block_f64(A[1].cpu(), B[1].cpu(), 0, 520, 0, 776)) <= TOL[torch.bfloat16]


def test_host_gemm_pipelined_transfers_match_device_path():
    """ops.HostGemm (pinned host operands, panel-pipelined copies) gives the bits of the device-resident call."""
    torch.manual_seed(21)
    M, N, K = 1024, 776, 520
    hA = torch.randn(M, K, dtype=torch.bfloat16).pin_memory()
    hB = torch.randn(K, N, dtype=torch.bfloat16).pin_memory()
    hC = torch.empty(M, N, dtype=torch.bfloat16).pin_memory()
    g = b2.HostGemm(M, N, K, torch.bfloat16, device=DEV, panels=4)
    for _ in range(2):  # second call reuses the staging buffers and events
        hC.zero_()
        g(hA, hB, hC)
        torch.cuda.synchronize()
        assert torch.equal(hC, b2.matmul(hA.to(DEV), hB.to(DEV)).cpu())
    assert rel_err(hC, oracle.gemm_block_f64(hA, hB, 0, M, 0, N)) <= TOL[torch.bfloat16]
    with pytest.raises(RuntimeError, match="pinned"):
        g(torch.zeros(M, K, dtype=torch.bfloat16), hB, hC)
    # back-to-back calls with different operands and no synchronisation in between: the calls overlap (the next
    # upload starts while the previous GEMMs / downloads are in flight), every result must still be its own
    sets = []
    for k in range(5):
        a = torch.randn(M, K, dtype=torch.bfloat16).pin_memory()
        b = torch.randn(K, N, dtype=torch.bfloat16).pin_memory()
        sets.append((a, b, torch.zeros(M, N, dtype=torch.bfloat16).pin_memory()))
    for a, b, c in sets:
        g(a, b, c)
    torch.cuda.synchronize()
    for a, b, c in sets:
        assert torch.equal(c, b2.matmul(a.to(DEV), b.to(DEV)).cpu())


def test_programmatic_dependent_launch_keeps_dependencies(monkeypatch):
    """GEMMs are launched so that the next one may become resident while the previous one drains (PDL).  A chain in
    which each GEMM reads the previous one's output, and write-after-write on one C, must give the bits that
    serialised launches (B2MM_PDL=0) give."""
    results = {}
    for pdl in ("0", "1"):
        monkeypatch.setenv("B2MM_PDL", pdl)
        for N, dt in ((1024, torch.bfloat16), (768, torch.float32)):
            torch.manual_seed(N)
            X = torch.randn(N, N, dtype=dt, device=DEV)
            W = torch.randn(N, N, dtype=dt, device=DEV) / N ** 0.5
            bufs = [torch.empty(N, N, dtype=dt, device=DEV) for _ in range(2)]
            cur = X
            for i in range(16):
                b2.matmul(cur, W, out=bufs[i % 2])
                cur = bufs[i % 2]
            C = torch.empty(N, N, dtype=dt, device=DEV)
            for i in range(16):
                b2.matmul(X if i % 2 else W, W, out=C)
            torch.cuda.synchronize()
            results[(pdl, N)] = (cur.clone(), C.clone())
    for N in (1024, 768):
        assert torch.equal(results[("0", N)][0], results[("1", N)][0])
        assert torch.equal(results[("0", N)][1], results[("1", N)][1])
        assert torch.isfinite(results[("1", N)][0]).all()
