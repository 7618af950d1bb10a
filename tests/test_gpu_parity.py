"""GPU parity tests (B200): the CUDA path, called through the C ABI, against the CPU oracle.

Bar (BASELINE.json north_star): activation int4 codes, max_abs, group sums and the int32 accumulators
bit-exact; fp32 outputs within NMSE <= 1e-3 of the oracle (asserted much tighter: <= 1e-9, only the
fp32 association differs).  Nothing here reads /root/reference."""
import os
import numpy as np
import pytest

from conftest import block_perm, make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FP32_NMSE_SPEC = 1e-3      # tolerance stated by north_star
FP32_NMSE_TIGHT = 1e-9     # what we actually hold


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _unpack_ref_blocks(blocks: np.ndarray):
    """block_rrs_int4_tc bytes [M][KT][160] -> codes int8 [M][K], max_abs f32 [M][KT], sum_q int16 [M][KT][8]."""
    M, KT, _ = blocks.shape
    qs = blocks[:, :, :128]
    lo = (qs & 0xF).astype(np.int8)
    hi = (qs >> 4).astype(np.int8)
    lo = np.where(lo >= 8, lo - 16, lo).astype(np.int8)
    hi = np.where(hi >= 8, hi - 16, hi).astype(np.int8)
    codes = np.stack([lo, hi], axis=-1).reshape(M, KT * 256)
    mx = np.ascontiguousarray(blocks[:, :, 128:132]).view(np.float32).reshape(M, KT)
    sq = np.ascontiguousarray(blocks[:, :, 132:148]).view(np.int16).reshape(M, KT, 8)
    return codes, mx, sq


# ------------------------------------------------------------------------------------------------
# K1: activation stage
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,K,use_perm", [(1, 256, False), (3, 2560, False), (7, 1024, True), (130, 512, True),
                                          (512, 2560, False)])
def test_act_quant_bit_exact(t4, orc, M, K, use_perm):
    rng = np.random.default_rng(100 + M + K)
    x = make_acts(rng, M, K)
    perm = block_perm(rng, K) if use_perm else None
    blocks = t4.fwht_quantize_tc(_dev(x), _dev(perm) if use_perm else None).cpu().numpy()
    codes, mx, sq = _unpack_ref_blocks(blocks)
    rc, rmx, rsq = orc.quantize_act(x, perm)
    assert np.array_equal(codes, rc)
    assert np.array_equal(mx.view(np.uint32), rmx.view(np.uint32))
    assert np.array_equal(sq, rsq)
    # and byte-for-byte against the oracle's block_rrs_int4_tc writer
    ob = orc.quantize_act_blocks(x, perm).reshape(M, K // 256, 160)
    assert np.array_equal(blocks[:, :, :148], ob[:, :, :148])


def test_act_quant_edge_cases(t4, orc):
    K = 1024
    x = np.zeros((8, K), dtype=np.float32)
    x[1, :256] = 1e-12
    x[2, 7] = 1000.0
    x[3, :] = np.linspace(-3, 3, K, dtype=np.float32)
    x[4, 300] = np.inf
    x[5, 10] = np.nan
    x[6, :] = 1e30
    x[7, :] = -7.0
    rc, rmx, rsq = orc.quantize_act(x)
    codes, mx, sq = _unpack_ref_blocks(t4.fwht_quantize_tc(_dev(x)).cpu().numpy())
    assert np.array_equal(codes, rc)
    assert np.array_equal(sq, rsq)
    assert np.array_equal(mx, rmx, equal_nan=True)                       # NaN payload bits may differ


def test_act_quant_arbitrary_perm(t4, orc):
    """The reference kernel accepts any index vector (act_row[perm[k]], tcq4_k32.cu:774-775)."""
    rng = np.random.default_rng(5)
    K = 768
    x = make_acts(rng, 4, K)
    perm = rng.permutation(K).astype(np.int32)
    codes, _, _ = _unpack_ref_blocks(t4.fwht_quantize_tc(_dev(x), _dev(perm)).cpu().numpy())
    assert np.array_equal(codes, orc.quantize_act(x, perm)[0])


# ------------------------------------------------------------------------------------------------
# K3: fused decode GEMV
# ------------------------------------------------------------------------------------------------
GEMV_SHAPES = [(1, 8, 256), (1, 64, 512), (2, 256, 1024), (3, 136, 768), (4, 1024, 2560), (1, 2560, 9728),
               (4, 9728, 2560), (8, 1024, 2560), (9, 264, 512), (16, 4096, 2560), (13, 2560, 4096), (7, 16384, 256)]


@pytest.mark.parametrize("M,N,K", GEMV_SHAPES)
@pytest.mark.parametrize("use_perm", [False, True])
def test_gemv_vs_oracle(t4, orc, M, N, K, use_perm):
    rng = np.random.default_rng(200 + M + N + K)
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    perm = block_perm(rng, K) if use_perm else None
    tiles = orc.quantize_tiles(orc.apply_perm_weights(W, perm) if use_perm else W)
    ref = orc.mul_mat(tiles, x, N, perm, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(tiles), _dev(x), N, _dev(perm) if use_perm else None, kernel=t4.KERNEL_GEMV)
    torch.cuda.synchronize()
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])             # int32 accumulators bit-exact
    e = nmse(y.cpu().numpy(), ref["C"])
    assert e <= FP32_NMSE_TIGHT <= FP32_NMSE_SPEC, e


def test_gemv_whole_rows_match_the_gemm_bitwise(t4, orc):
    """KT = 1: every output row is whole inside one warp, and the GEMV's fp32 step per (row, channel, k-tile)
    is the GEMM's (fma(float(isum) * (a7/127), S, acc)), so the two kernels agree bit for bit; against the oracle
    (which divides by 127 per group, tcq4_k32.cu:849) only the last ulp differs."""
    rng = np.random.default_rng(7)
    M, N, K = 2, 16384, 256
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    x = make_acts(rng, M, K)
    ws = t4.Workspace.for_shape(M, N, K)
    y = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMV).clone()
    y2 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM)
    assert torch.equal(y, y2)
    ref = orc.mul_mat(tiles, x, N)["C"]
    assert nmse(y.cpu().numpy(), ref) <= 1e-12


@pytest.mark.parametrize("M", [2, 11])
def test_gemv_zero_points(t4, orc, M):
    """Tiles with non-zero Z/zc: the format carries them even though the quantiser writes 0."""
    rng = np.random.default_rng(8)
    N, K = 64, 512
    t = orc.quantize_tiles(make_weights(rng, N, K)).reshape(-1, 1184).copy()
    t[:, 1040:1056] = np.frombuffer(rng.uniform(-0.01, 0.01, (t.shape[0], 8)).astype(np.float16).tobytes(),
                                    dtype=np.uint8).reshape(t.shape[0], 16)
    t[:, 1120:1184] = rng.integers(-127, 128, (t.shape[0], 64)).astype(np.int8).view(np.uint8)
    t = t.reshape(-1)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(t, x, N)["C"]
    y = t4.mul_mat(_dev(t), _dev(x), N, workspace=t4.Workspace.for_shape(M, N, K), kernel=t4.KERNEL_GEMV)
    assert nmse(y.cpu().numpy(), ref) <= FP32_NMSE_TIGHT


def test_gemv_back_to_back_launches_overlap_safely(t4, orc):
    """The GEMV is launched with programmatic stream serialization: a launch prefetches its weights while the
    previous one is still running and only then waits for it.  A dependent chain y1 = W1 x, y2 = W2 y1', ...
    (each input produced by the previous launch, one shared workspace) must equal the same chain run with a
    device synchronise between launches, eagerly and as one CUDA graph."""
    rng = np.random.default_rng(21)
    dims = [2560, 1024, 2560, 512, 1024, 2560]
    M = 3
    Ws = [_dev(orc.quantize_tiles(make_weights(rng, dims[i + 1], dims[i]))) for i in range(len(dims) - 1)]
    x0 = _dev(make_acts(rng, M, dims[0]))
    ws = t4.Workspace.for_shape(M, max(dims), max(dims))
    bufs = [torch.empty((M, d), device="cuda") for d in dims[1:]]

    def chain(sync):
        cur = x0
        for W, out, n in zip(Ws, bufs, dims[1:]):
            t4.mul_mat(W, cur, n, out=out, workspace=ws, kernel=t4.KERNEL_GEMV)
            if sync:
                torch.cuda.synchronize()
            cur = out
        return [b.clone() for b in bufs]

    want = chain(True)
    for _ in range(3):
        got = chain(False)
        torch.cuda.synchronize()
        for a, b in zip(got, want):
            assert torch.equal(a, b)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            chain(False)
    for b in bufs:
        b.zero_()
    g.replay()
    torch.cuda.synchronize()
    for a, b in zip(bufs, want):
        assert torch.equal(a, b)


def test_gemv_repeatable_and_tickets_self_clean(t4, orc):
    rng = np.random.default_rng(9)
    M, N, K = 1, 1024, 2560
    tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
    x = _dev(make_acts(rng, M, K))
    ws = t4.Workspace.for_shape(M, N, K)
    y0 = t4.mul_mat(tiles, x, N, workspace=ws, kernel=t4.KERNEL_GEMV).clone()
    for _ in range(5):
        y = t4.mul_mat(tiles, x, N, workspace=ws, kernel=t4.KERNEL_GEMV)
        assert torch.equal(y, y0)                                        # deterministic split-K


# ------------------------------------------------------------------------------------------------
# K2: tcgen05 GEMM
# ------------------------------------------------------------------------------------------------
GEMM_SHAPES = [(5, 128, 256), (128, 128, 256), (130, 256, 512), (64, 8, 256), (33, 136, 768), (257, 1024, 1024),
               (512, 9728, 2560)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_vs_oracle(t4, orc, M, N, K):
    rng = np.random.default_rng(300 + M + N + K)
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    big = M * N * K > 2e9
    Mo = 24 if big else M                      # oracle on a row subset at the full size
    rows = np.sort(rng.choice(M, Mo, replace=False)) if big else np.arange(M)
    ref = orc.mul_mat(tiles, x[rows], N, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(tiles), _dev(x), N, kernel=t4.KERNEL_GEMM)
    torch.cuda.synchronize()
    assert np.array_equal(isum.cpu().numpy()[rows], ref["isum"])        # TMEM accumulators bit-exact
    e = nmse(y.cpu().numpy()[rows], ref["C"])
    assert e <= FP32_NMSE_TIGHT <= FP32_NMSE_SPEC, e


GEMM2_SHAPES = [(129, 256, 512), (256, 128, 256), (300, 1024, 1024), (257, 264, 768), (130, 8, 256), (640, 2560, 1024),
                (512, 9728, 2560)]


@pytest.mark.parametrize("M,N,K", GEMM2_SHAPES)
def test_gemm2_pairs_vs_oracle(t4, orc, M, N, K):
    """The CTA-pair kernel (cta_group::2): TMEM accumulators bit-exact, fp32 as tight as the single-CTA kernel."""
    rng = np.random.default_rng(400 + M + N + K)
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    big = M * N * K > 2e9
    Mo = 24 if big else M
    rows = np.sort(rng.choice(M, Mo, replace=False)) if big else np.arange(M)
    ref = orc.mul_mat(tiles, x[rows], N, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(tiles), _dev(x), N, kernel=t4.KERNEL_GEMM2)
    torch.cuda.synchronize()
    assert np.array_equal(isum.cpu().numpy()[rows], ref["isum"])
    e = nmse(y.cpu().numpy()[rows], ref["C"])
    assert e <= FP32_NMSE_TIGHT <= FP32_NMSE_SPEC, e
    # and the non-debug instantiation agrees bit for bit with the single-CTA kernel
    ws = t4.Workspace.for_shape(M, N, K)
    y2 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM2).clone()
    y1 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM)
    assert torch.equal(y1, y2)


GEMM3_SHAPES = GEMM2_SHAPES + [(300, 256, 2048), (512, 2560, 4096), (17, 200, 2304), (1000, 392, 512), (256, 6144, 256)]


@pytest.mark.parametrize("M,N,K", GEMM3_SHAPES)
def test_gemm3_f16_pairs_vs_oracle(t4, orc, M, N, K):
    """The round-2 prefill kernel (cta_group::2 kind::f16 on exact fp16 integers, nibble operands widened in the SM): the
    fp32 TMEM accumulators hold the oracle's int32 sums EXACTLY, fp32 outputs as tight as the kind::i8 kernels'.  Shapes
    with few tiles and K >= 2048 also run the K split (two halves meeting in y through a TMA reduce-add)."""
    rng = np.random.default_rng(900 + M + N + K)
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    big = M * N * K > 2e9
    Mo = 24 if big else M
    rows = np.sort(rng.choice(M, Mo, replace=False)) if big else np.arange(M)
    ref = orc.mul_mat(tiles, x[rows], N, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(tiles), _dev(x), N, kernel=t4.KERNEL_GEMM3)
    torch.cuda.synchronize()
    assert np.array_equal(isum.cpu().numpy()[rows], ref["isum"])
    e = nmse(y.cpu().numpy()[rows], ref["C"])
    assert e <= FP32_NMSE_TIGHT <= FP32_NMSE_SPEC, e
    ws = t4.Workspace.for_shape(M, N, K)
    y3 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM3).clone()
    y3b = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM3).clone()
    assert torch.equal(y3, y3b)                      # deterministic, K split included (two addends commute)
    assert torch.equal(y3, y)                        # debug and plain instantiations agree
    y1 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM)
    if K < 2048:                                     # no K split: the same fp32 operation order as the kind::i8 kernels
        assert torch.equal(y1, y3)
    else:
        assert nmse(y3.cpu().numpy(), y1.cpu().numpy()) <= 1e-12


def test_gemm3_zero_points_perm_negative_scales(t4, orc):
    rng = np.random.default_rng(16)
    M, N, K = 200, 136, 512
    perm = block_perm(rng, K)
    t = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm)).reshape(-1, 1184).copy()
    t[:, 1040:1056] = np.frombuffer(rng.uniform(-0.01, 0.01, (t.shape[0], 8)).astype(np.float16).tobytes(),
                                    dtype=np.uint8).reshape(t.shape[0], 16)
    t[:, 1120:1184] = rng.integers(-127, 128, (t.shape[0], 64)).astype(np.int8).view(np.uint8)
    sc = t[:, 1056:1120].view(np.int8)
    sc[:, ::5] = -sc[:, ::5]
    sc[:, 7] = -127
    sc[:, 9] = 127
    t = t.reshape(-1)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(t, x, N, perm, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(t), _dev(x), N, _dev(perm), kernel=t4.KERNEL_GEMM3)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])
    assert nmse(y.cpu().numpy(), ref["C"]) <= 1e-8


def test_gemm3_extreme_codes_are_exact(t4, orc):
    """every weight nibble -8 or 7 with |sc| = 127 and saturated activation codes: the largest sums the format allows
    (|isum| up to 256 * 7 * 1016) must still be exact in the fp32 accumulator"""
    rng = np.random.default_rng(3)
    M, N, K = 130, 64, 512
    t = orc.quantize_tiles(make_weights(rng, N, K)).reshape(-1, 1184).copy()
    t[:, :1024] = rng.choice(np.array([0x88, 0x77, 0x87, 0x78], dtype=np.uint8), (t.shape[0], 1024))
    t[:, 1056:1120] = rng.choice(np.array([127, -127], dtype=np.int8), (t.shape[0], 64)).view(np.uint8)
    t = t.reshape(-1)
    x = np.where(rng.random((M, K)) < 0.5, -1.0, 1.0).astype(np.float32)      # FWHT of +-1 rows: many saturated codes
    x[::2] = 1.0                                                              # constant rows: one huge coefficient per block
    ref = orc.mul_mat(t, x, N, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(t), _dev(x), N, kernel=t4.KERNEL_GEMM3)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])
    assert nmse(y.cpu().numpy(), ref["C"]) <= FP32_NMSE_TIGHT


def test_gemm2_zero_points_and_perm(t4, orc):
    rng = np.random.default_rng(15)
    M, N, K = 200, 136, 512
    perm = block_perm(rng, K)
    t = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm)).reshape(-1, 1184).copy()
    t[:, 1040:1056] = np.frombuffer(rng.uniform(-0.01, 0.01, (t.shape[0], 8)).astype(np.float16).tobytes(),
                                    dtype=np.uint8).reshape(t.shape[0], 16)
    t[:, 1120:1184] = rng.integers(-127, 128, (t.shape[0], 64)).astype(np.int8).view(np.uint8)
    sc = t[:, 1056:1120].view(np.int8)
    sc[:, ::5] = -sc[:, ::5]
    t = t.reshape(-1)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(t, x, N, perm)["C"]
    y = t4.mul_mat(_dev(t), _dev(x), N, _dev(perm), workspace=t4.Workspace.for_shape(M, N, K), kernel=t4.KERNEL_GEMM2)
    assert nmse(y.cpu().numpy(), ref) <= 1e-8


def test_gemm_with_perm_and_negative_scale_codes(t4, orc):
    """perm path + sc < 0 (int8 allows it; the quantiser never writes it) exercise the unpack bias math."""
    rng = np.random.default_rng(11)
    M, N, K = 40, 128, 512
    perm = block_perm(rng, K)
    t = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm)).reshape(-1, 1184).copy()
    sc = t[:, 1056:1120].view(np.int8)
    sc[:, ::3] = -sc[:, ::3]
    sc[:, 5] = -127
    t = t.reshape(-1)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(t, x, N, perm, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(t), _dev(x), N, _dev(perm), kernel=t4.KERNEL_GEMM)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])
    assert nmse(y.cpu().numpy(), ref["C"]) <= FP32_NMSE_TIGHT


def test_gemm_zero_points(t4, orc):
    rng = np.random.default_rng(12)
    M, N, K = 20, 128, 512
    t = orc.quantize_tiles(make_weights(rng, N, K)).reshape(-1, 1184).copy()
    t[:, 1040:1056] = np.frombuffer(rng.uniform(-0.01, 0.01, (t.shape[0], 8)).astype(np.float16).tobytes(),
                                    dtype=np.uint8).reshape(t.shape[0], 16)
    t[:, 1120:1184] = rng.integers(-127, 128, (t.shape[0], 64)).astype(np.int8).view(np.uint8)
    t = t.reshape(-1)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(t, x, N)["C"]
    y = t4.mul_mat(_dev(t), _dev(x), N, workspace=t4.Workspace.for_shape(M, N, K), kernel=t4.KERNEL_GEMM)
    assert nmse(y.cpu().numpy(), ref) <= 1e-8


def test_gemm_equals_gemv_on_small_m(t4, orc):
    """Both kernels form the same int32 accumulators; fp32 differs only by association."""
    rng = np.random.default_rng(13)
    M, N, K = 3, 512, 1024
    tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
    x = _dev(make_acts(rng, M, K))
    y1, i1 = t4.mul_mat_debug(tiles, x, N, kernel=t4.KERNEL_GEMV)
    y2, i2 = t4.mul_mat_debug(tiles, x, N, kernel=t4.KERNEL_GEMM)
    assert torch.equal(i1, i2)
    assert nmse(y2.cpu().numpy(), y1.cpu().numpy()) <= FP32_NMSE_TIGHT


# ------------------------------------------------------------------------------------------------
# size-independent properties at BASELINE.json's full sizes (synthetic tiles, no oracle run)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K", [(1, 9728, 2560), (512, 9728, 2560), (64, 14336, 4096)])
def test_full_size_properties(t4, M, N, K):
    tiles = t4.random_tiles(N, K, seed=3)
    g = torch.Generator(device="cuda")
    g.manual_seed(4)
    x = torch.randn((M, K), device="cuda", generator=g)
    ws = t4.Workspace.for_shape(M, N, K)
    y = t4.mul_mat(tiles, x, N, workspace=ws).clone()
    assert torch.isfinite(y).all()
    # (1) power-of-two scaling of x scales y exactly (codes unchanged, a7 scales exactly)
    y4 = t4.mul_mat(tiles, x * 4.0, N, workspace=ws)
    assert torch.equal(y4, y * 4.0)
    # (2) rows are independent: a row computed alone through the GEMV has the same integer
    #     accumulators, so fp32 agrees to association error
    r = M // 2
    y1 = t4.mul_mat(tiles, x[r:r + 1].contiguous(), N, workspace=ws, kernel=t4.KERNEL_GEMV)
    assert nmse(y1.cpu().numpy(), y[r:r + 1].cpu().numpy()) <= FP32_NMSE_TIGHT
    # (3) N-sharding: concatenated shard outputs are bit-identical to the full output
    parts, kernels = [], {t4.select_kernel(M, N, K)}
    for rank in range(4):
        sw = t4.ShardedWeight(N, K, 4, rank, tiles_full=tiles)
        parts.append(sw.local(x, workspace=ws).clone())
        kernels.add(t4.select_kernel(M, sw.n_local, K))
    ycat = torch.cat(parts, dim=1)
    if kernels == {t4.KERNEL_GEMM}:                  # one kernel family, no K split: the same fp32 operation order
        assert torch.equal(ycat, y)
    else:
        assert nmse(ycat.cpu().numpy(), y.cpu().numpy()) <= FP32_NMSE_TIGHT


# ------------------------------------------------------------------------------------------------
# boundary: registry, ggml mirror, host entry, capture safety
# ------------------------------------------------------------------------------------------------
def test_registry_and_ggml_mirror(t4, orc):
    rng = np.random.default_rng(14)
    M, N, K = 2, 64, 512
    perm = block_perm(rng, K)
    tiles = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm))
    x = make_acts(rng, M, K)
    t4.clear_perms()
    t4.set_reorder_enabled(True)
    name = "blk.0.ffn_up.weight"
    assert not t4.has_perm(name)
    t4.register_perm(name, perm)
    assert t4.has_perm(name) and t4.get_perm_ptr(name) != 0
    y = t4.ggml_mul_mat(_dev(tiles), _dev(x), N, name=name)
    assert nmse(y.cpu().numpy(), orc.mul_mat(tiles, x, N, perm)["C"]) <= FP32_NMSE_TIGHT
    t4.set_reorder_enabled(False)                       # rrs.cu:102: lookups return nullptr when disabled
    assert not t4.has_perm(name)
    y2 = t4.ggml_mul_mat(_dev(tiles), _dev(x), N, name=name)
    assert nmse(y2.cpu().numpy(), orc.mul_mat(tiles, x, N, None)["C"]) <= FP32_NMSE_TIGHT
    t4.set_reorder_enabled(True)
    t4.register_perm(name, np.arange(K, dtype=np.int32))   # re-registration replaces (rrs.cu:88-90)
    y3 = t4.ggml_mul_mat(_dev(tiles), _dev(x), N, name=name)
    assert torch.equal(y3, y2)
    t4.clear_perms()
    assert not t4.has_perm(name)


def test_error_behaviour(t4):
    x = torch.zeros((1, 300), device="cuda")
    w = torch.zeros(1184, dtype=torch.uint8, device="cuda")
    with pytest.raises((t4.Tcq4Error, ValueError)):
        t4.mul_mat(w, x, 8)
    with pytest.raises(t4.Tcq4Error) as ei:
        t4.mul_mat(w, torch.zeros((1, 256), device="cuda"), 8, workspace=t4.Workspace(256))
    assert ei.value.status == 3


def test_host_entry_matches_device_entry(t4, orc):
    rng = np.random.default_rng(15)
    for M in (1, 37):
        N, K = 256, 512
        tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
        xh = torch.from_numpy(make_acts(rng, M, K)).pin_memory()
        yh = torch.empty((M, N), dtype=torch.float32).pin_memory()
        t4.mul_mat_host(tiles, xh, yh, N)
        yd = t4.mul_mat(tiles, xh.cuda(), N)
        assert torch.equal(yh, yd.cpu())


def test_pipelined_host_entry_matches_device_entry(t4, orc):
    """Queued host-buffer matmuls (copies and kernels of different calls overlap) give the device entry's bits,
    including when slots are reused with different shapes."""
    rng = np.random.default_rng(151)
    cases = []
    for i, (M, N, K) in enumerate([(3, 256, 512), (150, 512, 256), (1, 1024, 512), (300, 256, 1024), (17, 512, 512),
                                   (260, 1024, 256)]):
        tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
        xh = torch.from_numpy(make_acts(rng, M, K)).pin_memory()
        yh = torch.full((M, N), float("nan"), dtype=torch.float32).pin_memory()
        cases.append((tiles, xh, yh, N))
    for rep in range(2):
        for tiles, xh, yh, N in cases:
            t4.mul_mat_host_async(tiles, xh, yh, N)
        t4.host_sync()
        for tiles, xh, yh, N in cases:
            assert torch.equal(yh, t4.mul_mat(tiles, xh.cuda(), N).cpu())


@pytest.mark.parametrize("M", [1, 12, 40, 300])
def test_pipelined_host_fanout_matches_device_fanout(t4, orc, M):
    """tcq4_b200_mul_mat_fanout_host_async: a group of matrices that share x, x copied once, interleaved with single-matrix
    host calls on the same two-deep pipeline; results equal the device fan-out's bits and the oracle."""
    rng = np.random.default_rng(152 + M)
    K, Ns = 512, [512, 128, 128]
    tl = [_dev(orc.quantize_tiles(make_weights(rng, n, K))) for n in Ns]
    xh = torch.from_numpy(make_acts(rng, M, K)).pin_memory()
    yhs = [torch.full((M, n), float("nan"), dtype=torch.float32).pin_memory() for n in Ns]
    single = torch.full((M, Ns[0]), float("nan"), dtype=torch.float32).pin_memory()
    for rep in range(3):
        t4.mul_mat_fanout_host_async(tl, xh, yhs, Ns)
        t4.mul_mat_host_async(tl[0], xh, single, Ns[0])
    t4.host_sync()
    want = t4.Fanout(tl, Ns, K)(xh.cuda())
    torch.cuda.synchronize()
    for y, w, t, n in zip(yhs, want, tl, Ns):
        assert torch.equal(y, w.cpu())
        assert nmse(y.numpy(), orc.mul_mat(t.cpu().numpy(), xh.numpy(), n)["C"]) <= FP32_NMSE_TIGHT
    assert nmse(single.numpy(), yhs[0].numpy()) <= 1e-12


@pytest.mark.parametrize("M", [1, 3, 16, 64, 300])
def test_peer_destinations_strided_window(t4, orc, M):
    """tcq4_b200_mul_mat_peers: the shard's columns land at column col0 of EVERY destination matrix (here two
    buffers on the same GPU stand in for two GPUs' symmetric buffers), nothing else is touched, bits equal the
    plain entry's.  Covers the GEMV (M <= 16) and the pair GEMM (which also takes over 17 <= M <= 128)."""
    rng = np.random.default_rng(300 + M)
    N, K, ldc, col0 = 256, 512, 1024, 512
    tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
    x = _dev(make_acts(rng, M, K))
    want = t4.mul_mat(tiles, x, N)
    dst = [torch.full((M, ldc), -7.0, device="cuda") for _ in range(2)]
    t4.mul_mat_peers(tiles, x, [d.data_ptr() for d in dst], ldc, col0, N)
    torch.cuda.synchronize()
    for d in dst:
        assert torch.equal(d[:, col0:col0 + N], want)
        assert bool((d[:, :col0] == -7.0).all()) and bool((d[:, col0 + N:] == -7.0).all())


def test_cuda_graph_capture(t4, orc):
    """Launches are capture-safe: no allocation / sync on the path (ggml-cuda.cu:3791-3799)."""
    rng = np.random.default_rng(16)
    N, K = 512, 1024
    tiles = _dev(orc.quantize_tiles(make_weights(rng, N, K)))
    for M in (1, 48):
        x = _dev(make_acts(rng, M, K))
        ws = t4.Workspace.for_shape(M, N, K)
        out = torch.empty((M, N), device="cuda")
        eager = t4.mul_mat(tiles, x, N, workspace=ws).clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                t4.mul_mat(tiles, x, N, out=out, workspace=ws)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)


@pytest.mark.gpu
def test_gemm2_fused_activation_stage_in_a_child_process():
    """TCQ4_G2_FUSE=1 (read once at load, hence the child): the pair GEMM quantises the activations on its own warps behind
    a grid-wide counter; accumulators must stay bit-exact and the counters must re-arm themselves launch after launch."""
    import subprocess
    import sys
    code = r'''
import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import __graft_entry__ as graft
from conftest import make_acts, make_weights, block_perm
t4 = graft.load_package(); orc = graft.load_oracle().Oracle()
rng = np.random.default_rng(11)
for (M, N, K, use_perm) in [(300, 256, 1024, False), (129, 136, 512, True), (512, 1024, 2560, False)]:
    tiles = orc.quantize_tiles(make_weights(rng, N, K)); x = make_acts(rng, M, K)
    perm = block_perm(rng, K) if use_perm else None
    ref = orc.mul_mat(tiles, x, N, perm=perm, want_isum=True)
    ws = t4.Workspace.for_shape(M, N, K)
    td, xd = torch.from_numpy(tiles).cuda(), torch.from_numpy(x).cuda()
    pd = torch.from_numpy(perm).cuda() if use_perm else None
    for rep in range(3):
        y, isum = t4.mul_mat_debug(td, xd, N, perm=pd, kernel=t4.KERNEL_GEMM2, workspace=ws)
        torch.cuda.synchronize()
        assert np.array_equal(isum.cpu().numpy(), ref["isum"]), (M, N, K, rep)
        yy = y.cpu().numpy(); assert ((yy - ref["C"]) ** 2).sum() / (ref["C"] ** 2).sum() < 1e-9
print("FUSED_OK")
'''
    env = dict(os.environ, TCQ4_G2_FUSE="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=240)
    assert "FUSED_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
