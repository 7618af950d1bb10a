"""the device weight quantiser (quantize_tcq4_tile, ggml-quants.c:1528-1639, with
apply_channel_perm_to_weights fused), derive_tcq4_perms_from_imatrix (quantize.cpp:127-165) and GET_ROWS for TCQ4
tensors (getrows.cu:96-171 + 348-380).  Bar: byte-identical tiles, bit-identical rows."""
import numpy as np
import pytest

from conftest import block_perm, make_weights


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ---------------------------------------------------------------------------------------------------------------
# CPU: the oracle's GET_ROWS against its own pinned pieces, and the host-side perm derivation of the C ABI
# ---------------------------------------------------------------------------------------------------------------
def test_oracle_get_rows_is_dequant_then_fwht(orc):
    rng = np.random.default_rng(5)
    N, K = 24, 768
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    Wh = orc.dequant_matrix(tiles, N, K)                        # Hadamard-domain rows (pinned to the reference dequantiser)
    rows = np.array([0, 7, 8, 23, 7], dtype=np.int32)
    got = orc.get_rows(tiles, rows, K)
    want = np.stack([np.concatenate([orc.fwht(Wh[r, b * 256:(b + 1) * 256].copy()) for b in range(K // 256)]) for r in rows])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_oracle_get_rows_returns_the_weights(orc):
    """FWHT is its own inverse: GET_ROWS of quantised W is W up to the int4 quantisation error."""
    rng = np.random.default_rng(6)
    N, K = 16, 512
    W = make_weights(rng, N, K)
    got = orc.get_rows(orc.quantize_tiles(W), np.arange(N, dtype=np.int32), K)
    assert float(((got - W) ** 2).sum() / (W ** 2).sum()) < 2e-2


def test_ref_quantiser_pins_the_tiles_get_rows_reads(orc, ref):
    """the tile bytes GET_ROWS consumes are the reference quantiser's (memcmp), so the row values are pinned too"""
    rng = np.random.default_rng(8)
    W = make_weights(rng, 16, 512)
    assert np.array_equal(orc.quantize_tiles(W), ref.quantize_tiles(W))


def test_perm_from_imatrix_host(orc):
    t4 = pytest.importorskip("__graft_entry__").load_package()
    rng = np.random.default_rng(9)
    for K in (256, 2560, 4096):
        im = np.exp(rng.standard_normal(K)).astype(np.float32)         # exp(N(0,1))
        p = t4.perm_from_imatrix(im)
        assert np.array_equal(p, orc.perm_from_imatrix(im))
        blocks = p.reshape(K // 256, 256)
        assert np.array_equal(np.sort(blocks, axis=1), np.arange(K, dtype=np.int32).reshape(K // 256, 256))
        assert (np.diff(im[blocks], axis=1) <= 0).all()                # descending inside every block
    im = np.ones(300, dtype=np.float32)                                # ties keep index order; partial last block
    assert np.array_equal(t4.perm_from_imatrix(im), np.arange(300, dtype=np.int32))


def test_quantize_size_and_argument_checks():
    t4 = pytest.importorskip("__graft_entry__").load_package()
    assert t4.quantize_size(8, 256) == 1184
    assert t4.quantize_size(13, 512) == 2 * 2 * 1184
    assert t4.quantize_size(8, 100) == 0
    assert t4.LIB.tcq4_b200_quantize_tiles(None, None, 8, 100, None, None) == 1     # TCQ4_ERR_SHAPE before any CUDA call
    assert t4.LIB.tcq4_b200_get_rows(None, None, None, 1, 8, 100, 100, None, None) == 1


# ---------------------------------------------------------------------------------------------------------------
# GPU parity through the C ABI
# ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def t4():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import __graft_entry__ as graft
    return graft.load_package()


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,use_perm", [(8, 256, False), (64, 512, False), (13, 768, False), (40, 2560, True),
                                          (256, 1024, True)])
def test_quantiser_bytes_match_the_oracle(t4, orc, N, K, use_perm):
    rng = np.random.default_rng(N * 7 + K)
    W = make_weights(rng, N, K)
    W[:, ::53] *= 30.0                                                  # outlier columns
    if N > 8:
        W[3] = 0.0                                                      # an all-zero channel: scale < 1e-10 -> 1
        W[5, 256:] = 1e-12
    perm = block_perm(rng, K) if use_perm else None
    want = orc.quantize_tiles(orc.apply_perm_weights(W, perm) if use_perm else W)
    got = t4.quantize_tiles(_dev(W), _dev(perm) if use_perm else None).cpu().numpy()
    assert got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_quantiser_rounding_ties_and_clamps(t4, orc):
    """values that land exactly on .5 after scaling, +-absmax (q = +-7), and the -8 code"""
    K = 256
    W = np.zeros((8, K), dtype=np.float32)
    rng = np.random.default_rng(3)
    base = rng.integers(-7, 8, size=(8, K)).astype(np.float32)
    W[:] = base * 0.125
    W[0, :32] = np.linspace(-7, 7, 32, dtype=np.float32) * 0.5
    got = t4.quantize_tiles(_dev(W)).cpu().numpy()
    assert np.array_equal(got, orc.quantize_tiles(W))


@pytest.mark.gpu
def test_quantiser_and_get_rows_match_the_golden_fixtures(t4, orc):
    """the committed golden vectors are outputs of the reference's own compiled quantiser / dequantiser
    (tests/golden/make_golden.py): device tiles must equal them byte for byte, and GET_ROWS of the golden tiles must equal
    FWHT256 of the golden dequantised rows bit for bit"""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    from make_golden import CASES, make_w
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "tcq4_golden.npz"))
    for name, N, K, seed, std, special in CASES:
        W = make_w(N, K, seed, std, special)
        tiles = gold[f"{name}.tiles"]
        got = t4.quantize_tiles(_dev(W)).cpu().numpy()
        assert np.array_equal(got, tiles), name
        NP = (N + 7) // 8 * 8
        Wh = gold[f"{name}.dequant"].reshape(NP // 8, K // 256, 8, 256).transpose(0, 2, 1, 3).reshape(NP, K)
        want = np.stack([np.concatenate([orc.fwht(Wh[r, b * 256:(b + 1) * 256].copy()) for b in range(K // 256)])
                         for r in range(NP)])
        rows = np.arange(NP, dtype=np.int32)
        back = t4.get_rows(_dev(tiles), _dev(rows), NP, K).cpu().numpy()
        assert np.array_equal(back.view(np.uint32), want.view(np.uint32)), name


@pytest.mark.gpu
def test_quantised_on_device_then_mul_mat(t4, orc):
    """tiles produced on the device drive the MUL_MAT: same int32 accumulators as the oracle fed by ITS quantiser"""
    rng = np.random.default_rng(12)
    M, N, K = 3, 128, 1024
    W = make_weights(rng, N, K)
    x = rng.standard_normal((M, K)).astype(np.float32)
    tiles_d = t4.quantize_tiles(_dev(W))
    y, isum = t4.mul_mat_debug(tiles_d, _dev(x), N)
    ref = orc.mul_mat(orc.quantize_tiles(W), x, N, want_isum=True)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])


@pytest.mark.gpu
def test_quantiser_full_size_checksum(t4, orc):
    """BASELINE config 1 shape through both quantisers: one FNV-1a over 14.4 MB of tiles"""
    import torch
    N, K = 9728, 2560
    g = torch.Generator(device="cuda")
    g.manual_seed(1234)
    W = torch.randn((N, K), device="cuda", generator=g) * 0.02
    got = t4.quantize_tiles(W).cpu().numpy()
    want = orc.quantize_tiles(W.cpu().numpy())
    assert got.size == 14397440
    assert orc.fnv1a64(got) == orc.fnv1a64(want)


@pytest.mark.gpu
@pytest.mark.parametrize("N,K", [(16, 256), (64, 1024), (24, 2560)])
def test_get_rows_bits_match_the_oracle(t4, orc, N, K):
    import torch
    rng = np.random.default_rng(N + K)
    tiles = orc.quantize_tiles(make_weights(rng, N, K)).copy()
    # zero points too (never produced by the fork's quantiser, but part of the format)
    t = tiles.reshape(-1, 1184)
    t[:, 1040:1056] = np.frombuffer(np.full(8, 0.01, dtype=np.float16).tobytes(), dtype=np.uint8)
    t[:, 1120:1184] = rng.integers(-127, 128, size=(t.shape[0], 64)).astype(np.int8).view(np.uint8)
    rows = rng.integers(0, N, size=37).astype(np.int32)
    rows[:3] = [0, N - 1, N - 1]
    want = orc.get_rows(tiles, rows, K)
    got = t4.get_rows(_dev(tiles), _dev(rows), N, K).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    # strided destination, nothing else touched
    buf = torch.full((rows.size, K + 8), -1.0, device="cuda")
    t4.get_rows(_dev(tiles), _dev(rows), N, K, out=buf)
    assert np.array_equal(buf[:, :K].cpu().numpy().view(np.uint32), want.view(np.uint32))
    assert bool((buf[:, K:] == -1.0).all())


@pytest.mark.gpu
def test_get_rows_bad_index_is_flagged_not_read(t4, orc):
    import torch
    rng = np.random.default_rng(2)
    N, K = 16, 512
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    rows = np.array([1, 16, -1, 2], dtype=np.int32)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    got = t4.get_rows(_dev(tiles), _dev(rows), N, K, err_flag=flag).cpu().numpy()
    assert int(flag.item()) == 1
    assert not got[1].any() and not got[2].any()
    want = orc.get_rows(tiles, np.array([1, 2], dtype=np.int32), K)
    assert np.array_equal(got[[0, 3]].view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_get_rows_of_device_quantised_embedding_roundtrip(t4):
    """size-independent property at a full-size shape: quantise on the device, GET_ROWS every row, compare with W"""
    import torch
    N, K = 4096, 2560
    g = torch.Generator(device="cuda")
    g.manual_seed(77)
    W = torch.randn((N, K), device="cuda", generator=g) * 0.02
    tiles = t4.quantize_tiles(W)
    rows = torch.arange(N, dtype=torch.int32, device="cuda")
    back = t4.get_rows(tiles, rows, N, K)
    nmse = float(((back - W) ** 2).sum() / (W ** 2).sum())
    assert nmse < 2e-2
    # idempotence of the format: re-quantising the dequantised rows gives rows within one int4 step everywhere
    again = t4.get_rows(t4.quantize_tiles(back), rows, N, K)
    assert float(((again - back) ** 2).sum() / (back ** 2).sum()) < 2e-2


@pytest.mark.gpu
def test_ggml_get_rows_mirror_batched(t4, orc):
    """ggml_cuda_op_get_rows shape handling: index slice (i11, i12) reads tile-tensor slice (i11, i12) (getrows.cu:126)"""
    rng = np.random.default_rng(31)
    N, K, ne10, ne11, ne12 = 16, 512, 5, 2, 3
    slices = [orc.quantize_tiles(make_weights(rng, N, K)) for _ in range(ne11 * ne12)]
    idx = rng.integers(0, N, size=(ne12, ne11, ne10)).astype(np.int32)
    got = t4.ggml_get_rows(_dev(np.concatenate(slices)), _dev(idx), N, K).cpu().numpy()
    for i12 in range(ne12):
        for i11 in range(ne11):
            want = orc.get_rows(slices[i12 * ne11 + i11], idx[i12, i11], K)
            assert np.array_equal(got[i12, i11].view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_quantiser_exact_half_ties_everywhere(t4, orc):
    """every code boundary k + 0.5 (k = -8 .. 6) in both directions: the kernel's one-FRND roundf must round half away
    from zero exactly like roundf (ggml-quants.c:1620), including after the clamp to [-8, 7]"""
    K = 256
    W = np.zeros((8, K), dtype=np.float32)
    ties = np.arange(-8, 7, dtype=np.float32) + 0.5                       # 15 ties
    for c in range(8):
        for g in range(8):
            blk = np.zeros(32, dtype=np.float32)
            blk[0] = 7.0 * (c + 1)                                        # absmax -> scale = c + 1 exactly
            blk[1:16] = ties * (c + 1)
            blk[16:31] = -ties * (c + 1)
            W[c, g * 32:(g + 1) * 32] = blk
    # the quantiser sees FWHT(W): feed the inverse transform so that the transformed block is exactly `W`
    Wt = np.stack([np.concatenate([orc.fwht(W[c, b * 256:(b + 1) * 256].copy()) for b in range(K // 256)]) for c in range(8)])
    got = t4.quantize_tiles(_dev(Wt)).cpu().numpy()
    assert np.array_equal(got, orc.quantize_tiles(Wt))
