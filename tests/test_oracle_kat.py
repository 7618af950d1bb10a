"""Known-answer tests of the CPU oracle (no GPU, no reference needed).

The reference holds no automated test for this path; the pins reused here are
  * tests/test-imma-micro.cu:679-748  fragment <-> (channel, k) maps of mma.m16n8k32.s4
  * tests/test-imma-micro.cu:911-968  hand-computed group value 1.5*(2*64 + 0.5*32) = 216
  * verify-rrs.c                      FWHT self-inverse / norm preservation
  * ggml/src/ggml-quants.c:1381-1400  closed form of the pack order
"""
import numpy as np
import pytest

from conftest import make_acts, make_weights, nmse


def test_pack_closed_form(orc):
    """element (c,k) of a group lives in byte c*16 + (k/8)*4 + (k%8)/2, low nibble iff k even."""
    rng = np.random.default_rng(1)
    w = rng.integers(-8, 8, size=(8, 32)).astype(np.int8)
    p = orc.pack_group(w)
    for c in range(8):
        for k in range(32):
            b = int(p[c * 16 + (k // 8) * 4 + (k % 8) // 2])
            nib = (b >> 4) if (k & 1) else (b & 0xF)
            val = nib - 16 if nib >= 8 else nib
            assert val == w[c, k]
    assert np.array_equal(orc.unpack_group(p), w)


def test_pack_imma_b_fragment_lane_map(orc):
    """B fragment of mma.m16n8k32: lane L holds column L/4, k = (L%4)*8 .. +8, element i at bits 4i
    (test-imma-micro.cu:719-727, ggml-quants.c:1381-1400)."""
    w = np.zeros((8, 32), dtype=np.int8)
    for c in range(8):
        for k in range(32):
            w[c, k] = ((c * 3 + k) % 16) - 8
    p = orc.pack_group(w).view(np.uint32)
    for lane in range(32):
        c, s = lane // 4, lane % 4
        for i in range(8):
            nib = (int(p[lane]) >> (4 * i)) & 0xF
            val = nib - 16 if nib >= 8 else nib
            assert val == w[c, s * 8 + i]


def test_pack_extremes(orc):
    for v in (-8, -1, 0, 7):
        w = np.full((8, 32), v, dtype=np.int8)
        assert np.array_equal(orc.unpack_group(orc.pack_group(w)), w)


def test_fwht_properties(orc):
    """verify-rrs.c: FWHT(FWHT(x)) = x and the L2 norm is preserved (n = 8 there; 8 and 256 here)."""
    rng = np.random.default_rng(2)
    for n in (8, 256):
        x = rng.standard_normal(n).astype(np.float32)
        y = orc.fwht(x)
        assert abs(np.linalg.norm(y) - np.linalg.norm(x)) < 1e-4 * np.linalg.norm(x)
        assert np.max(np.abs(orc.fwht(y) - x)) < 1e-5 * max(1.0, np.max(np.abs(x)))


def test_fwht_matches_hadamard_matrix(orc):
    """The butterfly order h=1..128 with (x+y, x-y) is the natural-ordered (Sylvester) Hadamard."""
    H = np.array([[1.0]])
    while H.shape[0] < 256:
        H = np.block([[H, H], [H, -H]])
    rng = np.random.default_rng(3)
    x = rng.standard_normal(256).astype(np.float32)
    assert np.allclose(orc.fwht(x), (H @ x.astype(np.float64)) / 16.0, atol=1e-4)
    e = np.zeros(256, dtype=np.float32)
    e[5] = 16.0
    assert np.array_equal(orc.fwht(e), H[:, 5].astype(np.float32))


def test_fp16_roundtrip(orc):
    vals = [0.0, 1.0, -1.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 6.1e-5, 0.1, 3.14159, 1e-3, 2.0 ** -25,
            2.0 ** -24, 1.0009765625, 1.00048828125]
    a = np.array(vals, dtype=np.float32)
    mine = np.array([orc.fp32_to_fp16(float(v)) for v in a], dtype=np.uint16)
    with np.errstate(over="ignore"):
        ref = a.astype(np.float16).view(np.uint16)
    assert np.array_equal(mine, ref)
    for h in (0x0000, 0x0001, 0x03FF, 0x0400, 0x3C00, 0x7BFF, 0xBC00, 0x8001):
        assert orc.fp16_to_fp32(h) == float(np.array([h], dtype=np.uint16).view(np.float16)[0])


def test_quantiser_layout_and_sizes(orc):
    """type traits: 148 B per 256-element 'row', 1184 B per tile, [N/8][K/256] n-tile major."""
    rng = np.random.default_rng(4)
    N, K = 16, 512
    t = orc.quantize_tiles(make_weights(rng, N, K))
    assert t.size == (N // 8) * (K // 256) * 1184 == N * (K // 256) * 148
    tiles = t.reshape(-1, 1184)
    Z = tiles[:, 1040:1056]
    zc = tiles[:, 1120:1184]
    assert not Z.any() and not zc.any()                    # symmetric quantiser (ggml-quants.c:1603,1609)
    sc = tiles[:, 1056:1120].view(np.int8)
    assert sc.max() == 127 and sc.min() >= 0               # max group of every channel maps to 127


def test_quantiser_error_and_zero_group(orc):
    rng = np.random.default_rng(5)
    N, K = 8, 256
    W = make_weights(rng, N, K)
    t = orc.quantize_tiles(W)
    Wh = np.stack([orc.fwht(W[n]) for n in range(N)])
    D = orc.dequant_matrix(t, N, K)
    assert nmse(D, Wh) < 2e-2                               # probe value 9.3e-3 on uniform data
    # an all-zero row quantises to scale 1.0 / codes 0 and dequantises to exactly 0
    W[3] = 0.0
    D = orc.dequant_matrix(orc.quantize_tiles(W), N, K)
    assert not D[3].any()


def test_partial_row_group_zero_padded(orc):
    """nrow % 8 != 0: the last tile row is zero-padded (ggml-quants.c:1553-1569)."""
    rng = np.random.default_rng(6)
    W = make_weights(rng, 13, 256)
    t = orc.quantize_tiles(W)
    assert t.size == 2 * 1184
    Wpad = np.zeros((16, 256), dtype=np.float32)
    Wpad[:13] = W
    assert np.array_equal(t, orc.quantize_tiles(Wpad))


def test_hand_computed_group_216(orc):
    """test-imma-micro.cu:911-968 (hand-computed): q_a = 1, q_b = 2, a-scale 1.5, b_scale 2.0, b_zero 0.5,
    sum_a = 32 -> per group 1.5 * (2.0*64 + 0.5*32) = 216, zero-point term included.  Re-expressed on the
    per-channel tile: S=2, sc=127 -> b_scale 2; Z=0.5, zc=127 -> b_zero 0.5; max_abs 10.5 -> a7 1.5.
    One element has to carry max_abs (code 7), so group 7 is 1.5*(2*76 + 0.5*38) = 256.5 instead."""
    tile = np.zeros(1184, dtype=np.uint8)
    w = np.zeros((8, 32), dtype=np.int8)
    w[0, :] = 2
    for g in range(8):
        tile[g * 128:(g + 1) * 128] = orc.pack_group(w)
    tile[1024:1026] = np.array([2.0], dtype=np.float16).view(np.uint8)
    tile[1040:1042] = np.array([0.5], dtype=np.float16).view(np.uint8)
    tile[1056:1064] = 127                                  # sc[0][0..7]
    tile[1120:1128] = 127                                  # zc[0][0..7]
    v = np.full(256, 1.5, dtype=np.float32)                # code 1 at a7 = 1.5
    v[255] = 10.5                                          # pins max_abs = 10.5 (code 7)
    x = orc.fwht(v)[None, :]                               # FWHT is its own inverse
    r = orc.mul_mat(tile, x, 8, want_dots=True, want_act=True)
    assert r["max_abs"][0, 0] == pytest.approx(10.5, rel=1e-6)
    assert np.array_equal(r["codes"][0, :255], np.ones(255, dtype=np.int8)) and r["codes"][0, 255] == 7
    assert list(r["dots"][0, 0, 0]) == [64] * 7 + [76]
    assert list(r["sum_q"][0, 0]) == [32] * 7 + [38]
    assert r["C"][0, 0] == pytest.approx(7 * 216.0 + 256.5, rel=1e-6)
    assert not r["C"][0, 1:].any()


def test_activation_edge_cases(orc):
    K = 512
    x = np.zeros((3, K), dtype=np.float32)
    x[1, :256] = 1e-12                                      # below the 1e-10 threshold -> max_abs = 1.0
    x[2, 7] = 1000.0                                        # single huge outlier: FWHT spreads it evenly
    codes, mx, sq = orc.quantize_act(x)
    assert not codes[0].any() and np.all(mx[0] == 1.0)
    assert np.all(mx[1] == 1.0) and not codes[1].any()
    assert np.all(np.abs(codes[2, :256]) == 7) and mx[2, 0] == pytest.approx(1000.0 / 16.0)
    assert codes.min() >= -7 and codes.max() <= 7           # never -8 (tcq4_k32.cu:829)
    assert np.array_equal(sq, codes.reshape(3, 2, 8, 32).sum(-1).astype(np.int16))


def test_rintf_half_to_even(orc):
    """codes use rintf (half to even), not roundf: v*(7/max) = 2.5 -> 2, 3.5 -> 4."""
    v = np.zeros(256, dtype=np.float32)
    v[0] = 7.0
    v[1] = 2.5
    v[2] = 3.5
    v[3] = -2.5
    x = orc.fwht(v)[None, :]
    codes, mx, _ = orc.quantize_act(x)
    assert mx[0, 0] == 7.0
    assert list(codes[0, :4]) == [7, 2, 4, -2]


def test_identity_w_dot_x(orc):
    """C = W x = (W P^T H)(H P x): the O2 product approximates the true product at weight-quant error,
    O1 adds the activation int4 noise."""
    rng = np.random.default_rng(8)
    N, K, M = 64, 1024, 4
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K, outliers=False)
    t = orc.quantize_tiles(W)
    true = x @ W.T
    assert nmse(orc.mul_mat_f32(t, x, N), true) < 3e-2
    assert nmse(orc.mul_mat(t, x, N)["C"], true) < 8e-2


def test_perm_convention(orc):
    """weights W'[n][j] = W[n][perm[j]] (llama-quant.cpp:26-40) and activations a[i] = x[perm[i]]
    (tcq4_k32.cu:771-776) cancel: the permuted pair computes the same product."""
    from conftest import block_perm
    rng = np.random.default_rng(9)
    N, K, M = 16, 512, 2
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K, outliers=False)
    perm = block_perm(rng, K)
    tp = orc.quantize_tiles(orc.apply_perm_weights(W, perm))
    true = x @ W.T
    assert nmse(orc.mul_mat_f32(tp, x, N, perm), true) < 3e-2
    assert nmse(orc.mul_mat(tp, x, N, perm)["C"], true) < 8e-2


def test_perm_from_imatrix(orc):
    rng = np.random.default_rng(10)
    im = np.exp(rng.standard_normal(512)).astype(np.float32)
    perm = orc.perm_from_imatrix(im)
    for b in range(2):
        blk = perm[b * 256:(b + 1) * 256]
        assert sorted(blk) == list(range(b * 256, (b + 1) * 256))
        assert np.all(np.diff(im[blk]) <= 0)


def test_isum_consistency(orc):
    """isum[m][n][kt] = sum_g sc[n][kt][g] * dot_g -- the quantity the tcgen05 kernel accumulates."""
    rng = np.random.default_rng(11)
    N, K, M = 16, 512, 3
    t = orc.quantize_tiles(make_weights(rng, N, K))
    r = orc.mul_mat(t, make_acts(rng, M, K), N, want_dots=True, want_isum=True)
    sc = t.reshape(N // 8, K // 256, 1184)[:, :, 1056:1120].view(np.int8).reshape(N // 8, K // 256, 8, 8)
    sc_n = sc.transpose(0, 2, 1, 3).reshape(N, K // 256, 8).astype(np.int64)
    assert np.array_equal(r["isum"], (r["dots"] * sc_n[None]).sum(-1))
    assert np.abs(r["dots"]).max() <= 32 * 7 * 8
