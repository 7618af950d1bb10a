"""Pin the oracle: (1) against the committed golden fixtures generated from the reference's own CPU
code (tests/golden/make_golden.py), (2) live against oracle/_ref/libggml_ref.so when it is present."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden import CASES, make_w  # noqa: E402

from conftest import make_weights  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "tcq4_golden.npz"))


def test_golden_type_traits():
    assert list(GOLD["traits"]) == [148, 256]           # ggml.c:750-757


def test_golden_tiles_and_dequant(orc):
    for name, N, K, seed, std, special in CASES:
        W = make_w(N, K, seed, std, special)
        tiles = orc.quantize_tiles(W)
        assert np.array_equal(tiles, GOLD[f"{name}.tiles"]), name          # byte-identical tiles
        assert np.array_equal(orc.dequant_tiles(tiles), GOLD[f"{name}.dequant"]), name


def test_golden_fwht(orc):
    assert np.array_equal(orc.fwht(GOLD["fwht.in"]), GOLD["fwht.out"])   # bit-identical floats


def test_live_quantiser_memcmp(orc, ref):
    """quantize_tcq4_tile (ggml-quants.c:1528-1639) through ggml_quantize_chunk (ggml.c:7564-7565)."""
    rng = np.random.default_rng(21)
    for N, K, std in [(8, 256, 0.02), (16, 512, 1.0), (24, 768, 1e-4), (13, 256, 0.02), (64, 2560, 0.02),
                      (40, 1024, 30.0)]:
        W = make_weights(rng, N, K, std)
        W[0, : min(K, 40)] = 0.0
        if N > 3:
            W[3, 100] = 50.0 * std
        a = orc.quantize_tiles(W)
        assert np.array_equal(a, ref.quantize_tiles(W)), (N, K)
        assert np.array_equal(a, ref.quantize_tiles(W, via_chunk=False)), (N, K)
        assert np.array_equal(orc.dequant_tiles(a), ref.dequant_tiles(a)), (N, K)


def test_live_dequant_with_zero_points(orc, ref):
    """dequantize_row_tcq4_tile on tiles with non-zero Z/zc (the format allows it)."""
    rng = np.random.default_rng(22)
    t = orc.quantize_tiles(make_weights(rng, 16, 512)).reshape(-1, 1184).copy()
    t[:, 1040:1056] = np.frombuffer(rng.uniform(-0.01, 0.01, (t.shape[0], 8)).astype(np.float16).tobytes(),
                                    dtype=np.uint8).reshape(t.shape[0], 16)
    t[:, 1120:1184] = rng.integers(-127, 128, (t.shape[0], 64)).astype(np.int8).view(np.uint8)
    t = t.reshape(-1)
    assert np.array_equal(orc.dequant_tiles(t), ref.dequant_tiles(t))


def test_live_fwht_bit_equal(orc, ref):
    rng = np.random.default_rng(23)
    for scale in (1.0, 1e-3, 1e4):
        x = (rng.standard_normal(256) * scale).astype(np.float32)
        assert np.array_equal(orc.fwht(x), ref.fwht(x))


def test_live_type_traits(ref):
    tr = ref.type_traits()
    assert tr == {"name": "tcq4_tile", "type_size": 148, "blck_size": 256}
