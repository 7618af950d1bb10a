"""Batched decode, 17 <= M <= 64 (BASELINE configs[2]): the kernel that runs the weights through TMEM (tcq4_gemm4.cu) against
the CPU oracle -- the fp32 accumulators in TMEM hold the oracle's int32 sums exactly, the fp32 outputs follow the oracle's
operation order.  Same bar as test_gpu_parity.py."""
import numpy as np
import pytest

from conftest import block_perm, make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FP32_NMSE_TIGHT = 1e-9


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# (M, N, K): both token widths (Mt = 32 / 64), ragged N (not a multiple of 128), one k-tile, several blocks per CTA
# (N = 19200 > 148 blocks: stream-K, a block cut by a range boundary meets its other part in y), K halves (few blocks),
# K parts with the last arriver's ordered reduction (very few blocks, many k-tiles: (24, 128, 8192), (17, 384, 6144))
GEMM4_SHAPES = [(17, 128, 256), (32, 200, 512), (33, 1024, 1024), (48, 136, 2304), (64, 2560, 4096), (64, 19200, 512),
                (20, 8, 256), (64, 392, 2048), (40, 4096, 2560), (24, 128, 8192), (17, 384, 6144), (12, 1024, 1024)]


@pytest.mark.parametrize("M,N,K", GEMM4_SHAPES)
def test_gemm4_vs_oracle(t4, orc, M, N, K):
    rng = np.random.default_rng(4000 + M + N + K)
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    big = M * N * K > 1e9
    ref = orc.mul_mat(tiles, x if not big else x[:8], N, want_isum=True)
    rows = slice(0, 8) if big else slice(0, M)
    y, isum = t4.mul_mat_debug(_dev(tiles), _dev(x), N, kernel=t4.KERNEL_GEMM4)
    torch.cuda.synchronize()
    assert np.array_equal(isum.cpu().numpy()[rows], ref["isum"])
    assert nmse(y.cpu().numpy()[rows], ref["C"]) <= FP32_NMSE_TIGHT
    ws = t4.Workspace.for_shape(M, N, K)
    assert t4.select_kernel(M, N, K) == t4.KERNEL_GEMM4 or M <= 16
    ya = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws).clone()             # the automatic dispatch takes this kernel
    yb = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws).clone()
    assert torch.equal(ya, yb) and torch.equal(ya, y)                          # deterministic (a cut block has two parts: their sum commutes)
    y2 = t4.mul_mat(_dev(tiles), _dev(x), N, workspace=ws, kernel=t4.KERNEL_GEMM2)
    # blocks not cut by a range boundary follow the kind::i8 kernels' fp32 operation order exactly; cut ones differ in association
    assert torch.equal(y2, ya) or nmse(ya.cpu().numpy(), y2.cpu().numpy()) <= 1e-12


@pytest.mark.parametrize("M,N,K", [(50, 136, 512), (20, 136, 6144)])
def test_gemm4_zero_points_perm_negative_scales(t4, orc, M, N, K):
    rng = np.random.default_rng(41)
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
    y, isum = t4.mul_mat_debug(_dev(t), _dev(x), N, _dev(perm), kernel=t4.KERNEL_GEMM4)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])
    assert nmse(y.cpu().numpy(), ref["C"]) <= 1e-8


def test_gemm4_extreme_codes_are_exact(t4, orc):
    rng = np.random.default_rng(5)
    M, N, K = 64, 64, 512
    t = orc.quantize_tiles(make_weights(rng, N, K)).reshape(-1, 1184).copy()
    t[:, :1024] = rng.choice(np.array([0x88, 0x77, 0x87, 0x78], dtype=np.uint8), (t.shape[0], 1024))
    t[:, 1056:1120] = rng.choice(np.array([127, -127], dtype=np.int8), (t.shape[0], 64)).view(np.uint8)
    t = t.reshape(-1)
    x = np.where(rng.random((M, K)) < 0.5, -1.0, 1.0).astype(np.float32)
    x[::2] = 1.0
    ref = orc.mul_mat(t, x, N, want_isum=True)
    y, isum = t4.mul_mat_debug(_dev(t), _dev(x), N, kernel=t4.KERNEL_GEMM4)
    assert np.array_equal(isum.cpu().numpy(), ref["isum"])
    assert nmse(y.cpu().numpy(), ref["C"]) <= FP32_NMSE_TIGHT


@pytest.mark.parametrize("M,K,Ns", [(48, 1024, [1024, 256, 256]), (48, 2048, [1024, 256, 256]), (33, 512, [19200, 1032]),
                                    (20, 6144, [128, 136])])
def test_gemm4_strided_output_and_fanout(t4, orc, M, K, Ns):
    """A fan-out group of a decode batch (Q/K/V, gate/up) into one fused buffer (strided outputs): one activation launch and
    ONE GEMM launch whose blocks run over all matrices -- whole blocks, K halves (outputs cleared row by row by K1),
    stream-K over 158 blocks, K parts"""
    rng = np.random.default_rng(6 + K)
    x = make_acts(rng, M, K)
    tl = [orc.quantize_tiles(make_weights(rng, n, K)) for n in Ns]
    buf = torch.full((M, sum(Ns) + 8), float("nan"), device="cuda")
    outs, o = [], 0
    for n in Ns:
        outs.append(buf[:, o:o + n])
        o += n
    xd = _dev(x)
    g = t4.GgmlGraph([(_dev(t), xd, out, n, "") for t, n, out in zip(tl, Ns, outs)])
    n0 = t4.launch_count()
    g.compute()
    torch.cuda.synchronize()
    assert t4.launch_count() - n0 == 2
    first = [o_.clone() for o_ in outs]
    g.compute()                                           # a second call on the same workspace: counters / cleared outputs
    torch.cuda.synchronize()
    assert torch.isnan(buf[:, sum(Ns):]).all()            # nothing written outside the outputs
    for t, n, out, f in zip(tl, Ns, outs, first):
        assert torch.equal(out, f)
        rows = slice(0, 8) if n * K * M > 3e8 else slice(0, M)
        ref = orc.mul_mat(t, x[rows], n)["C"]
        assert nmse(out[rows].cpu().numpy(), ref) <= FP32_NMSE_TIGHT
