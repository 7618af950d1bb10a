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


def test_gemm4_fuzz_against_pair_gemm(t4):
    """60 seeded random shapes (M 9..64, N any multiple of 8 up to 40 000, K any multiple of 256 up to 8 192): the batched-decode
    GEMM against the kind::i8 pair GEMM on the same inputs -- the int32 sums of both are pinned to the oracle elsewhere, here the
    planner's corners are what is exercised (ragged N, blocks just above / below the SM count, K halves, K parts, one k-tile)."""
    rng = np.random.default_rng(77)
    shapes = [(int(rng.integers(9, 65)), int(rng.integers(1, 64)) * 8, int(rng.integers(1, 9)) * 256) for _ in range(30)]
    shapes += [(int(rng.integers(9, 65)), int(rng.integers(140, 160)) * 128 + int(rng.integers(0, 16)) * 8, 512) for _ in range(6)]
    shapes += [(int(rng.integers(9, 65)), int(rng.integers(1, 12)) * 128 - int(rng.integers(0, 15)) * 8, int(rng.integers(8, 33)) * 256)
               for _ in range(14)]
    shapes += [(int(rng.integers(9, 65)), int(rng.integers(300, 5000)) * 8, int(rng.integers(1, 5)) * 256) for _ in range(10)]
    for M, N, K in shapes:
        N = max(N, 8)
        tiles = t4.random_tiles(N, K, seed=M + N + K)
        x = torch.randn((M, K), device="cuda") * float(rng.uniform(0.1, 4.0))
        ws = t4.Workspace.for_shape(M, N, K)
        y4 = t4.mul_mat(tiles, x, N, workspace=ws, kernel=t4.KERNEL_GEMM4).clone()
        y4b = t4.mul_mat(tiles, x, N, workspace=ws, kernel=t4.KERNEL_GEMM4).clone()
        y2 = t4.mul_mat(tiles, x, N, workspace=ws, kernel=t4.KERNEL_GEMM2)
        assert torch.equal(y4, y4b), (M, N, K)
        e = nmse(y4.cpu().numpy(), y2.cpu().numpy())
        assert e <= 1e-12, (M, N, K, e)


def test_gemm4_fuzz_fanout_groups(t4):
    """random fan-out groups of 2..4 matrices into one strided buffer through the graph entry, against per-matrix calls"""
    rng = np.random.default_rng(78)
    for _ in range(12):
        M = int(rng.integers(9, 65))
        K = int(rng.integers(1, 13)) * 256
        Ns = [int(rng.integers(1, 300)) * 8 for _ in range(int(rng.integers(2, 5)))]
        if rng.random() < 0.3:
            Ns[0] = int(rng.integers(145, 152)) * 128
        tl = [t4.random_tiles(n, K, seed=n + K) for n in Ns]
        x = torch.randn((M, K), device="cuda")
        buf = torch.full((M, sum(Ns) + 8), float("nan"), device="cuda")
        outs, o = [], 0
        for n in Ns:
            outs.append(buf[:, o:o + n])
            o += n
        t4.GgmlGraph([(t, x, out, n, "") for t, n, out in zip(tl, Ns, outs)]).compute()
        torch.cuda.synchronize()
        assert torch.isnan(buf[:, sum(Ns):]).all()
        for t, n, out in zip(tl, Ns, outs):
            want = t4.mul_mat(t, x, n, kernel=t4.KERNEL_GEMM2)
            e = nmse(out.cpu().numpy(), want.cpu().numpy())
            assert e <= 1e-12, (M, K, Ns, n, e)


@pytest.mark.parametrize("M", [12, 40, 64])
def test_gemm4_chain_of_groups_in_a_cuda_graph(t4, M):
    """a decode-batch layer's worth of launches back to back under programmatic dependent launch, captured and replayed: every
    mode of the planner in one chain on ONE shared arena (whole blocks, K halves with the outputs cleared by K1, K parts with
    arrival counters, stream-K over > 148 blocks), each output checked against the pair GEMM after every replay"""
    rng = np.random.default_rng(91)
    shapes = [([1024, 256, 256], 1024), ([512], 4096), ([9600, 9600], 512), ([256], 6144), ([392], 2048)]
    nodes, ref = [], []
    for Ns, K in shapes:
        x = torch.randn((M, K), device="cuda")
        for n in Ns:
            w = t4.random_tiles(n, K, seed=n + K)
            out = torch.empty((M, n), device="cuda")
            nodes.append((w, x, out, n, ""))
    t4.reserve(M, 9600, 6144)
    g = t4.GgmlGraph(nodes)
    g.compute()
    torch.cuda.synchronize()
    cg = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        with torch.cuda.graph(cg, stream=st):
            g.compute(stream=st)
    torch.cuda.current_stream().wait_stream(st)
    for rep in range(3):
        for (_, x, out, _, _) in nodes:
            out.fill_(float("nan"))
        for x in {id(n[1]): n[1] for n in nodes}.values():
            x.normal_()
        for _ in range(2):                     # twice on the same inputs: counters and cleared outputs must come back clean
            cg.replay()
        torch.cuda.synchronize()
        for (w, x, out, n, _) in nodes:
            want = t4.mul_mat(w, x, n, kernel=t4.KERNEL_GEMM2)
            e = nmse(out.cpu().numpy(), want.cpu().numpy())
            assert e <= 1e-12, (M, n, x.shape[1], rep, e)
