"""Input-side fusion: RMSNorm * weight in front of Q/K/V and gate/up, SwiGLU in front of down, applied while the
activation stage loads x (tcq4_b200_mul_mat_fanout_input).  Two bars: (1) BIT-IDENTICAL to the unfused composition
tcq4_b200_apply_input -> plain fan-out (the fused load evaluates the same expression, ggml-cuda's norm.cu:75-150 /
unary.cu:255-267 order); (2) the transform itself against a float64 numpy restatement, and the fused MUL_MAT against the CPU
oracle fed with the transform's own output (so the oracle bar of the plain path carries over)."""
import numpy as np
import pytest

from conftest import block_perm, make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("M", [1, 8, 12, 48, 200])
@pytest.mark.parametrize("use_perm", [False, True])
def test_rms_norm_mul_fused_into_the_fanout(t4, orc, M, use_perm):
    rng = np.random.default_rng(500 + M)
    K, Ns, eps = 1024, [512, 128, 136], 1e-6
    perm = block_perm(rng, K) if use_perm else None
    tl = [orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, n, K), perm) if use_perm else make_weights(rng, n, K)) for n in Ns]
    x = make_acts(rng, M, K)
    nw = rng.uniform(0.5, 1.5, K).astype(np.float32)
    xd, nwd, pd = _dev(x), _dev(nw), (_dev(perm) if use_perm else None)
    rs = t4.rms_row_scale(xd, eps)
    want_rs = 1.0 / np.sqrt((x.astype(np.float64) ** 2).mean(axis=1) + eps)
    assert np.allclose(rs.cpu().numpy(), want_rs, rtol=2e-6)
    xn = t4.apply_input(t4.INPUT_RMS_NORM_MUL, xd, row_scale=rs, col_weight=nwd)
    want_xn = (rs.cpu().numpy()[:, None].astype(np.float32) * x) * nw[None, :]          # same two roundings, same order
    assert np.array_equal(xn.cpu().numpy(), want_xn)
    fused = t4.mul_mat_fanout_input([_dev(t) for t in tl], xd, Ns, t4.INPUT_RMS_NORM_MUL, row_scale=rs, col_weight=nwd, perm=pd)
    plain = t4.Fanout([_dev(t) for t in tl], Ns, K)(xn, perm=pd)
    torch.cuda.synchronize()
    for f, p, t, n in zip(fused, plain, tl, Ns):
        assert torch.equal(f, p)
        assert nmse(f.cpu().numpy(), orc.mul_mat(t, xn.cpu().numpy(), n, perm)["C"]) <= 1e-9


@pytest.mark.parametrize("M", [1, 8, 12, 48, 200])
def test_swiglu_fused_into_the_down_projection(t4, orc, M):
    rng = np.random.default_rng(600 + M)
    K, N = 1536, 264
    perm = block_perm(rng, K)
    tiles = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm))
    gate, up = make_acts(rng, M, K), make_acts(rng, M, K)
    gd, ud, pd = _dev(gate), _dev(up), _dev(perm)
    act = t4.apply_input(t4.INPUT_SWIGLU, gd, x2=ud)
    g64, u64 = gate.astype(np.float64), up.astype(np.float64)
    want = g64 / (1.0 + np.exp(-g64)) * u64
    assert np.allclose(act.cpu().numpy(), want, rtol=1e-5, atol=1e-7)          # device expf against numpy's, a few ulp
    fused = t4.mul_mat_fanout_input([_dev(tiles)], gd, [N], t4.INPUT_SWIGLU, x2=ud, perm=pd)[0]
    plain = t4.mul_mat(_dev(tiles), act, N, pd)
    assert torch.equal(fused, plain)
    assert nmse(fused.cpu().numpy(), orc.mul_mat(tiles, act.cpu().numpy(), N, perm)["C"]) <= 1e-9


def test_input_fusion_strided_outputs_and_bad_arguments(t4, orc):
    rng = np.random.default_rng(700)
    M, K, Ns = 40, 512, [256, 128]
    tl = [_dev(orc.quantize_tiles(make_weights(rng, n, K))) for n in Ns]
    x = _dev(make_acts(rng, M, K))
    nw = _dev(rng.uniform(0.5, 1.5, K).astype(np.float32))
    rs = t4.rms_row_scale(x, 1e-5)
    buf = torch.full((M, sum(Ns) + 8), float("nan"), device="cuda")
    outs = [buf[:, :256], buf[:, 256:384]]
    t4.mul_mat_fanout_input(tl, x, Ns, t4.INPUT_RMS_NORM_MUL, row_scale=rs, col_weight=nw, outs=outs)
    want = t4.Fanout(tl, Ns, K)(t4.apply_input(t4.INPUT_RMS_NORM_MUL, x, row_scale=rs, col_weight=nw))
    torch.cuda.synchronize()
    assert torch.equal(outs[0], want[0]) and torch.equal(outs[1], want[1]) and torch.isnan(buf[:, 384:]).all()
    with pytest.raises(t4.Tcq4Error):
        t4.mul_mat_fanout_input(tl, x, Ns, t4.INPUT_SWIGLU)                     # no second operand
    with pytest.raises(t4.Tcq4Error):
        t4.mul_mat_fanout_input(tl, x, Ns, t4.INPUT_RMS_NORM_MUL, row_scale=rs)  # no weight
    y = t4.mul_mat(tl[0], x, Ns[0])                                             # the pending transform never leaks into a later call
    assert torch.equal(y, t4.mul_mat(tl[0], x, Ns[0]))


def test_input_fusion_is_graph_capturable(t4, orc):
    """row scale + fused fan-out + fused SwiGLU down captured in one CUDA graph and replayed on new activations"""
    rng = np.random.default_rng(800)
    M, K, Kd, Ns, Nd = 48, 512, 768, [256, 128], 264
    tl = [_dev(orc.quantize_tiles(make_weights(rng, n, K))) for n in Ns]
    wd = _dev(orc.quantize_tiles(make_weights(rng, Nd, Kd)))
    x = torch.randn((M, K), device="cuda")
    nw = _dev(rng.uniform(0.5, 1.5, K).astype(np.float32))
    gate, up = torch.randn((M, Kd), device="cuda"), torch.randn((M, Kd), device="cuda")
    rs = torch.empty((M,), device="cuda")
    outs = [torch.empty((M, n), device="cuda") for n in Ns]
    yd = [torch.empty((M, Nd), device="cuda")]

    def step(stream=None):
        t4.LIB.tcq4_b200_rms_row_scale(x.data_ptr(), rs.data_ptr(), M, K, 1e-6, stream.cuda_stream if stream else 0)
        t4.mul_mat_fanout_input(tl, x, Ns, t4.INPUT_RMS_NORM_MUL, row_scale=rs, col_weight=nw, outs=outs, stream=stream)
        t4.mul_mat_fanout_input([wd], gate, [Nd], t4.INPUT_SWIGLU, x2=up, outs=yd, stream=stream)
    t4.reserve(M, max(Ns + [Nd]), max(K, Kd))
    step()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            step(st)
    torch.cuda.current_stream().wait_stream(st)
    for _ in range(2):
        x.normal_()
        gate.normal_()
        up.normal_()
        g.replay()
        torch.cuda.synchronize()
        xn = t4.apply_input(t4.INPUT_RMS_NORM_MUL, x, row_scale=t4.rms_row_scale(x, 1e-6), col_weight=nw)
        want = t4.Fanout(tl, Ns, K)(xn)
        assert torch.equal(outs[0], want[0]) and torch.equal(outs[1], want[1])
        assert torch.equal(yd[0], t4.mul_mat(wd, t4.apply_input(t4.INPUT_SWIGLU, gate, x2=up), Nd))
