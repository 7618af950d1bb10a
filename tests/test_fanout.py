"""Fan-out: several weight matrices over the same activations with ONE activation stage.  The bar: every output is
BIT-IDENTICAL to the separate MUL_MAT (which is itself pinned to the oracle in test_gpu_parity.py) at decode and prefill batch."""
import numpy as np
import pytest

from conftest import block_perm, make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def t4():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import __graft_entry__ as graft
    return graft.load_package()


@pytest.mark.parametrize("M", [1, 3, 8, 16, 17, 130, 300])
@pytest.mark.parametrize("use_perm", [False, True])
def test_fanout_equals_separate_mul_mats(t4, orc, M, use_perm):
    rng = np.random.default_rng(100 + M)
    K, Ns = 1024, [512, 128, 136]                       # widest first, a narrow one, one that is not a multiple of 128
    perm = block_perm(rng, K) if use_perm else None
    pd = torch.from_numpy(perm).cuda() if use_perm else None
    tiles = [orc.quantize_tiles(make_weights(rng, N, K)) for N in Ns]
    tds = [torch.from_numpy(t).cuda() for t in tiles]
    x = make_acts(rng, M, K)
    xd = torch.from_numpy(x).cuda()
    fan = t4.Fanout(tds, Ns, K)
    for rep in range(2):                                # second pass: the workspace holds stale codes of the first
        outs = fan(xd, perm=pd)
        for td, N, y in zip(tds, Ns, outs):
            assert torch.equal(y, t4.mul_mat(td, xd, N, perm=pd)), (N, rep)
        xd = xd * 0.5 + 0.25
    ref = orc.mul_mat(tiles[2], (x * 0.5 + 0.25).astype(np.float32), Ns[2], perm)
    assert nmse(outs[2].cpu().numpy(), ref["C"]) <= 1e-9


def test_fanout_in_a_graph_and_other_order(t4, orc):
    """narrow matrix first (the shared workspace layout is the widest one's), captured and replayed"""
    rng = np.random.default_rng(7)
    M, K, Ns = 200, 512, [64, 1024]
    tds = [torch.from_numpy(orc.quantize_tiles(make_weights(rng, N, K))).cuda() for N in Ns]
    xd = torch.from_numpy(make_acts(rng, M, K)).cuda()
    fan = t4.Fanout(tds, Ns, K)
    outs = fan(xd)
    want = [t4.mul_mat(td, xd, N).clone() for td, N in zip(tds, Ns)]
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            fan(xd, outs=outs)
    for o in outs:
        o.zero_()
    g.replay()
    torch.cuda.synchronize()
    for o, w in zip(outs, want):
        assert torch.equal(o, w)
