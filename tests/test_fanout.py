"""Fan-out: several weight matrices over the same activations with ONE activation stage.  The bar: at prefill batch
every output is BIT-IDENTICAL to the separate MUL_MAT (itself pinned to the oracle in test_gpu_parity.py); at decode batch the
int32 accumulators are the same and the fp32 sum over 256-blocks may be cut elsewhere (NMSE <= 1e-12), and the oracle bar holds."""
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
            sep = t4.mul_mat(td, xd, N, perm=pd)
            if M > 16:
                assert torch.equal(y, sep), (N, rep)            # same GEMM launches, minus the repeated activation kernel
            else:
                # one segmented GEMV: same int32 accumulators per 256-block, but a channel's blocks may be cut between
                # warps at other places than in the separate launch, so the fp32 sum over blocks can differ in the last bits
                assert nmse(y.cpu().numpy(), sep.cpu().numpy()) <= 1e-12, (N, rep)
        xd = xd * 0.5 + 0.25
    x2 = (x * 0.5 + 0.25).astype(np.float32)
    for i in range(len(Ns)):
        ref = orc.mul_mat(tiles[i], x2, Ns[i], perm)
        assert nmse(outs[i].cpu().numpy(), ref["C"]) <= 1e-9, i


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


def test_fanout_decode_one_launch_per_matrix_in_a_child_process():
    """TCQ4_FANOUT_NOSEG=1 (read once at load, hence the child): decode batches through one GEMV per matrix instead of the
    single segmented launch -- the path taken with more than four matrices or when sum(N) does not plan as a GEMV."""
    import os
    import subprocess
    import sys
    if os.environ.get("TCQ4_FANOUT_NOSEG"):
        pytest.skip("already the child")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TCQ4_FANOUT_NOSEG="1")
    out = subprocess.run([sys.executable, "-m", "pytest", "tests/test_fanout.py", "-m", "gpu", "-q", "-x", "-k",
                          "equals_separate and not (17 or 130 or 300)"], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and " passed" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]


@pytest.mark.parametrize("M", [2, 96])
def test_ggml_mirror_same_perm_fans_out_and_different_perms_fall_back(t4, orc, M):
    """perms looked up by tensor name (rrs.cu:266-270): equal perms -> one fan-out call, different perms -> node by node"""
    rng = np.random.default_rng(900 + M)
    K, Ns, names = 512, [256, 64], ["blk.0.attn_q.weight", "blk.0.attn_k.weight"]
    p0, p1 = block_perm(rng, K), block_perm(rng, K)
    xd = torch.from_numpy(make_acts(rng, M, K)).cuda()
    t4.set_reorder_enabled(True)
    try:
        for perms in ([p0, p0.copy()], [p0, p1]):
            t4.clear_perms()
            tds = []
            for N, nm, p in zip(Ns, names, perms):
                t4.register_perm(nm, p)
                tds.append(torch.from_numpy(orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), p))).cuda())
            c0 = t4.launch_count()
            outs = t4.ggml_mul_mat_fanout(tds, xd, Ns, names)
            launches = t4.launch_count() - c0
            same = perms[1] is not p1
            if M <= 16:
                assert launches == (1 if same else 2)           # one segmented GEMV vs one GEMV per node
            else:
                assert launches == (2 if same else 4)           # one activation kernel + ONE segmented GEMM launch vs two of each
            for td, N, nm, y in zip(tds, Ns, names, outs):
                sep = t4.ggml_mul_mat(td, xd, N, name=nm)
                assert nmse(y.cpu().numpy(), sep.cpu().numpy()) <= 1e-12, (nm, same)
    finally:
        t4.clear_perms()
