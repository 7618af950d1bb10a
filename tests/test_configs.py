"""BASELINE.json configs[2..4] as parity cases, scaled where the scalar oracle would take minutes:
C3 Llama-3-8B with imatrix channel reordering, decode batch 1-64; C4 Qwen3-32B prefill N-sharded over 2/4/8 ranks;
C5 Llama-3-70B column-parallel over 8 ranks with a batch sweep.  Every case runs the PRODUCT chain end to end:
perm derivation -> device quantiser with the perm fused -> registry by tensor name -> ggml MUL_MAT mirror."""
import numpy as np
import pytest

from conftest import make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def t4():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import __graft_entry__ as graft
    return graft.load_package()


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ---- C3: Llama-3-8B (hidden 4096, ffn 14336, kv 1024), synthetic imatrix exp(N(0,1)) seed 42, M = 1 .. 64 ----------------
@pytest.mark.parametrize("name,N,K", [("blk.0.attn_k.weight", 1024, 4096), ("blk.0.ffn_up.weight", 512, 4096),
                                      ("blk.0.ffn_down.weight", 256, 14336)])
def test_c3_llama3_8b_imatrix_reorder_decode_batches(t4, orc, name, N, K):
    rng = np.random.default_rng(42)
    im = np.exp(rng.standard_normal(K)).astype(np.float32)
    perm = t4.perm_from_imatrix(im)                                       # quantize.cpp:127-165
    assert np.array_equal(perm, orc.perm_from_imatrix(im))
    W = make_weights(rng, N, K)
    tiles_d = t4.quantize_tiles(_dev(W), _dev(perm))                      # llama-quant.cpp:26-40 + ggml-quants.c:1528-1639
    tiles = orc.quantize_tiles(orc.apply_perm_weights(W, perm))
    assert np.array_equal(tiles_d.cpu().numpy(), tiles)
    t4.clear_perms()
    t4.set_reorder_enabled(True)
    t4.register_perm(name, perm)                                          # llama-model.cpp:7192-7197
    try:
        for M in (1, 2, 4, 8, 16, 32, 64):
            x = make_acts(rng, M, K)
            ref = orc.mul_mat(tiles, x, N, perm, want_isum=True)
            y = t4.ggml_mul_mat(tiles_d, _dev(x), N, name=name)           # perm looked up by src0->name (rrs.cu:266-270)
            assert nmse(y.cpu().numpy(), ref["C"]) <= 1e-9, M
            _, isum = t4.mul_mat_debug(tiles_d, _dev(x), N, _dev(perm))
            assert np.array_equal(isum.cpu().numpy(), ref["isum"]), M     # int32 accumulators bit-exact at every batch
            # what the reordering is for: the W4A4 result tracks the F32 product of the ORIGINAL weights
            exact = x @ W.T
            assert nmse(y.cpu().numpy(), exact) < 5e-2, M
    finally:
        t4.clear_perms()


# ---- C4: Qwen3-32B (hidden 5120, ffn 25600), prefill, N-sharded over 2 / 4 / 8 ranks ---------------------------------
@pytest.mark.parametrize("world", [2, 4, 8])
def test_c4_qwen3_32b_prefill_n_shards(t4, orc, world):
    """full K, N cut to 2048 of the 25600 channels and M to 384 tokens so the oracle finishes in seconds; the shard
    arithmetic (byte ranges at multiples of 128 channels) is the full-size one"""
    rng = np.random.default_rng(320 + world)
    M, N, K = 384, 2048, 5120
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    td, xd = _dev(tiles), _dev(x)
    ws = t4.Workspace.for_shape(M, N, K)
    full = t4.mul_mat(td, xd, N, workspace=ws).clone()
    parts = [t4.ShardedWeight(N, K, world, r, tiles_full=td).local(xd, workspace=ws).clone() for r in range(world)]
    assert sum(p.shape[1] for p in parts) == N
    assert torch.equal(torch.cat(parts, dim=1), full)                     # column blocks are bit-identical to the whole
    rows = [0, 129, M - 1]
    ref = orc.mul_mat(tiles, x[rows], N)
    assert nmse(full[rows].cpu().numpy(), ref["C"]) <= 1e-9


def test_c4_full_size_shard_ranges(t4):
    for N, K in [(25600, 5120), (5120, 25600), (8192, 5120), (1024, 5120)]:
        for world in (2, 4, 8):
            edges = [t4.shard_range(N, K, world, r, 128) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == N
            for (a0, a1, off, ln), (b0, _, boff, _) in zip(edges, edges[1:]):
                assert a1 == b0 and off + ln == boff and ln == (a1 - a0) // 8 * (K // 256) * 1184
            assert all(e[0] % 128 == 0 for e in edges)


# ---- C5: Llama-3-70B (hidden 8192, ffn 28672), column-parallel over 8 ranks, batch sweep ----------------------------
@pytest.mark.parametrize("M", [1, 2, 3, 8, 16, 17, 64, 128, 129, 512])
def test_c5_llama3_70b_column_parallel_batch_sweep(t4, orc, M):
    """N = 1024 of the 28672 channels (128 per rank), full K = 8192; oracle on up to 32 rows of the batch"""
    rng = np.random.default_rng(700 + M)
    N, K, world = 1024, 8192, 8
    W = make_weights(rng, N, K)
    x = make_acts(rng, M, K)
    tiles = orc.quantize_tiles(W)
    td, xd = _dev(tiles), _dev(x)
    ws = t4.Workspace.for_shape(M, N, K)
    parts = [t4.ShardedWeight(N, K, world, r, tiles_full=td).local(xd, workspace=ws).clone() for r in range(world)]
    y = torch.cat(parts, dim=1)
    rows = sorted(set(np.linspace(0, M - 1, min(M, 32)).astype(int).tolist()))
    ref = orc.mul_mat(tiles, x[rows], N, want_isum=True)
    assert nmse(y[rows].cpu().numpy(), ref["C"]) <= 1e-9
    _, isum = t4.mul_mat_debug(td, xd, N, workspace=ws)
    assert np.array_equal(isum[rows].cpu().numpy(), ref["isum"])


@pytest.mark.parametrize("M", [1, 64, 1024, 8192])
def test_c5_full_size_properties(t4, M):
    """the real 70B matrices (synthetic tiles, no oracle run): exact scaling, finite outputs, row independence"""
    N, K = 8192, 8192
    tiles = t4.random_tiles(N, K, seed=70)
    g = torch.Generator(device="cuda")
    g.manual_seed(M)
    x = torch.randn((M, K), device="cuda", generator=g)
    ws = t4.Workspace.for_shape(M, N, K)
    y = t4.mul_mat(tiles, x, N, workspace=ws).clone()
    assert torch.isfinite(y).all()
    assert torch.equal(t4.mul_mat(tiles, x * 0.5, N, workspace=ws), y * 0.5)
    r = M // 2
    y1 = t4.mul_mat(tiles, x[r:r + 1].contiguous(), N, kernel=t4.KERNEL_GEMV)
    assert nmse(y1.cpu().numpy(), y[r:r + 1].cpu().numpy()) <= 1e-9
