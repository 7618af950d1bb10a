"""MUL_MAT_ID for TCQ4 expert tensors (reference: ggml_cuda_mul_mat_id, ggml-cuda.cu:2287-2436, which sorts rows by
expert on the host and calls the TCQ4 MUL_MAT once per expert).  The oracle is the MUL_MAT oracle applied per (token, slot) with
the expert the ids name.  With more than 8 tokens the routing runs on the device (three launches, capturable; round 2) and
every row equals the plain MUL_MAT of its expert up to the fp32 summation order over 256-blocks; with TCQ4_MOE_HOSTSORT=1 (the
reference's scheme, kept for comparison; child process) rows are BIT-IDENTICAL to the plain MUL_MAT of the same expert and batch."""
import os

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


def _experts(orc, rng, n_expert, N, K, perm=None):
    Ws = [make_weights(rng, N, K) for _ in range(n_expert)]
    tiles = [orc.quantize_tiles(orc.apply_perm_weights(W, perm) if perm is not None else W) for W in Ws]
    return tiles, _dev(np.stack(tiles))                    # device: [n_expert][tile bytes]


def _oracle(orc, tiles, x, ids, N, perm=None):
    n_tokens, x_slots, K = x.shape
    n_used = ids.shape[1]
    ref = np.zeros((n_tokens, n_used, N), np.float32)
    for t in range(n_tokens):
        for s in range(n_used):
            ref[t, s] = orc.mul_mat(tiles[int(ids[t, s])], x[t, s % x_slots][None, :], N, perm)["C"][0]
    return ref


@pytest.mark.parametrize("x_slots_full", [False, True])
def test_decode_one_token_device_routed(t4, orc, x_slots_full):
    """n_tokens == 1: one GEMV per slot, the expert id is read by the kernel (no host round trip)"""
    rng = np.random.default_rng(7 + x_slots_full)
    n_expert, n_used, N, K = 8, 4, 512, 1024
    perm = np.concatenate([b * 256 + rng.permutation(256) for b in range(K // 256)]).astype(np.int32)
    tiles, td = _experts(orc, rng, n_expert, N, K, perm)
    x = make_acts(rng, n_used if x_slots_full else 1, K)[None]          # [1][x_slots][K]
    ids = rng.permutation(n_expert)[:n_used].astype(np.int32)[None]     # [1][n_used]
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    y = t4.mul_mat_id(td, _dev(x), _dev(ids), N, perm=_dev(perm), err_flag=err)
    assert err.item() == 0
    assert nmse(y.cpu().numpy(), _oracle(orc, tiles, x, ids, N, perm)) <= 1e-9
    for s in range(n_used):
        # same int32 accumulators; the grouped launch may cut a channel's 256-blocks between warps elsewhere (fp32 sum order)
        plain = t4.mul_mat(td[int(ids[0, s])], _dev(x[0, s % x.shape[1]][None]), N, perm=_dev(perm))
        assert nmse(y[0, s].cpu().numpy(), plain[0].cpu().numpy()) <= 1e-12, s


def test_decode_is_graph_capturable_and_follows_device_ids(t4, orc):
    rng = np.random.default_rng(11)
    n_expert, n_used, N, K = 6, 2, 256, 512
    tiles, td = _experts(orc, rng, n_expert, N, K)
    x = make_acts(rng, 1, K)[None]
    xd = _dev(x)
    ids_d = torch.zeros((1, n_used), dtype=torch.int32, device="cuda")
    out = torch.empty((1, n_used, N), dtype=torch.float32, device="cuda")
    t4.reserve(1, N, K)
    t4.mul_mat_id(td, xd, ids_d, N, out=out)                            # warm-up outside capture
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            t4.mul_mat_id(td, xd, ids_d, N, out=out)
    for pick in ([5, 1], [0, 3], [2, 2]):
        ids = np.array([pick], np.int32)
        ids_d.copy_(_dev(ids))                                          # the router's output changes ON THE DEVICE only
        g.replay()
        torch.cuda.synchronize()
        assert nmse(out.cpu().numpy(), _oracle(orc, tiles, x, ids, N)) <= 1e-9, pick


@pytest.mark.parametrize("x_slots_full", [False, True])
def test_decode_batch_of_tokens_device_routed(t4, orc, x_slots_full):
    """2..8 tokens: one grouped launch per token, still no host round trip (ids with a row stride)"""
    rng = np.random.default_rng(17 + x_slots_full)
    n_expert, n_used, n_tokens, N, K = 6, 3, 5, 128, 768
    tiles, td = _experts(orc, rng, n_expert, N, K)
    wide = np.stack([rng.permutation(n_expert) for _ in range(n_tokens)]).astype(np.int32)
    ids_view = _dev(wide)[:, :n_used]                                   # stride(0) = n_expert
    x = np.stack([make_acts(rng, n_used if x_slots_full else 1, K) for _ in range(n_tokens)])
    c0 = t4.launch_count()
    y = t4.mul_mat_id(td, _dev(x), ids_view, N)
    if not __import__("os").environ.get("TCQ4_MOE_NOGROUP"):
        assert t4.launch_count() - c0 == n_tokens
    assert nmse(y.cpu().numpy(), _oracle(orc, tiles, x, wide[:, :n_used], N)) <= 1e-9


def test_decode_bad_expert_id_sets_flag(t4, orc):
    rng = np.random.default_rng(12)
    tiles, td = _experts(orc, rng, 2, 64, 256)
    x = _dev(make_acts(rng, 1, 256)[None])
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    t4.mul_mat_id(td, x, _dev(np.array([[1, 7]], np.int32)), 64, err_flag=err)
    torch.cuda.synchronize()
    assert err.item() == 1


@pytest.mark.parametrize("x_slots_full", [False, True])
def test_prefill_sorted_by_expert(t4, orc, x_slots_full):
    """n_tokens > 8: rows binned by expert (device-routed chunks, or host counting sort with TCQ4_MOE_HOSTSORT=1); expert 0 gets
    20 rows, others fewer, one none"""
    rng = np.random.default_rng(21 + x_slots_full)
    n_expert, n_used, n_tokens, N, K = 5, 2, 24, 256, 512
    tiles, td = _experts(orc, rng, n_expert, N, K)
    ids = np.zeros((n_tokens, n_used), np.int32)
    for t in range(n_tokens):
        ids[t, 0] = 0 if t < 20 else 1                                  # 20 rows for expert 0
        ids[t, 1] = 1 + (t % 3)                                         # experts 1..3; expert 4 never used
    x = np.stack([make_acts(rng, n_used if x_slots_full else 1, K) for _ in range(n_tokens)])
    y = t4.mul_mat_id(td, _dev(x), _dev(ids), N)
    assert nmse(y.cpu().numpy(), _oracle(orc, tiles, x, ids, N)) <= 1e-9
    # bit-identical to the plain MUL_MAT of each expert over its rows in (token, slot) order
    for e in range(n_expert):
        rows = [(t, s) for t in range(n_tokens) for s in range(n_used) if ids[t, s] == e]
        if not rows:
            continue
        xe = np.stack([x[t, s % x.shape[1]] for t, s in rows])
        plain = t4.mul_mat(td[e], _dev(xe), N)
        got = torch.stack([y[t, s] for t, s in rows])
        if os.environ.get("TCQ4_MOE_HOSTSORT"):
            assert torch.equal(got, plain), e
        else:
            assert nmse(got.cpu().numpy(), plain.cpu().numpy()) <= 1e-12, e
    y2 = t4.mul_mat_id(td, _dev(x), _dev(ids), N)
    assert torch.equal(y, y2)                                           # deterministic whatever slot a row lands in


def test_prefill_strided_ids_and_ggml_mirror_with_registered_perm(t4, orc):
    """ids as a view of a wider [n_tokens][n_expert] top-k tensor (ggml nb[1] > n_used * 4); perm looked up by tensor name"""
    rng = np.random.default_rng(31)
    n_expert, n_used, n_tokens, N, K = 4, 2, 9, 128, 512
    name = "blk.3.ffn_up_exps.weight"
    perm = np.concatenate([b * 256 + rng.permutation(256) for b in range(K // 256)]).astype(np.int32)
    tiles, td = _experts(orc, rng, n_expert, N, K, perm)
    wide = np.stack([rng.permutation(n_expert) for _ in range(n_tokens)]).astype(np.int32)
    wide_d = _dev(wide)
    ids_view = wide_d[:, :n_used]                                       # stride(0) = n_expert
    x = np.stack([make_acts(rng, 1, K) for _ in range(n_tokens)])
    ref = _oracle(orc, tiles, x, wide[:, :n_used], N, perm)
    y = t4.mul_mat_id(td, _dev(x), ids_view, N, perm=_dev(perm))
    assert nmse(y.cpu().numpy(), ref) <= 1e-9
    t4.clear_perms()
    t4.set_reorder_enabled(True)
    t4.register_perm(name, perm)
    try:
        y2 = t4.ggml_mul_mat_id(td, _dev(x), ids_view, N, name=name)
        assert torch.equal(y2, y)
    finally:
        t4.clear_perms()


def test_prefill_bad_expert_id_is_an_error(t4, orc):
    """more than 8 tokens.  Host-sorted path: the call fails (the ids are read on the host).  Device-routed path: the ids never
    reach the host -- the flag is raised and the row skipped, as in the decode path; every other row is computed."""
    rng = np.random.default_rng(41)
    tiles, td = _experts(orc, rng, 2, 64, 256)
    xs = np.stack([make_acts(rng, 1, 256) for _ in range(9)])
    x = _dev(xs)
    good = [[0], [1], [1], [0], [0], [1], [0], [1]]
    if os.environ.get("TCQ4_MOE_HOSTSORT"):
        with pytest.raises(Exception, match="expert id"):
            t4.mul_mat_id(td, x, _dev(np.array(good + [[2]], np.int32)), 64)
        with pytest.raises(Exception):
            t4.mul_mat_id(td, x, _dev(np.array([[-1]] + good, np.int32)), 64)
        return
    for bad_ids in (good + [[2]], [[-1]] + good):
        ids = np.array(bad_ids, np.int32)
        err = torch.zeros(1, dtype=torch.int32, device="cuda")
        out = torch.full((9, 1, 64), 123.0, device="cuda")
        t4.mul_mat_id(td, x, _dev(ids), 64, out=out, err_flag=err)
        torch.cuda.synchronize()
        assert err.item() == 1
        ok = [t for t in range(9) if 0 <= ids[t, 0] < 2]
        ref = _oracle(orc, tiles, xs[ok], ids[ok], 64)
        assert nmse(out[ok].cpu().numpy(), ref) <= 1e-9
        bad = [t for t in range(9) if not 0 <= ids[t, 0] < 2]
        assert (out[bad] == 123.0).all()                                # the skipped row is not written


def test_prefill_many_tokens_several_chunks_graph_capture(t4, orc):
    """200 tokens x 4 slots over 16 experts with a skewed router: experts with more than 64 rows span several chunks, some get
    none; a channel permutation; the whole op captured in a CUDA graph and replayed after the ids changed ON THE DEVICE."""
    rng = np.random.default_rng(51)
    n_expert, n_used, n_tokens, N, K = 16, 4, 200, 136, 512
    perm = np.concatenate([b * 256 + rng.permutation(256) for b in range(K // 256)]).astype(np.int32)
    tiles, td = _experts(orc, rng, n_expert, N, K, perm)
    pr = np.array([30, 20, 10, 8, 6, 5, 4, 3, 3, 3, 2, 2, 2, 1, 1, 0], np.float64)
    pr /= pr.sum()

    def route():
        return np.stack([rng.choice(n_expert, n_used, replace=False, p=pr) for _ in range(n_tokens)]).astype(np.int32)
    ids1, ids2 = route(), route()
    assert np.bincount(ids1.ravel(), minlength=n_expert).max() > 128
    x = np.stack([make_acts(rng, 1, K) for _ in range(n_tokens)])
    xd, pd = _dev(x), _dev(perm)
    ids_d = _dev(ids1)
    out = torch.empty((n_tokens, n_used, N), device="cuda")
    err = torch.zeros(1, dtype=torch.int32, device="cuda")
    t4.mul_mat_id(td, xd, ids_d, N, perm=pd, out=out, err_flag=err)      # warm-up: grows the arena outside the capture
    torch.cuda.synchronize()
    n0 = t4.launch_count()
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            t4.mul_mat_id(td, xd, ids_d, N, perm=pd, out=out, err_flag=err, stream=st)
    torch.cuda.current_stream().wait_stream(st)
    if not os.environ.get("TCQ4_MOE_HOSTSORT"):
        assert t4.launch_count() - n0 == 3                                # routing, activation stage, grouped GEMM
    for ids in (ids1, ids2):
        ids_d.copy_(_dev(ids))
        out.fill_(float("nan"))
        g.replay()
        torch.cuda.synchronize()
        rows = rng.choice(n_tokens, 24, replace=False)                    # the scalar oracle on a sample of the tokens
        ref = _oracle(orc, tiles, x[rows], ids[rows], N, perm)
        assert nmse(out[rows].cpu().numpy(), ref) <= 1e-9
        assert not torch.isnan(out).any()
        assert err.item() == 0


def test_prefill_host_sorted_scheme_in_a_child_process():
    """TCQ4_MOE_HOSTSORT=1 (read once at load, hence the child): the reference's scheme -- ids to the host, counting sort, one
    MUL_MAT per expert -- with its bit-identity to the per-expert plain MUL_MAT."""
    import subprocess
    import sys
    if os.environ.get("TCQ4_MOE_HOSTSORT"):
        pytest.skip("already the child")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TCQ4_MOE_HOSTSORT="1")
    out = subprocess.run([sys.executable, "-m", "pytest", "tests/test_moe.py", "-m", "gpu", "-q", "-x", "-k",
                          "prefill and not child and not graph_capture"],
                         cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and " passed" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]


def test_decode_per_slot_fallback_in_a_child_process():
    """TCQ4_MOE_NOGROUP=1 (read once at load, hence the child): one GEMV per slot instead of the grouped launch -- the path
    taken when the slots' activation rows do not fit the GEMV's shared memory.  Same decode tests, same bar."""
    import os
    import subprocess
    import sys
    if os.environ.get("TCQ4_MOE_NOGROUP"):
        pytest.skip("already the child")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TCQ4_MOE_NOGROUP="1")
    out = subprocess.run([sys.executable, "-m", "pytest", "tests/test_moe.py", "-m", "gpu", "-q", "-x", "-k", "decode and not child"],
                         cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and " passed" in out.stdout, out.stdout[-3000:] + out.stderr[-2000:]


def test_prefill_fuzz_device_routed_against_per_expert_calls(t4):
    """random (experts, slots, tokens, shapes, one row per token or per slot): the device-routed op against one plain MUL_MAT per
    expert over that expert's rows"""
    rng = np.random.default_rng(61)
    for _ in range(10):
        n_expert = int(rng.integers(1, 40))
        n_used = int(rng.integers(1, min(n_expert, 6) + 1))
        n_tokens = int(rng.integers(9, 180))
        N, K = int(rng.integers(1, 40)) * 8, int(rng.integers(1, 5)) * 256
        x_slots = n_used if rng.random() < 0.5 else 1
        td = torch.stack([t4.random_tiles(N, K, seed=1000 + e) for e in range(n_expert)])
        ids = np.stack([rng.choice(n_expert, n_used, replace=False) for _ in range(n_tokens)]).astype(np.int32)
        x = torch.randn((n_tokens, x_slots, K), device="cuda")
        y = t4.mul_mat_id(td, x, _dev(ids), N)
        for e in range(n_expert):
            rows = [(t, s) for t in range(n_tokens) for s in range(n_used) if ids[t, s] == e]
            if not rows:
                continue
            xe = torch.stack([x[t, s % x_slots] for t, s in rows])
            plain = t4.mul_mat(td[e], xe, N, kernel=t4.KERNEL_GEMM2)
            got = torch.stack([y[t, s] for t, s in rows])
            assert nmse(got.cpu().numpy(), plain.cpu().numpy()) <= 1e-12, (n_expert, n_used, n_tokens, N, K, e)
