"""tcq4_b200_mul_mat_static -- the allocation-free entry point (scratch in a file-scope __device__ array of the library).

No counterpart in the reference.  Collected LAST on purpose:
the new entry is checked through the C ABI against the existing tcq4_b200_mul_mat on two calls in a
row -- same kernels and workspace layout, so the outputs must be bit-identical -- once on the decode GEMV and once on
the activation kernel + tcgen05 GEMM pair, which is the path that actually uses the scratch."""
import numpy as np
import pytest

from conftest import make_acts, make_weights

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch


@pytest.mark.parametrize("M", [1, 8, 130, 300])
def test_static_entry_matches_arena_entry_twice(t4, orc, torch_cuda, M):
    torch = torch_cuda
    rng = np.random.default_rng(100 + M)
    N, K = 384, 1024
    tiles_d = torch.from_numpy(orc.quantize_tiles(make_weights(rng, N, K))).cuda()
    out = torch.empty((M, N), dtype=torch.float32, device="cuda")
    for call in range(2):                                   # the array keeps the previous call's contents
        x = torch.from_numpy(make_acts(rng, M, K)).cuda()
        want = t4.mul_mat(tiles_d, x, N)
        torch.cuda.synchronize()
        out.fill_(float("nan"))
        torch.cuda.synchronize()
        a0 = torch.cuda.memory_allocated()
        got = t4.mul_mat_static(tiles_d, x, N, out=out)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == a0          # nothing allocated through torch on the call path
        assert got.data_ptr() == out.data_ptr()
        assert torch.equal(got, want), f"call {call}: static-scratch entry differs from tcq4_b200_mul_mat"
    # and against the oracle, within the tolerance every parity test uses (spec: 1e-3 NMSE; gate 1e-9)
    ref = orc.mul_mat(tiles_d.cpu().numpy(), x.cpu().numpy(), N)["C"]
    err = float(((got.cpu().numpy().astype(np.float64) - ref) ** 2).sum() / (ref.astype(np.float64) ** 2).sum())
    assert err < 1e-9, err


def test_static_entry_with_perm(t4, orc, torch_cuda):
    from conftest import block_perm
    torch = torch_cuda
    rng = np.random.default_rng(77)
    M, N, K = 40, 256, 512
    perm = block_perm(rng, K)
    tiles_d = torch.from_numpy(orc.quantize_tiles(make_weights(rng, N, K))).cuda()
    perm_d = torch.from_numpy(perm).cuda()
    for _ in range(2):
        x = torch.from_numpy(make_acts(rng, M, K)).cuda()
        want = t4.mul_mat(tiles_d, x, N, perm=perm_d)
        got = t4.mul_mat_static(tiles_d, x, N, perm=perm_d)
        torch.cuda.synchronize()
        assert torch.equal(got, want)


def test_static_entry_rejects_oversize(t4, torch_cuda):
    torch = torch_cuda
    M, N, K = 4096, 9728, 2560                              # needs ~30 MB of scratch; the array holds 8 MiB
    assert t4.workspace_size(M, N, K) > (8 << 20)
    tiles_d = torch.zeros(t4.tile_bytes(N, K), dtype=torch.uint8, device="cuda")
    x = torch.zeros((M, K), dtype=torch.float32, device="cuda")
    with pytest.raises(t4.Tcq4Error) as ei:
        t4.mul_mat_static(tiles_d, x, N)
    assert "static array holds" in str(ei.value)
