"""Epilogue fusion: y = W x + bias + residual inside the MUL_MAT (tcq4_b200_mul_mat_add).  Oracle: the MUL_MAT
oracle followed by the two fp32 adds, as ggml_add would do them; the fused kernels differ in summation order only."""
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


@pytest.mark.parametrize("M", [1, 3, 8, 16, 40, 64, 129, 300])  # GEMV dp4a / IMMA loops, the batched-decode GEMM (9..64), the pair GEMM
@pytest.mark.parametrize("N,K", [(136, 512), (1024, 1024), (256, 4096)])      # a ragged last column tile, a wide shallow matrix,
                                                                              # a narrow deep one (K parts with the last arriver)
def test_bias_and_residual_fused(t4, orc, M, N, K):
    rng = np.random.default_rng(M * 7 + N)
    perm = block_perm(rng, K)
    tiles = orc.quantize_tiles(orc.apply_perm_weights(make_weights(rng, N, K), perm))
    x = make_acts(rng, M, K)
    res = rng.standard_normal((M, N)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    ref = orc.mul_mat(tiles, x, N, perm)["C"]
    td, xd, pd = (torch.from_numpy(a).cuda() for a in (tiles, x, perm))
    rd, bd = torch.from_numpy(res).cuda(), torch.from_numpy(bias).cuda()
    plain = t4.mul_mat(td, xd, N, perm=pd)
    for use_res, use_bias in ((True, True), (True, False), (False, True), (False, False)):
        want = ref + (res if use_res else 0) + (bias[None, :] if use_bias else 0)
        y = t4.mul_mat_add(td, xd, N, residual=rd if use_res else None, bias=bd if use_bias else None, perm=pd)
        assert nmse(y.cpu().numpy(), want) <= 1e-9, (use_res, use_bias)
        # against this library's own MUL_MAT followed by torch adds: only the fp32 summation order may differ
        sep = plain + (rd if use_res else 0) + (bd[None, :] if use_bias else 0)
        assert float((y - sep).abs().max()) <= 1e-5 * float(sep.abs().max() + 1.0), (use_res, use_bias)
        if not use_res and not use_bias:
            assert torch.equal(y, plain)


@pytest.mark.parametrize("M", [2, 24, 64])
@pytest.mark.parametrize("N,K", [(256, 512), (256, 6144), (19200, 256)])
def test_residual_in_place(t4, orc, M, N, K):
    """the residual stream updated in place: out is the residual tensor (ggml_add inplace); also where the batched-decode GEMM
    would otherwise cut blocks (K parts, stream-K): with a residual in y the parts must not meet there through red.add"""
    rng = np.random.default_rng(5 + M)
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    x = make_acts(rng, M, K)
    res = rng.standard_normal((M, N)).astype(np.float32)
    td, xd, h = torch.from_numpy(tiles).cuda(), torch.from_numpy(x).cuda(), torch.from_numpy(res).cuda()
    t4.mul_mat_add(td, xd, N, residual=h, out=h)
    rows = slice(0, 4) if N * K * M > 2e8 else slice(0, M)
    assert nmse(h[rows].cpu().numpy(), orc.mul_mat(tiles, x[rows], N)["C"] + res[rows]) <= 1e-9
    h2 = torch.from_numpy(res).cuda()
    t4.mul_mat_add(td, xd, N, residual=h2, out=h2)
    assert torch.equal(h, h2)                                    # deterministic


@pytest.mark.gpu
def test_epilogue_add_through_the_f16_pair_gemm_in_a_child_process():
    """TCQ4_AUTO_GEMM3=1 (read once at load, hence the child) routes single matrices to the kind::f16 pair GEMM as well: its
    bias / residual instantiation (incl. the K split meeting a pre-added residual) must pass the same checks."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TCQ4_AUTO_GEMM3="1")
    out = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", os.path.join(root, "tests", "test_epilogue_add.py"),
                          "-k", "not child_process"], cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
