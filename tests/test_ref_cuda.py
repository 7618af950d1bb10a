"""O3 of SURVEY 8(c): the FORK's own CUDA kernels (tcq4_k32.cu compiled unmodified for sm_100a with the fork's
-use_fast_math into oracle/_ref/libtcq4_refcuda.so by `make -C oracle refcuda`) run on this box against the oracle.

  * tcq4_rrs_fwht_quantize_tc (tcq4_k32.cu:181-277): its int4 activation codes against the oracle's -- a mismatch RATE, not
    equality: the fork computes 7/max with fast-math reciprocals, so codes may differ by one step at rounding ties
    (SURVEY H3).  Bars: every difference is +-1, fewer than 1 % of the codes differ, scales agree to 1e-6 relative.
  * tcq4_rrs_fused_gemm_smallM (the live kernel, :733-889) and tcq4_rrs_fwht_quantize_tc + tcq4_rrs_gemm_imma (:1217-1412):
    fp32 outputs against the oracle within the spec tolerance (NMSE <= 1e-3), and this library's output on the same inputs
    within 1e-9 of the oracle beside them.
Skipped when the prebuilt library is absent (it is built where /root/reference exists and travels with the snapshot)."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import block_perm, make_acts, make_weights, nmse

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libtcq4_refcuda.so")


@pytest.fixture(scope="module")
def refcuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libtcq4_refcuda.so absent (built by `make -C oracle refcuda` where /root/reference exists)")
    lib = C.CDLL(LIB)
    vp, ci = C.c_void_p, C.c_int
    f = {}
    f["smallM"] = lib._Z26tcq4_rrs_fused_gemm_smallMPKfPKiPKvPfiiiP11CUstream_st
    f["smallM"].argtypes = [vp, vp, vp, vp, ci, ci, ci, vp]
    f["quant_tc"] = lib._Z25tcq4_rrs_fwht_quantize_tcPKfPviiP11CUstream_st
    f["quant_tc"].argtypes = [vp, vp, ci, ci, vp]
    f["imma"] = lib._Z18tcq4_rrs_gemm_immaPKvS0_PfiiiP11CUstream_st
    f["imma"].argtypes = [vp, vp, vp, ci, ci, ci, vp]
    for v in f.values():
        v.restype = None
    return f


def _codes(blocks):
    """block_rrs_int4_tc bytes [M][KT][160] -> int codes [M][KT][256] (qs[t] = q[2t] | q[2t+1] << 4, two's complement),
    scales [M][KT], group sums [M][KT][8]"""
    b = blocks.reshape(blocks.shape[0], blocks.shape[1], 160)
    qs = b[:, :, :128].astype(np.int16)
    lo, hi = qs & 0xF, qs >> 4
    q = np.stack([lo, hi], axis=-1).reshape(b.shape[0], b.shape[1], 256)
    q = np.where(q >= 8, q - 16, q)
    scale = b[:, :, 128:132].copy().view(np.float32)[:, :, 0]
    sums = b[:, :, 132:148].copy().view(np.int16)
    return q, scale, sums


@pytest.mark.parametrize("M,K", [(64, 1024), (8, 9728), (300, 2560)])
def test_fork_activation_kernel_code_mismatch_rate(t4, orc, refcuda, M, K):
    rng = np.random.default_rng(M + K)
    x = make_acts(rng, M, K)
    xd = torch.from_numpy(x).cuda()
    fork = torch.zeros((M, K // 256, 160), dtype=torch.uint8, device="cuda")
    refcuda["quant_tc"](xd.data_ptr(), fork.data_ptr(), K, M, None)
    torch.cuda.synchronize()
    want = orc.quantize_act_blocks(x)                                   # the oracle's blocks (pinned to the CPU reference)
    own = t4.fwht_quantize_tc(xd).cpu().numpy()
    assert np.array_equal(own.reshape(-1), np.asarray(want).reshape(-1))  # this library: byte-identical
    qf, sf, _ = _codes(fork.cpu().numpy())
    qo, so, _ = _codes(np.asarray(want).reshape(M, K // 256, 160))
    diff = qf - qo
    rate = float((diff != 0).mean())
    print(f"fork activation kernel vs oracle, M={M} K={K}: {int((diff != 0).sum())} of {diff.size} codes differ ({rate:.2e}); "
          f"max |diff| = {int(np.abs(diff).max())}; scale rel err {float(np.abs(sf / so - 1).max()):.2e}")
    assert np.abs(diff).max() <= 1                                      # only rounding ties move, by one step
    assert rate < 1e-2
    assert np.abs(sf / so - 1).max() < 1e-6


@pytest.mark.parametrize("M,N,K,use_perm", [(1, 256, 1024, False), (32, 512, 2560, True), (130, 256, 512, False)])
def test_fork_kernels_and_this_library_against_the_oracle(t4, orc, refcuda, M, N, K, use_perm):
    rng = np.random.default_rng(M * 3 + N)
    perm = block_perm(rng, K) if use_perm else None
    W = make_weights(rng, N, K)
    tiles = orc.quantize_tiles(orc.apply_perm_weights(W, perm) if use_perm else W)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(tiles, x, N, perm)["C"]
    td, xd = torch.from_numpy(tiles).cuda(), torch.from_numpy(x).cuda()
    pd = torch.from_numpy(perm).cuda() if use_perm else None
    y_live = torch.zeros((M, N), device="cuda")
    refcuda["smallM"](xd.data_ptr(), pd.data_ptr() if use_perm else None, td.data_ptr(), y_live.data_ptr(), M, N, K, None)
    torch.cuda.synchronize()
    e_live = nmse(y_live.cpu().numpy(), ref)
    e_own = nmse(t4.mul_mat(td, xd, N, perm=pd).cpu().numpy(), ref)
    msg = f"M={M} N={N} K={K} perm={use_perm}: NMSE vs oracle: fork live kernel {e_live:.2e}, this library {e_own:.2e}"
    if not use_perm:                                                    # the fork's IMMA path takes no permutation
        act = torch.zeros((M * (K // 256) * 160,), dtype=torch.uint8, device="cuda")
        y_imma = torch.zeros((M, N), device="cuda")
        refcuda["quant_tc"](xd.data_ptr(), act.data_ptr(), K, M, None)
        refcuda["imma"](act.data_ptr(), td.data_ptr(), y_imma.data_ptr(), M, N, K, None)
        torch.cuda.synchronize()
        e_imma = nmse(y_imma.cpu().numpy(), ref)
        msg += f", fork IMMA path {e_imma:.2e}"
        assert e_imma <= 1e-3
    print(msg)
    assert e_live <= 1e-3                                               # the spec's tolerance (north_star)
    assert e_own <= 1e-9
