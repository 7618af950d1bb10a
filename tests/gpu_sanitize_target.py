"""compute-sanitizer target (not a pytest file): a small-shape parity subset through every kernel family -- K1 (both staging
formats), K3 (dp4a and IMMA loops, segmented, pre-pass), the kind::i8 pair GEMM (narrow and wide NTL, bias / residual), the
kind::f16 pair GEMM (fan-out segments, K split), the device quantiser and GET_ROWS -- each checked against the oracle so a
silent corruption would also show.   compute-sanitizer --tool memcheck|racecheck|synccheck python tests/gpu_sanitize_target.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import block_perm, make_acts, make_weights, nmse  # noqa: E402

t4 = graft.load_package()
orc = graft.load_oracle().Oracle()
rng = np.random.default_rng(5)
which = sys.argv[1] if len(sys.argv) > 1 else "all"


def check(name, y, ref, tol=1e-9):
    e = nmse(y.cpu().numpy(), ref)
    print(f"{name}: nmse {e:.2e}", flush=True)
    assert e <= tol, name


def mm(M, N, K, kernel, perm=False, name=""):
    p = block_perm(rng, K) if perm else None
    W = make_weights(rng, N, K)
    tiles = orc.quantize_tiles(orc.apply_perm_weights(W, p) if perm else W)
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(tiles, x, N, p)["C"]
    td, xd = torch.from_numpy(tiles).cuda(), torch.from_numpy(x).cuda()
    pd = torch.from_numpy(p).cuda() if perm else None
    ws = t4.Workspace.for_shape(M, N, K)
    for rep in range(2):                       # back-to-back launches (dependent-launch chains)
        y = t4.mul_mat(td, xd, N, perm=pd, workspace=ws, kernel=kernel)
    torch.cuda.synchronize()
    check(f"{name} M={M} N={N} K={K}", y, ref)


if which in ("all", "gemv"):
    mm(1, 136, 512, t4.KERNEL_GEMV, name="gemv dp4a")
    mm(3, 64, 768, t4.KERNEL_GEMV, perm=True, name="gemv dp4a perm")
    mm(8, 136, 512, t4.KERNEL_GEMV, name="gemv imma")
    mm(16, 64, 1024, t4.KERNEL_GEMV, name="gemv imma 16")
if which in ("all", "gemm"):
    mm(40, 136, 512, t4.KERNEL_GEMM, name="single-CTA gemm")
if which in ("all", "gemm2"):
    mm(130, 40, 512, t4.KERNEL_GEMM2, name="pair gemm i8 narrow")
    mm(300, 264, 768, t4.KERNEL_GEMM2, perm=True, name="pair gemm i8 wide")
if which in ("all", "gemm3"):
    mm(130, 40, 512, t4.KERNEL_GEMM3, name="pair gemm f16")
    mm(300, 392, 2048, t4.KERNEL_GEMM3, perm=True, name="pair gemm f16 K split")
    # fan-out: three matrices, one launch
    K, Ns, M = 512, [136, 64, 200], 140
    Ws = [make_weights(rng, n, K) for n in Ns]
    tl = [orc.quantize_tiles(w) for w in Ws]
    x = make_acts(rng, M, K)
    fo = t4.Fanout([torch.from_numpy(t).cuda() for t in tl], Ns, K)
    outs = fo(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    for t, n, o in zip(tl, Ns, outs):
        check(f"fan-out f16 N={n}", o, orc.mul_mat(t, x, n)["C"])
    # decode fan-out (segmented GEMV)
    x1 = make_acts(rng, 2, K)
    outs = fo(torch.from_numpy(x1).cuda())
    torch.cuda.synchronize()
    for t, n, o in zip(tl, Ns, outs):
        check(f"fan-out gemv N={n}", o, orc.mul_mat(t, x1, n)["C"])
if which in ("all", "misc"):
    # bias / residual epilogue, device quantiser, GET_ROWS, stand-alone activation stage
    M, N, K = 130, 136, 512
    W = make_weights(rng, N, K)
    tiles = orc.quantize_tiles(W)
    x = make_acts(rng, M, K)
    res = rng.standard_normal((M, N)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    y = t4.mul_mat_add(torch.from_numpy(tiles).cuda(), torch.from_numpy(x).cuda(), N, residual=torch.from_numpy(res).cuda(),
                       bias=torch.from_numpy(bias).cuda())
    torch.cuda.synchronize()
    check("bias + residual", y, orc.mul_mat(tiles, x, N)["C"] + res + bias[None, :])
    q = t4.quantize_tiles(torch.from_numpy(W).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(q.cpu().numpy(), tiles), "device quantiser"
    blocks = t4.fwht_quantize_tc(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(blocks.cpu().numpy().reshape(-1), orc.quantize_act_blocks(x)), "activation stage"
    print("quantiser / activation stage byte-identical", flush=True)
print("SANITIZE_TARGET_OK", flush=True)
