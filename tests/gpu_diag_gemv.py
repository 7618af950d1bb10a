"""GPU diagnostic (not a pytest file): GEMV phase timeline and the back-to-back (PDL) chain check.
Usage on the box:  python tests/gpu_diag_gemv.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import make_acts, make_weights  # noqa: E402

t4 = graft.load_package()
orc = graft.load_oracle().Oracle()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def phases(M, N, K):
    tiles = t4.random_tiles(N, K, seed=1)
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    for _ in range(3):
        t4.mul_mat_debug(tiles, x, N, kernel=t4.KERNEL_GEMV, workspace=ws)
    torch.cuda.synchronize()
    raw = np.zeros(160 * 16 + 16 * 8 * 3, dtype=np.uint64)
    t4.diag_lib().tcq4_b200_debug_gemv_phases(raw.ctypes.data, raw.nbytes)
    buf = raw[:160 * 16].reshape(160, 16)
    it = raw[160 * 16:].reshape(16, 8, 3).astype(np.int64)
    b = buf[:148, :6].astype(np.int64)
    t0 = b[:, 0].min()
    rel = b - t0
    names = ["start", "prefetch", "dep", "act", "stream", "end"]
    print(f"[phases M={M} N={N} K={K}] (ns since first CTA start; min / median / max over CTAs)")
    for i, n in enumerate(names):
        print(f"   {n:9s} {rel[:, i].min():7d} {int(np.median(rel[:, i])):7d} {rel[:, i].max():7d}")
    it0 = it - b[0, 0]
    for wi in (0, 5, 15):
        print(f"   cta0 warp{wi} tiles (wait_begin, ready, done): " + " ".join(
            f"[{it0[wi, i, 0]},{it0[wi, i, 1]},{it0[wi, i, 2]}]" for i in range(7) if it[wi, i, 2] > b[0, 0]))
    d = np.diff(b, axis=1)
    print("   durations median:", {names[i + 1]: int(np.median(d[:, i])) for i in range(5)})


def chain_check():
    rng = np.random.default_rng(21)
    dims = [2560, 1024, 2560, 512, 1024, 2560]
    M = 3
    Ws = [dev(orc.quantize_tiles(make_weights(rng, dims[i + 1], dims[i]))) for i in range(len(dims) - 1)]
    x0 = dev(make_acts(rng, M, dims[0]))
    ws = t4.Workspace.for_shape(M, max(dims), max(dims))
    bufs = [torch.empty((M, d), device="cuda") for d in dims[1:]]

    def chain(sync):
        cur = x0
        for W, out, n in zip(Ws, bufs, dims[1:]):
            t4.mul_mat(W, cur, n, out=out, workspace=ws, kernel=t4.KERNEL_GEMV)
            if sync:
                torch.cuda.synchronize()
            cur = out
        torch.cuda.synchronize()
        return [b.clone() for b in bufs]

    want = chain(True)
    want2 = chain(True)
    print("sync chain repeatable:", [bool(torch.equal(a, b)) for a, b in zip(want, want2)])
    for rep in range(3):
        got = chain(False)
        for i, (a, b) in enumerate(zip(got, want)):
            if not torch.equal(a, b):
                diff = (a - b).abs()
                bad = (diff > 0).nonzero()
                print(f"   rep {rep} stage {i} (N={dims[i + 1]} K={dims[i]}): {bad.shape[0]} differ, max abs {diff.max().item():.3e},"
                      f" rel {(diff.max() / b.abs().max()).item():.3e}; first {bad[:6].tolist()}")
            else:
                print(f"   rep {rep} stage {i}: equal")


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    chain_check()
    for (M, N, K) in [(1, 9728, 2560), (1, 2560, 9728), (1, 1024, 2560), (4, 9728, 2560), (16, 9728, 2560)]:
        phases(M, N, K)
