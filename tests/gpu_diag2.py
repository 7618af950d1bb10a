"""GPU diagnostic (not a pytest file): localise fp32 differences of the tcgen05 GEMM against the oracle.
Usage on the box:  python tests/gpu_diag2.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import make_acts, make_weights, nmse  # noqa: E402

t4 = graft.load_package()
orc = graft.load_oracle().Oracle()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def localise(M, N, K, seed, debug):
    rng = np.random.default_rng(seed)
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    x = make_acts(rng, M, K)
    Mo = min(M, 32)
    rows = np.sort(rng.choice(M, Mo, replace=False))
    ref = orc.mul_mat(tiles, x[rows], N, want_isum=True)
    td, xd = dev(tiles), dev(x)
    ws = t4.Workspace.for_shape(M, N, K)
    for rep in range(3):
        if debug:
            y, isum = t4.mul_mat_debug(td, xd, N, kernel=t4.KERNEL_GEMM, workspace=ws)
        else:
            y = t4.mul_mat(td, xd, N, workspace=ws, kernel=t4.KERNEL_GEMM)
        torch.cuda.synchronize()
        y = y.cpu().numpy()[rows]
        e = nmse(y, ref["C"])
        rel = np.abs(y - ref["C"]) / (np.abs(ref["C"]) + 1e-3 * np.abs(ref["C"]).mean())
        bad = rel > 1e-4
        print(f"[M={M} N={N} K={K} debug={debug} rep={rep}] nmse={e:.3e} bad={bad.sum()}/{bad.size}", flush=True)
        if bad.sum():
            print("   rows (m)      :", {int(rows[i]): int(bad[i].sum()) for i in range(Mo) if bad[i].sum()})
            print("   by n//128     :", {i: int(bad[:, i * 128:(i + 1) * 128].sum()) for i in range(N // 128)
                                         if bad[:, i * 128:(i + 1) * 128].sum()})
            print("   by (n%128)//8 :", [int(bad[:, [n for n in range(N) if (n % 128) // 8 == j]].sum()) for j in range(16)])
            print("   by n%8        :", [int(bad[:, i::8].sum()) for i in range(8)])
            idx = np.argwhere(bad)[:8]
            for i, n in idx:
                print(f"   m={rows[i]} n={n}: got {y[i, n]:.6g} want {ref['C'][i, n]:.6g} rel={rel[i, n]:.3e}")


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    localise(512, 9728, 2560, 300 + 512 + 9728 + 2560, True)
    localise(512, 9728, 2560, 300 + 512 + 9728 + 2560, False)
    localise(128, 9728, 2560, 1, False)
    localise(512, 9728, 1024, 2, False)
    localise(512, 2048, 2560, 3, False)
    localise(1024, 1024, 2560, 4, False)
