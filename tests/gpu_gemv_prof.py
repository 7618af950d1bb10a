"""GPU diagnostic (not a pytest file): GEMV phase stamps of the PRODUCT arithmetic (no isum dump): TCQ4_GV_PROF=1
selects the stamped instantiation.  Usage on the box:  python tests/gpu_gemv_prof.py"""
import os
import sys

os.environ["TCQ4_GV_PROF"] = "1"
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()


def phases(M, N, K):
    tiles = t4.random_tiles(N, K, seed=1)
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    out = torch.empty((M, N), device="cuda")
    for _ in range(3):
        t4.mul_mat(tiles, x, N, out=out, workspace=ws, kernel=t4.KERNEL_GEMV)
    torch.cuda.synchronize()
    raw = np.zeros(160 * 16 + 16 * 8 * 3, dtype=np.uint64)
    t4.diag_lib().tcq4_b200_debug_gemv_phases(raw.ctypes.data, raw.nbytes)
    b8 = raw[:160 * 16].reshape(160, 16)[:148].astype(np.int64)
    b = b8[:, :6]
    it = raw[160 * 16:].reshape(16, 8, 3).astype(np.int64)
    t0 = b[:, 0].min()
    rel = b - t0
    ends = rel[:, 5][rel[:, 5] > 0]          # CTA 1 of a cluster returns before the last stamp
    names = ["start", "prefetch", "dep", "act", "stream", "end"]
    print(f"[phases M={M} N={N} K={K}] ns since first CTA start; min / median / max over CTAs")
    for i, n in enumerate(names[:5]):
        print(f"   {n:9s} {rel[:, i].min():7d} {int(np.median(rel[:, i])):7d} {rel[:, i].max():7d}")
    print(f"   {'end':9s} {ends.min():7d} {int(np.median(ends)):7d} {ends.max():7d}")
    a1 = b8[:, 6] - t0
    cb = (b8[:, 7] - t0)[rel[:, 5] > 0]
    print(f"   (x loaded {int(np.median(b8[:, 8] - t0))}; quantised {int(np.median(b8[:, 9] - t0))})")
    print(f"   (first act block stored {int(np.median(a1))}; CTA 0: partials landed {int(np.median(cb))})")
    it0 = it - b[0, 0]
    for wi in (0, 15):
        print(f"   cta0 warp{wi} tiles (wait_begin, ready, done): " + " ".join(
            f"[{it0[wi, i, 0]},{it0[wi, i, 1]},{it0[wi, i, 2]}]" for i in range(7) if it[wi, i, 2] > b[0, 0]))


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    for shp in [(1, 9728, 2560), (1, 2560, 9728), (1, 4096, 2560), (1, 1024, 2560), (4, 9728, 2560), (8, 9728, 2560),
                (16, 9728, 2560)]:
        phases(*shp)
