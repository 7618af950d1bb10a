"""GPU diagnostic (not a pytest file): timeline of one cluster of the pair GEMM.  Run with TCQ4_G2_PROF=1.
Usage on the box:  TCQ4_G2_PROF=1 python tests/gpu_diag_gemm2.py [M N K]"""
import os
import sys

import numpy as np
import torch

os.environ.setdefault("TCQ4_G2_PROF", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 9728, 2560)
tiles = t4.random_tiles(N, K, seed=1)
x = torch.randn((M, K), device="cuda")
ws = t4.Workspace.for_shape(M, N, K)
out = torch.empty((M, N), device="cuda")
for _ in range(3):
    t4.mul_mat(tiles, x, N, out=out, workspace=ws, kernel=t4.KERNEL_GEMM2)
torch.cuda.synchronize()
raw = np.zeros(2 * 6 * 32 * 8 + 4096 * 6, dtype=np.int64)
rc = t4.diag_lib().tcq4_b200_debug_gemm2_timeline(raw.ctypes.data, raw.nbytes)
buf = raw[:2 * 6 * 32 * 8].reshape(2, 6, 32, 8)
cta = raw[2 * 6 * 32 * 8:].reshape(4096, 6)
cta = cta[cta[:, 1] > 0]
if len(cta):
    t_first = cta[:, 1].min()
    d = cta[:, 1:] - t_first
    print(f"{len(cta)} CTAs; kernel span {int(d[:, 4].max())} ns")
    ph = np.diff(cta[:, 1:], axis=1)
    for i, nm in enumerate(["setup", "to last tile's loop end", "store y", "final sync"]):
        print(f"   {nm:10s} median {int(np.median(ph[:, i])):6d} ns   p90 {int(np.percentile(ph[:, i], 90)):6d}   max {int(ph[:, i].max()):6d}")
    gaps = []
    for sm in np.unique(cta[:, 0]):
        c = cta[cta[:, 0] == sm]
        c = c[np.argsort(c[:, 1])]
        gaps += list(c[1:, 1] - c[:-1, 5])
    if gaps:
        print(f"   gap between a CTA's exit and the next CTA's entry on the same SM: median {int(np.median(gaps))} ns, p90 {int(np.percentile(gaps, 90))}, max {int(max(gaps))}")
        per_sm = [int((cta[:, 0] == sm).sum()) for sm in np.unique(cta[:, 0])]
        print(f"   CTAs per SM: min {min(per_sm)} max {max(per_sm)}")
KT = min(3 * (K // 256), 32)
names = {0: ("producer", ["empty_ok", "issued"]),
         1: ("mma", ["top", "accempty_ok", "bfull0_ok", "bfull1_ok", "committed"]),
         2: ("unpack w0", ["top", "full_ok", "bfree0_ok", "done0", "bfree1_ok", "done1"]),
         3: ("unpack w7", ["top", "full_ok", "bfree0_ok", "done0", "bfree1_ok", "done1"]),
         4: ("epi w0", ["top", "accfull_ok", "handback", "end"]),
         5: ("epi w7", ["top", "accfull_ok", "handback", "end"])}
print(f"rc={rc} M={M} N={N} K={K}  (SM clocks relative to the CTA's first producer stamp)")
for cta in range(2):
    t0 = buf[cta, 0, 0, 0]
    print(f"--- CTA rank {cta}")
    for role in range(6):
        if cta == 1 and role == 1:
            continue
        nm, evs = names[role]
        print(f"  {nm}: " + ", ".join(evs))
        for kt in range(KT):
            print(f"     kt{kt:2d} " + " ".join(f"{int(buf[cta, role, kt, e] - t0):7d}" for e in range(len(evs))))
