"""GPU diagnostic (not a pytest file): timeline of one cluster of the kind::f16 pair GEMM (diagnostics library, TCQ4_G3_PROF=1).
Usage on the box:  python tests/gpu_diag_gemm3.py [M N K]"""
import os
import sys

import numpy as np
import torch

os.environ["TCQ4_G3_PROF"] = "1"
os.environ["TCQ4_B200_USE_DIAG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 9728, 2560)
import ctypes as C
d = t4.diag_lib()
tiles = t4.random_tiles(N, K, seed=1)
x = torch.randn((M, K), device="cuda")
out = torch.empty((M, N), device="cuda")
ws = t4.Workspace.for_shape(M, N, K)
for _ in range(3):
    t4.mul_mat(tiles, x, N, out=out, workspace=ws, kernel=t4.KERNEL_GEMM3)
torch.cuda.synchronize()
raw = np.zeros(4 * 64 * 8, dtype=np.int64)
rc = d.tcq4_b200_debug_gemm3_timeline(C.c_void_p(raw.ctypes.data), C.c_size_t(raw.nbytes))
buf = raw.reshape(4, 64, 8)
names = {0: ("mma", ["top", "accempty_ok", "bfull_ok", "afull0_ok", "afull1_ok", "afull2_ok", "afull3_ok"]),
         1: ("act w0", ["top", "raw_ok", "afree(q0)_ok", "q0 done", "-", "afree(q2)_ok", "q2 done"]),
         2: ("wgt w0", ["top", "raw_ok", "bfree_ok", "done"]),
         3: ("epi w0", ["top", "accfull_ok", "handback", "end"])}
t0 = buf[0, 0, 0]
KT = min(4 * (K // 256), 64)
print(f"rc={rc} M={M} N={N} K={K}  (SM clocks relative to the MMA warp's first stamp)")
for role in range(4):
    nm, evs = names[role]
    print(f"  {nm}: " + ", ".join(evs))
    for kt in range(KT):
        print(f"     kt{kt:2d} " + " ".join(f"{int(buf[role, kt, e] - t0):7d}" for e in range(len(evs))))
