"""GPU diagnostic (not a pytest file): timeline of one CTA of the batched-decode GEMM (diagnostics library, TCQ4_G4_PROF=1),
and what-if timings with parts of the CUDA-core work switched off (TCQ4_G4_EXP).
Usage on the box:  python tests/gpu_diag_gemm4.py [M N K]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if os.environ.get("G4_CHILD") != "1":
    args = sys.argv[1:4] if len(sys.argv) > 3 else ["64", "28672", "8192"]
    for exp in os.environ.get("G4_EXPS", "0,1,2,4,3,7,8,15").split(","):
        env = dict(os.environ, G4_CHILD="1", TCQ4_G4_EXP=exp, TCQ4_B200_USE_DIAG="1")
        if exp in os.environ.get("G4_PROF_EXPS", "0").split(","):
            env["TCQ4_G4_PROF"] = "1"
        subprocess.run([sys.executable, os.path.abspath(__file__)] + args, cwd=ROOT, env=env)
    sys.exit(0)

import numpy as np
import torch
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_g3_bench as b
import ctypes as C
t4 = b.t4
M, N, K = (int(a) for a in sys.argv[1:4])
b.M = M
b.NBUF = max(3, min(12, int(400e6 / (N * K * 0.578))))
exp = os.environ.get("TCQ4_G4_EXP", "0")
b.one(N, K, t4.KERNEL_GEMM4, f"gemm4 exp={exp}")
if os.environ.get("TCQ4_G4_PROF"):
    d = t4.diag_lib()
    raw = np.zeros(4 * 64 * 8, dtype=np.int64)
    rc = d.tcq4_b200_debug_gemm4_timeline(C.c_void_p(raw.ctypes.data), C.c_size_t(raw.nbytes))
    buf = raw.reshape(4, 64, 8)
    names = {0: ("mma", ["top", "accempty_ok", "opfull_ok", "issued"]),
             1: ("cw 0", ["top", "wfull_ok", "afull_ok", "opfree_ok", "w cvt+st", "act sts", "wait::st", "published"]),
             2: ("cw15", ["top", "wfull_ok", "afull_ok", "opfree_ok", "w cvt+st", "act sts", "wait::st", "published"]),
             3: ("epi0", ["top", "snfull_ok", "accfull_ok", "ld done", "math end"])}
    t0 = buf[0, 0, 0]
    KT = min(K // 256, 40)
    print(f"rc={rc} M={M} N={N} K={K}  (SM clocks relative to the MMA warp's first stamp)")
    for role in range(4):
        nm, evs = names[role]
        print(f"  {nm}: " + ", ".join(evs))
        for kt in range(KT):
            print(f"     kt{kt:2d} " + " ".join(f"{int(buf[role, kt, e] - t0):7d}" for e in range(len(evs))))
