"""GPU diagnostic (not a pytest file): tcgen05.mma kind::i8 cta_group::2 issue loop with background shared-memory traffic.
The env var is read per call: python tests/gpu_umma_bg.py"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
for N in (192, 256):
    for bg in ("0,0", "2,1", "4,1", "8,1", "12,1", "4,2", "8,2", "12,2"):
        os.environ["TCQ4_UMMA2_BG"] = bg
        tops, cyc = C.c_double(), C.c_double()
        rc = t4.diag_lib().tcq4_b200_measure_umma2(N, 2000, C.byref(tops), C.byref(cyc))
        w, m = bg.split(",")
        print(f"M=256 N={N}: background {w} warps of {'STS.128' if m == '1' else 'LDS.128' if m == '2' else 'nothing'}: rc={rc} "
              f"{tops.value:7.1f} TOPS {cyc.value:6.1f} cyc/MMA", flush=True)
