"""GPU diagnostic (not a pytest file): does tcgen05.mma kind::f16 accumulate small integers exactly?"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
d = t4.diag_lib()
for pattern in (0, 1, 2, 3):
    for seed in (1, 2, 3):
        bad, mx = C.c_int(), C.c_int()
        rc = d.tcq4_b200_measure_f16_exact(pattern, seed, C.byref(bad), C.byref(mx))
        print(f"kind::f16 integer exactness: pattern={pattern} seed={seed} rc={rc} mismatches={bad.value} of 32768, max |sum|={mx.value}", flush=True)
