"""Measure the tcgen05 tensor-pipe peak per kind / N / layout (not a pytest file)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
names = {0: "i8", 1: "bf16", 2: "fp8"}
for kind in (0, 1, 2):
    for N in (64, 128, 256):
        for swz in (0, 1):
            tops, cyc = t4.measure_umma(kind, N, swz, 4000)
            print(f"kind={names[kind]:5s} N={N:3d} swizzle={'128B' if swz else 'none'}: {tops:8.1f} TOPS  {cyc:7.1f} cyc/MMA", flush=True)
import ctypes as C
for N in (64, 128, 256):
    tops, cyc = C.c_double(), C.c_double()
    rc = t4.diag_lib().tcq4_b200_measure_umma2(N, 4000, C.byref(tops), C.byref(cyc))
    print(f"kind=i8 cta_group::2 M=256 N={N:3d}: rc={rc} {tops.value:8.1f} TOPS  {cyc.value:7.1f} cyc/MMA", flush=True)
