"""GPU diagnostic (not a pytest file): L2 -> SM fabric rate (tcq4_b200_measure_fabric) and the cta_group::2 kind::i8
issue rate per N (tcq4_b200_measure_umma2).  python tests/gpu_fabric_peak.py"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
d = t4.diag_lib()
print("# L2 -> SM fabric: delivered GB/s over all SMs; unit 32 KB out of a 32 MB (L2-resident) region, 4 units in flight per SM")
for unit, region in ((32768, 32 << 20), (16384, 32 << 20), (32768, 4 << 30)):
    for cluster, mc in ((1, 0), (2, 0), (2, 1), (4, 0), (4, 1)):
        for pattern in (0, 1):
            gbs, ctas = C.c_double(), C.c_int()
            rc = d.tcq4_b200_measure_fabric(cluster, mc, pattern, unit, region, 3000, C.byref(gbs), C.byref(ctas))
            print(f"unit={unit:6d} region={region >> 20:5d} MB cluster={cluster} multicast={mc} pattern={'same' if pattern else 'distinct'}: "
                  f"rc={rc} ctas={ctas.value} delivered {gbs.value:8.1f} GB/s", flush=True)
print("# tcgen05.mma.cta_group::2.kind::i8 M=256 issue loop")
for N in (64, 96, 128, 160, 192, 224, 256):
    tops, cyc = C.c_double(), C.c_double()
    rc = d.tcq4_b200_measure_umma2(N, 4000, C.byref(tops), C.byref(cyc))
    print(f"M=256 N={N:3d}: rc={rc} {tops.value:7.1f} TOPS {cyc.value:6.1f} cyc/MMA", flush=True)
