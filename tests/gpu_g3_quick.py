"""GPU diagnostic (not a pytest file): quick timing of the kind::f16 pair GEMM on a few shapes (graph of NBUF MUL_MATs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import gpu_g3_bench as b
t4 = b.t4
for M in (512, 4096):
    b.M = M
    for N, K in ((9728, 2560), (2560, 9728), (6144, 2560)):
        b.one(N, K, t4.KERNEL_GEMM3, "gemm3 (f16)")
b.M = 512
b.fan([4096, 1024, 1024], 2560, "qkv")
b.fan([9728, 9728], 2560, "gate/up")
