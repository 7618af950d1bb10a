"""GPU diagnostic (not a pytest file): per-shape time of the two pair GEMMs (graph of NBUF MUL_MATs), for the dispatch rule."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
t4 = b.t4
for M in (64, 128, 256, 512, 1024, 4096):
    b.M = M
    for N, K in ((1024, 2560), (4096, 2560), (2560, 4096), (9728, 2560), (2560, 9728), (14336, 4096), (4096, 14336)):
        if M * N * K > 3e11:
            continue
        a = b.one(N, K, t4.KERNEL_GEMM2, "gemm2")
        c = b.one(N, K, t4.KERNEL_GEMM3, "gemm3")
        print(f"   -> M={M} N={N} K={K}: gemm2 {a:.1f} us, gemm3 {c:.1f} us, gemm3/gemm2 = {c / a:.2f}", flush=True)
