"""GPU diagnostic (not a pytest file): where should the batched-decode GEMM take over from the GEMV?  Per-MUL_MAT time of both
for M = 2..16 on the Llama-3-8B / Qwen3-4B shapes (graph of NBUF MUL_MATs over distinct weight buffers)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
t4 = b.t4
SHAPES = [(14336, 4096), (4096, 14336), (6144, 4096), (4096, 4096), (9728, 2560), (2560, 9728), (28672, 8192), (8192, 28672)]
for M in (int(a) for a in (sys.argv[1:] or ["2", "4", "6", "8", "12", "16"])):
    b.M = M
    for N, K in SHAPES:
        b.NBUF = max(3, min(12, int(400e6 / (N * K * 0.578))))
        try:
            g = b.one(N, K, t4.KERNEL_GEMV, "gemv ")
        except t4.Tcq4Error:
            g = float("nan")
        h = b.one(N, K, t4.KERNEL_GEMM4, "gemm4")
        a = b.one(N, K, t4.KERNEL_AUTO, "auto ")
        print(f"M={M} N={N} K={K}: gemv {g:.2f} gemm4 {h:.2f} auto {a:.2f} us", flush=True)
