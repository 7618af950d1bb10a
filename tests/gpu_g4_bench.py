"""GPU diagnostic (not a pytest file): batched decode, 17 <= M <= 64 -- per-MUL_MAT time (activation stage + GEMM) of the
TMEM-weights kernel (KERNEL_GEMM4) against the kernels that served these M in round 1, inside a CUDA graph of NBUF
back-to-back MUL_MATs over distinct weight buffers (L2-cold weights).   python tests/gpu_g4_bench.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
t4 = b.t4

SHAPES = [(14336, 4096, "llama3-8b gate/up"), (4096, 14336, "llama3-8b down"), (6144, 4096, "llama3-8b qkv"),
          (4096, 4096, "llama3-8b o"), (28672, 8192, "llama3-70b gate/up"), (8192, 28672, "llama3-70b down"),
          (9728, 2560, "qwen3-4b gate/up"), (2560, 9728, "qwen3-4b down")]
if __name__ == "__main__":
    for M in (int(a) for a in (sys.argv[1:] or ["64", "32", "17"])):
        b.M = M
        for N, K, nm in SHAPES:
            b.NBUF = max(3, min(12, int(400e6 / (N * K * 0.578))))
            floor = N * K * 0.578 / 7.0e6      # us at ~7 TB/s
            print(f"-- {nm}: weight stream floor {floor:.1f} us")
            for name, k in (("gemm4", t4.KERNEL_GEMM4), ("gemm2", t4.KERNEL_GEMM2), ("gemm ", t4.KERNEL_GEMM)):
                b.one(N, K, k, name)
