"""GPU diagnostic (not a pytest file): batched decode, 9 <= M <= 64 -- per-MUL_MAT time (activation stage + GEMM) of the
TMEM-weights kernel (KERNEL_GEMM4) against the kernels that served these M before it, inside a CUDA graph of NBUF
back-to-back MUL_MATs over distinct weight buffers (L2-cold weights), and the eager split activation kernel / GEMM from the
library's profile tap.   python tests/gpu_g4_bench.py [M ...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
t4 = b.t4

SHAPES = [(14336, 4096, "llama3-8b gate/up"), (4096, 14336, "llama3-8b down"), (6144, 4096, "llama3-8b qkv"),
          (4096, 4096, "llama3-8b o"), (28672, 8192, "llama3-70b gate/up"), (8192, 28672, "llama3-70b down"),
          (9728, 2560, "qwen3-4b gate/up"), (2560, 9728, "qwen3-4b down")]


def split(M, N, K):
    """eager: (GEMM us, activation-kernel us) per MUL_MAT from the profile tap (CUDA events around each launch)"""
    tiles = t4.random_tiles(N, K, seed=99)
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    out = torch.empty((M, N), device="cuda")
    for _ in range(3):
        t4.mul_mat(tiles, x, N, out=out, workspace=ws, kernel=t4.KERNEL_GEMM4)
    torch.cuda.synchronize()
    t4.profile_begin(64)
    for _ in range(20):
        t4.mul_mat(tiles, x, N, out=out, workspace=ws, kernel=t4.KERNEL_GEMM4)
    main_ms, pre_ms, n = t4.profile_end(t4.KERNEL_GEMM4)
    return main_ms * 1e3 / n, pre_ms * 1e3 / n


if __name__ == "__main__":
    only4 = os.environ.get("G4_ONLY") is not None
    print("# per MUL_MAT, us: graph = activation kernel + GEMM inside a CUDA graph (PDL overlap); eager split from the profile tap")
    for M in (int(a) for a in (sys.argv[1:] or ["64", "32", "17"])):
        b.M = M
        for N, K, nm in SHAPES:
            b.NBUF = max(3, min(12, int(400e6 / (N * K * 0.578))))
            floor = N * K * 0.578 / 7.0e6      # us at ~7 TB/s
            g4 = b.one(N, K, t4.KERNEL_GEMM4, "gemm4")
            g2 = gm = gv = float("nan")
            if not only4:
                g2 = b.one(N, K, t4.KERNEL_GEMM2, "gemm2")
                if M <= 16:
                    try:
                        gv = b.one(N, K, t4.KERNEL_GEMV, "gemv ")
                    except t4.Tcq4Error:
                        pass
            mu, pu = split(M, N, K)
            print(f"ROW M={M:2d} {nm:20s} N={N:5d} K={K:5d} floor {floor:5.1f} | gemm4 {g4:6.2f} (eager: act {pu:5.2f} + gemm {mu:6.2f}) | "
                  f"gemm2 {g2:6.2f} | gemv {gv:6.2f} | x floor {g4 / floor:4.1f}", flush=True)
