"""GPU diagnostic (not a pytest file): BASELINE configs[3..4] shapes (Qwen3-32B, Llama-3-70B) through every kernel --
index arithmetic at full size, timings, and row-independence checks (a row computed alone by the GEMV agrees with the
same row inside the large GEMM to fp32 association error)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()


def nmse(a, b):
    return float(((a - b) ** 2).sum() / (b ** 2).sum())


def run(M, N, K):
    tiles = t4.random_tiles(N, K, seed=9)
    g = torch.Generator(device="cuda")
    g.manual_seed(M + N)
    x = torch.randn((M, K), device="cuda", generator=g)
    ws = t4.Workspace.for_shape(M, N, K)
    y = torch.empty((M, N), device="cuda")
    t4.mul_mat(tiles, x, N, out=y, workspace=ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        t4.mul_mat(tiles, x, N, out=y, workspace=ws)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 3
    errs = []
    for r in sorted({0, M // 2, M - 1}):
        y1 = t4.mul_mat(tiles, x[r:r + 1].contiguous(), N, kernel=t4.KERNEL_GEMV)
        errs.append(nmse(y1, y[r:r + 1]))
    ok = bool(torch.isfinite(y).all()) and max(errs) < 1e-9
    print(f"M={M} N={N} K={K} kernel={t4.select_kernel(M, N, K)}: {us:.1f} us, {2.0 * M * N * K / us / 1e6:.0f} TOP/s, "
          f"{(N * K * 0.578125 + 4 * M * (N + K)) / us / 1e3:.0f} GB/s, rows vs GEMV nmse max {max(errs):.1e} {'OK' if ok else 'FAIL'}",
          flush=True)
    return ok


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    allok = True
    # Llama-3-70B: hidden 8192, ffn 28672, kv 1024; Qwen3-32B: hidden 5120, ffn 25600
    for shp in [(1, 28672, 8192), (1, 8192, 28672), (8, 28672, 8192), (16, 8192, 8192), (64, 28672, 8192), (512, 28672, 8192),
                (4096, 25600, 5120), (4096, 5120, 25600), (8192, 28672, 8192), (8192, 8192, 28672), (8192, 1024, 8192)]:
        allok &= run(*shp)
    print("ALL OK" if allok else "SOME FAILED")
