"""GPU diagnostic (not a pytest file): HBM roofline of the device quantiser (K4) and GET_ROWS (K5).
Usage on the box:  python tests/gpu_quant_bench.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()
try:
    HBM = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:  # noqa: BLE001
    HBM = 6550.0


def timeit(fn, reps):
    for _ in range(3):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def quant(N, K, nbuf=4):
    Ws = [torch.randn((N, K), device="cuda") * 0.02 for _ in range(nbuf)]      # 4 x 100 MB > 126 MB L2
    outs = [torch.empty(t4.quantize_size(N, K), dtype=torch.uint8, device="cuda") for _ in range(nbuf)]

    def fn(i):
        t4.LIB.tcq4_b200_quantize_tiles(Ws[i % nbuf].data_ptr(), outs[i % nbuf].data_ptr(), N, K, None, None)
    s = timeit(fn, 20)
    b = N * K * 4 + t4.quantize_size(N, K)
    print(f"quantize_tiles N={N} K={K}: {s * 1e6:.1f} us, {b / s / 1e9:.0f} GB/s algorithmic = {b / s / 1e9 / HBM:.2f} of HBM "
          f"({N * K / s / 1e9:.1f} G weights/s)", flush=True)


def getrows(V, K, n):
    tiles = t4.random_tiles(V, K, seed=5)
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    idxs = [torch.randint(0, V, (n,), device="cuda", generator=g, dtype=torch.int32) for _ in range(8)]
    out = torch.empty((n, K), device="cuda")

    def fn(i):
        t4.get_rows(tiles, idxs[i % 8], V, K, out=out)
    s = timeit(fn, 20)
    b = n * K * 4 + n * (K // 256) * 148 + n * 4          # rows written, the selected rows' share of the tiles, indices
    print(f"get_rows V={V} K={K} n={n}: {s * 1e6:.1f} us, {b / s / 1e9:.0f} GB/s algorithmic = {b / s / 1e9 / HBM:.2f} of HBM",
          flush=True)


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), "HBM peak", HBM, flush=True)
    quant(9728, 2560)
    quant(2560, 9728)
    quant(28672, 8192, nbuf=2)
    getrows(151936, 2560, 512)
    getrows(151936, 2560, 4096)
    getrows(151936, 2560, 32768)
