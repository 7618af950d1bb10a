"""GPU diagnostics (not a pytest file): run each kernel on small cases, print where it differs from the
oracle, then time the headline shapes.  Usage on the box:  python tests/gpu_diag.py [quick|perf]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import make_acts, make_weights, nmse  # noqa: E402

t4 = graft.load_package()
orc = graft.load_oracle().Oracle()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def report(tag, isum, ref_isum, y, ref_y):
    ok = np.array_equal(isum, ref_isum)
    print(f"[{tag}] isum exact={ok}  fp32 nmse={nmse(y, ref_y):.3e}", flush=True)
    if ok:
        return True
    bad = isum != ref_isum
    M, N, KT = bad.shape
    print(f"   mismatches {bad.sum()} / {bad.size}")
    print("   by kt      :", bad.sum(axis=(0, 1)).tolist())
    print("   by m%8     :", [int(bad[i::8].sum()) for i in range(8)])
    print("   by m//32   :", [int(bad[i * 32:(i + 1) * 32].sum()) for i in range((M + 31) // 32)])
    print("   by n%8     :", [int(bad[:, i::8].sum()) for i in range(8)])
    print("   by n//16   :", [int(bad[:, i * 16:(i + 1) * 16].sum()) for i in range(min(16, (N + 15) // 16))])
    idx = np.argwhere(bad)[:6]
    for m, n, kt in idx:
        print(f"   (m={m}, n={n}, kt={kt}): got {isum[m, n, kt]}  want {ref_isum[m, n, kt]}")
    # hypotheses
    if M > 1 and np.array_equal(isum[:, :, :], ref_isum[::-1][:, :, :]):
        print("   -> rows reversed")
    return False


def case(M, N, K, kernel, seed=0, perm=None):
    rng = np.random.default_rng(seed)
    tiles = orc.quantize_tiles(make_weights(rng, N, K))
    x = make_acts(rng, M, K)
    ref = orc.mul_mat(tiles, x, N, perm, want_isum=True)
    y, isum = t4.mul_mat_debug(dev(tiles), dev(x), N, dev(perm) if perm is not None else None, kernel=kernel)
    torch.cuda.synchronize()
    return report(f"{'GEMV' if kernel == 1 else 'GEMM'} M={M} N={N} K={K}", isum.cpu().numpy(), ref["isum"],
                  y.cpu().numpy(), ref["C"])


def time_calls(fn, iters, warm=3):
    for _ in range(warm):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3     # us


def perf(M, N, K, nbuf=12, iters=48):
    """time one MUL_MAT over nbuf distinct weight buffers (> L2) so weights come from HBM"""
    bufs = [t4.random_tiles(N, K, seed=i) for i in range(nbuf)]
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    out = torch.empty((M, N), device="cuda")
    us = time_calls(lambda i: t4.mul_mat(bufs[i % nbuf], x, N, out=out, workspace=ws), iters)
    # same through one CUDA graph of nbuf launches
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for i in range(nbuf):
                t4.mul_mat(bufs[i], x, N, out=out, workspace=ws)
    us_g = time_calls(lambda i: g.replay(), 8) / nbuf
    wbytes = bufs[0].numel()
    alg = wbytes + 4 * M * K + 4 * M * N
    ops = 2.0 * M * N * K
    print(f"[perf] M={M} N={N} K={K} kernel={t4.select_kernel(M, N, K)}: eager {us:.2f} us, graph {us_g:.2f} us/launch"
          f" -> {alg / us_g / 1e3:.1f} GB/s alg, {ops / us_g / 1e6:.2f} TOPS", flush=True)
    return us_g


if __name__ == "__main__":
    mode = sys.argv[1] if len(sys.argv) > 1 else "quick"
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    t0 = time.time()
    ok = True
    ok &= case(1, 8, 256, t4.KERNEL_GEMV)
    ok &= case(2, 256, 1024, t4.KERNEL_GEMV, 1)
    ok &= case(4, 1024, 2560, t4.KERNEL_GEMV, 2)
    ok &= case(128, 128, 256, t4.KERNEL_GEMM, 3)
    ok &= case(128, 128, 512, t4.KERNEL_GEMM, 4)
    ok &= case(200, 264, 768, t4.KERNEL_GEMM, 5)
    print("diag cases done in %.1fs ok=%s" % (time.time() - t0, ok), flush=True)
    if mode == "perf":
        for (M, N, K) in [(1, 9728, 2560), (1, 2560, 9728), (1, 4096, 2560), (1, 1024, 2560), (4, 9728, 2560),
                          (512, 9728, 2560), (512, 2560, 9728), (512, 4096, 2560), (128, 9728, 2560),
                          (16, 9728, 2560), (4096, 9728, 2560)]:
            perf(M, N, K)
