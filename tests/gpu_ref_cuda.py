"""GPU diagnostic (not a pytest file): the FORK's own CUDA kernels (tcq4_k32.cu compiled unmodified for sm_100a into
oracle/_ref/libtcq4_refcuda.so by `make -C oracle refcuda`) timed on this box next to this library, same inputs.
  live path : tcq4_rrs_fused_gemm_smallM                      (what ggml_cuda_rrs_mul_mat calls, rrs.cu:241-344)
  IMMA path : tcq4_rrs_fwht_quantize_tc + tcq4_rrs_gemm_imma  (the fork's tensor-core kernels, benchmark-only there)
Outputs are compared by NMSE (the fork builds with -use_fast_math, so codes can differ at rounding ties).
Usage on the box:  python tests/gpu_ref_cuda.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import __graft_entry__ as graft  # noqa: E402
from conftest import make_acts, make_weights  # noqa: E402

t4 = graft.load_package()
orc = graft.load_oracle().Oracle()
ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libtcq4_refcuda.so"))
vp, ci = C.c_void_p, C.c_int
smallM = ref._Z26tcq4_rrs_fused_gemm_smallMPKfPKiPKvPfiiiP11CUstream_st
smallM.argtypes = [vp, vp, vp, vp, ci, ci, ci, vp]
smallM.restype = None
quant_tc = ref._Z25tcq4_rrs_fwht_quantize_tcPKfPviiP11CUstream_st
quant_tc.argtypes = [vp, vp, ci, ci, vp]
quant_tc.restype = None
imma = ref._Z18tcq4_rrs_gemm_immaPKvS0_PfiiiP11CUstream_st
imma.argtypes = [vp, vp, vp, ci, ci, ci, vp]
imma.restype = None


def timeit(fn, n_buf, reps=3):
    for i in range(n_buf):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for i in range(n_buf):
            fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_buf)


def nmse(a, b):
    a, b = a.double(), b.double()
    return float(((a - b) ** 2).sum() / (b ** 2).sum())


def run(M, N, K, n_buf=10):
    rng = np.random.default_rng(7)
    w0 = torch.from_numpy(orc.quantize_tiles(make_weights(rng, N, K))).cuda()       # the fork's CPU quantiser
    bufs = [w0] + [t4.random_tiles(N, K, seed=i) for i in range(1, n_buf)]         # rotate > L2 worth of weights
    x = torch.from_numpy(make_acts(rng, M, K)).cuda()
    ws = t4.Workspace.for_shape(M, N, K)
    y_own = torch.empty((M, N), device="cuda")
    y_live = torch.empty((M, N), device="cuda")
    y_imma = torch.empty((M, N), device="cuda")
    act = torch.empty((M * (K // 256) * 160,), dtype=torch.uint8, device="cuda")
    s = torch.cuda.current_stream().cuda_stream

    def own(i):
        t4.mul_mat(bufs[i], x, N, out=y_own, workspace=ws)

    def live(i):
        smallM(x.data_ptr(), None, bufs[i].data_ptr(), y_live.data_ptr(), M, N, K, s)

    def tc(i):
        quant_tc(x.data_ptr(), act.data_ptr(), K, M, s)
        imma(act.data_ptr(), bufs[i].data_ptr(), y_imma.data_ptr(), M, N, K, s)

    t_own = timeit(own, n_buf)
    t_live = timeit(live, n_buf, reps=1 if M > 64 else 3)
    own(0); live(0)
    torch.cuda.synchronize()
    line = (f"M={M:5d} N={N} K={K}: this library {t_own:9.1f} us | fork live kernel (smallM) {t_live:10.1f} us "
            f"({t_live / t_own:6.1f}x) nmse {nmse(y_live, y_own):.2e}")
    if M % 16 == 0 and N % 128 == 0:
        t_tc = timeit(tc, n_buf)
        tc(0)
        torch.cuda.synchronize()
        line += f" | fork IMMA path (quantize_tc + gemm_imma) {t_tc:9.1f} us ({t_tc / t_own:5.1f}x) nmse {nmse(y_imma, y_own):.2e}"
    print(line, flush=True)


if __name__ == "__main__":
    print(t4.version(), torch.cuda.get_device_name(0), flush=True)
    for (M, N, K) in [(1, 9728, 2560), (8, 9728, 2560), (16, 9728, 2560), (512, 9728, 2560), (512, 2560, 9728),
                      (512, 4096, 2560), (4096, 9728, 2560)]:
        run(M, N, K)
