"""GPU diagnostic (not a pytest file): time the pair GEMM at one shape (env TCQ4_G2_XFLAGS / TCQ4_G2_NTL experiments)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 9728, 2560)
kern = int(sys.argv[4]) if len(sys.argv) > 4 else t4.KERNEL_AUTO      # 0 auto, 1 GEMV, 2 GEMM (single CTA), 3 GEMM2 (pairs)
nbuf = 10
bufs = [t4.random_tiles(N, K, seed=i) for i in range(nbuf)]
x = torch.randn((M, K), device="cuda")
ws = t4.Workspace.for_shape(M, N, K)
out = torch.empty((M, N), device="cuda")
for i in range(nbuf): t4.mul_mat(bufs[i], x, N, out=out, workspace=ws, kernel=kern)
torch.cuda.synchronize()
t4.profile_begin(4 * nbuf)
for r in range(4):
    for i in range(nbuf): t4.mul_mat(bufs[i], x, N, out=out, workspace=ws, kernel=kern)
ms, pre, n = t4.profile_end()
us = ms * 1e3 / n
print(f"M={M} N={N} K={K} kernel={kern} ntl={os.environ.get('TCQ4_G2_NTL','auto')}: "
      f"main {us:.1f} us ({2.0*M*N*K/us/1e6:.0f} TOPS), act {pre*1e3/n:.1f} us", flush=True)
