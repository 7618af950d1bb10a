"""Profiling target (not a pytest file): a handful of launches of each kernel at the headline shape.
  python tests/gpu_prof_target.py [gemv|gemm|both] [M]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()
which = sys.argv[1] if len(sys.argv) > 1 else "both"
N, K = 9728, 2560
nbuf = 10
bufs = [t4.random_tiles(N, K, seed=i) for i in range(nbuf)]
for M in ([1] if which == "gemv" else [int(sys.argv[2]) if len(sys.argv) > 2 else 512] if which == "gemm" else [1, 512]):
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    out = torch.empty((M, N), device="cuda")
    for i in range(nbuf):
        t4.mul_mat(bufs[i], x, N, out=out, workspace=ws)
    torch.cuda.synchronize()
print("done")
