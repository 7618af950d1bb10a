"""GPU diagnostic (not a pytest file): a few batched-decode MUL_MATs over distinct weight buffers, for ncu captures of the
TMEM-weights kernel.   python tests/gpu_g4_one.py [M N K]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
M, N, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (64, 28672, 8192)
bufs = [t4.random_tiles(N, K, seed=i) for i in range(3)]
x = torch.randn((M, K), device="cuda")
ws = t4.Workspace.for_shape(M, N, K)
out = torch.empty((M, N), device="cuda")
for r in range(2):
    for b in bufs:
        t4.mul_mat(b, x, N, out=out, workspace=ws)
torch.cuda.synchronize()
print("ok", t4.select_kernel(M, N, K))
