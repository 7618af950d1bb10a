"""GPU diagnostic (not a pytest file): one pass over the round-2 additions for an ncu capture -- a Llama-3-8B decode-batch layer
(M = 32) through the graph entry (activation kernel + batched-decode GEMM per fan-out group), one device-routed MUL_MAT_ID
(512 tokens, 128 experts), one fused RMSNorm fan-out and one fused SwiGLU down projection at M = 512."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
import bench
t4 = graft.load_package()
PARTS = os.environ.get("R02_PARTS", "layer,moe,fusion").split(",")
M = 32
xs = {src: torch.randn((M, K), device="cuda") for _, _, K, src in bench.LLAMA3_8B_LAYER}
nodes = []
for rep in range(2):
    for j, (nm, N, K, src) in enumerate(bench.LLAMA3_8B_LAYER):
        nodes.append((t4.random_tiles(N, K, seed=rep * 8 + j), xs[src], torch.empty((M, N), device="cuda"), N, ""))
t4.reserve(M, 14336, 14336)
g = t4.GgmlGraph(nodes)
for _ in range(2 if "layer" in PARTS else 0):
    g.compute()
torch.cuda.synchronize()
rng = np.random.default_rng(0)
n_expert, n_used, N, K, n_tokens = 128, 8, 768, 2048, 512
w = torch.stack([t4.random_tiles(N, K, seed=e) for e in range(n_expert)])
ids = torch.from_numpy(np.stack([rng.choice(n_expert, n_used, replace=False) for _ in range(n_tokens)]).astype(np.int32)).cuda()
x = torch.randn((n_tokens, 1, K), device="cuda")
for _ in range(2 if "moe" in PARTS else 0):
    t4.mul_mat_id(w, x, ids, N)
torch.cuda.synchronize()
Mp, Kq = 512, 2560
xq = torch.randn((Mp, Kq), device="cuda")
nw = torch.rand((Kq,), device="cuda") + 0.5
qkv = [t4.random_tiles(n, Kq, seed=50 + i) for i, n in enumerate((4096, 1024, 1024))]
for _ in range(2 if "fusion" in PARTS else 0):
    rs = t4.rms_row_scale(xq, 1e-6)
    t4.mul_mat_fanout_input(qkv, xq, [4096, 1024, 1024], t4.INPUT_RMS_NORM_MUL, row_scale=rs, col_weight=nw)
gate, up = torch.randn((Mp, 9728), device="cuda"), torch.randn((Mp, 9728), device="cuda")
down = t4.random_tiles(2560, 9728, seed=60)
for _ in range(2 if "fusion" in PARTS else 0):
    t4.mul_mat_fanout_input([down], gate, [2560], t4.INPUT_SWIGLU, x2=up)
torch.cuda.synchronize()
print("ok")
