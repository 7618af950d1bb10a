"""GPU diagnostic (not a pytest file): MUL_MAT_ID with many tokens -- device-routed path (routing kernel + activation stage +
grouped batched-decode GEMM, three launches) against the reference's host-sorted scheme (TCQ4_MOE_HOSTSORT=1: two stream
synchronisations, one MUL_MAT per expert).  Qwen3-30B-A3B-like expert shapes.   python tests/gpu_moe_prefill_bench.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if os.environ.get("MOE_CHILD") != "1":
    for hs in ("", "1"):
        env = dict(os.environ, MOE_CHILD="1")
        if hs:
            env["TCQ4_MOE_HOSTSORT"] = "1"
        subprocess.run([sys.executable, os.path.abspath(__file__)], cwd=ROOT, env=env)
    sys.exit(0)
import time
import numpy as np
import torch
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
t4 = graft.load_package()
mode = "host-sorted" if os.environ.get("TCQ4_MOE_HOSTSORT") else "device-routed"
rng = np.random.default_rng(1)
n_expert, n_used = 128, 8
for name, N, K in (("ffn_up_exps", 768, 2048), ("ffn_down_exps", 2048, 768)):
    w = torch.stack([t4.random_tiles(N, K, seed=e) for e in range(n_expert)])
    for n_tokens in (16, 64, 512, 2048):
        ids = torch.from_numpy(np.stack([rng.choice(n_expert, n_used, replace=False) for _ in range(n_tokens)]).astype(np.int32)).cuda()
        x = torch.randn((n_tokens, 1, K), device="cuda")
        out = torch.empty((n_tokens, n_used, N), device="cuda")
        for _ in range(3):
            t4.mul_mat_id(w, x, ids, N, out=out)
        torch.cuda.synchronize()
        reps = 10
        t0 = time.perf_counter()
        for _ in range(reps):
            t4.mul_mat_id(w, x, ids, N, out=out)
        torch.cuda.synchronize()
        us = (time.perf_counter() - t0) / reps * 1e6
        wbytes = n_expert * N * K * 0.578
        print(f"{mode:13s} {name:13s} N={N} K={K} tokens={n_tokens:5d} rows={n_tokens * n_used:6d}: {us:8.1f} us per op (wall, eager)  "
              f"= {2 * n_tokens * n_used * N * K / us / 1e6:6.1f} TOP/s, expert weights {wbytes / 1e6:.0f} MB", flush=True)
