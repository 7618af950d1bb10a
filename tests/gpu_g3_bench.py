"""GPU diagnostic (not a pytest file): per-shape time of the prefill GEMMs inside a CUDA graph of NBUF back-to-back MUL_MATs
over distinct weight buffers -- the kind::f16 pair GEMM (KERNEL_GEMM3) against the round-1 kind::i8 pair GEMM (KERNEL_GEMM2),
per node and as fan-out groups.   python tests/gpu_g3_bench.py [M]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 512
NBUF = int(os.environ.get("NBUF", "12"))


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


def graph_of(fn):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    return g


def one(N, K, kernel, name):
    bufs = [t4.random_tiles(N, K, seed=i) for i in range(NBUF)]
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    out = torch.empty((M, N), device="cuda")

    def chain():
        for b in bufs:
            t4.mul_mat(b, x, N, out=out, workspace=ws, kernel=kernel)
    us = timeit(graph_of(chain).replay) / NBUF
    print(f"M={M} N={N:6d} K={K:5d} {name}: {us:7.2f} us per MUL_MAT (act + gemm) = {2 * M * N * K / us / 1e6:7.1f} TOP/s", flush=True)
    return us


def fan(Ns, K, name):
    bufs = [[t4.random_tiles(N, K, seed=i * 8 + j) for j, N in enumerate(Ns)] for i in range(NBUF)]
    fo = [t4.Fanout(b, Ns, K) for b in bufs]
    x = torch.randn((M, K), device="cuda")
    outs = [torch.empty((M, N), device="cuda") for N in Ns]
    t4.reserve(M, max(Ns), K)

    def chain():
        for f in fo:
            f(x, outs=outs)
    us = timeit(graph_of(chain).replay) / NBUF
    print(f"M={M} fan-out {name} N={Ns} K={K}: {us:7.2f} us per group = {2 * M * sum(Ns) * K / us / 1e6:7.1f} TOP/s", flush=True)
    return us


if __name__ == "__main__":
    tot = {}
    for name, k in (("gemm2 (i8 hi/lo)", t4.KERNEL_GEMM2), ("gemm3 (f16)", t4.KERNEL_GEMM3)):
        tot[name] = 0.0
        for N, K, mult in ((4096, 2560, 1), (1024, 2560, 2), (2560, 4096, 1), (9728, 2560, 2), (2560, 9728, 1)):
            tot[name] += mult * one(N, K, k, name)
        print(f"  per Qwen3-4B layer, per node, {name}: {tot[name]:.1f} us -> {36 * tot[name] / 1e3:.2f} ms/step", flush=True)
    a = fan([4096, 1024, 1024], 2560, "qkv")
    b = fan([9728, 9728], 2560, "gate/up")
    o = one(2560, 4096, t4.KERNEL_GEMM3, "gemm3 (f16)")
    d = one(2560, 9728, t4.KERNEL_GEMM3, "gemm3 (f16)")
    print(f"  per layer with fan-out: {a + b + o + d:.1f} us -> {36 * (a + b + o + d) / 1e3:.2f} ms/step", flush=True)
    if M >= 512:
        for N, K in ((9728, 2560), (28672, 8192)):
            if M * N * K > 4096 * 9728 * 2560 * 3 and M < 4096:
                continue
        M = 4096
        one(9728, 2560, t4.KERNEL_GEMM2, "gemm2 (i8 hi/lo)")
        one(9728, 2560, t4.KERNEL_GEMM3, "gemm3 (f16)")
