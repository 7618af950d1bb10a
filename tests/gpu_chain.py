"""GPU diagnostic (not a pytest file): time a chain of back-to-back GEMV launches, eager and as a CUDA graph.
   TCQ4_NO_PDL=1 python tests/gpu_chain.py   disables programmatic dependent launch (on by default) for comparison."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()

def run(M, N, K, nbuf=int(os.environ.get('NBUF', '24'))):
    bufs = [t4.random_tiles(N, K, seed=i) for i in range(nbuf)]
    x = torch.randn((M, K), device="cuda")
    ws = t4.Workspace.for_shape(M, N, K)
    outs = [torch.empty((M, N), device="cuda") for _ in range(2)]
    def chain():
        for i in range(nbuf):
            t4.mul_mat(bufs[i], x, N, out=outs[i & 1], workspace=ws)
    def timeit(fn, reps):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * nbuf)
    us_e = timeit(chain, 10)
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            chain()
    us_g = timeit(g.replay, 10)
    wb = bufs[0].numel()
    print(f"M={M} N={N} K={K} pdl={'off' if os.environ.get('TCQ4_NO_PDL') else 'on'}: eager {us_e:.2f} us/launch, graph {us_g:.2f} us/launch "
          f"-> {wb / us_g / 1e3:.0f} GB/s", flush=True)

def layer_chain(M, layers=int(os.environ.get('LAYERS', '36'))):
    """the bench's decode chain: 7 matmuls per layer, distinct weights per layer, inputs independent of outputs"""
    shapes = [(4096, 2560), (1024, 2560), (1024, 2560), (2560, 4096), (9728, 2560), (9728, 2560), (2560, 9728)]
    W = [[t4.random_tiles(n, k, seed=l * 8 + j) for j, (n, k) in enumerate(shapes)] for l in range(layers)]
    x = {k: torch.randn((M, k), device="cuda") for k in (2560, 4096, 9728)}
    y = {n: torch.empty((M, n), device="cuda") for n in (4096, 1024, 2560, 9728)}
    ws = t4.Workspace.for_shape(M, 9728, 9728)
    fan = os.environ.get("FANOUT")                 # q/k/v and gate/up share their activations: one activation stage each
    if fan:
        yk, yv, yu = torch.empty_like(y[1024]), torch.empty_like(y[1024]), torch.empty_like(y[9728])
        qkv = [t4.Fanout([W[l][0], W[l][1], W[l][2]], [4096, 1024, 1024], 2560) for l in range(layers)]
        gu = [t4.Fanout([W[l][4], W[l][5]], [9728, 9728], 2560) for l in range(layers)]
        t4.reserve(M, 9728, 9728)
    def chain():
        for l in range(layers):
            if fan:
                qkv[l](x[2560], outs=[y[4096], yk, yv])
                t4.mul_mat(W[l][3], x[4096], 2560, out=y[2560])
                gu[l](x[2560], outs=[y[9728], yu])
                t4.mul_mat(W[l][6], x[9728], 2560, out=y[2560])
                continue
            for j, (n, k) in enumerate(shapes):
                t4.mul_mat(W[l][j], x[k], n, out=y[n], workspace=ws)
    for _ in range(2): chain()
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            chain()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"layer chain M={M} layers={layers}: {ms:.3f} ms/step, {ms * 1e3 / (layers * 7):.2f} us/launch", flush=True)

if os.environ.get("LAYER_CHAIN"):
    layer_chain(int(os.environ.get("CHAIN_M", "1")))
elif os.environ.get("SWEEP"):
    # BASELINE configs[2]: Llama-3-8B gate/up and down projections, decode batch 1..64
    for n, k in [(14336, 4096), (4096, 14336)]:
        for m in (1, 2, 3, 4, 5, 8, 12, 16, 17, 24, 32, 48, 64):
            run(m, n, k, nbuf=8)
elif os.environ.get("CHAIN_M"):
    m = int(os.environ["CHAIN_M"])
    for n, k in [(4096, 2560), (1024, 2560), (2560, 4096), (9728, 2560), (2560, 9728)]:
        run(m, n, k, nbuf=12)
else:
    for shp in [(1, 9728, 2560), (1, 2560, 9728), (1, 4096, 2560), (1, 1024, 2560), (8, 9728, 2560)]:
        run(*shp)
