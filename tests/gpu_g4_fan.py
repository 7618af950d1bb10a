"""GPU diagnostic (not a pytest file): fan-out groups of a decode batch (Q/K/V, gate/up) -- one activation launch + one
launch of the batched-decode GEMM over all matrices of the group, in a graph over distinct weights.
TCQ4_NO_GEMM4=1 gives the path these M took before (kind::f16 pair GEMM with a segmented tile list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
for M in (int(a) for a in (sys.argv[1:] or ["64", "32", "17"])):
    b.M = M
    b.NBUF = 6
    b.fan([4096, 1024, 1024], 4096, "llama3-8b q/k/v")
    b.fan([14336, 14336], 4096, "llama3-8b gate+up")
    b.fan([4096, 512, 512], 2560, "qwen3-4b q/k/v") if False else None
    b.fan([9728, 9728], 2560, "qwen3-4b gate+up")
    b.NBUF = 3
    b.fan([8192, 1024, 1024], 8192, "llama3-70b q/k/v")
    b.fan([28672, 28672], 8192, "llama3-70b gate+up")
