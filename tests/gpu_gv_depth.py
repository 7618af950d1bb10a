"""GPU diagnostic (not a pytest file): decode GEMV ring depth (TCQ4_GV_DEPTH, read at load: one child per value) on large matrices."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if os.environ.get("GVD_CHILD") != "1":
    for d in ("", "3", "4", "5", "6", "8"):
        env = dict(os.environ, GVD_CHILD="1")
        if d:
            env["TCQ4_GV_DEPTH"] = d
        print("== TCQ4_GV_DEPTH =", d or "(plan)", flush=True)
        subprocess.run([sys.executable, os.path.abspath(__file__)] + sys.argv[1:], cwd=ROOT, env=env)
    sys.exit(0)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_g3_bench as b
t4 = b.t4
for M in (int(a) for a in (sys.argv[1:] or ["1"])):
    b.M = M
    b.NBUF = 6
    b.fan([14336, 14336], 4096, "llama3-8b gate+up")
    b.one(4096, 14336, t4.KERNEL_AUTO, "llama3-8b down")
    b.fan([4096, 1024, 1024], 4096, "llama3-8b qkv")
    b.NBUF = 3
    b.one(28672, 8192, t4.KERNEL_AUTO, "llama3-70b gate")
    b.one(8192, 28672, t4.KERNEL_AUTO, "llama3-70b down")
    b.NBUF = 12
    b.fan([9728, 9728], 2560, "qwen3-4b gate+up")
    b.one(2560, 9728, t4.KERNEL_AUTO, "qwen3-4b down")
