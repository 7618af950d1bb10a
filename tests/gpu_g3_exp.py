"""GPU diagnostic (not a pytest file): what-if timings of the kind::f16 pair GEMM with parts of its CUDA-core work switched
off (results are wrong on purpose; PROF instantiation of the diagnostics library): which role bounds the k-tile period?"""
import os, subprocess, sys
code = r"""
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), 'tests'))
import torch
import gpu_g3_bench as b
t4 = b.t4
b.M = int(os.environ.get('EXP_M', '4096'))
b.one(9728, 2560, t4.KERNEL_GEMM3, 'gemm3 exp=' + os.environ.get('TCQ4_G3_EXP', '0'))
"""
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for exp in ("0", "1", "2", "4", "6", "7"):
    env = dict(os.environ, TCQ4_G3_EXP=exp, TCQ4_G3_PROF="1", TCQ4_B200_USE_DIAG="1")
    subprocess.run([sys.executable, "-c", code], cwd=root, env=env)
