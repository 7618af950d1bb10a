"""cta_group::2 M=256 MMA issue loop: kind::i8 against kind::f16 with A in shared memory / in TMEM (not a pytest file).
The kind is read from the environment once per process, so each variant runs in a child."""
import os, subprocess, sys
code = r'''
import os, sys, ctypes as C
sys.path.insert(0, os.getcwd())
import __graft_entry__ as graft
t4 = graft.load_package()
for N in (96, 128, 160, 192, 224, 256):
    tops, cyc = C.c_double(), C.c_double()
    rc = t4.diag_lib().tcq4_b200_measure_umma2(N, 4000, C.byref(tops), C.byref(cyc))
    print(f"kind={os.environ.get('TCQ4_UMMA2_KIND','0')} bg={os.environ.get('TCQ4_UMMA2_BG','-')} M=256 N={N:3d}: rc={rc} {tops.value:8.1f} TOPS  {cyc.value:7.1f} cyc/MMA", flush=True)
'''
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for kind in ("0", "1", "2"):
    for bg in (None, "12,1"):
        env = dict(os.environ, TCQ4_UMMA2_KIND=kind)
        if bg:
            env["TCQ4_UMMA2_BG"] = bg
        subprocess.run([sys.executable, "-c", code], cwd=root, env=env)
