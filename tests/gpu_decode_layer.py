"""GPU diagnostic (not a pytest file): the four launches of a Llama-3-8B decode layer one by one (graph of NBUF launches over
distinct weights), M on the command line: q/k/v fan-out, o, gate+up fan-out, down -- time and weight GB/s of each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpu_g3_bench as b
t4 = b.t4
for M in (int(a) for a in (sys.argv[1:] or ["1"])):
    b.M = M
    b.NBUF = 6
    tot = 0.0
    for kind, Ns, K in (("fan", [4096, 1024, 1024], 4096), ("one", [4096], 4096), ("fan", [14336, 14336], 4096), ("one", [4096], 14336)):
        us = b.fan(Ns, K, "llama3-8b") if kind == "fan" else b.one(Ns[0], K, t4.KERNEL_AUTO, "auto")
        mb = sum(Ns) * K * 0.578 / 1e6
        print(f"   M={M} N={Ns} K={K}: {us:6.2f} us, {mb:5.1f} MB of weights = {mb / us * 1e-3 * 1e6 / 1e3:6.0f} GB/s", flush=True)
        tot += us
    print(f"M={M}: layer total {tot:.1f} us", flush=True)
