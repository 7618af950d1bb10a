"""GPU diagnostic (not a pytest file): what the channel permutation costs in the Llama-3-8B decode-batch leg (bench.py configs[2])."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft, bench
t4 = graft.load_package()
Ms = tuple(int(a) for a in sys.argv[1:]) or (1, 2, 4, 8, 32)
r = bench.batch_decode_leg(t4, torch, 6650.0, Ms)
print("with perms   ", [round(x["us_per_layer"], 1) for x in r["per_batch"]], flush=True)
orig = t4.register_perm
t4.register_perm = lambda name, perm: None          # same leg, no permutation registered
r = bench.batch_decode_leg(t4, torch, 6650.0, Ms)
print("without perms", [round(x["us_per_layer"], 1) for x in r["per_batch"]], flush=True)
