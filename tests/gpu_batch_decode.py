"""GPU diagnostic (not a pytest file): bench.py's configs[2] leg (Llama-3-8B layer, imatrix permutation, decode batch) for the
batch sizes on the command line -- used with the GEMV's experiment switches (TCQ4_GV_PRE_M, TCQ4_GV_PRE_BLOCKS, ...)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
import bench
t4 = graft.load_package()
Ms = tuple(int(a) for a in sys.argv[1:]) or (1, 2, 4, 8)
r = bench.batch_decode_leg(t4, torch, 6650.0, Ms)
for row in r["per_batch"]:
    print(f"M={row['M']:2d}: {row['us_per_layer']:7.2f} us per layer, {row['launches_per_layer']:.0f} launches, {row['weight_gbs']:6.0f} GB/s of weights", flush=True)
