"""Measure warp-level pipe costs and the TMEM read rate (not a pytest file)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
t4 = graft.load_package()
for kind, name in enumerate(["mma.sync.m16n8k32.s8", "dp4a", "shfl.xor", "redux.add", "cvt.rn.f32.s32", "mad.lo.s32"]):
    v = C.c_double()
    rc = t4.diag_lib().tcq4_b200_measure_pipe(kind, 2000, C.byref(v))
    print(f"{name:24s} rc={rc} {v.value:8.2f} cycles per warp-instruction per SM (16 warps resident)", flush=True)
for cols in (8, 16, 32):
    for warps in (4, 8, 16, 32):
        v = C.c_double()
        rc = t4.diag_lib().tcq4_b200_measure_tmem_ld(cols, warps, 4000, C.byref(v))
        print(f"tcgen05.ld.32x32b.x{cols:<2d} {warps:2d} warps rc={rc} {v.value:8.1f} B/clk/SM", flush=True)
