"""GPU diagnostic (not a pytest file): MUL_MAT_ID timing on a Qwen3-30B-A3B-shaped expert block (128 experts, 8 used per token,
hidden 2048, expert ffn 768).  Decode (1 token): gate/up/down as one CUDA graph, ids changed on the device between replays.
Prefill (512 tokens): the host-sorted path, eager (it synchronises the stream to read the ids, like the reference)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

t4 = graft.load_package()
E, U, H, F = 128, 8, 2048, 768
LAYERS = int(os.environ.get("LAYERS", "8"))


def experts(N, K, seed):
    return torch.stack([t4.random_tiles(N, K, seed=seed * 1000 + e) for e in range(E)])


gate = [experts(F, H, 3 * l) for l in range(LAYERS)]
up = [experts(F, H, 3 * l + 1) for l in range(LAYERS)]
down = [experts(H, F, 3 * l + 2) for l in range(LAYERS)]
bytes_per_layer = U * (2 * t4.tile_bytes(F, H) + t4.tile_bytes(H, F))

# ---- decode --------------------------------------------------------------------------------------------------------
x = torch.randn((1, 1, H), device="cuda")
h = torch.randn((1, U, F), device="cuda")
ids = torch.randperm(E, device="cuda")[:U].to(torch.int32).reshape(1, U).contiguous()
yg = torch.empty((1, U, F), device="cuda")
yu = torch.empty((1, U, F), device="cuda")
yd = torch.empty((1, U, H), device="cuda")
t4.reserve(1, H, H)


def decode_step():
    for l in range(LAYERS):
        t4.mul_mat_id(gate[l], x, ids, F, out=yg)
        t4.mul_mat_id(up[l], x, ids, F, out=yu)
        t4.mul_mat_id(down[l], h, ids, H, out=yd)


decode_step()
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        decode_step()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 20
new_ids = [torch.randperm(E, device="cuda")[:U].to(torch.int32).reshape(1, U) for _ in range(reps)]
torch.cuda.synchronize()
e0.record()
for r in range(reps):
    ids.copy_(new_ids[r])
    g.replay()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"decode: {LAYERS} layers x (gate, up, down) x {U} experts: {ms * 1e3:.1f} us/step, {ms * 1e3 / (LAYERS * 3):.2f} us per MUL_MAT_ID, "
      f"{ms * 1e3 / (LAYERS * 3 * U):.2f} us per expert GEMV, {LAYERS * bytes_per_layer / ms / 1e6:.0f} GB/s of expert weights", flush=True)

# ---- prefill -------------------------------------------------------------------------------------------------------
T = int(os.environ.get("TOKENS", "512"))
xp = torch.randn((T, 1, H), device="cuda")
hp = torch.randn((T, U, F), device="cuda")
idp = torch.stack([torch.randperm(E, device="cuda")[:U] for _ in range(T)]).to(torch.int32)
og = torch.empty((T, U, F), device="cuda")
od = torch.empty((T, U, H), device="cuda")
for _ in range(2):
    t4.mul_mat_id(gate[0], xp, idp, F, out=og)
    t4.mul_mat_id(down[0], hp, idp, H, out=od)
torch.cuda.synchronize()
t0 = time.perf_counter()
for l in range(LAYERS):
    t4.mul_mat_id(gate[l], xp, idp, F, out=og)
    t4.mul_mat_id(up[l], xp, idp, F, out=og)
    t4.mul_mat_id(down[l], hp, idp, H, out=od)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
ops = 2.0 * T * U * (2 * F * H + H * F) * LAYERS
print(f"prefill: {T} tokens x {U} of {E} experts, {LAYERS} layers x (gate, up, down): {dt * 1e3:.2f} ms wall "
      f"({dt * 1e3 / (LAYERS * 3):.3f} ms per MUL_MAT_ID, {ops / dt / 1e12:.1f} useful TOP/s)", flush=True)
