#!/usr/bin/env python
"""bench.py -- TCQ4_K32 W4A4 MUL_MAT on B200: Qwen3-4B's 252 linear matmuls (BASELINE.json configs[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W]            own arm (CUDA, through the C ABI)
  python bench.py --impl reference [--steps K] [--warmup W]      the reference's CPU path (oracle/_ref)
  torchrun ... bench.py --gpus N ...                             N-sharded (column-parallel) + NCCL all-gather

A step = one 512-token prefill pass over the 36 x 7 TCQ4_K32 matmuls of Qwen3-4B (q,k,v,o,gate,up,down),
synthetic tiles, 252 distinct weight buffers (2.1 GB >> 126 MB L2, so every weight byte comes from HBM).
`value` = prefill tokens/s of that matmul chain (linear layers only -- all this path covers); the same
JSON line also carries the batch-1 decode pass (`decode`), the per-kernel roofline legs, the end-to-end
number through the host-buffer C-ABI call, and the CPU baseline.  Nothing here reads /root/reference.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Qwen3-4B (hidden 2560, ffn 9728, 32 q heads x 128, 8 kv heads x 128, 36 layers): (name, N, K)
QWEN3_4B_LAYER = [("attn_q", 4096, 2560), ("attn_k", 1024, 2560), ("attn_v", 1024, 2560), ("attn_output", 2560, 4096),
                  ("ffn_gate", 9728, 2560), ("ffn_up", 9728, 2560), ("ffn_down", 2560, 9728)]
QWEN3_4B_LAYERS = 36
PREFILL_M = 512
METRIC = "qwen3_4b_tcq4_k32_mul_mat_prefill512_tokens_per_s"
UNIT = "tok/s"
TILE_BYTES = 1184


def alg_bytes(M, N, K):
    """N*K*1184/2048 weight bytes + fp32 activations in + fp32 outputs out."""
    return N * K * TILE_BYTES // 2048 + 4 * M * K + 4 * M * N


def layer_ops(M):
    return sum(2 * M * N * K for _, N, K in QWEN3_4B_LAYER)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# --------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md recipe)
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = f"/tmp/tcq4_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "20"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1]))
                    mx.append(float(c[2]))
                except ValueError:
                    continue
                for i, nme in enumerate(names):
                    if c[5 + i].lower().startswith("active"):
                        reasons.add(nme)
            os.unlink(self.path)
        except Exception:  # noqa: BLE001
            pass
        if sm:
            busy = [v for v in sm if v > 0.5 * max(sm)] or sm
            out.update(sm_mhz=statistics.median(busy), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# --------------------------------------------------------------------------------------------------
# CPU side: the reference's own CPU code (oracle/_ref), used ONLY as the reported baseline
# --------------------------------------------------------------------------------------------------
def cpu_random_tiles(N, K, seed):
    """Same distribution as the device generator: random nibbles, sc in [1,127], S ~ U(1e-3,1e-2) fp16, Z = zc = 0."""
    import numpy as np
    rng = np.random.default_rng(seed)
    nt = (N // 8) * (K // 256)
    t = np.zeros((nt, TILE_BYTES), dtype=np.uint8)
    t[:, :1024] = rng.integers(0, 256, (nt, 1024), dtype=np.uint8)
    t[:, 1024:1040] = (rng.random((nt, 8)) * 9e-3 + 1e-3).astype(np.float16).view(np.uint8).reshape(nt, 16)
    t[:, 1056:1120] = rng.integers(1, 128, (nt, 64), dtype=np.uint8)
    return t.reshape(-1)


class CpuLayer:
    """One Qwen3-4B layer's 7 matmuls on the host cores: dequantize_row_tcq4_tile -> F32, ggml_fwht_impl
    on the activations, ggml CPU backend F32 MUL_MAT.  A bounded sample of the workload:
    tokens/s = M / (36 * t_layer)."""

    def __init__(self, M):
        import numpy as np
        from oracle import pyoracle
        self.np = np
        self.rb = pyoracle.RefBaseline()
        self.M = M
        self.tiles = [cpu_random_tiles(N, K, 1000 + i) for i, (_, N, K) in enumerate(QWEN3_4B_LAYER)]
        rng = np.random.default_rng(5)
        self.x = {K: rng.standard_normal((M, K)).astype(np.float32) for K in {k for _, _, k in QWEN3_4B_LAYER}}

    def run(self):
        parts = {"dequant_s": 0.0, "fwht_s": 0.0, "matmul_s": 0.0}
        t0 = time.perf_counter()
        for (_, N, K), tl in zip(QWEN3_4B_LAYER, self.tiles):
            _, t = self.rb.mul_mat(tl, self.x[K], N)
            for k in parts:
                parts[k] += t[k]
        return time.perf_counter() - t0, parts


def cpu_baseline_leg(M=PREFILL_M, reps=2):
    layer = CpuLayer(M)
    layer.run()                                  # warm-up: thread pool, page faults
    best, parts = None, None
    for _ in range(reps):
        t, p = layer.run()
        if best is None or t < best:
            best, parts = t, p
    return {"value": M / (best * QWEN3_4B_LAYERS), "unit": UNIT, "cores": layer.rb.threads, "kind": layer.rb.kind,
            "sample": f"1 of {QWEN3_4B_LAYERS} layers (7 matmuls, M={M}), best of {reps}, scaled x{QWEN3_4B_LAYERS}",
            "layer_s": best, "parts_s": {k: round(v, 4) for k, v in parts.items()}}


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path (there is no CPU TCQ4 MUL_MAT in
    the fork, README.md:54; north_star defines the CPU baseline as quantiser output -> dequant -> F32 MUL_MAT)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    layer = CpuLayer(PREFILL_M)
    for _ in range(max(1, min(args.warmup, 2))):
        layer.run()                                        # warm-up: thread pool, page faults (one layer each)
    ts = []
    for _ in range(args.steps):                            # a step = all 36 layers (the same layer's tiles 36 times: the
        t0 = time.perf_counter()                           # arithmetic of a step, a bounded sample of its weights)
        for _l in range(QWEN3_4B_LAYERS):
            layer.run()
        ts.append(time.perf_counter() - t0)
    t_step = sum(ts) / len(ts)
    val = PREFILL_M / t_step
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": base_config(),
            "detail": {"arm": "reference CPU path (oracle/_ref: dequantize_row_tcq4_tile + ggml_fwht_impl + ggml CPU F32 MUL_MAT)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": layer.rb.threads, "kind": layer.rb.kind,
                             "sample": f"each step = {QWEN3_4B_LAYERS} x (7 matmuls, M={PREFILL_M}) over one layer's synthetic tiles"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def base_config():
    """`config` of BOTH arms (key-identical; arm-specific detail goes under `detail`)."""
    return {"workload": "Qwen3-4B TCQ4_K32 252 linear matmuls, 512-token prefill (BASELINE.json configs[1])",
            "M": PREFILL_M, "layers": QWEN3_4B_LAYERS, "matmuls_per_layer": len(QWEN3_4B_LAYER)}


# --------------------------------------------------------------------------------------------------
# own arm
# --------------------------------------------------------------------------------------------------
class Chain:
    """The 252 matmuls of Qwen3-4B, N-sharded over `world` ranks (shard = contiguous tile byte range)."""

    def __init__(self, t4, torch, M, world, rank, weights=None, gather=True, mode="fused", per_node=False, barrier="layer"):
        self.t4, self.torch, self.M, self.world, self.rank = t4, torch, M, world, rank
        self.gather = gather                           # False: the sharded MUL_MATs alone (SURVEY 8e reports both)
        self.barrier = barrier                         # cross-GPU barrier per "layer" (a consumer of layer l's outputs may start) or per "step"
        self.ops = []                                  # (sharded weight, K, N)
        if weights is None:
            weights = []
            for layer in range(QWEN3_4B_LAYERS):
                for j, (_, N, K) in enumerate(QWEN3_4B_LAYER):
                    n0, n1, _, _ = t4.shard_range(N, K, world, rank, 64)
                    tl = t4.random_tiles(n1 - n0, K, seed=(layer * 16 + j) * 64 + rank)
                    weights.append(t4.ShardedWeight(N, K, world, rank, tiles_local=tl, align=64))
        self.weights = weights
        g = torch.Generator(device="cuda")
        g.manual_seed(17)
        Ks = sorted({K for _, _, K in QWEN3_4B_LAYER})
        Ns = sorted({N for _, N, _ in QWEN3_4B_LAYER})
        self.x = {K: torch.randn((M, K), device="cuda", generator=g) for K in Ks}
        self.y_local = {N: torch.empty((M, t4.shard_range(N, 256, world, rank, 64)[1] - t4.shard_range(N, 256, world, rank, 64)[0]),
                                       device="cuda") for N in Ns}
        wsz = max(t4.workspace_size(M, w.n_local, w.K) for w in weights)
        self.ws = t4.Workspace(wsz)
        # N > 1 with the gather: every matmul keeps its own column block until its all-gather (second stream) has
        # read it, so the gather of matmul i overlaps the kernels of matmul i+1
        self.y_each = None
        self.comm = None
        self.sym = None                                # N -> (symmetric [M][N] buffer, handle, peer pointers)
        self.mode = mode
        self.gather_kind = "none"
        if world > 1 and gather:
            try:
                # fused path: every rank's epilogue stores its column block straight into every rank's C
                import torch.distributed as dist
                import torch.distributed._symmetric_memory as symm_mem
                self.sym = {}                          # one symmetric C per matmul of a layer (every node has its own dst)
                for j, (_, N, _) in enumerate(QWEN3_4B_LAYER):
                    buf = symm_mem.empty((M, N), dtype=torch.float32, device="cuda")
                    hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
                    self.sym[j] = (buf, hdl, [int(p) for p in hdl.buffer_ptrs])
                self.mc = {j: int(getattr(self.sym[j][1], "multicast_ptr", 0) or 0) for j in self.sym}
                self.pbar = t4.PeerBarrier()           # tcq4_b200_peer_barrier: one 32-thread kernel per barrier
                if mode == "multicast" and not all(self.mc.values()):
                    raise RuntimeError("no multicast (NVLS) address for the symmetric buffers")
                self.mode = mode
                if mode == "multicast":
                    self.gather_kind = ("epilogue stores through the NVSwitch multicast address of the symmetric C (multimem.st, "
                                        f"tcq4_b200_mul_mat_multicast) + one cross-GPU barrier per {barrier}")
                elif mode == "fused":
                    self.gather_kind = ("fused peer stores (tcq4_b200_mul_mat_peers over symmetric memory) + one cross-GPU "
                                        f"barrier per {barrier}")
                else:
                    self.copy_streams = [torch.cuda.Stream() for _ in range(world)]
                    # the MUL_MATs of this variant are ordinary single-destination calls, so they go through the graph-level
                    # entry like the single-GPU step: one GgmlGraph per run of nodes sharing src1 (q/k/v, o, gate/up, down),
                    # each node's dst = this rank's column window of the symmetric C
                    Lr = len(QWEN3_4B_LAYER)
                    fan = {"attn_q": 0, "attn_k": 0, "attn_v": 0, "ffn_gate": 1, "ffn_up": 1}
                    self.groups = []
                    i = 0
                    while i < len(weights):
                        j = i + 1
                        gi = fan.get(QWEN3_4B_LAYER[i % Lr][0])
                        while gi is not None and j < len(weights) and j // Lr == i // Lr and fan.get(QWEN3_4B_LAYER[j % Lr][0]) == gi:
                            j += 1
                        nodes = []
                        for t in range(i, j):
                            w = weights[t]
                            if w.n_local:
                                out = self.sym[t % Lr][0][:, w.n0:w.n1]
                                nodes.append((w.tiles, self.x[w.K], out, w.n_local, f"blk.{t // Lr}.{QWEN3_4B_LAYER[t % Lr][0]}.weight"))
                        self.groups.append((t4.GgmlGraph(nodes) if nodes else None, list(range(i, j))))
                        i = j
                    t4.reserve(M, max(max(w.n_local for w in weights), 8), max(K for _, _, K in QWEN3_4B_LAYER))
                    t4.reserve(M, max(8, sum(w.n_local for w in weights[4:6])), 2560)
                    self.gather_kind = ("MUL_MAT into the local symmetric C + copy-engine pushes of the column block to every "
                                        f"peer on side streams (tcq4_b200_push_columns) + one cross-GPU barrier per {barrier}")
            except Exception as e:  # noqa: BLE001
                print(f"[bench] symmetric memory unavailable ({type(e).__name__}: {e}); NCCL all-gather on a second stream",
                      file=sys.stderr)
                self.sym = None
                self.y_each = [torch.empty((M, w.n_local), device="cuda") for w in weights]
                self.comm = torch.cuda.Stream()
                self.gather_kind = "NCCL all-gather of every output on a second stream (overlaps the next matmul)"
        self.graph = None
        self.launches_per_step = None
        # single GPU: the step is ONE call of the graph-level entry (tcq4_b200_ggml_graph_compute) over the 252 MUL_MAT nodes in
        # execution order, every node with its own dst as in a ggml graph; the library fans out the runs that share src1
        self.ggml = None
        if (world == 1 or not gather) and not per_node:      # (N > 1 without the exchange: the same graph over the shards)
            Lr = len(QWEN3_4B_LAYER)
            self.y_node = {QWEN3_4B_LAYER[j][0]: torch.empty((M, max(8, weights[j].n_local)), device="cuda") for j in range(Lr)}
            nodes = []
            for i, w in enumerate(weights):
                name = QWEN3_4B_LAYER[i % Lr][0]
                if w.n_local:
                    nodes.append((w.tiles, self.x[w.K], self.y_node[name], w.n_local, f"blk.{i // Lr}.{name}.weight"))
            self.ggml = t4.GgmlGraph(nodes)
            t4.reserve(M, max(8, max(w.n_local for w in weights)), max(K for _, _, K in QWEN3_4B_LAYER))
            t4.reserve(M, max(8, sum(w.n_local for w in weights[4:6])), 2560)

    def weight_bytes(self):
        return sum(w.tiles.numel() for w in self.weights)

    def step_eager(self):
        torch = self.torch
        last = None
        L = len(QWEN3_4B_LAYER)
        any_hdl = self.sym[0][1] if self.sym is not None else None
        if self.sym is not None and self.mode in ("multicast", "fused"):
            for i, w in enumerate(self.weights):
                buf, hdl, ptrs = self.sym[i % L]
                if w.n_local:
                    if self.mode == "multicast":
                        self.t4.mul_mat_multicast(w.tiles, self.x[w.K], self.mc[i % L], w.N, w.n0, w.n_local)
                    else:
                        self.t4.mul_mat_peers(w.tiles, self.x[w.K], ptrs, w.N, w.n0, w.n_local)
                last = buf
                if self.barrier == "layer" and (i + 1) % L == 0:
                    self.pbar()                      # every rank's stores of this layer have landed everywhere
            if self.barrier != "layer":
                self.pbar()
            return last
        if self.sym is not None:
            main = torch.cuda.current_stream()
            for g, idxs in self.groups:
                if g is not None:
                    g.compute()
                ev = torch.cuda.Event()
                ev.record(main)
                for r in range(self.world):
                    if r != self.rank:
                        cs = self.copy_streams[r]
                        cs.wait_event(ev)
                        for t in idxs:
                            w = self.weights[t]
                            if w.n_local:
                                ptrs = self.sym[t % L][2]
                                self.t4.push_columns(ptrs[r], ptrs[self.rank], self.M, w.N, w.n0, w.n_local, stream=cs)
                last = self.sym[idxs[-1] % L][0]
                if (self.barrier == "layer" and (idxs[-1] + 1) % L == 0) or idxs[-1] + 1 == len(self.weights):
                    for r in range(self.world):
                        if r != self.rank:
                            main.wait_stream(self.copy_streams[r])
                    self.pbar()
            return last
        if self.ggml is not None:
            self.ggml.compute()
            return self.y_node[QWEN3_4B_LAYER[-1][0]]
        if self.comm is None:
            for w in self.weights:
                last = w.local(self.x[w.K], out=self.y_local[w.N], workspace=self.ws)
            return last
        main = torch.cuda.current_stream()
        for i, w in enumerate(self.weights):
            yl = w.local(self.x[w.K], out=self.y_each[i], workspace=self.ws)
            ev = torch.cuda.Event()
            ev.record(main)
            self.comm.wait_event(ev)
            with torch.cuda.stream(self.comm):
                last = w.gather(yl)                      # NCCL all-gather + re-layout into C[m*N + n]
        main.wait_stream(self.comm)                      # every output of the step is gathered before the step ends
        return last

    def capture(self):
        torch = self.torch
        c0 = self.t4.launch_count()
        self.step_eager()
        self.launches_per_step = self.t4.launch_count() - c0
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                with torch.cuda.graph(g, stream=s):
                    self.step_eager()
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            self.graph = g
        except Exception as e:  # noqa: BLE001  (NCCL capture can be refused; eager is still correct)
            print(f"[bench] CUDA graph capture failed ({type(e).__name__}: {e}); running eager", file=sys.stderr)
            self.graph = None
            torch.cuda.synchronize()

    def step(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.step_eager()


def per_node_leg(t4, torch, pre, dec, steps, warmup):
    """The same chains with one tcq4_b200_mul_mat call per MUL_MAT node (no graph-level fusion), reported beside the headline;
    the outputs of the last layer are compared (prefill: bit-identical unless a K split applies to the node; decode: the fp32
    sum over 256-blocks is cut between warps at other places in the segmented launch, same int32 accumulators)."""
    out = {"what": "one tcq4_b200_mul_mat call per MUL_MAT node (the round-1 headline path); the headline above goes through "
                   "tcq4_b200_ggml_graph_compute, which fans out consecutive nodes sharing src1"}
    L = len(QWEN3_4B_LAYER)
    for key, base, M, n in (("prefill", pre, pre.M, steps), ("decode", dec, 1, max(steps, 20))):
        pn = Chain(t4, torch, M, 1, 0, weights=base.weights, per_node=True)
        pn.capture()
        ms = timed_steps(torch, None, pn, n, max(warmup, 3), 1)
        base.step()
        torch.cuda.synchronize()
        same, worst = True, 0.0
        for j, (name, N, K) in enumerate(QWEN3_4B_LAYER):
            ref = t4.mul_mat(base.weights[-L + j].tiles, base.x[K], N)
            got = base.y_node[name]
            same = same and bool(torch.equal(ref, got))
            worst = max(worst, float(((ref - got).double() ** 2).sum() / (ref.double() ** 2).sum()))
        out[key] = {"ms_per_step": ms, "value": M / (ms * 1e-3), "unit": UNIT, "launches_per_step": int(pn.launches_per_step),
                    "graph_outputs_bit_identical_to_per_node": same, "nmse_graph_vs_per_node": worst}
        del pn
    return out


# BASELINE.json configs[3] and configs[4]: one transformer layer's seven TCQ4 matmuls (name, N, K), N-sharded over the ranks
BIG_CONFIGS = {
    "qwen3_32b_prefill4096": {"Ms": [4096], "worlds": (1, 2, 4, 8),
                              "layer": [("attn_q", 8192, 5120), ("attn_k", 1024, 5120), ("attn_v", 1024, 5120), ("attn_output", 5120, 8192),
                                        ("ffn_gate", 25600, 5120), ("ffn_up", 25600, 5120), ("ffn_down", 5120, 25600)]},
    "llama3_70b_column_parallel": {"Ms": [1, 8, 64, 512, 4096, 8192], "worlds": (1, 8),
                                   "layer": [("attn_q", 8192, 8192), ("attn_k", 1024, 8192), ("attn_v", 1024, 8192), ("attn_output", 8192, 8192),
                                             ("ffn_gate", 28672, 8192), ("ffn_up", 28672, 8192), ("ffn_down", 8192, 28672)]},
}


def configs_leg(t4, torch, dist, world, rank, mode):
    """Per config and batch size: one layer's seven N-sharded MUL_MATs, (a) into this rank's own C only, (b) with the exchange
    of the fp32 column blocks (`mode`) and one cross-GPU barrier per layer.  ms per layer (max over ranks), useful TOP/s per GPU,
    outbound exchange bytes per GPU and layer, and the NVLink time those bytes would need at 900 GB/s per direction."""
    out = {}
    symm_mem = None
    if world > 1:
        import torch.distributed._symmetric_memory as symm_mem
        pbar = t4.PeerBarrier()
        copy_streams = [torch.cuda.Stream() for _ in range(world)]
    for cname, cfg in BIG_CONFIGS.items():
        if world not in cfg["worlds"]:
            continue
        res = []
        for M in cfg["Ms"]:
            ws_, xs, cs, fans = [], {}, [], None
            for j, (_, N, K) in enumerate(cfg["layer"]):
                n0, n1, _, _ = t4.shard_range(N, K, world, rank, 64)
                ws_.append(t4.ShardedWeight(N, K, world, rank, tiles_local=t4.random_tiles(n1 - n0, K, seed=977 + j * 8 + rank), align=64))
                if K not in xs:
                    xs[K] = torch.randn((M, K), device="cuda")
                if world > 1:
                    buf = symm_mem.empty((M, N), dtype=torch.float32, device="cuda")
                    hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
                    cs.append((buf, hdl, [int(p) for p in hdl.buffer_ptrs], int(getattr(hdl, "multicast_ptr", 0) or 0)))
                else:
                    cs.append((torch.empty((M, N), device="cuda"), None, None, 0))
            t4.reserve(M, max(w.n_local for w in ws_) or 8, max(K for _, _, K in cfg["layer"]))

            def layer(exchange):
                main = torch.cuda.current_stream()
                for j, w in enumerate(ws_):
                    buf, hdl, ptrs, mc = cs[j]
                    if not w.n_local:
                        continue
                    if world == 1:
                        t4.mul_mat(w.tiles, xs[w.K], w.N, out=buf)
                    elif not exchange or mode == "dma":
                        t4.mul_mat_peers(w.tiles, xs[w.K], [ptrs[rank]], w.N, w.n0, w.n_local)
                        if exchange:                       # copy engines push the finished column block while the next matmul runs
                            ev = torch.cuda.Event()
                            ev.record(main)
                            for r in range(world):
                                if r != rank:
                                    copy_streams[r].wait_event(ev)
                                    t4.push_columns(ptrs[r], ptrs[rank], M, w.N, w.n0, w.n_local, stream=copy_streams[r])
                    elif mode == "multicast" and mc:
                        t4.mul_mat_multicast(w.tiles, xs[w.K], mc, w.N, w.n0, w.n_local)
                    else:
                        t4.mul_mat_peers(w.tiles, xs[w.K], ptrs, w.N, w.n0, w.n_local)
                if world > 1 and exchange:
                    if mode == "dma":
                        for r in range(world):
                            if r != rank:
                                main.wait_stream(copy_streams[r])
                    pbar()

            def time_it(exchange, reps=5):
                for _ in range(2):
                    layer(exchange)
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(reps):
                    layer(exchange)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                if world > 1:
                    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    ms = float(t.item())
                return ms

            ms_local = time_it(False)
            ms_x = time_it(True) if world > 1 else None
            ops_gpu = sum(2 * M * w.n_local * w.K for w in ws_)
            out_bytes = sum(4 * M * w.n_local for w in ws_) * (1 if (mode == "multicast") else (world - 1)) if world > 1 else 0      # per GPU, outbound
            r = {"M": M, "ms_per_layer_shard_local": ms_local, "useful_tops_per_gpu_shard_local": ops_gpu / (ms_local * 1e-3) / 1e12,
                 "weight_gbs_per_gpu_shard_local": sum(w.tiles.numel() for w in ws_) / (ms_local * 1e-3) / 1e9}
            if world > 1:
                r.update({"ms_per_layer_with_exchange": ms_x, "useful_tops_per_gpu_with_exchange": ops_gpu / (ms_x * 1e-3) / 1e12,
                          "exchange_outbound_bytes_per_gpu": out_bytes,
                          "nvlink_bound_ms_at_900GBs": out_bytes / 900e9 * 1e3,
                          "exchange_gbs_per_gpu_if_serial": out_bytes / (ms_x * 1e-3) / 1e9})
            res.append(r)
            del ws_, xs, cs
            torch.cuda.empty_cache()
        out[cname] = {"shapes": [f"{n}: N={N} K={K}" for n, N, K in cfg["layer"]], "exchange": mode if world > 1 else "none",
                      "timing": "eager calls, CUDA events over 5 layer passes, max over ranks", "per_batch": res}
    return out


# BASELINE.json configs[2]: Llama-3-8B with imatrix channel reordering, decode batch 1-64.  One layer's seven TCQ4 matmuls
LLAMA3_8B_LAYER = [("attn_q", 4096, 4096, "attn_norm"), ("attn_k", 1024, 4096, "attn_norm"), ("attn_v", 1024, 4096, "attn_norm"),
                   ("attn_output", 4096, 4096, "kqv_out"), ("ffn_gate", 14336, 4096, "ffn_norm"), ("ffn_up", 14336, 4096, "ffn_norm"),
                   ("ffn_down", 4096, 14336, "ffn_act")]


def batch_decode_leg(t4, torch, peak_gbs, Ms=(1, 2, 4, 8, 16, 32, 64)):
    """Llama-3-8B layer shapes with a synthetic-imatrix channel permutation per activation tensor (tensors fed by the same
    activations share it, as derive_tcq4_perms_from_imatrix gives them), batch sizes 1..64: the ggml-graph entry over three
    layers of distinct weights (378 MB, L2-cold), replayed as one CUDA graph; us per layer, launches per layer, weight GB/s."""
    import numpy as np
    rng = np.random.default_rng(2024)
    n_layers = 3
    t4.set_reorder_enabled(True)
    perms = {}
    for _, _, K, src in LLAMA3_8B_LAYER:
        if src not in perms:
            perms[src] = t4.perm_from_imatrix(rng.gamma(2.0, 1.0, K).astype(np.float32))
    weights = []
    for l in range(n_layers):
        for j, (nm, N, K, src) in enumerate(LLAMA3_8B_LAYER):
            name = f"blk.{l}.{nm}.weight"
            t4.register_perm(name, perms[src])
            weights.append((name, N, K, src, t4.random_tiles(N, K, seed=5000 + l * 8 + j)))
    wbytes = sum(w[4].numel() for w in weights) / n_layers
    res = []
    for M in Ms:
        xs = {src: torch.randn((M, K), device="cuda") for _, _, K, src in LLAMA3_8B_LAYER}
        nodes = [(w, xs[src], torch.empty((M, N), device="cuda"), N, name) for (name, N, K, src, w) in weights]
        t4.reserve(M, 14336, 14336)
        g = t4.GgmlGraph(nodes)
        n0 = t4.launch_count()
        g.compute()
        launches = (t4.launch_count() - n0) / n_layers
        torch.cuda.synchronize()
        cg = torch.cuda.CUDAGraph()
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            with torch.cuda.graph(cg, stream=st):
                g.compute(stream=st)
        torch.cuda.current_stream().wait_stream(st)
        for _ in range(3):
            cg.replay()
        torch.cuda.synchronize()
        reps, trials = 20, []
        for _ in range(5):      # 5 bursts of 20 replays, median burst: a burst is 3-7 ms, short enough for one host hiccup to show
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                cg.replay()
            e1.record()
            torch.cuda.synchronize()
            trials.append(e0.elapsed_time(e1) * 1e3 / reps / n_layers)
        trials.sort()
        us = trials[len(trials) // 2]
        res.append({"M": M, "us_per_layer": us, "us_per_layer_min": trials[0], "us_per_layer_max": trials[-1],
                    "launches_per_layer": launches, "weight_gbs": wbytes / (us * 1e-6) / 1e9,
                    "frac_hbm": wbytes / (us * 1e-6) / 1e9 / peak_gbs, "tok_per_s_per_layer": M / (us * 1e-6)})
        del nodes, g, cg
    t4.clear_perms()
    return {"shapes": [f"{n}: N={N} K={K}" for n, N, K, _ in LLAMA3_8B_LAYER], "perm": "synthetic imatrix, per activation tensor",
            "timing": "tcq4_b200_ggml_graph_compute over 3 layers of distinct weights in one CUDA graph; median of 5 bursts of 20 replays",
            "weight_bytes_per_layer": wbytes, "per_batch": res}


def timed_steps(torch, dist, chain, steps, warmup, world):
    """W untimed steps, then exactly K steps between barrier + synchronize, CUDA events, max over ranks."""
    for _ in range(warmup):
        chain.step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        chain.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    return ms / steps


def kernel_leg(t4, torch, chain, steps, kernel, M):
    """Per-launch CUDA events around the main kernel over an eager pass of the same steps (roofline leg)."""
    n_launch = len(chain.weights) * steps
    t4.profile_begin(n_launch)
    for _ in range(steps):
        if chain.ggml is not None:
            chain.ggml.compute()
        else:
            for w in chain.weights:
                w.local(chain.x[w.K], out=chain.y_local[w.N], workspace=chain.ws)
    main_ms, pre_ms, n = t4.profile_end(kernel)
    return main_ms, pre_ms, n


def e2e_leg(t4, torch, chain, steps, M):
    """Same chain through the host-buffer C-ABI entries (tcq4_b200_mul_mat_host_async / tcq4_b200_mul_mat_fanout_host_async +
    tcq4_b200_host_sync): per MUL_MAT node -- per group of consecutive nodes that share src1 (Q/K/V, gate/up), as the graph-level
    entry groups them -- an H2D copy of x from pinned memory, the MUL_MATs, a D2H copy of every y into pinned memory; the
    library overlaps the copies of neighbouring calls with the kernels; one sync per step.  Returns (seconds per step, H2D
    bytes, D2H bytes) with the bytes counted from the tensors actually copied."""
    xh = {K: v.cpu().pin_memory() for K, v in chain.x.items()}
    yh = [torch.empty((M, w.n_local), dtype=torch.float32).pin_memory() for w in chain.weights[:len(QWEN3_4B_LAYER)]]
    L = len(QWEN3_4B_LAYER)
    share = {"attn_q": 0, "attn_k": 0, "attn_v": 0, "ffn_gate": 1, "ffn_up": 1}       # nodes fed by the same activations
    groups, i = [], 0
    while i < len(chain.weights):
        j = i + 1
        nm = QWEN3_4B_LAYER[i % L][0]
        while (j < len(chain.weights) and j // L == i // L and nm in share and QWEN3_4B_LAYER[j % L][0] in share
               and share[QWEN3_4B_LAYER[j % L][0]] == share[nm]):
            j += 1
        groups.append(list(range(i, j)))
        i = j
    h2d = sum(4 * M * chain.weights[g[0]].K for g in groups)
    d2h = sum(4 * M * w.n_local for w in chain.weights)

    def one():
        for g in groups:
            ws = [chain.weights[t] for t in g]
            if len(g) == 1:
                t4.mul_mat_host_async(ws[0].tiles, xh[ws[0].K], yh[g[0] % L], ws[0].n_local)
            else:
                t4.mul_mat_fanout_host_async([w.tiles for w in ws], xh[ws[0].K], [yh[t % L] for t in g], [w.n_local for w in ws])
        t4.host_sync()              # every y of the step is in host memory here
    one()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps, h2d, d2h


def load_traffic():
    """per-launch DRAM traffic (dram__bytes_read + write) of the dominant kernels from the committed ncu capture of this
    round, written by tests/ncu_summary.py --traffic; null when the file is absent (never a hard-coded constant)"""
    p = os.path.join(ROOT, "profiles", "traffic_r02.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
        return {}


def run_own_arm(args):
    import torch
    import __graft_entry__ as graft

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"          # keep NCCL's version banner off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    t4 = graft.load_package()
    peaks = load_peaks()
    M = PREFILL_M

    # ---- prefill chain (the headline step) -------------------------------------------------------------
    pre = Chain(t4, torch, M, world, rank)
    pre.capture()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    torch.cuda.profiler.start()     # cudaProfilerStart: `ncu --profile-from-start off` sees only the timed regions
    ms_step = timed_steps(torch, dist, pre, args.steps, args.warmup, world)
    # ---- decode chain (batch 1) over the same weights, timed in the same clock-sampled window --------------
    dec = Chain(t4, torch, 1, world, rank, weights=pre.weights)
    dec.capture()
    dec_steps = max(args.steps, 20)
    ms_dec = timed_steps(torch, dist, dec, dec_steps, max(args.warmup, 3), world)
    torch.cuda.profiler.stop()
    clocks = sampler.stop() if rank == 0 else None
    variants = None
    exchange_rule = None
    if world > 1 and pre.sym is not None:
        L = len(QWEN3_4B_LAYER)

        def check(chain):
            # one-off check (not timed): after a step, the symmetric C of EVERY matmul of the last layer equals the NCCL
            # all-gather of the shard outputs, bitwise, on every rank
            for j in chain.sym:
                chain.sym[j][0].fill_(float("nan"))
            torch.cuda.synchronize()
            dist.barrier()
            chain.step_eager()
            torch.cuda.synchronize()
            dist.barrier()
            ok = True
            for j in range(L):
                wl = chain.weights[-L + j]
                want = wl.forward(chain.x[wl.K], workspace=chain.ws)
                torch.cuda.synchronize()
                ok = ok and bool(torch.equal(chain.sym[j][0], want))
            t = torch.tensor([1 if ok else 0], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            dist.barrier()
            return bool(t.item())

        variants = {}
        timed = {}
        for key, mode in (("fused_peer_stores", "fused"), ("switch_multicast_stores", "multicast"), ("copy_engine_pushes", "dma")):
            try:
                if mode == "fused":
                    p_, d_ = pre, dec
                else:
                    p_ = Chain(t4, torch, M, world, rank, weights=pre.weights, mode=mode)
                    if p_.sym is None or p_.mode != mode:
                        raise RuntimeError(f"{mode} unavailable")
                    p_.capture()
                    d_ = Chain(t4, torch, 1, world, rank, weights=pre.weights, mode=mode)
                    d_.capture()
                ms_p = ms_step if mode == "fused" else timed_steps(torch, dist, p_, args.steps, args.warmup, world)
                ms_d = ms_dec if mode == "fused" else timed_steps(torch, dist, d_, dec_steps, max(args.warmup, 3), world)
                variants[key] = {"ms_per_step": ms_p, "decode_ms_per_step": ms_d, "matches_nccl_all_7_outputs": check(p_),
                                 "sync": "one cross-GPU barrier per layer (36 per step)"}
                timed[key] = (ms_p, ms_d, p_.gather_kind)
            except Exception as e:  # noqa: BLE001
                variants[key] = {"unavailable": f"{type(e).__name__}: {e}"}
        # ONE variant per run, chosen by rule (not the fastest of three per leg): the switch's multicast when the fabric offers
        # it (outbound traffic 1x instead of (G-1)x), else the fused peer stores
        # (round 2: the copy-engine variant is the rule's first choice -- its MUL_MATs are ordinary single-destination calls and so
        # take the graph-level fan-out like the single-GPU step, and its transfers use no SM; the store variants follow)
        # G = 2: one push per output; G > 2: the copy engines would move every block G-1 times (and at G = 8 issue 1 764 copies per
        # step), the switch replicates ONE multicast store stream
        order = ("copy_engine_pushes", "switch_multicast_stores", "fused_peer_stores") if world == 2 else \
                ("switch_multicast_stores", "copy_engine_pushes", "fused_peer_stores")
        exchange_rule = next(k for k in order if k in timed)
        ms_step, ms_dec, pre.gather_kind = timed[exchange_rule]
        # the same variant with ONE barrier per step (round 1's placement), for comparison
        try:
            p1 = Chain(t4, torch, M, world, rank, weights=pre.weights, mode={"switch_multicast_stores": "multicast", "copy_engine_pushes": "dma",
                                                                            "fused_peer_stores": "fused"}[exchange_rule], barrier="step")
            p1.capture()
            variants[exchange_rule]["ms_per_step_one_barrier_per_step"] = timed_steps(torch, dist, p1, args.steps, args.warmup, world)
            del p1
        except Exception as e:  # noqa: BLE001
            variants[exchange_rule]["one_barrier_per_step_error"] = f"{type(e).__name__}: {e}"
    # ---- N > 1: the same chains without the fp32 all-gather (the sharded MUL_MATs alone; SURVEY 8e asks for both) ----
    ms_step_ng = ms_dec_ng = None
    if world > 1:
        pre_ng = Chain(t4, torch, M, world, rank, weights=pre.weights, gather=False)
        pre_ng.capture()
        ms_step_ng = timed_steps(torch, dist, pre_ng, args.steps, args.warmup, world)
        dec_ng = Chain(t4, torch, 1, world, rank, weights=pre.weights, gather=False)
        dec_ng.capture()
        ms_dec_ng = timed_steps(torch, dist, dec_ng, dec_steps, max(args.warmup, 3), world)

    # ---- per-kernel roofline legs (N-shard-local kernels; events on the launching stream) ----------------
    gm_ms, act_ms, gm_n = kernel_leg(t4, torch, pre_ng if world > 1 else pre, args.steps, t4.KERNEL_AUTO, M)
    gv_ms, _, gv_n = kernel_leg(t4, torch, dec_ng if world > 1 else dec, min(dec_steps, 10), t4.KERNEL_GEMV, 1)
    ops_local = sum(2 * M * w.n_local * w.K for w in pre.weights)                 # per step, this rank
    bytes_local_dec = sum(alg_bytes(1, w.n_local, w.K) for w in dec.weights)
    i8_peak = 2.0 * peaks["bf16_tflops"]          # dense int8 = 2 x dense bf16 on the same tensor pipe
    try:
        umma_i8, _ = t4.measure_umma(0, 256, 0, 4000)    # bare tcgen05.mma kind::i8 128x256x32 issue loop, all SMs
    except Exception:  # noqa: BLE001
        umma_i8 = None
    gv_gbs_eager = bytes_local_dec * (min(dec_steps, 10)) / (gv_ms * 1e-3) / 1e9 if gv_n else 0.0
    # Dominant prefill kernel: its share of the step comes from the eager pass (per-launch events around the activation kernel
    # and the GEMM launches of every call: share = GEMM time / (GEMM + activation time), host launch gaps hit both alike); the
    # duration it is charged with is that share of the TIMED region (CUDA-graph replays): kernel share x step.
    ms_pre_local = ms_step_ng if (world > 1 and ms_step_ng) else ms_step
    gemm_share = gm_ms / max(1e-9, gm_ms + act_ms)
    gemm_launches_per_step = gm_n / max(1, args.steps)
    gm_tops = ops_local / (ms_pre_local * 1e-3 * gemm_share) / 1e12
    gemm_avg_us = ms_pre_local * 1e3 * gemm_share / max(1.0, gemm_launches_per_step)
    # the decode kernel's launch duration inside the TIMED region (CUDA-graph replays, events around the steps): the step
    # is nothing but these launches, so step time / launches is the per-launch duration including every gap between them
    ms_dec_local = ms_dec_ng if (world > 1 and ms_dec_ng) else ms_dec
    gv_gbs = bytes_local_dec / (ms_dec_local * 1e-3) / 1e9
    dec_launches = int((dec_ng if world > 1 else dec).launches_per_step or len(dec.weights))
    traffic = load_traffic()

    # ---- BASELINE configs[3] / configs[4] on this many GPUs ------------------------------------------------
    big_configs = None
    if not args.no_configs:
        try:
            big_configs = configs_leg(t4, torch, dist, world, rank,
                                      {"switch_multicast_stores": "multicast", "copy_engine_pushes": "dma"}.get(exchange_rule, "peers"))
        except Exception as e:  # noqa: BLE001
            big_configs = {"error": f"{type(e).__name__}: {e}"}
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if world == 1:      # configs[2]: Llama-3-8B decode batch 1-64 (one GPU)
            try:
                big_configs["llama3_8b_decode_batch"] = batch_decode_leg(t4, torch, peaks["hbm_gbs"])
            except Exception as e:  # noqa: BLE001
                big_configs["llama3_8b_decode_batch"] = {"error": f"{type(e).__name__}: {e}"}

    # ---- end to end through the host-buffer C-ABI call ---------------------------------------------------
    e2e_s, h2d, d2h = e2e_leg(t4, torch, pre, max(1, min(args.steps, 2)), M)
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    if rank == 0:
        total_ops = QWEN3_4B_LAYERS * layer_ops(M)
        total_dec_bytes = QWEN3_4B_LAYERS * sum(alg_bytes(1, N, K) for _, N, K in QWEN3_4B_LAYER)
        line = {
            "metric": METRIC, "value": M / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": base_config(),
            "detail": {"weight_bytes_per_gpu": pre.weight_bytes(),
                       "parallelism": (f"N-shard x{world}; gather: {pre.gather_kind}" if world > 1 else "single GPU"),
                       "data_plane": ("symmetric memory (peer-mapped stores / NVSwitch multicast / copy engines); NCCL only for the "
                                      "bootstrap and the reference all-gather check" if world > 1 else "none (one GPU)"),
                       "api": ("one tcq4_b200_ggml_graph_compute call per step over the 252 MUL_MAT nodes (consecutive nodes sharing "
                               "src1 fan out: 4 activation + 4 GEMM launches per layer)" if world == 1 else
                               "one tcq4_b200_mul_mat_peers / _multicast call per MUL_MAT node"),
                       "l2": "252 distinct weight buffers (2.1 GB total) >> 126 MB L2; activations stay L2-resident",
                       "cuda_graph": pre.graph is not None},
            "tops": total_ops / (ms_step * 1e-3) / 1e12,
            "decode": {"value": 1.0 / (ms_dec * 1e-3), "unit": UNIT, "ms_per_step": ms_dec, "steps": dec_steps,
                       "gbs": total_dec_bytes / (ms_dec * 1e-3) / 1e9, "frac_hbm": total_dec_bytes / (ms_dec * 1e-3) / 1e9
                       / (peaks["hbm_gbs"] * world), "cuda_graph": dec.graph is not None},
            "roofline": {"bound": "tensor", "kernel": "tcq4_gemm3_kernel (tcgen05.mma.cta_group::2 kind::f16 on exact fp16 integers; fan-out groups) "
                                                       "and tcq4_gemm2_kernel (kind::i8; single matrices)",
                         "achieved": gm_tops, "peak": i8_peak, "unit": "TOP/s", "frac": gm_tops / i8_peak,
                         "traffic": traffic.get("prefill_gemm_bytes_per_launch"),
                         "traffic_note": traffic.get("prefill_note", "no ncu capture of this round in profiles/traffic_r02.json"),
                         "peak_source": f"2 x bf16_tflops of MEASURED_PEAKS.json ({peaks['source']}, burst): dense int8 rate of the pipe",
                         "peak_umma_i8_measured": umma_i8, "frac_of_umma_i8": (gm_tops / umma_i8) if umma_i8 else None,
                         "ops": "useful 2*M*N*K; the tensor pipe runs them as fp16 MACs with fp32 accumulators holding the exact integer "
                                "(kind::f16 = half the int8 rate: the exact per-group scale caps USEFUL ops at 50 % of the int8 peak)",
                         "launches_per_step": gemm_launches_per_step, "avg_launch_us": gemm_avg_us,
                         "kernel_share_of_step": gemm_share,
                         "act_kernel_share_of_step": 1.0 - gemm_share,
                         "timed": "kernel share (eager per-launch events) x step time of the timed region (graph replays)"},
            "roofline_decode": {"bound": "hbm", "kernel": "tcq4_gemv_kernel", "achieved": gv_gbs, "peak": peaks["hbm_gbs"],
                                "unit": "GB/s", "frac": gv_gbs / peaks["hbm_gbs"],
                                "traffic": traffic.get("decode_gemv_bytes_per_launch"),
                                "traffic_note": traffic.get("decode_note", "no ncu capture of this round in profiles/traffic_r02.json"),
                                "launches": dec_launches * dec_steps,
                                "avg_launch_us": ms_dec_local * 1e3 / max(1, dec_launches),
                                "timed": "decode steps of the timed region (graph replays): step time / launches, gaps included",
                                "eager_event_gbs": gv_gbs_eager,
                                "peak_source": f"hbm_gbs ({peaks['source']})"},
            "e2e": {"value": M / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_s * 1e3, "api": "tcq4_b200_mul_mat_host_async per node, tcq4_b200_mul_mat_fanout_host_async per group of nodes sharing src1 (x crosses PCIe once per group; pinned host x and y, copies inside the timed region) + tcq4_b200_host_sync per step"},
            "gpu_launches": int(pre.launches_per_step * args.steps),
            "clocks": clocks,
        }
        if world > 1:
            line["detail"]["exchange_variants"] = variants
            line["detail"]["exchange_rule"] = f"value and decode use {exchange_rule} (rule: copy engines at 2 GPUs, switch multicast above, peer stores as the fallback; never best-of-three)"
            line["sharded_mul_mat_only"] = {
                "note": "same chains without the fp32 all-gather of the N-sharded outputs (max over ranks)",
                "value": M / (ms_step_ng * 1e-3), "unit": UNIT, "ms_per_step": ms_step_ng,
                "tops": total_ops / (ms_step_ng * 1e-3) / 1e12,
                "decode": {"value": 1.0 / (ms_dec_ng * 1e-3), "unit": UNIT, "ms_per_step": ms_dec_ng}}
        if world == 1:
            try:
                line["per_node"] = per_node_leg(t4, torch, pre, dec, args.steps, args.warmup)
            except Exception as e:  # noqa: BLE001
                line["per_node"] = {"error": f"{type(e).__name__}: {e}"}
        line["configs"] = big_configs
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline_leg()
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"}
        emit(line)
    if world > 1:
        # Captured graphs hold NCCL kernels; tearing the communicator down under them can block for minutes.
        # Everything is measured and printed: synchronise, meet once more, and leave without the teardown.
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line of the contract, on the process's original stdout."""
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    # Everything else any library prints to fd 1 (NCCL's version banner, for one) goes to stderr: stdout carries
    # exactly one JSON line.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the BASELINE configs[3]/[4] layer timings")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_own_arm(args)


if __name__ == "__main__":
    main()
