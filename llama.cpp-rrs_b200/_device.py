"""Torch-facing half of the binding: device memory, streams, torch.distributed plumbing.

Mirrors the reference operator interface for this path:
  mul_mat / ggml_mul_mat      <- ggml_cuda_rrs_mul_mat            (rrs.cu:241-344)
  fwht_quantize_tc            <- tcq4_rrs_(perm_)fwht_quantize_tc (tcq4_k32.cuh:79-94)
  register_perm / clear_perms / set_reorder_enabled / get_perm    (rrs.cu:82-127)
  ShardedWeight               <- N-sharded MUL_MAT + all-gather (no reference counterpart)
"""
from __future__ import annotations

import ctypes as C

import torch

from . import (GGML_TYPE_F32, GGML_TYPE_TCQ4_K32, KERNEL_AUTO, LIB, TILE_BYTES, TILE_CHANNELS, TILE_K,
               ACT_BLOCK_BYTES, TensorView, _check, quantize_size, shard_range, tile_bytes, workspace_size)

__all__ = ["Workspace", "mul_mat", "mul_mat_id", "ggml_mul_mat_id", "Fanout", "ggml_mul_mat_fanout", "GgmlGraph", "mul_mat_add", "mul_mat_static", "mul_mat_debug", "mul_mat_peers", "mul_mat_multicast", "push_columns", "PeerBarrier", "mul_mat_host", "mul_mat_host_async", "mul_mat_fanout_host_async", "rms_row_scale", "apply_input", "mul_mat_fanout_input", "INPUT_RMS_NORM_MUL", "INPUT_SWIGLU", "host_sync", "fwht_quantize_tc", "register_perm",
           "clear_perms", "set_reorder_enabled", "get_perm_ptr", "has_perm", "reserve", "ggml_mul_mat",
           "ggml_supports_op", "ShardedWeight", "random_tiles", "quantize_tiles", "get_rows", "ggml_get_rows"]


def _stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def _req_cuda(t: torch.Tensor, dtype, what: str):
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{what} must be a contiguous CUDA tensor of dtype {dtype}")


class Workspace:
    """Caller-owned scratch arena for tcq4_b200_mul_mat_ws (the ggml shim would take it from ctx.pool())."""

    def __init__(self, nbytes: int, device=None):
        nbytes = max(int(nbytes), 256)
        self.buf = torch.zeros(nbytes + 256, dtype=torch.uint8, device=device or "cuda")
        off = (-self.buf.data_ptr()) % 256
        self.ptr = self.buf.data_ptr() + off
        self.nbytes = nbytes
        # torch.zeros already cleared the split-K tickets; init again through the ABI for parity of use
        _check(LIB.tcq4_b200_workspace_init(self.ptr, self.nbytes, _stream_ptr()))

    @classmethod
    def for_shape(cls, M: int, N: int, K: int, device=None) -> "Workspace":
        return cls(workspace_size(M, N, K), device)


def mul_mat(w_tiles: torch.Tensor, x: torch.Tensor, N: int, perm: torch.Tensor | None = None,
            out: torch.Tensor | None = None, workspace: Workspace | None = None, kernel: int = KERNEL_AUTO,
            stream=None) -> torch.Tensor:
    """y[M][N] = W4A4(w_tiles [N][K], x [M][K]).  Asynchronous on the current stream."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    if w_tiles.numel() != tile_bytes(N, K):
        raise ValueError(f"w_tiles has {w_tiles.numel()} bytes, expected {tile_bytes(N, K)} for N={N} K={K}")
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=x.device)
    _req_cuda(out, torch.float32, "out")
    pp = 0
    if perm is not None:
        _req_cuda(perm, torch.int32, "perm")
        pp = perm.data_ptr()
    if workspace is None:
        _check(LIB.tcq4_b200_mul_mat(w_tiles.data_ptr(), x.data_ptr(), out.data_ptr(), M, N, K, pp,
                                     _stream_ptr(stream)))
    else:
        _check(LIB.tcq4_b200_mul_mat_ws(w_tiles.data_ptr(), x.data_ptr(), out.data_ptr(), M, N, K, pp,
                                        workspace.ptr, workspace.nbytes, kernel, _stream_ptr(stream)))
    return out


def mul_mat_static(w_tiles: torch.Tensor, x: torch.Tensor, N: int, perm: torch.Tensor | None = None,
                   out: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """tcq4_b200_mul_mat_static: the same MUL_MAT with its scratch in the library's file-scope device array -- no
    cudaMalloc on the call path (the output must be passed in by callers that may not allocate either).  Calls that
    overlap on the GPU race on the shared array."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    if w_tiles.numel() != tile_bytes(N, K):
        raise ValueError(f"w_tiles has {w_tiles.numel()} bytes, expected {tile_bytes(N, K)} for N={N} K={K}")
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=x.device)
    _req_cuda(out, torch.float32, "out")
    pp = 0
    if perm is not None:
        _req_cuda(perm, torch.int32, "perm")
        pp = perm.data_ptr()
    _check(LIB.tcq4_b200_mul_mat_static(w_tiles.data_ptr(), x.data_ptr(), out.data_ptr(), M, N, K, pp, _stream_ptr(stream)))
    return out


def mul_mat_debug(w_tiles: torch.Tensor, x: torch.Tensor, N: int, perm: torch.Tensor | None = None,
                  kernel: int = KERNEL_AUTO, workspace: Workspace | None = None):
    """Verification tap: returns (y, isum int32 [M][N][K/256]) -- the integer accumulators."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    ws = workspace or Workspace.for_shape(M, N, K, x.device)
    y = torch.empty((M, N), dtype=torch.float32, device=x.device)
    isum = torch.zeros((M, N, K // TILE_K), dtype=torch.int32, device=x.device)
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_debug(w_tiles.data_ptr(), x.data_ptr(), y.data_ptr(), M, N, K, pp, ws.ptr,
                                       ws.nbytes, kernel, isum.data_ptr(), _stream_ptr()))
    return y, isum


def mul_mat_peers(w_tiles: torch.Tensor, x: torch.Tensor, peer_ptrs, ldc: int, col0: int, N: int,
                  perm: torch.Tensor | None = None, stream=None) -> None:
    """N-sharded MUL_MAT whose epilogue stores its N columns at column col0 of every [M][ldc] fp32 matrix in
    `peer_ptrs` (device pointers, normally one peer-mapped symmetric buffer per GPU): the all-gather fused away."""
    import ctypes as _C
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    arr = (_C.c_void_p * len(peer_ptrs))(*[int(p) for p in peer_ptrs])
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_peers(w_tiles.data_ptr(), x.data_ptr(), arr, len(peer_ptrs), ldc, col0, M, N, K, pp,
                                       _stream_ptr(stream)))


def mul_mat_multicast(w_tiles: torch.Tensor, x: torch.Tensor, mc_ptr: int, ldc: int, col0: int, N: int,
                      perm: torch.Tensor | None = None, stream=None) -> None:
    """N-sharded MUL_MAT storing through the multicast (NVLS) address of the symmetric C: one multimem.st per element
    group, replicated by the switch into every GPU's copy."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_multicast(w_tiles.data_ptr(), x.data_ptr(), int(mc_ptr), ldc, col0, M, N, K, pp,
                                           _stream_ptr(stream)))


def push_columns(dst_ptr: int, src_ptr: int, M: int, ldc: int, col0: int, ncols: int, stream=None) -> None:
    """One 2-D async copy of a column window of a row-major [M][ldc] fp32 matrix into the same window of another
    (peer-mapped) matrix: the copy-engine half of the N-shard exchange."""
    _check(LIB.tcq4_b200_push_columns(int(dst_ptr), int(src_ptr), M, ldc, col0, ncols, _stream_ptr(stream)))


class PeerBarrier:
    """tcq4_b200_peer_barrier over a symmetric int32 flag buffer (torch symmetric memory supplies the peer mapping)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.buf = symm_mem.empty((16,), dtype=torch.int32, device="cuda")
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, group)
        torch.cuda.synchronize()
        self.hdl.barrier()                      # everybody's flags are zero before the first use
        self.ptrs = (C.c_void_p * self.world)(*[int(p) for p in self.hdl.buffer_ptrs])

    def __call__(self, stream=None):
        _check(LIB.tcq4_b200_peer_barrier(self.ptrs, self.world, self.rank, _stream_ptr(stream)))


def mul_mat_host(w_tiles: torch.Tensor, x_host: torch.Tensor, y_host: torch.Tensor, N: int,
                 perm: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """End-to-end call with HOST (pinned) x and y; weights resident on the device.  Synchronous."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    if x_host.is_cuda or y_host.is_cuda or x_host.dtype != torch.float32 or y_host.dtype != torch.float32:
        raise ValueError("x_host / y_host must be float32 CPU tensors (pinned for full speed)")
    M, K = x_host.shape
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_host(w_tiles.data_ptr(), x_host.data_ptr(), y_host.data_ptr(), M, N, K, pp,
                                      _stream_ptr(stream)))
    return y_host


def mul_mat_host_async(w_tiles: torch.Tensor, x_host: torch.Tensor, y_host: torch.Tensor, N: int,
                       perm: torch.Tensor | None = None) -> None:
    """Queue one host-buffer MUL_MAT on the library's copy/compute pipeline; y_host is valid after host_sync()."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    if x_host.is_cuda or y_host.is_cuda or x_host.dtype != torch.float32 or y_host.dtype != torch.float32:
        raise ValueError("x_host / y_host must be float32 CPU tensors (pinned for asynchronous copies)")
    M, K = x_host.shape
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_host_async(w_tiles.data_ptr(), x_host.data_ptr(), y_host.data_ptr(), M, N, K, pp))


def mul_mat_fanout_host_async(w_tiles: list, x_host: torch.Tensor, y_hosts: list, Ns: list,
                              perm: torch.Tensor | None = None) -> None:
    """Queue a fan-out group (matrices that share x: Q/K/V, gate/up) on the host pipeline: x crosses PCIe once."""
    for w in w_tiles:
        _req_cuda(w, torch.uint8, "w_tiles")
    if x_host.is_cuda or x_host.dtype != torch.float32 or any(y.is_cuda or y.dtype != torch.float32 for y in y_hosts):
        raise ValueError("x_host / y_hosts must be float32 CPU tensors (pinned for asynchronous copies)")
    M, K = x_host.shape
    n = len(w_tiles)
    wp = (C.c_void_p * n)(*[w.data_ptr() for w in w_tiles])
    yp = (C.c_void_p * n)(*[y.data_ptr() for y in y_hosts])
    Np = (C.c_int64 * n)(*Ns)
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_mul_mat_fanout_host_async(n, wp, Np, yp, x_host.data_ptr(), M, K, pp))


def host_sync() -> None:
    _check(LIB.tcq4_b200_host_sync())


def fwht_quantize_tc(x: torch.Tensor, perm: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """Stand-alone activation stage -> block_rrs_int4_tc bytes [M][K/256][160]."""
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    out = torch.empty((M, K // TILE_K, ACT_BLOCK_BYTES), dtype=torch.uint8, device=x.device)
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_fwht_quantize_tc(x.data_ptr(), out.data_ptr(), pp, K, M, _stream_ptr(stream)))
    return out


def quantize_tiles(w: torch.Tensor, perm: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """Device weight quantiser (quantize_tcq4_tile, ggml-quants.c:1528-1639): F32 [nrow][K] -> uint8 tile bytes
    [ceil(nrow/8) * K/256 * 1184]; perm (int32[K], device) quantises W[:, perm] (llama-quant.cpp:26-40)."""
    _req_cuda(w, torch.float32, "w")
    nrow, K = w.shape
    out = torch.empty(quantize_size(nrow, K), dtype=torch.uint8, device=w.device)
    pp = perm.data_ptr() if perm is not None else 0
    _check(LIB.tcq4_b200_quantize_tiles(w.data_ptr(), out.data_ptr(), nrow, K, pp, _stream_ptr(stream)))
    return out


def get_rows(w_tiles: torch.Tensor, rows: torch.Tensor, nrows_src: int, K: int, out: torch.Tensor | None = None,
             err_flag: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """GET_ROWS of a TCQ4 tensor (getrows.cu:96-171 + 348-380): out[i] = FWHT256(dequant(row rows[i])), F32 [n][K]."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(rows, torch.int32, "rows")
    n = rows.numel()
    if out is None:
        out = torch.empty((n, K), dtype=torch.float32, device=w_tiles.device)
    ef = err_flag.data_ptr() if err_flag is not None else 0
    _check(LIB.tcq4_b200_get_rows(w_tiles.data_ptr(), rows.data_ptr(), out.data_ptr(), n, nrows_src, K, out.stride(0), ef,
                                  _stream_ptr(stream)))
    return out


def reserve(max_M: int, max_N: int, max_K: int):
    _check(LIB.tcq4_b200_reserve(max_M, max_N, max_K))


# ---- channel-permutation registry (ggml_backend_cuda_tcq4_*, ggml-cuda.cu:4148-4158) ------------
def register_perm(name: str, perm) -> None:
    p = torch.as_tensor(perm, dtype=torch.int32).contiguous().cpu()
    LIB.ggml_backend_cuda_tcq4_register_perm(name.encode(), p.data_ptr(), int(p.numel()))


def clear_perms() -> None:
    LIB.ggml_backend_cuda_tcq4_clear_perms()


def set_reorder_enabled(enabled: bool) -> None:
    LIB.ggml_backend_cuda_tcq4_set_reorder_enabled(bool(enabled))


def get_perm_ptr(name: str) -> int:
    return int(LIB.tcq4_b200_get_perm(name.encode()) or 0)


def has_perm(name: str) -> bool:
    return bool(LIB.tcq4_b200_has_perm(name.encode()))


# ---- ggml operator mirror ---------------------------------------------------------------------
def _view(t: torch.Tensor | None, gtype: int, ne, nb, name: bytes | None, data_ptr: int | None = None) -> TensorView:
    v = TensorView()
    v.type = gtype
    for i in range(4):
        v.ne[i] = ne[i] if i < len(ne) else 1
        v.nb[i] = nb[i] if i < len(nb) else nb[-1]
    v.data = data_ptr if data_ptr is not None else (t.data_ptr() if t is not None else 0)
    v.name = name
    return v


def _views(w_tiles, x, y, N, name):
    M, K = x.shape
    row = (K // TILE_K) * 148
    src0 = _view(w_tiles, GGML_TYPE_TCQ4_K32, [K, N, 1, 1], [148, row, row * N, row * N], name)
    src1 = _view(x, GGML_TYPE_F32, [K, M, 1, 1], [4, 4 * K, 4 * K * M, 4 * K * M], b"src1")
    ldc = y.stride(0) if (y is not None and y.dim() == 2 and M > 1) else N      # a column window of a wider matrix is a strided view
    dst = _view(y, GGML_TYPE_F32, [N, M, 1, 1], [4, 4 * ldc, 4 * ldc * M, 4 * ldc * M], b"dst")
    return src0, src1, dst


def ggml_supports_op(K: int, N: int, M: int, cc: int = 100) -> bool:
    src0 = _view(None, GGML_TYPE_TCQ4_K32, [K, N, 1, 1], [148, (K // TILE_K) * 148] * 2, b"w", 0)
    src1 = _view(None, GGML_TYPE_F32, [K, M, 1, 1], [4, 4 * K, 4 * K * M, 4 * K * M], b"x", 0)
    dst = _view(None, GGML_TYPE_F32, [N, M, 1, 1], [4, 4 * N, 4 * N * M, 4 * N * M], b"y", 0)
    return bool(LIB.tcq4_b200_ggml_supports_op(C.byref(src0), C.byref(src1), C.byref(dst), cc))


def ggml_get_rows(w_tiles: torch.Tensor, idx: torch.Tensor, nrows: int, K: int, err_flag: torch.Tensor | None = None,
                  stream=None) -> torch.Tensor:
    """ggml_cuda_op_get_rows for a TCQ4 src0: idx int32 [ne12][ne11][ne10] (torch order), w_tiles holds one tile tensor
    per (i12, i11) slice back to back; returns F32 [ne12][ne11][ne10][K]."""
    _req_cuda(idx, torch.int32, "idx")
    while idx.dim() < 3:
        idx = idx.unsqueeze(0)
    ne12, ne11, ne10 = idx.shape
    row = (K // TILE_K) * 148
    out = torch.empty((ne12, ne11, ne10, K), dtype=torch.float32, device=w_tiles.device)
    src0 = _view(w_tiles, GGML_TYPE_TCQ4_K32, [K, nrows, ne11, ne12], [148, row, row * nrows, row * nrows * ne11], b"emb")
    src1 = _view(idx, 26, [ne10, ne11, ne12, 1], [4, 4 * ne10, 4 * ne10 * ne11, 4 * ne10 * ne11 * ne12], b"idx")
    dst = _view(out, GGML_TYPE_F32, [K, ne10, ne11, ne12], [4, 4 * K, 4 * K * ne10, 4 * K * ne10 * ne11], b"dst")
    ef = err_flag.data_ptr() if err_flag is not None else 0
    _check(LIB.tcq4_b200_ggml_get_rows(C.byref(src0), C.byref(src1), C.byref(dst), ef, _stream_ptr(stream)))
    return out


def mul_mat_add(w_tiles: torch.Tensor, x: torch.Tensor, N: int, residual: torch.Tensor | None = None,
                bias: torch.Tensor | None = None, perm: torch.Tensor | None = None, out: torch.Tensor | None = None,
                stream=None) -> torch.Tensor:
    """y = W x + bias + residual in the MUL_MAT's epilogue (tcq4_b200_mul_mat_add); out may be the residual tensor itself."""
    _req_cuda(w_tiles, torch.uint8, "w_tiles")
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=x.device)
    for t, what, shape in ((residual, "residual", (M, N)), (bias, "bias", (N,))):
        if t is not None:
            _req_cuda(t, torch.float32, what)
            if tuple(t.shape) != shape:
                raise ValueError(f"{what} must have shape {shape}")
    _check(LIB.tcq4_b200_mul_mat_add(w_tiles.data_ptr(), x.data_ptr(), out.data_ptr(), M, N, K,
                                     perm.data_ptr() if perm is not None else 0,
                                     residual.data_ptr() if residual is not None else 0,
                                     bias.data_ptr() if bias is not None else 0, _stream_ptr(stream)))
    return out


class Fanout:
    """Several weight matrices over the same activations (Q/K/V, gate/up): tcq4_b200_mul_mat_fanout.  The pointer
    tables are built once; call with x (and optionally the outputs of the first call) as often as needed."""

    def __init__(self, w_tiles: list, Ns: list, K: int):
        for w, N in zip(w_tiles, Ns):
            _req_cuda(w, torch.uint8, "w_tiles")
            if w.numel() != tile_bytes(N, K):
                raise ValueError(f"w_tiles has {w.numel()} bytes, expected {tile_bytes(N, K)} for N={N} K={K}")
        self.n, self.K, self.Ns, self._keep = len(Ns), K, list(Ns), list(w_tiles)
        self._w = (C.c_void_p * self.n)(*[w.data_ptr() for w in w_tiles])
        self._N = (C.c_int64 * self.n)(*Ns)

    def __call__(self, x: torch.Tensor, outs: list | None = None, perm: torch.Tensor | None = None, stream=None) -> list:
        _req_cuda(x, torch.float32, "x")
        M, K = x.shape
        if K != self.K:
            raise ValueError("x has the wrong K")
        if outs is None:
            outs = [torch.empty((M, N), dtype=torch.float32, device=x.device) for N in self.Ns]
        y = (C.c_void_p * self.n)(*[o.data_ptr() for o in outs])
        pp = perm.data_ptr() if perm is not None else 0
        _check(LIB.tcq4_b200_mul_mat_fanout(self.n, self._w, self._N, y, x.data_ptr(), M, K, pp, _stream_ptr(stream)))
        return outs


INPUT_RMS_NORM_MUL, INPUT_SWIGLU = 1, 2


def rms_row_scale(x: torch.Tensor, eps: float, stream=None) -> torch.Tensor:
    """rsqrt(mean(x^2) + eps) per row, in the reduction order of ggml-cuda's rms_norm_f32."""
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    out = torch.empty((M,), dtype=torch.float32, device=x.device)
    _check(LIB.tcq4_b200_rms_row_scale(x.data_ptr(), out.data_ptr(), M, K, float(eps), _stream_ptr(stream)))
    return out


def apply_input(mode: int, x: torch.Tensor, x2=None, row_scale=None, col_weight=None, stream=None) -> torch.Tensor:
    """The input transform on its own (what the fused activation-stage load computes): [M][K] -> [M][K]."""
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    out = torch.empty_like(x)
    p = lambda t: t.data_ptr() if t is not None else 0
    _check(LIB.tcq4_b200_apply_input(mode, x.data_ptr(), p(x2), p(row_scale), p(col_weight), out.data_ptr(), M, K, _stream_ptr(stream)))
    return out


def mul_mat_fanout_input(w_tiles: list, x: torch.Tensor, Ns: list, mode: int, x2=None, row_scale=None, col_weight=None,
                         perm=None, outs=None, stream=None) -> list:
    """tcq4_b200_mul_mat_fanout_input: the fan-out (or one matrix) with the op that produced x fused into the activation stage."""
    _req_cuda(x, torch.float32, "x")
    M, K = x.shape
    n = len(Ns)
    if outs is None:
        outs = [torch.empty((M, N), dtype=torch.float32, device=x.device) for N in Ns]
    wp = (C.c_void_p * n)(*[w.data_ptr() for w in w_tiles])
    yp = (C.c_void_p * n)(*[o.data_ptr() for o in outs])
    Np = (C.c_int64 * n)(*Ns)
    ld = (C.c_int64 * n)(*[o.stride(0) for o in outs])
    p = lambda t: t.data_ptr() if t is not None else 0
    _check(LIB.tcq4_b200_mul_mat_fanout_input(n, wp, Np, yp, ld, x.data_ptr(), mode, p(x2), p(row_scale), p(col_weight), M, K,
                                              p(perm), _stream_ptr(stream)))
    return outs


def ggml_mul_mat_fanout(w_tiles: list, x: torch.Tensor, Ns: list, names: list, stream=None) -> list:
    """tcq4_b200_ggml_mul_mat_fanout: consecutive MUL_MAT nodes with one src1; perms looked up by each src0->name."""
    M, K = x.shape
    n = len(Ns)
    outs = [torch.empty((M, N), dtype=torch.float32, device=x.device) for N in Ns]
    views = [_views(w, x, o, N, nm.encode()) for w, o, N, nm in zip(w_tiles, outs, Ns, names)]
    s0 = (C.POINTER(TensorView) * n)(*[C.pointer(v[0]) for v in views])
    ds = (C.POINTER(TensorView) * n)(*[C.pointer(v[2]) for v in views])
    _check(LIB.tcq4_b200_ggml_mul_mat_fanout(n, s0, C.byref(views[0][1]), ds, _stream_ptr(stream)))
    return outs


class GgmlGraph:
    """tcq4_b200_ggml_graph_compute: the TCQ4 MUL_MAT nodes of a ggml graph in execution order, as the backend's graph
    loop sees them (ggml-cuda.cu:3481-3688).  nodes: list of (w_tiles, x, out, N, name); consecutive nodes that share x
    are fanned out by the library.  The tensor views are built once; compute() is one C call."""

    def __init__(self, nodes):
        self.keep = nodes
        self.n = len(nodes)
        self.views = [_views(w, x, o, N, (nm or "").encode()) for (w, x, o, N, nm) in nodes]
        P = C.POINTER(TensorView)
        self.s0 = (P * self.n)(*[C.pointer(v[0]) for v in self.views])
        self.s1 = (P * self.n)(*[C.pointer(v[1]) for v in self.views])
        self.ds = (P * self.n)(*[C.pointer(v[2]) for v in self.views])

    def compute(self, stream=None):
        _check(LIB.tcq4_b200_ggml_graph_compute(self.n, self.s0, self.s1, self.ds, _stream_ptr(stream)))


def mul_mat_id(w_experts: torch.Tensor, x: torch.Tensor, ids: torch.Tensor, N: int, perm: torch.Tensor | None = None,
               out: torch.Tensor | None = None, err_flag: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """MUL_MAT_ID (ggml_cuda_mul_mat_id, ggml-cuda.cu:2287-2436): w_experts uint8 [n_expert][tile bytes of N x K],
    x F32 [n_tokens][1 or n_used][K], ids int32 [n_tokens][n_used] -> y F32 [n_tokens][n_used][N] with
    y[t][s] = W[ids[t][s]] x[t][s % x_slots]."""
    _req_cuda(w_experts, torch.uint8, "w_experts")
    _req_cuda(x, torch.float32, "x")
    if not ids.is_cuda or ids.dtype != torch.int32 or ids.dim() != 2:
        raise ValueError("ids must be a 2-D CUDA tensor of dtype torch.int32 (rows may be strided)")
    n_tokens, x_slots, K = x.shape
    n_used = ids.shape[1]
    n_expert = w_experts.shape[0]
    if ids.shape[0] != n_tokens or (n_used > 1 and ids.stride(1) != 1):
        raise ValueError("ids must be [n_tokens][n_used] with contiguous slots")
    if w_experts.dim() != 2 or w_experts.shape[1] != tile_bytes(N, K) or w_experts.stride(1) != 1:
        raise ValueError(f"w_experts must be [n_expert][{tile_bytes(N, K)}] bytes")
    if out is None:
        out = torch.empty((n_tokens, n_used, N), dtype=torch.float32, device=x.device)
    _req_cuda(out, torch.float32, "out")
    pp = 0
    if perm is not None:
        _req_cuda(perm, torch.int32, "perm")
        pp = perm.data_ptr()
    ef = err_flag.data_ptr() if err_flag is not None else 0
    ids_stride = ids.stride(0) if n_tokens > 1 else n_used          # a size-1 dimension may carry any stride
    _check(LIB.tcq4_b200_mul_mat_id(w_experts.data_ptr(), n_expert, w_experts.stride(0), x.data_ptr(), x_slots,
                                    ids.data_ptr(), ids_stride, n_used, n_tokens, out.data_ptr(), N, K, pp, ef,
                                    _stream_ptr(stream)))
    return out


def ggml_mul_mat_id(w_experts: torch.Tensor, x: torch.Tensor, ids: torch.Tensor, N: int, name: str = "",
                    err_flag: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """The same through the ggml operator mirror: src0 [K, N, n_expert], src1 [K, ne11, n_tokens], ids [n_used, n_tokens],
    dst [N, n_used, n_tokens]; the perm is looked up by src0->name."""
    n_tokens, ne11, K = x.shape
    n_used = ids.shape[1]
    n_expert = w_experts.shape[0]
    row = (K // TILE_K) * 148
    ist = ids.stride(0) if n_tokens > 1 else n_used
    out = torch.empty((n_tokens, n_used, N), dtype=torch.float32, device=x.device)
    src0 = _view(w_experts, GGML_TYPE_TCQ4_K32, [K, N, n_expert, 1], [148, row, w_experts.stride(0), w_experts.stride(0) * n_expert],
                 name.encode())
    src1 = _view(x, GGML_TYPE_F32, [K, ne11, n_tokens, 1], [4, 4 * K, 4 * K * ne11, 4 * K * ne11 * n_tokens], b"x")
    idv = _view(ids, 26, [n_used, n_tokens, 1, 1], [4, 4 * ist, 4 * ist * n_tokens, 4 * ist * n_tokens], b"ids")
    dst = _view(out, GGML_TYPE_F32, [N, n_used, n_tokens, 1], [4, 4 * N, 4 * N * n_used, 4 * N * n_used * n_tokens], b"dst")
    ef = err_flag.data_ptr() if err_flag is not None else 0
    _check(LIB.tcq4_b200_ggml_mul_mat_id(C.byref(src0), C.byref(src1), C.byref(idv), C.byref(dst), ef, _stream_ptr(stream)))
    return out


def ggml_mul_mat(w_tiles: torch.Tensor, x: torch.Tensor, N: int, name: str = "", out: torch.Tensor | None = None,
                 workspace: Workspace | None = None, stream=None) -> torch.Tensor:
    """ggml_cuda_rrs_mul_mat(ctx, src0, src1, dst): perm is looked up by src0->name (rrs.cu:266-270)."""
    M, K = x.shape
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=x.device)
    src0, src1, dst = _views(w_tiles, x, out, N, name.encode())
    wp, wb = (workspace.ptr, workspace.nbytes) if workspace is not None else (0, 0)
    _check(LIB.tcq4_b200_ggml_mul_mat(C.byref(src0), C.byref(src1), C.byref(dst), wp, wb, _stream_ptr(stream)))
    return out


# ---- synthetic tiles for throughput runs (weights synthesised directly as tiles) -----
def random_tiles(N: int, K: int, seed: int = 0, device="cuda") -> torch.Tensor:
    """Uniform random nibbles, sc in [1,127], S in fp16 U(1e-3,1e-2), Z = zc = 0."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    nt = (N // TILE_CHANNELS) * (K // TILE_K)
    t = torch.zeros((nt, TILE_BYTES), dtype=torch.uint8, device=device)
    t[:, :1024] = torch.randint(0, 256, (nt, 1024), dtype=torch.uint8, device=device, generator=g)
    S = (torch.rand((nt, 8), device=device, generator=g) * 9e-3 + 1e-3).to(torch.float16)
    t[:, 1024:1040] = S.view(torch.uint8).reshape(nt, 16)
    sc = torch.randint(1, 128, (nt, 64), dtype=torch.int16, device=device, generator=g).to(torch.uint8)
    t[:, 1056:1120] = sc
    return t.reshape(-1)


# ---- N-sharded MUL_MAT across the GPUs of one box ------------------------------------------------
class ShardedWeight:
    """Rows [n0,n1) of an [N][K] TCQ4 tensor on this rank (a contiguous byte range, no repacking).

    forward(): local MUL_MAT on the replicated activations, then an all-gather of the fp32 column
    blocks over NCCL/NVLink on the compute stream.  The gather lands in [world][M][N/world] and one
    strided copy puts it into ggml's dst layout C[m*N + n]; for M == 1 the shard is already contiguous.
    """

    def __init__(self, N: int, K: int, world: int, rank: int, tiles_full: torch.Tensor | None = None,
                 tiles_local: torch.Tensor | None = None, align: int = 128):
        self.N, self.K, self.world, self.rank = N, K, world, rank
        self.n0, self.n1, off, ln = shard_range(N, K, world, rank, align)
        self.ranges = [shard_range(N, K, world, r, align)[:2] for r in range(world)]
        self.even = all((b - a) == (self.ranges[0][1] - self.ranges[0][0]) for a, b in self.ranges)
        if tiles_local is not None:
            self.tiles = tiles_local
        else:
            self.tiles = tiles_full[off:off + ln].contiguous()
        assert self.tiles.numel() == ln

    @property
    def n_local(self) -> int:
        return self.n1 - self.n0

    def local(self, x: torch.Tensor, perm=None, out=None, workspace=None) -> torch.Tensor:
        if self.n_local == 0:
            return torch.empty((x.shape[0], 0), dtype=torch.float32, device=x.device)
        return mul_mat(self.tiles, x, self.n_local, perm, out=out, workspace=workspace)

    def gather(self, yl: torch.Tensor) -> torch.Tensor:
        """All-gather the per-rank column blocks [M][n_local] into ggml's dst layout [M][N]."""
        import torch.distributed as dist
        M = yl.shape[0]
        if self.world == 1:
            return yl
        if self.even:
            flat = torch.empty((self.world * M, self.n_local), dtype=yl.dtype, device=yl.device)
            dist.all_gather_into_tensor(flat, yl.contiguous())      # rank-major concatenation (NCCL and gloo agree)
            if M == 1:
                return flat.reshape(1, self.N)
            return flat.view(self.world, M, self.n_local).permute(1, 0, 2).reshape(M, self.N)
        # uneven shards (N not a multiple of world * align): pad every block to the widest one, one collective
        n_max = max(b - a for a, b in self.ranges)
        mine = torch.zeros((M, n_max), dtype=yl.dtype, device=yl.device)
        if self.n_local:
            mine[:, :self.n_local] = yl
        flat = torch.empty((self.world * M, n_max), dtype=yl.dtype, device=yl.device)
        dist.all_gather_into_tensor(flat, mine)
        blocks = flat.view(self.world, M, n_max)
        return torch.cat([blocks[r, :, :b - a] for r, (a, b) in enumerate(self.ranges)], dim=1)

    def forward(self, x: torch.Tensor, perm=None, workspace=None) -> torch.Tensor:
        return self.gather(self.local(x, perm, workspace=workspace))
