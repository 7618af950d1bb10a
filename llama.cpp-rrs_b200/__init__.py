"""tcq4_b200 -- host-side binding of the B200-native TCQ4_K32 W4A4 MUL_MAT path.

The product is the C-ABI shared library ``libtcq4_b200.so`` (include/tcq4_b200.h) built from
``csrc/``; this module is the thin Python mirror of the reference's operator interface for this
path (``ggml_cuda_rrs_mul_mat`` and the ``ggml_backend_cuda_tcq4_*`` registry,
/root/reference/ggml/src/ggml-cuda/rrs.cu:82-127, 241-344).  PyTorch is used for device memory,
streams and torch.distributed plumbing only: no torch type crosses the C boundary.

There is NO CPU fallback and nothing here touches ``oracle/``: if the CUDA library is missing the
import fails loudly, and compute calls without an sm_100 device raise ``Tcq4Error``.

The directory name ``llama.cpp-rrs_b200`` is not a Python identifier; load it with
``__graft_entry__.load_package()`` (registers it as module ``tcq4_b200``).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libtcq4_b200.so")
# diagnostics (include/tcq4_b200_diag.h): a superset build of the product with the instrumented kernel instantiations
# and the microbenchmarks.  Never loaded by the product path; TCQ4_B200_USE_DIAG=1 makes the box-side diagnostic
# scripts (tests/gpu_*.py) run the whole module on it.
DIAG_LIB_PATH = os.path.join(HERE, "libtcq4_b200_diag.so")
if os.environ.get("TCQ4_B200_USE_DIAG") == "1":
    LIB_PATH = DIAG_LIB_PATH

TILE_BYTES = 1184
TILE_K = 256
TILE_CHANNELS = 8
ACT_BLOCK_BYTES = 160
GGML_TYPE_F32 = 0
GGML_TYPE_TCQ4_K32 = 42

KERNEL_AUTO, KERNEL_GEMV, KERNEL_GEMM, KERNEL_GEMM2, KERNEL_GEMM3, KERNEL_GEMM4 = 0, 1, 2, 3, 4, 5

STATUS = {0: "OK", 1: "ERR_SHAPE", 2: "ERR_ARG", 3: "ERR_WORKSPACE", 4: "ERR_CUDA", 5: "ERR_NO_DEVICE"}

# every symbol include/tcq4_b200.h declares (tests check the library exports all of them)
EXPORTS = [
    "tcq4_b200_version", "tcq4_b200_last_error", "tcq4_b200_supports", "tcq4_b200_select_kernel",
    "tcq4_b200_workspace_size", "tcq4_b200_mul_mat_ws", "tcq4_b200_workspace_init", "tcq4_b200_mul_mat",
    "tcq4_b200_reserve", "tcq4_b200_mul_mat_static", "tcq4_b200_mul_mat_peers", "tcq4_b200_mul_mat_multicast", "tcq4_b200_push_columns", "tcq4_b200_peer_barrier", "tcq4_b200_mul_mat_host", "tcq4_b200_mul_mat_host_async", "tcq4_b200_host_sync", "tcq4_b200_fwht_quantize_tc", "tcq4_b200_mul_mat_debug",
    "ggml_backend_cuda_tcq4_register_perm", "ggml_backend_cuda_tcq4_clear_perms",
    "ggml_backend_cuda_tcq4_set_reorder_enabled", "tcq4_b200_get_perm", "tcq4_b200_has_perm",
    "tcq4_b200_get_reorder_enabled", "tcq4_b200_shard_range", "tcq4_b200_ggml_supports_op",
    "tcq4_b200_ggml_mul_mat", "tcq4_b200_launch_count",
    "tcq4_b200_profile_begin", "tcq4_b200_profile_end",
    "tcq4_b200_quantize_size", "tcq4_b200_quantize_tiles", "tcq4_b200_perm_from_imatrix", "tcq4_b200_get_rows", "tcq4_b200_ggml_get_rows",
    "tcq4_b200_mul_mat_id", "tcq4_b200_ggml_mul_mat_id", "tcq4_b200_mul_mat_fanout", "tcq4_b200_mul_mat_fanout_ld", "tcq4_b200_mul_mat_fanout_host_async", "tcq4_b200_rms_row_scale", "tcq4_b200_apply_input", "tcq4_b200_mul_mat_fanout_input", "tcq4_b200_ggml_mul_mat_fanout", "tcq4_b200_ggml_graph_compute", "tcq4_b200_mul_mat_add", "tcq4_b200_quant_type_for",
    "tcq4_b200_gguf_open", "tcq4_b200_gguf_close", "tcq4_b200_gguf_reorder_enabled", "tcq4_b200_gguf_n_tensors",
    "tcq4_b200_gguf_n_perms", "tcq4_b200_gguf_tensor_info", "tcq4_b200_gguf_perm", "tcq4_b200_gguf_register_perms",
    "tcq4_b200_gguf_read_tensor", "tcq4_b200_gguf_load_tensor", "tcq4_b200_gguf_write",
]


class Tcq4Error(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"tcq4_b200: {STATUS.get(status, status)}: {msg}")
        self.status = status


class TensorView(C.Structure):
    """tcq4_tensor_view (include/tcq4_b200.h): the ggml_tensor fields ggml_cuda_rrs_mul_mat reads."""
    _fields_ = [("type", C.c_int), ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4), ("data", C.c_void_p),
                ("name", C.c_char_p)]


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU or PyTorch fallback for this path.")
    lib = C.CDLL(LIB_PATH)
    i64, vp, sz, ci = C.c_int64, C.c_void_p, C.c_size_t, C.c_int
    lib.tcq4_b200_version.restype = C.c_char_p
    lib.tcq4_b200_last_error.restype = C.c_char_p
    lib.tcq4_b200_supports.argtypes = [i64, i64, i64, ci]
    lib.tcq4_b200_select_kernel.argtypes = [i64, i64, i64]
    lib.tcq4_b200_workspace_size.restype = sz
    lib.tcq4_b200_workspace_size.argtypes = [i64, i64, i64]
    lib.tcq4_b200_mul_mat_ws.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp, sz, ci, vp]
    lib.tcq4_b200_workspace_init.argtypes = [vp, sz, vp]
    lib.tcq4_b200_mul_mat.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp]
    lib.tcq4_b200_reserve.argtypes = [i64, i64, i64]
    lib.tcq4_b200_mul_mat_static.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_multicast.argtypes = [vp, vp, vp, i64, i64, i64, i64, i64, vp, vp]
    lib.tcq4_b200_push_columns.argtypes = [vp, vp, i64, i64, i64, i64, vp]
    lib.tcq4_b200_peer_barrier.argtypes = [vp, ci, ci, vp]
    lib.tcq4_b200_mul_mat_peers.argtypes = [vp, vp, vp, ci, i64, i64, i64, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_host.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_host_async.argtypes = [vp, vp, vp, i64, i64, i64, vp]
    lib.tcq4_b200_host_sync.argtypes = []
    lib.tcq4_b200_fwht_quantize_tc.argtypes = [vp, vp, vp, i64, i64, vp]
    lib.tcq4_b200_mul_mat_debug.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp, sz, ci, vp, vp]
    lib.ggml_backend_cuda_tcq4_register_perm.restype = None
    lib.ggml_backend_cuda_tcq4_register_perm.argtypes = [C.c_char_p, vp, ci]
    lib.ggml_backend_cuda_tcq4_clear_perms.restype = None
    lib.ggml_backend_cuda_tcq4_set_reorder_enabled.restype = None
    lib.ggml_backend_cuda_tcq4_set_reorder_enabled.argtypes = [C.c_bool]
    lib.tcq4_b200_get_perm.restype = vp
    lib.tcq4_b200_get_perm.argtypes = [C.c_char_p]
    lib.tcq4_b200_has_perm.argtypes = [C.c_char_p]
    lib.tcq4_b200_shard_range.argtypes = [i64, i64, ci, ci, i64, C.POINTER(i64), C.POINTER(i64),
                                          C.POINTER(sz), C.POINTER(sz)]
    lib.tcq4_b200_ggml_supports_op.argtypes = [C.POINTER(TensorView)] * 3 + [ci]
    lib.tcq4_b200_ggml_mul_mat.argtypes = [C.POINTER(TensorView)] * 3 + [vp, sz, vp]
    lib.tcq4_b200_launch_count.restype = C.c_uint64
    lib.tcq4_b200_profile_begin.argtypes = [ci]
    lib.tcq4_b200_gguf_open.argtypes = [C.c_char_p, C.POINTER(vp)]
    lib.tcq4_b200_gguf_close.argtypes = [vp]
    lib.tcq4_b200_gguf_close.restype = None
    lib.tcq4_b200_gguf_reorder_enabled.argtypes = [vp]
    lib.tcq4_b200_gguf_n_tensors.argtypes = [vp]
    lib.tcq4_b200_gguf_n_tensors.restype = i64
    lib.tcq4_b200_gguf_n_perms.argtypes = [vp]
    lib.tcq4_b200_gguf_n_perms.restype = i64
    lib.tcq4_b200_gguf_tensor_info.argtypes = [vp, i64, C.POINTER(C.c_char_p), C.POINTER(ci), C.POINTER(i64 * 4),
                                               C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.tcq4_b200_gguf_perm.argtypes = [vp, i64, C.POINTER(C.c_char_p), C.POINTER(vp), C.POINTER(i64)]
    lib.tcq4_b200_gguf_register_perms.argtypes = [vp]
    lib.tcq4_b200_gguf_read_tensor.argtypes = [vp, i64, vp, sz]
    lib.tcq4_b200_gguf_load_tensor.argtypes = [vp, i64, vp, sz, vp]
    lib.tcq4_b200_gguf_write.argtypes = [C.c_char_p, ci, C.POINTER(C.c_char_p), C.POINTER(ci), C.POINTER(i64), C.POINTER(i64),
                                         C.POINTER(vp), ci, C.POINTER(C.c_char_p), C.POINTER(vp), C.POINTER(i64), ci]
    lib.tcq4_b200_quantize_size.restype = sz
    lib.tcq4_b200_quantize_size.argtypes = [i64, i64]
    lib.tcq4_b200_quantize_tiles.argtypes = [vp, vp, i64, i64, vp, vp]
    lib.tcq4_b200_perm_from_imatrix.argtypes = [vp, i64, vp]
    lib.tcq4_b200_ggml_get_rows.argtypes = [C.POINTER(TensorView)] * 3 + [vp, vp]
    lib.tcq4_b200_get_rows.argtypes = [vp, vp, vp, i64, i64, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_id.argtypes = [vp, i64, C.c_size_t, vp, i64, vp, i64, i64, i64, vp, i64, i64, vp, vp, vp]
    lib.tcq4_b200_ggml_mul_mat_id.argtypes = [C.POINTER(TensorView)] * 4 + [vp, vp]
    lib.tcq4_b200_mul_mat_fanout.argtypes = [ci, vp, vp, vp, vp, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_fanout_ld.argtypes = [ci, vp, vp, vp, vp, vp, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_fanout_host_async.argtypes = [ci, vp, vp, vp, vp, i64, i64, vp]
    lib.tcq4_b200_rms_row_scale.argtypes = [vp, vp, i64, i64, C.c_float, vp]
    lib.tcq4_b200_apply_input.argtypes = [ci, vp, vp, vp, vp, vp, i64, i64, vp]
    lib.tcq4_b200_mul_mat_fanout_input.argtypes = [ci, vp, vp, vp, vp, vp, ci, vp, vp, vp, i64, i64, vp, vp]
    lib.tcq4_b200_mul_mat_add.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp, vp, vp]
    lib.tcq4_b200_quant_type_for.argtypes = [C.c_char_p, ci, i64, i64, ci]
    lib.tcq4_b200_ggml_mul_mat_fanout.argtypes = [ci, vp, C.POINTER(TensorView), vp, vp]
    lib.tcq4_b200_ggml_graph_compute.argtypes = [ci, vp, vp, vp, vp]
    lib.tcq4_b200_profile_end.argtypes = [ci, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(ci)]
    return lib


LIB = _load()

# every symbol include/tcq4_b200_diag.h declares
DIAG_EXPORTS = ["tcq4_b200_measure_umma", "tcq4_b200_measure_umma2", "tcq4_b200_measure_pipe", "tcq4_b200_measure_tmem_ld",
                "tcq4_b200_measure_fabric", "tcq4_b200_measure_f16_exact", "tcq4_b200_debug_gemv_phases", "tcq4_b200_debug_gemm2_timeline", "tcq4_b200_debug_gemm3_timeline", "tcq4_b200_debug_gemm4_timeline"]
_DIAG = None


def diag_lib() -> C.CDLL:
    """libtcq4_b200_diag.so (diagnostics only: microbenchmarks and timeline readers)."""
    global _DIAG
    if _DIAG is None:
        lib = LIB if LIB_PATH == DIAG_LIB_PATH else C.CDLL(DIAG_LIB_PATH)
        ci, vp, sz, dp = C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_double)
        lib.tcq4_b200_measure_umma.argtypes = [ci, ci, ci, ci, dp, dp]
        lib.tcq4_b200_measure_umma2.argtypes = [ci, ci, dp, dp]
        lib.tcq4_b200_measure_pipe.argtypes = [ci, ci, dp]
        lib.tcq4_b200_measure_tmem_ld.argtypes = [ci, ci, ci, dp]
        lib.tcq4_b200_measure_fabric.argtypes = [ci, ci, ci, ci, sz, ci, dp, C.POINTER(ci)]
        lib.tcq4_b200_measure_f16_exact.argtypes = [ci, C.c_uint, C.POINTER(ci), C.POINTER(ci)]
        lib.tcq4_b200_debug_gemv_phases.argtypes = [vp, sz]
        lib.tcq4_b200_debug_gemm2_timeline.argtypes = [vp, sz]
        lib.tcq4_b200_debug_gemm3_timeline.argtypes = [vp, sz]
        lib.tcq4_b200_debug_gemm4_timeline.argtypes = [vp, sz]
        _DIAG = lib
    return _DIAG


def _check(status: int):
    if status != 0:
        raise Tcq4Error(status, LIB.tcq4_b200_last_error().decode(errors="replace"))


def version() -> str:
    return LIB.tcq4_b200_version().decode()


def launch_count() -> int:
    return int(LIB.tcq4_b200_launch_count())


def measure_umma(kind: int, N: int = 256, swizzle: int = 0, iters: int = 2000):
    """(TOPS, cycles per MMA) of a bare tcgen05.mma loop: kind 0 = i8, 1 = bf16, 2 = fp8."""
    tops, cyc = C.c_double(), C.c_double()
    _check(diag_lib().tcq4_b200_measure_umma(kind, N, swizzle, iters, C.byref(tops), C.byref(cyc)))
    return tops.value, cyc.value


def profile_begin(max_launches: int) -> None:
    _check(LIB.tcq4_b200_profile_begin(max_launches))


def profile_end(kernel: int = KERNEL_AUTO):
    """(main-kernel ms summed, activation-kernel ms summed, launches) since profile_begin."""
    a, b, n = C.c_double(), C.c_double(), C.c_int()
    _check(LIB.tcq4_b200_profile_end(kernel, C.byref(a), C.byref(b), C.byref(n)))
    return a.value, b.value, n.value


def supports(M: int, N: int, K: int, cc: int = 100) -> bool:
    return bool(LIB.tcq4_b200_supports(M, N, K, cc))


def select_kernel(M: int, N: int, K: int) -> int:
    return int(LIB.tcq4_b200_select_kernel(M, N, K))


def workspace_size(M: int, N: int, K: int) -> int:
    return int(LIB.tcq4_b200_workspace_size(M, N, K))


def tile_bytes(N: int, K: int) -> int:
    """ggml_nbytes of an [N][K] TCQ4_K32 tensor (ggml.c:750-757: N * K/256 * 148)."""
    return (N // TILE_CHANNELS) * (K // TILE_K) * TILE_BYTES


def quantize_size(nrow: int, K: int) -> int:
    """bytes of the tile tensor tcq4_b200_quantize_tiles writes: ceil(nrow/8) * K/256 * 1184."""
    return int(LIB.tcq4_b200_quantize_size(nrow, K))


def perm_from_imatrix(imatrix):
    """derive_tcq4_perms_from_imatrix for one tensor (tools/quantize/quantize.cpp:127-165); numpy in, numpy int32 out."""
    import numpy as np
    im = np.ascontiguousarray(imatrix, dtype=np.float32)
    out = np.empty(im.shape[0], dtype=np.int32)
    _check(LIB.tcq4_b200_perm_from_imatrix(im.ctypes.data, im.shape[0], out.ctypes.data))
    return out


class GgufFile:
    """GGUF file with TCQ4 tensors and permutations (tcq4_b200_gguf_*): host-side mirror of the loader's TCQ4 handling
    (src/llama-model.cpp:7129-7205).  Works without a GPU except register_perms() and load()."""

    def __init__(self, path: str):
        self._h = None
        h = C.c_void_p()
        _check_plain(LIB.tcq4_b200_gguf_open(path.encode(), C.byref(h)), f"cannot parse {path}")
        self._h = h
        self.tensors = []                      # (name, ggml type, ne tuple, file offset, nbytes)
        for i in range(LIB.tcq4_b200_gguf_n_tensors(h)):
            name, typ, ne = C.c_char_p(), C.c_int(), (C.c_int64 * 4)()
            off, nb = C.c_uint64(), C.c_uint64()
            LIB.tcq4_b200_gguf_tensor_info(h, i, C.byref(name), C.byref(typ), C.byref(ne), C.byref(off), C.byref(nb))
            self.tensors.append((name.value.decode(), typ.value, tuple(ne), off.value, nb.value))
        self.perms = {}                        # tensor name -> numpy int32[K]
        import numpy as np
        for i in range(LIB.tcq4_b200_gguf_n_perms(h)):
            name, ptr, K = C.c_char_p(), C.c_void_p(), C.c_int64()
            LIB.tcq4_b200_gguf_perm(h, i, C.byref(name), C.byref(ptr), C.byref(K))
            self.perms[name.value.decode()] = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_int32)), (K.value,)).copy()
        r = LIB.tcq4_b200_gguf_reorder_enabled(h)
        self.reorder_enabled = None if r < 0 else bool(r)

    def read(self, i: int):
        import numpy as np
        out = np.empty(self.tensors[i][4], dtype=np.uint8)
        _check_plain(LIB.tcq4_b200_gguf_read_tensor(self._h, i, out.ctypes.data, out.size), "read_tensor")
        return out

    def load(self, i: int, stream_ptr: int = 0):
        """tensor bytes -> a new CUDA uint8 tensor (device bytes == file bytes)"""
        import torch
        out = torch.empty(self.tensors[i][4], dtype=torch.uint8, device="cuda")
        _check_plain(LIB.tcq4_b200_gguf_load_tensor(self._h, i, out.data_ptr(), out.numel(), stream_ptr), "load_tensor")
        return out

    def register_perms(self):
        _check_plain(LIB.tcq4_b200_gguf_register_perms(self._h), "register_perms")

    def close(self):
        if self._h:
            LIB.tcq4_b200_gguf_close(self._h)
            self._h = None

    def __del__(self):
        self.close()


def _check_plain(status: int, what: str):
    if status != 0:
        raise Tcq4Error(status, what)


def gguf_write(path: str, tensors, perms=None, reorder_enabled=None) -> None:
    """tensors: list of (name, ggml type (0 F32, 1 F16, 42 TCQ4_K32), ne0, ne1, numpy bytes); perms: {tensor name: int32[K]}."""
    import numpy as np
    perms = perms or {}
    n = len(tensors)
    names = (C.c_char_p * n)(*[t[0].encode() for t in tensors])
    types = (C.c_int * n)(*[t[1] for t in tensors])
    ne0 = (C.c_int64 * n)(*[t[2] for t in tensors])
    ne1 = (C.c_int64 * n)(*[t[3] for t in tensors])
    keep = [np.ascontiguousarray(t[4]) for t in tensors]
    data = (C.c_void_p * n)(*[a.ctypes.data for a in keep])
    m = len(perms)
    pk = [np.ascontiguousarray(v, dtype=np.int32) for v in perms.values()]
    pn = (C.c_char_p * m)(*[k.encode() for k in perms])
    pp = (C.c_void_p * m)(*[a.ctypes.data for a in pk])
    pK = (C.c_int64 * m)(*[a.size for a in pk])
    flag = -1 if reorder_enabled is None else int(bool(reorder_enabled))
    _check_plain(LIB.tcq4_b200_gguf_write(path.encode(), n, names, types, ne0, ne1, data, m, pn, pp, pK, flag), "gguf_write")


def shard_range(N: int, K: int, world: int, rank: int, align: int = 128):
    n0, n1 = C.c_int64(), C.c_int64()
    off, ln = C.c_size_t(), C.c_size_t()
    _check(LIB.tcq4_b200_shard_range(N, K, world, rank, align, C.byref(n0), C.byref(n1), C.byref(off), C.byref(ln)))
    return n0.value, n1.value, off.value, ln.value


from ._device import *  # noqa: E402,F401,F403  (torch-facing half: memory, streams, distributed)
