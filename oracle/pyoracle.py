"""ctypes bindings for the CPU oracle and the compiled reference -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module (see oracle/tcq4_oracle.h).  The product package never does.

  Oracle   -> oracle/liboracle.so            scalar C restatement (numpy in / numpy out)
  RefCPU   -> oracle/_ref/libggml_ref.so     the reference's own CPU code compiled from
                                             /root/reference by oracle/Makefile (`make ref`)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TILE_BYTES = 1184
ACT_BLOCK_BYTES = 160

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_i16p = C.POINTER(C.c_int16)
_i8p = C.POINTER(C.c_int8)
_u8p = C.POINTER(C.c_uint8)


def _ptr(a, typ):
    if a is None:
        return C.cast(None, typ)
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(typ)


def build_oracle(force: bool = False) -> str:
    so = os.path.join(HERE, "liboracle.so")
    src = os.path.join(HERE, "tcq4_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    return so


def build_ref() -> str | None:
    """Build oracle/_ref/*.so from /root/reference when it is present; else use the prebuilt files."""
    if os.path.isdir("/root/reference/ggml/src"):
        subprocess.check_call(["make", "-C", HERE, "ref", "-j8"], stdout=subprocess.DEVNULL)
        subprocess.check_call(["make", "-C", HERE, "refcuda"], stdout=subprocess.DEVNULL)
    so = os.path.join(HERE, "_ref", "libggml_ref.so")
    return so if os.path.exists(so) else None


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.orc_fp32_to_fp16.restype = C.c_uint16
        L.orc_fp32_to_fp16.argtypes = [C.c_float]
        L.orc_fp16_to_fp32.restype = C.c_float
        L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
        L.orc_fwht.argtypes = [_f32p, C.c_int]
        L.orc_pack_group.argtypes = [_i8p, _u8p]
        L.orc_unpack_group.argtypes = [_u8p, _i8p]
        L.orc_quantize_tiles.restype = C.c_size_t
        L.orc_quantize_tiles.argtypes = [_f32p, C.c_void_p, C.c_int64, C.c_int64]
        L.orc_dequant_tiles.argtypes = [C.c_void_p, _f32p, C.c_int64]
        L.orc_get_rows.argtypes = [C.c_void_p, _i32p, C.c_int64, C.c_int64, _f32p]
        L.orc_apply_perm_weights.argtypes = [_f32p, _f32p, C.c_int64, C.c_int64, _i32p]
        L.orc_perm_from_imatrix.argtypes = [_f32p, C.c_int64, _i32p]
        L.orc_quantize_act_row.argtypes = [_f32p, _i32p, C.c_int64, _i8p, _f32p, _i16p]
        L.orc_quantize_act_blocks.argtypes = [_f32p, _i32p, C.c_int64, C.c_int64, C.c_void_p]
        L.orc_mul_mat.argtypes = [C.c_void_p, _f32p, _i32p, C.c_int64, C.c_int64, C.c_int64, _f32p,
                                  _i32p, _i32p, _i8p, _f32p, _i16p]
        L.orc_mul_mat_f32.argtypes = [C.c_void_p, _f32p, _i32p, C.c_int64, C.c_int64, C.c_int64, _f32p]
        L.orc_fnv1a64.restype = C.c_uint64
        L.orc_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]

    # --- scalars -------------------------------------------------------------------------
    def fp32_to_fp16(self, f: float) -> int:
        return int(self.lib.orc_fp32_to_fp16(C.c_float(f)))

    def fp16_to_fp32(self, h: int) -> float:
        return float(self.lib.orc_fp16_to_fp32(C.c_uint16(h)))

    # --- building blocks ----------------------------------------------------------------
    def fwht(self, x: np.ndarray) -> np.ndarray:
        y = np.ascontiguousarray(x, dtype=np.float32).copy()
        self.lib.orc_fwht(_ptr(y, _f32p), y.size)
        return y

    def pack_group(self, w: np.ndarray) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.int8)
        assert w.shape == (8, 32)
        out = np.zeros(128, dtype=np.uint8)
        self.lib.orc_pack_group(_ptr(w, _i8p), _ptr(out, _u8p))
        return out

    def unpack_group(self, p: np.ndarray) -> np.ndarray:
        p = np.ascontiguousarray(p, dtype=np.uint8)
        assert p.size == 128
        out = np.zeros((8, 32), dtype=np.int8)
        self.lib.orc_unpack_group(_ptr(p, _u8p), _ptr(out, _i8p))
        return out

    def quantize_tiles(self, W: np.ndarray) -> np.ndarray:
        """W: [N][K] float32 (K % 256 == 0) -> uint8 tile bytes [ceil(N/8) * K/256 * 1184]."""
        W = np.ascontiguousarray(W, dtype=np.float32)
        N, K = W.shape
        assert K % 256 == 0
        ntile = ((N + 7) // 8) * (K // 256)
        out = np.zeros(ntile * TILE_BYTES, dtype=np.uint8)
        ret = self.lib.orc_quantize_tiles(_ptr(W, _f32p), out.ctypes.data, N, K)
        assert ret == N * (K // 256) * 148
        return out

    def dequant_tiles(self, tiles: np.ndarray) -> np.ndarray:
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        nt = tiles.size // TILE_BYTES
        y = np.zeros(nt * 2048, dtype=np.float32)
        self.lib.orc_dequant_tiles(tiles.ctypes.data, _ptr(y, _f32p), nt * 256)
        return y

    def dequant_matrix(self, tiles: np.ndarray, N: int, K: int) -> np.ndarray:
        """Tile bytes -> Hadamard-domain weight matrix [N][K] float32."""
        y = self.dequant_tiles(tiles).reshape(N // 8, K // 256, 8, 256)
        return np.ascontiguousarray(y.transpose(0, 2, 1, 3)).reshape(N, K)

    def get_rows(self, tiles: np.ndarray, rows: np.ndarray, K: int) -> np.ndarray:
        """GET_ROWS: [len(rows)][K] float32 = FWHT256(dequant(selected rows))."""
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        out = np.zeros((rows.size, K), dtype=np.float32)
        self.lib.orc_get_rows(tiles.ctypes.data, _ptr(rows, _i32p), rows.size, K, _ptr(out, _f32p))
        return out

    def apply_perm_weights(self, W: np.ndarray, perm: np.ndarray) -> np.ndarray:
        W = np.ascontiguousarray(W, dtype=np.float32)
        perm = np.ascontiguousarray(perm, dtype=np.int32)
        out = np.empty_like(W)
        self.lib.orc_apply_perm_weights(_ptr(W, _f32p), _ptr(out, _f32p), W.shape[0], W.shape[1], _ptr(perm, _i32p))
        return out

    def perm_from_imatrix(self, im: np.ndarray) -> np.ndarray:
        im = np.ascontiguousarray(im, dtype=np.float32)
        perm = np.zeros(im.size, dtype=np.int32)
        self.lib.orc_perm_from_imatrix(_ptr(im, _f32p), im.size, _ptr(perm, _i32p))
        return perm

    def quantize_act(self, x: np.ndarray, perm: np.ndarray | None = None):
        """x: [M][K] -> (codes int8 [M][K], max_abs f32 [M][KT], sum_q int16 [M][KT][8])."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        KT = K // 256
        codes = np.zeros((M, K), dtype=np.int8)
        mx = np.zeros((M, KT), dtype=np.float32)
        sq = np.zeros((M, KT, 8), dtype=np.int16)
        p = None if perm is None else np.ascontiguousarray(perm, dtype=np.int32)
        for m in range(M):
            self.lib.orc_quantize_act_row(_ptr(x[m], _f32p), _ptr(p, _i32p), K,
                                          _ptr(codes[m], _i8p), _ptr(mx[m], _f32p), _ptr(sq[m], _i16p))
        return codes, mx, sq

    def quantize_act_blocks(self, x: np.ndarray, perm: np.ndarray | None = None) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        out = np.zeros(M * (K // 256) * ACT_BLOCK_BYTES, dtype=np.uint8)
        p = None if perm is None else np.ascontiguousarray(perm, dtype=np.int32)
        self.lib.orc_quantize_act_blocks(_ptr(x, _f32p), _ptr(p, _i32p), M, K, out.ctypes.data)
        return out

    def mul_mat(self, tiles: np.ndarray, x: np.ndarray, N: int, perm: np.ndarray | None = None,
                want_dots: bool = False, want_isum: bool = False, want_act: bool = False):
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        KT = K // 256
        assert tiles.size == (N // 8) * KT * TILE_BYTES
        Cout = np.zeros((M, N), dtype=np.float32)
        dots = np.zeros((M, N, KT, 8), dtype=np.int32) if want_dots else None
        isum = np.zeros((M, N, KT), dtype=np.int32) if want_isum else None
        codes = np.zeros((M, K), dtype=np.int8) if want_act else None
        mx = np.zeros((M, KT), dtype=np.float32) if want_act else None
        sq = np.zeros((M, KT, 8), dtype=np.int16) if want_act else None
        p = None if perm is None else np.ascontiguousarray(perm, dtype=np.int32)
        self.lib.orc_mul_mat(tiles.ctypes.data, _ptr(x, _f32p), _ptr(p, _i32p), M, N, K, _ptr(Cout, _f32p),
                             _ptr(dots, _i32p), _ptr(isum, _i32p), _ptr(codes, _i8p), _ptr(mx, _f32p),
                             _ptr(sq, _i16p))
        res = {"C": Cout}
        if want_dots:
            res["dots"] = dots
        if want_isum:
            res["isum"] = isum
        if want_act:
            res.update(codes=codes, max_abs=mx, sum_q=sq)
        return res

    def mul_mat_f32(self, tiles: np.ndarray, x: np.ndarray, N: int, perm: np.ndarray | None = None) -> np.ndarray:
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        x = np.ascontiguousarray(x, dtype=np.float32)
        M, K = x.shape
        Cout = np.zeros((M, N), dtype=np.float32)
        p = None if perm is None else np.ascontiguousarray(perm, dtype=np.int32)
        self.lib.orc_mul_mat_f32(tiles.ctypes.data, _ptr(x, _f32p), _ptr(p, _i32p), M, N, K, _ptr(Cout, _f32p))
        return Cout

    def fnv1a64(self, a: np.ndarray) -> int:
        a = np.ascontiguousarray(a)
        return int(self.lib.orc_fnv1a64(a.ctypes.data, a.nbytes))


GGML_TYPE_F32 = 0
GGML_TYPE_TCQ4_K32 = 42  # ggml/include/ggml.h:432


class RefCPU:
    """The reference's own CPU functions (oracle/_ref/libggml_ref.so)."""

    def __init__(self, so: str | None = None):
        so = so or os.path.join(HERE, "_ref", "libggml_ref.so")
        if not os.path.exists(so):
            raise FileNotFoundError(so)
        self.lib = C.CDLL(so)
        L = self.lib
        L.quantize_tcq4_tile.restype = C.c_size_t
        L.quantize_tcq4_tile.argtypes = [_f32p, C.c_void_p, C.c_int64, C.c_int64, _f32p]
        L.ggml_quantize_chunk.restype = C.c_size_t
        L.ggml_quantize_chunk.argtypes = [C.c_int, _f32p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, _f32p]
        L.dequantize_row_tcq4_tile.argtypes = [C.c_void_p, _f32p, C.c_int64]
        L.ggml_fwht_impl.argtypes = [_f32p, C.c_int]
        L.ggml_type_size.restype = C.c_size_t
        L.ggml_type_size.argtypes = [C.c_int]
        L.ggml_blck_size.restype = C.c_int64
        L.ggml_blck_size.argtypes = [C.c_int]
        L.ggml_row_size.restype = C.c_size_t
        L.ggml_row_size.argtypes = [C.c_int, C.c_int64]
        L.ggml_type_name.restype = C.c_char_p
        L.ggml_type_name.argtypes = [C.c_int]

    # ---- the reference's own GGUF reader / writer (ggml/src/gguf.cpp), to pin the product's tcq4_b200_gguf_* ----------
    class _GgufInit(C.Structure):
        _fields_ = [("no_alloc", C.c_bool), ("ctx", C.c_void_p)]

    class _GgmlInit(C.Structure):
        _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]

    def _gguf_sigs(self):
        L = self.lib
        if getattr(self, "_gguf_ready", False):
            return L
        vp, i64, sz = C.c_void_p, C.c_int64, C.c_size_t
        L.gguf_init_from_file.restype = vp
        L.gguf_init_from_file.argtypes = [C.c_char_p, self._GgufInit]
        L.gguf_init_empty.restype = vp
        L.gguf_free.argtypes = [vp]
        for f, rt in (("gguf_get_n_tensors", i64), ("gguf_get_n_kv", i64), ("gguf_get_data_offset", sz)):
            getattr(L, f).restype = rt
            getattr(L, f).argtypes = [vp]
        for f, rt in (("gguf_get_tensor_name", C.c_char_p), ("gguf_get_tensor_type", C.c_int), ("gguf_get_tensor_offset", sz),
                      ("gguf_get_tensor_size", sz), ("gguf_get_key", C.c_char_p), ("gguf_get_kv_type", C.c_int),
                      ("gguf_get_arr_type", C.c_int), ("gguf_get_arr_n", sz), ("gguf_get_arr_data", vp),
                      ("gguf_get_val_bool", C.c_bool)):
            getattr(L, f).restype = rt
            getattr(L, f).argtypes = [vp, i64]
        L.gguf_find_key.restype = i64
        L.gguf_find_key.argtypes = [vp, C.c_char_p]
        L.gguf_set_val_bool.argtypes = [vp, C.c_char_p, C.c_bool]
        L.gguf_set_arr_data.argtypes = [vp, C.c_char_p, C.c_int, vp, sz]
        L.gguf_add_tensor.argtypes = [vp, vp]
        L.gguf_write_to_file.restype = C.c_bool
        L.gguf_write_to_file.argtypes = [vp, C.c_char_p, C.c_bool]
        L.ggml_init.restype = vp
        L.ggml_init.argtypes = [self._GgmlInit]
        L.ggml_free.argtypes = [vp]
        L.ggml_new_tensor_2d.restype = vp
        L.ggml_new_tensor_2d.argtypes = [vp, C.c_int, i64, i64]
        L.ggml_set_name.restype = vp
        L.ggml_set_name.argtypes = [vp, C.c_char_p]
        L.ggml_get_data.restype = vp
        L.ggml_get_data.argtypes = [vp]
        L.ggml_nbytes.restype = sz
        L.ggml_nbytes.argtypes = [vp]
        self._gguf_ready = True
        return L

    def gguf_read(self, path: str) -> dict:
        """Parse with the reference's gguf_init_from_file: tensors [(name, type, abs offset, size)], perms, reorder flag
        (the scan of src/llama-model.cpp:7133-7166)."""
        L = self._gguf_sigs()
        g = L.gguf_init_from_file(path.encode(), self._GgufInit(True, None))
        assert g, "reference gguf_init_from_file failed"
        try:
            base = L.gguf_get_data_offset(g)
            tensors = [(L.gguf_get_tensor_name(g, i).decode(), L.gguf_get_tensor_type(g, i), base + L.gguf_get_tensor_offset(g, i),
                        L.gguf_get_tensor_size(g, i)) for i in range(L.gguf_get_n_tensors(g))]
            perms, reorder = {}, None
            k = L.gguf_find_key(g, b"tcq4.reorder.enabled")
            if k >= 0:
                reorder = bool(L.gguf_get_val_bool(g, k))
            for i in range(L.gguf_get_n_kv(g)):
                key = L.gguf_get_key(g, i).decode()
                if key.startswith("tcq4.") and len(key) > 10 and key.endswith(".perm") and L.gguf_get_kv_type(g, i) == 9 \
                        and L.gguf_get_arr_type(g, i) == 5:
                    n = L.gguf_get_arr_n(g, i)
                    ptr = C.cast(L.gguf_get_arr_data(g, i), C.POINTER(C.c_int32))
                    perms[key[5:-5]] = np.ctypeslib.as_array(ptr, (n,)).copy()
            return dict(tensors=tensors, perms=perms, reorder=reorder)
        finally:
            L.gguf_free(g)

    def gguf_write(self, path: str, tensors, perms=None, reorder=None) -> None:
        """Write with the reference's gguf writer: tensors [(name, ggml type, ne0, ne1, numpy bytes)], the TCQ4 keys as
        src/llama-quant.cpp:840-854 sets them."""
        L = self._gguf_sigs()
        total = sum(int(np.asarray(t[4]).nbytes) for t in tensors) + 4096 * (len(tensors) + 4)
        ctx = L.ggml_init(self._GgmlInit(total, None, False))
        g = L.gguf_init_empty()
        try:
            if reorder is not None:
                L.gguf_set_val_bool(g, b"tcq4.reorder.enabled", bool(reorder))
            keep = []
            for name, p in (perms or {}).items():
                a = np.ascontiguousarray(p, dtype=np.int32)
                keep.append(a)
                L.gguf_set_arr_data(g, f"tcq4.{name}.perm".encode(), 5, a.ctypes.data, a.size)
            for name, typ, ne0, ne1, data in tensors:
                t = L.ggml_new_tensor_2d(ctx, typ, ne0, ne1)
                L.ggml_set_name(t, name.encode())
                raw = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
                assert L.ggml_nbytes(t) == raw.size, (name, L.ggml_nbytes(t), raw.size)
                C.memmove(L.ggml_get_data(t), raw.ctypes.data, raw.size)
                L.gguf_add_tensor(g, t)
            assert L.gguf_write_to_file(g, path.encode(), False)
        finally:
            L.gguf_free(g)
            L.ggml_free(ctx)

    def type_traits(self):
        L = self.lib
        return dict(name=L.ggml_type_name(GGML_TYPE_TCQ4_K32).decode(),
                    type_size=int(L.ggml_type_size(GGML_TYPE_TCQ4_K32)),
                    blck_size=int(L.ggml_blck_size(GGML_TYPE_TCQ4_K32)))

    def quantize_tiles(self, W: np.ndarray, via_chunk: bool = True) -> np.ndarray:
        W = np.ascontiguousarray(W, dtype=np.float32)
        N, K = W.shape
        ntile = ((N + 7) // 8) * (K // 256)
        out = np.zeros(ntile * TILE_BYTES, dtype=np.uint8)
        if via_chunk:   # ggml/src/ggml.c:7564-7565
            ret = self.lib.ggml_quantize_chunk(GGML_TYPE_TCQ4_K32, _ptr(W, _f32p), out.ctypes.data, 0, N, K,
                                               C.cast(None, _f32p))
        else:
            ret = self.lib.quantize_tcq4_tile(_ptr(W, _f32p), out.ctypes.data, N, K, C.cast(None, _f32p))
        assert ret == N * (K // 256) * 148, ret
        return out

    def dequant_tiles(self, tiles: np.ndarray) -> np.ndarray:
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        nt = tiles.size // TILE_BYTES
        y = np.zeros(nt * 2048, dtype=np.float32)
        self.lib.dequantize_row_tcq4_tile(tiles.ctypes.data, _ptr(y, _f32p), nt * 256)
        return y

    def fwht(self, x: np.ndarray) -> np.ndarray:
        n = x.size
        buf = np.zeros(n + 16, dtype=np.float32)          # 64-byte aligned view (AVX builds need it)
        off = (-buf.ctypes.data // 4) % 16
        y = buf[off:off + n]
        y[:] = x
        self.lib.ggml_fwht_impl(_ptr(y, _f32p), n)
        return y.copy()


class RefBaseline:
    """bench.py's CPU baseline: the reference's CPU-side TCQ4 path timed step by step --
    dequantize_row_tcq4_tile -> F32 Hadamard-domain weights, ggml_fwht_impl on (permuted) activations,
    ggml CPU backend F32 MUL_MAT -- all from oracle/_ref (the reference's own code).  Falls back to
    the scalar port (Oracle.mul_mat_f32) when oracle/_ref is absent; `kind` says which."""

    def __init__(self, threads: int | None = None):
        self.threads = int(threads or len(os.sched_getaffinity(0)))
        self.kind = "port"
        self.ref = None
        self.drv = None
        so = os.path.join(HERE, "_ref", "libref_baseline.so")
        try:
            build_ref()
            self.ref = RefCPU()
            self.drv = C.CDLL(so)
            self.drv.refbase_mul_mat_f32.argtypes = [_f32p, _f32p, _f32p, C.c_int64, C.c_int64, C.c_int64, C.c_int]
            self.kind = "reference"
        except Exception:  # noqa: BLE001
            self.orc = Oracle()

    # -- step (2): tiles -> F32 [N][K], threaded over tile ranges (ctypes releases the GIL) ----------
    def dequant(self, tiles: np.ndarray, N: int, K: int) -> np.ndarray:
        from concurrent.futures import ThreadPoolExecutor
        tiles = np.ascontiguousarray(tiles, dtype=np.uint8)
        nt = tiles.size // TILE_BYTES
        flat = np.empty(nt * 2048, dtype=np.float32)
        nchunk = min(self.threads, max(1, nt // 64))
        bounds = np.linspace(0, nt, nchunk + 1).astype(np.int64)

        def work(i):
            a, b = int(bounds[i]), int(bounds[i + 1])
            if b > a:
                self.ref.lib.dequantize_row_tcq4_tile(tiles.ctypes.data + a * TILE_BYTES,
                                                      _ptr(flat[a * 2048:b * 2048], _f32p), (b - a) * 256)
        with ThreadPoolExecutor(self.threads) as ex:
            list(ex.map(work, range(nchunk)))
        y = flat.reshape(N // 8, K // 256, 8, 256)
        return np.ascontiguousarray(y.transpose(0, 2, 1, 3)).reshape(N, K)

    # -- step (3): x -> FWHT256(perm(x)) on 64-byte aligned rows --------------------------------------
    def rotate(self, x: np.ndarray, perm: np.ndarray | None = None) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if perm is not None:
            x = np.ascontiguousarray(x[:, perm])
        M, K = x.shape
        buf = np.zeros(M * K + 16, dtype=np.float32)
        off = (-buf.ctypes.data // 4) % 16
        y = buf[off:off + M * K]
        y[:] = x.reshape(-1)
        base = y.ctypes.data
        for b in range(M * K // 256):
            self.ref.lib.ggml_fwht_impl(C.cast(base + b * 1024, _f32p), 256)
        return y.reshape(M, K)

    # -- step (4): ggml CPU backend F32 MUL_MAT -------------------------------------------------------
    def matmul(self, Wf: np.ndarray, xr: np.ndarray) -> np.ndarray:
        N, K = Wf.shape
        M = xr.shape[0]
        y = np.empty((M, N), dtype=np.float32)
        rc = self.drv.refbase_mul_mat_f32(_ptr(Wf, _f32p), _ptr(np.ascontiguousarray(xr), _f32p), _ptr(y, _f32p),
                                          M, N, K, self.threads)
        assert rc == 0, rc
        return y

    def mul_mat(self, tiles: np.ndarray, x: np.ndarray, N: int, perm: np.ndarray | None = None):
        """Full per-call CPU path; returns (y, {'dequant_s','fwht_s','matmul_s'})."""
        import time
        K = x.shape[1]
        if self.kind != "reference":
            t0 = time.perf_counter()
            y = self.orc.mul_mat_f32(tiles, x, N, perm)
            return y, {"dequant_s": 0.0, "fwht_s": 0.0, "matmul_s": time.perf_counter() - t0}
        t0 = time.perf_counter()
        Wf = self.dequant(tiles, N, K)
        t1 = time.perf_counter()
        xr = self.rotate(x, perm)
        t2 = time.perf_counter()
        y = self.matmul(Wf, xr)
        t3 = time.perf_counter()
        return y, {"dequant_s": t1 - t0, "fwht_s": t2 - t1, "matmul_s": t3 - t2}
