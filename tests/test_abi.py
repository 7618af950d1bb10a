"""The C-ABI library: loads, exports every symbol include/tcq4_b200.h declares, host logic works and
compute entry points fail LOUDLY without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "tcq4_b200.h")).read()
    return sorted(set(re.findall(r"^TCQ4_B200_API[^;(]*?\b(\w+)\s*\(", src, flags=re.M)))


def test_exports_match_header(t4):
    syms = _declared_symbols()
    assert len(syms) >= 20
    assert sorted(t4.EXPORTS) == syms
    lib = C.CDLL(t4.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), s


def test_diag_library_is_separate(t4):
    """measure_* / debug_* live in libtcq4_b200_diag.so (include/tcq4_b200_diag.h), not in the product library"""
    src = open(os.path.join(ROOT, "include", "tcq4_b200_diag.h")).read()
    syms = sorted(set(re.findall(r"^TCQ4_B200_API[^;(]*?\b(\w+)\s*\(", src, flags=re.M)))
    assert sorted(t4.DIAG_EXPORTS) == syms
    prod = C.CDLL(os.path.join(os.path.dirname(t4.DIAG_LIB_PATH), "libtcq4_b200.so"))
    diag = C.CDLL(t4.DIAG_LIB_PATH)
    for s in syms:
        assert hasattr(diag, s), s
        assert not hasattr(prod, s), f"{s} must not be exported by the product library"


def test_supports_and_select(t4):
    assert t4.supports(1, 9728, 2560) and t4.supports(512, 9728, 2560)
    assert not t4.supports(1, 9728, 2500)           # K % 256 (rrs.cu:258)
    assert not t4.supports(1, 9732, 2560)           # N % 8   (rrs.cu:259)
    assert not t4.supports(1, 9728, 2560, cc=86)    # sm_100a only
    assert not t4.supports(0, 8, 256)
    assert t4.select_kernel(1, 9728, 2560) == t4.KERNEL_GEMV
    assert t4.select_kernel(4, 2560, 9728) == t4.KERNEL_GEMV
    assert t4.select_kernel(8, 9728, 2560) == t4.KERNEL_GEMV
    assert t4.select_kernel(16, 9728, 2560) == t4.KERNEL_GEMM4     # 9 <= M <= 16 and <= 16 k-tiles per CTA: weights through TMEM
    assert t4.select_kernel(17, 9728, 2560) == t4.KERNEL_GEMM4     # batched decode: weights through TMEM
    assert t4.select_kernel(65, 9728, 2560) == t4.KERNEL_GEMM
    assert t4.select_kernel(128, 9728, 2560) == t4.KERNEL_GEMM
    assert t4.select_kernel(512, 9728, 2560) == t4.KERNEL_GEMM2
    assert t4.select_kernel(64, 2560, 9728) == t4.KERNEL_GEMM4
    assert t4.select_kernel(100, 2560, 9728) == t4.KERNEL_GEMM2     # deep or narrow matrices: the pair GEMM
    assert t4.select_kernel(100, 1024, 2560) == t4.KERNEL_GEMM2
    assert t4.select_kernel(4, 1024, 65536) == t4.KERNEL_GEMM4     # codes do not fit shared memory: a tcgen05 GEMM


def test_ggml_supports_op(t4):
    assert t4.ggml_supports_op(2560, 9728, 512)
    assert not t4.ggml_supports_op(2560 + 32, 9728, 512)


def test_workspace_size_monotone(t4):
    a = t4.workspace_size(1, 9728, 2560)
    b = t4.workspace_size(512, 9728, 2560)
    assert 0 < a < b and b % 256 == 0
    assert t4.workspace_size(1, 9728, 2500) == 0
    assert t4.tile_bytes(9728, 2560) == 14397440


def test_shard_range_covers_and_aligns(t4):
    for N, K, world in [(9728, 2560, 8), (25600, 5120, 4), (28672, 8192, 8), (1024, 2560, 8), (8, 256, 2)]:
        prev = 0
        total = 0
        for r in range(world):
            n0, n1, off, ln = t4.shard_range(N, K, world, r, 128)
            assert n0 == prev and n0 % 8 == 0 and (n1 % 128 == 0 or n1 == N)
            assert off == n0 * (K // 256) * 148 and ln == (n1 - n0) * (K // 256) * 148
            prev = n1
            total += ln
        assert prev == N and total == t4.tile_bytes(N, K)
    with pytest.raises(t4.Tcq4Error):
        t4.shard_range(9728, 2560, 0, 0)


def test_registry_without_device_is_safe(t4):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    t4.register_perm("blk.0.attn_q.weight", list(range(256)))
    assert not t4.has_perm("blk.0.attn_q.weight")
    t4.clear_perms()


def test_no_cpu_fallback(t4):
    """Without an sm_100 device compute calls return TCQ4_ERR_NO_DEVICE; nothing is computed on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    buf = (C.c_char * 4096)()
    addr = (C.addressof(buf) + 255) & ~255
    st = t4.LIB.tcq4_b200_mul_mat_ws(addr, addr, addr, 1, 8, 256, None, addr, 1 << 20, 0, None)
    assert st == 5, t4.LIB.tcq4_b200_last_error()
    assert b"no CPU fallback" in t4.LIB.tcq4_b200_last_error() or b"no fallback" in t4.LIB.tcq4_b200_last_error()
    st = t4.LIB.tcq4_b200_fwht_quantize_tc(addr, addr, None, 256, 1, None)
    assert st == 5


def test_product_never_imports_oracle():
    """The product path must not route through the oracle (oracle/tcq4_oracle.h rules)."""
    pkg = os.path.join(ROOT, "llama.cpp-rrs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "pyoracle" not in txt and "liboracle" not in txt and "tcq4_oracle" not in txt, f


def test_integration_shim_compiles_against_the_forks_headers():
    """INTEGRATION.md sections 2-4 (the shim that replaces ggml_cuda_rrs_mul_mat, the supports_op case, the registry names)
    as a translation unit against the fork's own ggml.h and ggml-cuda/common.cuh; also pins GGML_TYPE_TCQ4_K32 == 42.  Only
    where /root/reference exists (this container); the shim text in INTEGRATION.md must be the TU's text."""
    import subprocess
    if not os.path.isdir("/root/reference/ggml/src/ggml-cuda"):
        pytest.skip("no reference tree on this machine")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "shimcheck"], stdout=subprocess.DEVNULL)
    assert os.path.exists(os.path.join(ROOT, "oracle", "_ref", "obj", "shim_check.o"))
    tu = open(os.path.join(ROOT, "oracle", "shim_check.cu")).read()
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    body = tu[tu.index("static tcq4_tensor_view tcq4_view"):tu.index("// the MUL_MAT case of")].strip()
    assert body in doc, "INTEGRATION.md section 2 and oracle/shim_check.cu have drifted apart"
