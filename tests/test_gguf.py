"""GGUF files with TCQ4_K32 tensors and channel permutations (tcq4_b200_gguf_*).
Pins: (1) a golden file written by the REFERENCE's gguf writer + quantiser (tests/golden/make_golden_gguf.py) must parse
to the same tensors / permutations / flag; (2) live, when oracle/_ref is present: files written by the product parse
identically through the reference's own gguf_init_from_file, and vice versa."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden_gguf import OUT as GOLDEN_GGUF, inputs  # noqa: E402


@pytest.fixture(scope="module")
def t4():
    import __graft_entry__ as graft
    return graft.load_package()


def test_golden_gguf_written_by_the_reference_parses(t4, orc):
    g = t4.GgufFile(GOLDEN_GGUF)
    ws, norm = inputs()
    assert g.reorder_enabled is True
    names = [t[0] for t in g.tensors]
    assert names == ["blk.0.attn_q.weight", "blk.0.attn_norm.weight", "blk.0.ffn_down.weight"]
    for name, N, K, W, perm in ws:
        i = names.index(name)
        _, typ, ne, off, nbytes = g.tensors[i]
        assert typ == 42 and ne[:2] == (K, N) and nbytes == N * (K // 256) * 148 and off % 32 == 0
        # the bytes in the file are the reference quantiser's output for the permuted weights == the oracle's
        assert np.array_equal(g.read(i), orc.quantize_tiles(orc.apply_perm_weights(W, perm)))
        assert np.array_equal(g.perms[name], perm)
    i = names.index("blk.0.attn_norm.weight")
    assert g.tensors[i][1] == 0 and g.tensors[i][4] == 512 * 4
    assert np.array_equal(g.read(i).view(np.float32), norm)
    g.close()


def test_writer_reader_roundtrip_and_key_rules(t4, orc, tmp_path):
    rng = np.random.default_rng(1)
    W = (rng.standard_normal((8, 256)) * 0.02).astype(np.float32)
    tiles = orc.quantize_tiles(W)
    p = np.arange(256, dtype=np.int32)[::-1].copy()
    path = str(tmp_path / "a.gguf")
    t4.gguf_write(path, [("w", 42, 256, 8, tiles), ("f", 0, 4, 2, np.arange(8, dtype=np.float32))],
                  perms={"w": p, "blk.1.ffn_up.weight": np.arange(512, dtype=np.int32)}, reorder_enabled=False)
    g = t4.GgufFile(path)
    assert g.reorder_enabled is False
    assert [t[0] for t in g.tensors] == ["w", "f"]
    assert np.array_equal(g.read(0), tiles) and np.array_equal(g.read(1).view(np.float32), np.arange(8, dtype=np.float32))
    assert set(g.perms) == {"w", "blk.1.ffn_up.weight"} and np.array_equal(g.perms["w"], p)
    # no flag key -> None; keys of 10 characters or fewer are not permutations (llama-model.cpp:7146: strlen(key) > 10)
    path2 = str(tmp_path / "b.gguf")
    t4.gguf_write(path2, [("w", 42, 256, 8, tiles)], perms={"": p})          # key "tcq4..perm" has length 10
    g2 = t4.GgufFile(path2)
    assert g2.reorder_enabled is None and g2.perms == {}


def test_bad_files_are_rejected(t4, tmp_path):
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"GGML" + b"\0" * 64)
    with pytest.raises(t4.Tcq4Error):
        t4.GgufFile(str(bad))
    with pytest.raises(t4.Tcq4Error):
        t4.GgufFile(str(tmp_path / "missing.gguf"))
    trunc = tmp_path / "trunc.gguf"
    trunc.write_bytes(open(GOLDEN_GGUF, "rb").read()[:-100])                # tensor data cut short
    with pytest.raises(t4.Tcq4Error):
        t4.GgufFile(str(trunc))
    with pytest.raises(t4.Tcq4Error):                                       # TCQ4 tensors need K % 256 == 0, N % 8 == 0
        t4.gguf_write(str(tmp_path / "c.gguf"), [("w", 42, 200, 8, np.zeros(1184, dtype=np.uint8))])


def test_product_writer_against_the_reference_reader_and_back(t4, orc, ref, tmp_path):
    """live pin (needs oracle/_ref): same file, two parsers"""
    ws, norm = inputs()
    tensors, perms = [], {}
    for name, N, K, W, perm in ws:
        tensors.append((name, 42, K, N, orc.quantize_tiles(orc.apply_perm_weights(W, perm))))
        perms[name] = perm
    tensors.append(("blk.0.attn_norm.weight", 0, 512, 1, norm))
    mine = str(tmp_path / "mine.gguf")
    t4.gguf_write(mine, tensors, perms=perms, reorder_enabled=True)
    r = ref.gguf_read(mine)                                                 # the reference's gguf_init_from_file
    g = t4.GgufFile(mine)
    assert r["reorder"] is True and set(r["perms"]) == set(perms)
    for k in perms:
        assert np.array_equal(r["perms"][k], perms[k])
    assert [(t[0], t[1], t[3], t[4]) for t in g.tensors] == [(n, ty, off, sz) for n, ty, off, sz in r["tensors"]]
    theirs = str(tmp_path / "theirs.gguf")
    ref.gguf_write(theirs, tensors, perms, reorder=True)                    # the reference's writer
    g2 = t4.GgufFile(theirs)
    for i in range(len(tensors)):
        assert g2.tensors[i][:3] == g.tensors[i][:3] and np.array_equal(g2.read(i), g.read(i))
    assert all(np.array_equal(g2.perms[k], perms[k]) for k in perms)


@pytest.mark.gpu
def test_load_register_and_run_from_the_golden_file(t4, orc):
    """the whole caller side on the device: upload tiles verbatim, register the permutations by tensor name, MUL_MAT through
    the ggml mirror (perm looked up by src0->name) == the oracle on the original weights' quantisation"""
    import torch
    g = t4.GgufFile(GOLDEN_GGUF)
    t4.clear_perms()
    g.register_perms()
    ws, _ = inputs()
    names = [t[0] for t in g.tensors]
    rng = np.random.default_rng(5)
    for name, N, K, W, perm in ws:
        tiles_d = g.load(names.index(name))
        assert t4.has_perm(name)
        x = rng.standard_normal((3, K)).astype(np.float32)
        y = t4.ggml_mul_mat(tiles_d, torch.from_numpy(x).cuda(), N, name=name)
        want = orc.mul_mat(orc.quantize_tiles(orc.apply_perm_weights(W, perm)), x, N, perm=perm)["C"]
        got = y.cpu().numpy()
        assert float(((got - want) ** 2).sum() / (want ** 2).sum()) < 1e-9
        # and the permutation does what it is for: the result approximates W . x of the UNPERMUTED weights
        exact = x @ W.T
        assert float(((got - exact) ** 2).sum() / (exact ** 2).sum()) < 5e-2
    t4.clear_perms()


def test_quant_type_selection_rules():
    """llama_tensor_get_type under LLAMA_FTYPE_MOSTLY_TCQ4_K32 (src/llama-quant.cpp:204-526, 885-898): which tensors become
    TCQ4_K32 and what the others fall back to"""
    import __graft_entry__ as graft
    t4 = graft.load_package()
    F32, F16, Q5_0, Q8_0, Q4_K, Q6_K, TCQ4 = 0, 1, 6, 8, 12, 14, 42
    f = lambda name, nd, K, N, has_output=1: t4.LIB.tcq4_b200_quant_type_for(name.encode(), nd, K, N, has_output)
    assert f("blk.0.attn_q.weight", 2, 2560, 4096) == TCQ4
    assert f("blk.3.ffn_down_exps.weight", 3, 768, 2048) == TCQ4          # expert tensors are 3-D
    assert f("blk.0.attn_k.weight", 2, 2560, 1020) == Q4_K                # rows not a multiple of 8 (:481-483, :515)
    assert f("blk.0.ffn_up.weight", 2, 2432, 4096) == F16                 # K % 256 != 0: Q4_K needs it too -> F16 (:518-520)
    assert f("token_embd.weight", 2, 2560, 151936) == Q4_K                # :280-285
    assert f("token_embd.weight", 2, 4096, 128256, has_output=0) == Q6_K  # tied embeddings follow the output rule (:232)
    assert f("output.weight", 2, 4096, 128256) == Q6_K                    # 4096 % 256 == 0 (:237-253; blck_size 256, ggml.c:750-753)
    assert f("output.weight", 2, 2560, 151936) == Q6_K                    # Qwen3-4B: 2560 % 256 == 0 -> Q6_K
    assert f("token_embd.weight", 2, 2560, 151936, has_output=0) == Q6_K  # ... and its tied embedding
    assert f("output.weight", 2, 2432, 32000) == Q8_0                     # K % 256 != 0 -> Q8_0 (:243-244)
    UNCHANGED = -2                                                        # not quantised: keeps its source type (:885-898)
    assert f("blk.0.attn_norm.weight", 1, 2560, 1) == UNCHANGED
    assert f("blk.0.attn_q_norm.weight", 2, 128, 32) == UNCHANGED         # :891
    assert f("blk.0.ffn_gate_inp.weight", 2, 2048, 128) == UNCHANGED      # MoE router (:898)
    assert f("blk.0.attn_q.bias", 1, 4096, 1) == UNCHANGED
