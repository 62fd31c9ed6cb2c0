"""GGUF container + block-quant codecs against gguf-py (the reference's own Python reader / dequantisers)."""
import numpy as np
import pytest

from distributed_llm_pipeline_b200.gguf import (GGMLType, GGUFWriter, dequantize, open_gguf, quantize, random_blocks,
                                                 write_synthetic_gguf)
from distributed_llm_pipeline_b200.gguf.quants import pack_scale_min_k4, unpack_scale_min_k4
from distributed_llm_pipeline_b200.models.config import config_from_gguf, preset

gguf = pytest.importorskip("gguf")
from gguf import quants as GQ  # noqa: E402

TYPES = [(GGMLType.Q8_0, 3e-5), (GGMLType.Q4_0, 1e-2), (GGMLType.Q4_K, 1e-2), (GGMLType.Q5_K, 3e-3), (GGMLType.Q6_K, 1e-3),
         (GGMLType.F16, 1e-6), (GGMLType.BF16, 1e-4), (GGMLType.Q4_1, 1e-2), (GGMLType.Q5_0, 3e-3), (GGMLType.Q5_1, 3e-3),
         (GGMLType.Q2_K, 0.2), (GGMLType.Q3_K, 5e-2), (GGMLType.IQ4_NL, 1e-2), (GGMLType.IQ4_XS, 1e-2), (GGMLType.MXFP4, 2e-2),
         (GGMLType.TQ1_0, 0.7), (GGMLType.TQ2_0, 0.7)]


@pytest.mark.parametrize("t,tol", TYPES)
def test_quant_roundtrip_matches_gguf_py(t, tol):
    rng = np.random.default_rng(0)
    x = (rng.standard_normal((16, 512)) * 0.05).astype(np.float32)
    raw = quantize(x, t)
    mine = dequantize(raw, t, x.shape)
    ref = GQ.dequantize(raw.reshape(16, -1), gguf.GGMLQuantizationType(int(t)))
    assert np.array_equal(mine, ref.reshape(x.shape))          # bit-exact decode
    nmse = ((mine - x) ** 2).sum() / (x ** 2).sum()
    assert nmse < tol


@pytest.mark.parametrize("t", [GGMLType.Q4_0, GGMLType.Q4_1, GGMLType.Q5_0, GGMLType.Q5_1, GGMLType.Q8_0])
def test_legacy_encoders_are_bit_exact_with_gguf_py(t):
    rng = np.random.default_rng(3)
    x = (rng.standard_normal((8, 256)) * 0.05).astype(np.float32)
    ref = GQ.quantize(x, gguf.GGMLQuantizationType(int(t)))
    assert np.array_equal(quantize(x, t).reshape(ref.shape), ref)


@pytest.mark.parametrize("t,fields", [(GGMLType.Q2_K, [(80, 0.01), (82, 0.02)]), (GGMLType.Q3_K, [(108, 0.01)])])
def test_low_bit_k_decoders_on_arbitrary_blocks(t, fields):
    """Every bit pattern of the packed scale / quant fields decodes as gguf-py does (fp16 fields pinned to finite values)."""
    from distributed_llm_pipeline_b200.gguf.constants import BLOCK_GEOMETRY
    rng = np.random.default_rng(4)
    nbytes = BLOCK_GEOMETRY[t][1]
    raw = rng.integers(0, 256, size=(32, nbytes), dtype=np.uint8)
    for col, val in fields:
        raw[:, col:col + 2] = np.array([val], np.float16).view(np.uint8)
    ref = GQ.dequantize(raw.reshape(1, -1), gguf.GGMLQuantizationType(int(t))).reshape(-1)
    assert np.allclose(dequantize(raw, t), ref, rtol=0, atol=1e-7)


@pytest.mark.parametrize("t,f16_cols", [(GGMLType.IQ4_NL, [0]), (GGMLType.IQ4_XS, [0]), (GGMLType.TQ1_0, [52]), (GGMLType.TQ2_0, [64]),
                                        (GGMLType.MXFP4, [])])
def test_4bit_ternary_mx_decoders_on_arbitrary_blocks(t, f16_cols):
    """Random bytes in every packed field (scales pinned finite; every MX exponent 0..253) decode exactly as gguf-py does."""
    from distributed_llm_pipeline_b200.gguf.constants import BLOCK_GEOMETRY
    rng = np.random.default_rng(5)
    nbytes = BLOCK_GEOMETRY[t][1]
    raw = rng.integers(0, 256, size=(254, nbytes), dtype=np.uint8)
    for col in f16_cols:
        raw[:, col:col + 2] = np.array([0.0123], np.float16).view(np.uint8)
    if t == GGMLType.MXFP4:
        raw[:, 0] = np.arange(254)
    ref = GQ.dequantize(raw.reshape(1, -1), gguf.GGMLQuantizationType(int(t))).reshape(-1)
    assert np.array_equal(dequantize(raw, t), ref)


@pytest.mark.parametrize("name", ["IQ2_XXS", "IQ2_XS", "IQ2_S", "IQ3_XXS", "IQ3_S", "IQ1_S", "IQ1_M"])
def test_codebook_decoders_on_arbitrary_blocks(name):
    """The lattice code-book types: random bytes in every field (code indices, sign rows, sub-block scales, high bits) decode exactly as
    gguf-py does; only the fp16 super-scale is kept finite (for IQ1_M it is spread over the top nibbles of the four scale words)."""
    from distributed_llm_pipeline_b200.gguf.constants import BLOCK_GEOMETRY
    t = GGMLType[name]
    rng = np.random.default_rng(11)
    nbytes = BLOCK_GEOMETRY[t][1]
    raw = rng.integers(0, 256, size=(200, nbytes), dtype=np.uint8)
    if name == "IQ1_M":
        raw[:, 55] &= 0x3F
    else:
        raw[:, 1] &= 0x3B
    ref = GQ.dequantize(raw.reshape(1, -1), gguf.GGMLQuantizationType(int(t))).reshape(-1)
    out = dequantize(raw, t)
    assert np.isfinite(out).all() and np.array_equal(out, ref)


@pytest.mark.parametrize("t", [GGMLType.TQ1_0, GGMLType.TQ2_0])
def test_ternary_encoders_are_bit_exact_with_gguf_py(t):
    rng = np.random.default_rng(6)
    x = rng.integers(-1, 2, (8, 512)).astype(np.float32) * 0.25
    ref = GQ.quantize(x, gguf.GGMLQuantizationType(int(t)))
    assert np.array_equal(quantize(x, t).reshape(ref.shape), ref)
    assert np.array_equal(dequantize(quantize(x, t), t, x.shape), x)      # ternary data is represented exactly


@pytest.mark.parametrize("t", [GGMLType.Q8_0, GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K])
def test_random_blocks_are_valid(t):
    rng = np.random.default_rng(1)
    raw = random_blocks(t, 8192, rng)
    mine = dequantize(raw, t)
    ref = GQ.dequantize(raw.reshape(1, -1), gguf.GGMLQuantizationType(int(t))).reshape(-1)
    assert np.array_equal(mine, ref)
    assert np.isfinite(mine).all() and 0.005 < mine.std() < 0.2


def test_scale_min_pack_roundtrip():
    rng = np.random.default_rng(2)
    sc = rng.integers(0, 64, (10, 8), dtype=np.uint8)
    mn = rng.integers(0, 64, (10, 8), dtype=np.uint8)
    s2, m2 = unpack_scale_min_k4(pack_scale_min_k4(sc, mn))
    assert np.array_equal(sc, s2) and np.array_equal(mn, m2)


@pytest.mark.parametrize("ftype", ["F16", "Q8_0", "Q4_K_M", "Q5_1", "Q3_K", "IQ4_XS", "TQ1_0"])
def test_synthetic_file_readable_by_reference_reader(tmp_path, ftype):
    cfg = preset("tiny")
    path = write_synthetic_gguf(str(tmp_path / "m.gguf"), cfg, ftype)
    mine = open_gguf(path)
    ref = gguf.GGUFReader(path)
    assert {t.name for t in ref.tensors} == set(mine.tensors)
    for t in ref.tensors:
        a = mine.array(t.name)
        b = GQ.dequantize(t.data, t.tensor_type).reshape(a.shape)
        assert np.array_equal(a, b), t.name
    c2 = config_from_gguf(mine.metadata, mine.tensors)
    assert (c2.n_embd, c2.n_layer, c2.n_head, c2.n_head_kv, c2.n_ff, c2.vocab_size) == \
           (cfg.n_embd, cfg.n_layer, cfg.n_head, cfg.n_head_kv, cfg.n_ff, cfg.vocab_size)
    if ftype == "Q4_K_M":
        assert mine.tensors["output.weight"].ggml_type == GGMLType.Q6_K
        assert mine.tensors["blk.0.ffn_gate.weight"].ggml_type == GGMLType.Q4_K


def test_writer_kv_types_and_alignment(tmp_path):
    w = GGUFWriter(str(tmp_path / "kv.gguf"), alignment=64)
    w.add("general.architecture", "llama")
    w.add("a.int", 7)
    w.add("a.float", 0.5)
    w.add("a.bool", True)
    w.add("a.strs", ["x", "yz"])
    w.add("a.arr", np.arange(5, dtype=np.int32))
    w.add_array("t0", np.arange(64, dtype=np.float32).reshape(2, 32))
    w.add_array("t1", np.ones((4, 32), np.float32), GGMLType.Q8_0)
    w.write()
    f = open_gguf(str(tmp_path / "kv.gguf"))
    assert f.metadata["a.int"] == 7 and f.metadata["a.bool"] is True and f.metadata["a.strs"] == ["x", "yz"]
    assert list(f.metadata["a.arr"]) == [0, 1, 2, 3, 4]
    assert f.files[0].alignment == 64 and f.files[0].data_offset % 64 == 0
    assert np.array_equal(f.array("t0"), np.arange(64, dtype=np.float32).reshape(2, 32))
    assert np.allclose(f.array("t1"), 1.0, atol=1e-2)


def test_truncated_file_is_rejected(tmp_path):
    path = write_synthetic_gguf(str(tmp_path / "m.gguf"), preset("tiny"), "F16")
    data = open(path, "rb").read()
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(data[: len(data) // 2])
    with pytest.raises(ValueError):
        open_gguf(str(bad))


def test_loader_modes_mmap_read_direct_mlock(tmp_path, monkeypatch):
    """`--no-mmap`, `--direct-io`, `--mlock` (llama-mmap.cpp): every mode yields the same bytes; a refused mlock only warns."""
    import warnings
    from distributed_llm_pipeline_b200.gguf.reader import GGUFFile
    path = write_synthetic_gguf(str(tmp_path / "m.gguf"), preset("tiny"), "Q8_0")
    base = GGUFFile(path)
    assert base.load_mode == "mmap" and not base.mlocked
    want = {n: np.array(base.raw(n)) for n in base.tensors}
    rd = GGUFFile(path, use_mmap=False)
    assert rd.load_mode == "read" and all(np.array_equal(rd.raw(n), want[n]) for n in want)
    di = GGUFFile(path, direct_io=True)
    assert di.load_mode in ("direct", "read")                 # tmpfs refuses O_DIRECT: buffered fallback
    assert all(np.array_equal(di.raw(n), want[n]) for n in want) and list(di.metadata) == list(base.metadata)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        ml = GGUFFile(path, mlock=True)
        assert ml.mlocked or any("mlock" in str(x.message) for x in w)
    assert all(np.array_equal(ml.raw(n), want[n]) for n in want)
    monkeypatch.setenv("DLP_NO_MMAP", "1")                    # what --no-mmap sets for every loader in the process
    assert open_gguf(path).files[0].load_mode == "read"
