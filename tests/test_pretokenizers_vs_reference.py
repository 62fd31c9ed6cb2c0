"""Byte-level BPE with every pre-tokenizer (`tokenizer.ggml.pre`) against the reference's own tokenizer: a tiny model file gets a BPE
vocabulary trained here and one pre-tokenizer name after the other; `llama-perplexity` tokenizes a text with it (the token ids are in the
logits file it writes) and `utils/tokenizer.py` must produce the same ids.  (llama-vocab.cpp:279-480 regex table, :520-700 BPE session)"""
import collections
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline import reference as ref_build  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGUFWriter, open_gguf, write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import TokenType, ValueType  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.utils.tokenizer import PRE_TOKENIZERS, Tokenizer, _bytes_to_unicode  # noqa: E402

TEXT = ("Hello world's best 12345 tests, we'll see: foo_bar(3.14) != 42!!  Two  spaces\n\nNew paragraph\twith tab, I'M SHOUTING and CamelCaseWords. "
        "你好世界 こんにちは 1234567 e-mail@example.com $9.99 <tag> [x] {y} 'quoted' don't can't I'd ... end\n") * 6


def train_bpe(text: str, n_merges: int):
    """A tiny byte-level BPE trained on white-space words (the merges only have to be plausible, not optimal)."""
    b2u = _bytes_to_unicode()
    words = collections.Counter(" " + w for w in text.split())
    seqs = {w: [b2u[b] for b in w.encode("utf-8")] for w in words}
    merges = []
    for _ in range(n_merges):
        pairs = collections.Counter()
        for w, s in seqs.items():
            for a, b in zip(s, s[1:]):
                pairs[(a, b)] += words[w]
        if not pairs:
            break
        (a, b), _n = max(pairs.items(), key=lambda kv: (kv[1], kv[0]))
        merges.append((a, b))
        for w, s in seqs.items():
            out, i = [], 0
            while i < len(s):
                if i + 1 < len(s) and s[i] == a and s[i + 1] == b:
                    out.append(a + b)
                    i += 2
                else:
                    out.append(s[i])
                    i += 1
            seqs[w] = out
    return merges


@pytest.fixture(scope="module")
def setup(tmp_path_factory):
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-perplexity"):
        pytest.skip(f"reference binaries unavailable: {why}")
    scratch = str(tmp_path_factory.mktemp("pretok"))
    env = ref_build.cpu_env(scratch)
    b2u = _bytes_to_unicode()
    merges = train_bpe(TEXT, 180)
    toks = ["<unk>", "<s>", "</s>"] + [b2u[b] for b in range(256)] + [a + b for a, b in merges]
    assert len(set(toks)) == len(toks)
    cfg = preset("tiny", vocab_size=len(toks))
    base = write_synthetic_gguf(os.path.join(scratch, "base.gguf"), cfg, "F16", seed=1, vocab=False)
    text_path = os.path.join(scratch, "text.txt")
    open(text_path, "w", encoding="utf-8").write(TEXT)
    return scratch, env, base, toks, merges, text_path


def with_vocab(base: str, dst: str, toks, merges, pre: str) -> str:
    f = open_gguf(base)
    w = GGUFWriter(dst)
    for k, v in f.metadata.items():
        if not k.startswith("tokenizer."):
            w.add(k, v, f.files[0].metadata_types[k]) if not isinstance(v, (list, np.ndarray)) else w.add(k, v)
    w.add("tokenizer.ggml.model", "gpt2")
    w.add("tokenizer.ggml.pre", pre)
    w.add("tokenizer.ggml.tokens", toks, etype=ValueType.STRING)
    w.add("tokenizer.ggml.token_type", np.asarray([int(TokenType.CONTROL)] * 3 + [int(TokenType.NORMAL)] * (len(toks) - 3), np.int32), etype=ValueType.INT32)
    w.add("tokenizer.ggml.merges", [f"{a} {b}" for a, b in merges], etype=ValueType.STRING)
    w.add("tokenizer.ggml.bos_token_id", 1, ValueType.UINT32)
    w.add("tokenizer.ggml.eos_token_id", 2, ValueType.UINT32)
    w.add("tokenizer.ggml.unknown_token_id", 0, ValueType.UINT32)
    w.add("tokenizer.ggml.add_bos_token", False)
    for n, t in f.tensors.items():
        w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
    w.write()
    return dst


NAMES = ["gpt-2", "llama-bpe", "qwen2", "starcoder", "falcon", "deepseek-coder", "deepseek-v3", "chatglm-bpe", "tekken", "gpt-4o", "bloom", "viking",
         "seed-coder", "command-r", "smollm", "olmo", "stablelm2", "refact", "mpt", "bailingmoe2", "exaone-moe", "chameleon", "falcon3", "kormo",
         "trillion", "jina-v2-code", "solar-open"]


@pytest.mark.parametrize("pre", NAMES)
def test_bpe_pretokenizer_matches_reference(setup, pre):
    assert pre in PRE_TOKENIZERS
    scratch, env, base, toks, merges, text_path = setup
    model = with_vocab(base, os.path.join(scratch, f"m-{pre}.gguf"), toks, merges, pre)
    out = os.path.join(scratch, f"{pre}.logits")
    n_ctx = 32
    p = subprocess.run([ref_build.exe("llama-perplexity"), "-m", model, "-f", text_path, "-c", str(n_ctx), "-b", str(n_ctx), "--chunks", "1000", "-ngl", "0",
                        "-t", "4", "--save-all-logits", out], env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and os.path.exists(out), (p.stderr or p.stdout)[-800:]
    b = open(out, "rb").read()
    nc, nv, nch = np.frombuffer(b, np.int32, 3, 8)
    ref_ids = np.frombuffer(b, np.int32, nch * nc, 20).tolist()
    mine = Tokenizer(open_gguf(model).metadata).encode(TEXT, add_bos=False)
    assert len(mine) >= len(ref_ids) > 200
    assert mine[:len(ref_ids)] == ref_ids, next((i, mine[i - 3:i + 3], ref_ids[i - 3:i + 3]) for i in range(len(ref_ids)) if mine[i] != ref_ids[i])


def _ref_tokens(setup, model, name, text_path=None):
    scratch, env = setup[0], setup[1]
    text_path = text_path or setup[5]
    out = os.path.join(scratch, f"{name}.logits")
    p = subprocess.run([ref_build.exe("llama-perplexity"), "-m", model, "-f", text_path, "-c", "32", "-b", "32", "--chunks", "1000", "-ngl", "0", "-t", "4",
                        "--save-all-logits", out], env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and os.path.exists(out), (p.stderr or p.stdout)[-800:]
    b = open(out, "rb").read()
    nc, nv, nch = np.frombuffer(b, np.int32, 3, 8)
    return np.frombuffer(b, np.int32, nch * nc, 20).tolist()


def _model_with(setup, name, meta: dict, n_vocab: int):
    scratch = setup[0]
    base = write_synthetic_gguf(os.path.join(scratch, f"base-{name}.gguf"), preset("tiny", vocab_size=n_vocab), "F16", seed=1, vocab=False)
    f = open_gguf(base)
    dst = os.path.join(scratch, f"m-{name}.gguf")
    w = GGUFWriter(dst)
    for k, v in f.metadata.items():
        if not k.startswith("tokenizer."):
            w.add(k, v, f.files[0].metadata_types[k]) if not isinstance(v, (list, np.ndarray)) else w.add(k, v)
    for k, (v, kw) in meta.items():
        w.add(k, v, **kw)
    for n, t in f.tensors.items():
        w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
    w.write()
    return dst


def test_sentencepiece_vocab_matches_reference(setup):
    """The SPM tokenizer (score-driven bigram merges, byte fallback, space prefix) on the synthetic vocabulary every test model carries."""
    scratch = setup[0]
    model = write_synthetic_gguf(os.path.join(scratch, "spm.gguf"), preset("tiny"), "F16", seed=1)
    ref_ids = _ref_tokens(setup, model, "spm")
    tok = Tokenizer(open_gguf(model).metadata)
    mine = tok.encode(TEXT)                      # add_bos as the file says: the reference tool does the same
    assert len(ref_ids) > 200 and mine[:len(ref_ids)] == ref_ids


def test_wordpiece_vocab_matches_reference(setup):
    """WPM: NFD + lower-casing, punctuation / CJK as words of their own, greedy longest match, whole-word UNK, [CLS] ... [SEP]."""
    words = sorted({w.lower().strip(".,:!()'") for w in TEXT.split() if w.isascii() and w.strip(".,:!()'").isalpha()})
    pieces = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + ["▁" + w for w in words[:40]] + ["▁" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"] + \
        list("abcdefghijklmnopqrstuvwxyz0123456789") + ["▁" + c for c in ",.:!()'$<>[]{}_-@=?"] + ["▁你", "▁好", "▁世", "▁界", "ing", "ed", "s", "er"]
    pieces = list(dict.fromkeys(pieces))
    types = [int(TokenType.CONTROL)] * 5 + [int(TokenType.NORMAL)] * (len(pieces) - 5)
    meta = {"tokenizer.ggml.model": ("bert", {}), "tokenizer.ggml.tokens": (pieces, {"etype": ValueType.STRING}),
            "tokenizer.ggml.token_type": (np.asarray(types, np.int32), {"etype": ValueType.INT32}),
            "tokenizer.ggml.unknown_token_id": (1, {"vtype": ValueType.UINT32}), "tokenizer.ggml.cls_token_id": (2, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.bos_token_id": (2, {"vtype": ValueType.UINT32}), "tokenizer.ggml.seperator_token_id": (3, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.eos_token_id": (3, {"vtype": ValueType.UINT32}), "tokenizer.ggml.padding_token_id": (0, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.mask_token_id": (4, {"vtype": ValueType.UINT32})}
    model = _model_with(setup, "wpm", meta, len(pieces))
    ref_ids = _ref_tokens(setup, model, "wpm")
    mine = Tokenizer(open_gguf(model).metadata).encode(TEXT)
    assert len(ref_ids) > 200 and mine[:len(ref_ids)] == ref_ids


def test_unigram_vocab_matches_reference(setup):
    """UGM (t5): Viterbi over token scores, unknown characters, space prefix / collapsing."""
    words = collections.Counter(w for w in TEXT.split() if w.isascii())
    pieces = ["<pad>", "</s>", "<unk>"] + ["▁" + w for w, _ in words.most_common(30)] + ["▁"] + ["▁" + c for c in "abcdefghijklmnopqrstuvwxyz"] + \
        list("abcdefghijklmnopqrstuvwxyz0123456789.,:!'()_-") + ["ing", "er", "ed", "es", "th", "he", "in", "an", "on", "re", "st"]
    pieces = list(dict.fromkeys(pieces))
    rng = np.random.default_rng(0)
    scores = np.concatenate([[0.0, 0.0, 0.0], -(2.0 + 8.0 * rng.random(len(pieces) - 3))]).astype(np.float32)
    types = [int(TokenType.CONTROL), int(TokenType.CONTROL), int(TokenType.UNKNOWN)] + [int(TokenType.NORMAL)] * (len(pieces) - 3)
    meta = {"tokenizer.ggml.model": ("t5", {}), "tokenizer.ggml.tokens": (pieces, {"etype": ValueType.STRING}),
            "tokenizer.ggml.scores": (scores, {"etype": ValueType.FLOAT32}),
            "tokenizer.ggml.token_type": (np.asarray(types, np.int32), {"etype": ValueType.INT32}),
            "tokenizer.ggml.unknown_token_id": (2, {"vtype": ValueType.UINT32}), "tokenizer.ggml.eos_token_id": (1, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.padding_token_id": (0, {"vtype": ValueType.UINT32}), "tokenizer.ggml.add_bos_token": (False, {}),
            "tokenizer.ggml.add_eos_token": (False, {}), "tokenizer.ggml.add_space_prefix": (True, {}),     # the perplexity tool refuses add_eos
            "tokenizer.ggml.remove_extra_whitespaces": (True, {})}
    model = _model_with(setup, "ugm", meta, len(pieces))
    ref_ids = _ref_tokens(setup, model, "ugm")
    mine = Tokenizer(open_gguf(model).metadata).encode(TEXT)
    assert len(ref_ids) > 200 and mine[:len(ref_ids)] == ref_ids, next((i, mine[max(0, i - 3):i + 3], ref_ids[max(0, i - 3):i + 3]) for i in range(len(ref_ids)) if mine[i] != ref_ids[i])


def test_unigram_with_precompiled_charsmap_matches_reference(setup):
    """A sentencepiece-trained unigram model with its `precompiled_charsmap` in the file: the reference normalises through the double-array
    trie (llama-vocab.cpp:1057-1126) and so must `CharsMap`; the text is full of compatibility characters and has a user-defined symbol."""
    spm = pytest.importorskip("sentencepiece")
    from sentencepiece import sentencepiece_model_pb2 as pb
    scratch = setup[0]
    corpus = os.path.join(scratch, "ugm-corpus.txt")
    open(corpus, "w").write(TEXT.replace("你好世界 こんにちは ", ""))
    prefix = os.path.join(scratch, "ugmcm")
    spm.SentencePieceTrainer.train(input=corpus, model_prefix=prefix, vocab_size=150, model_type="unigram", character_coverage=1.0, bos_id=-1, eos_id=1,
                                   unk_id=2, pad_id=0, hard_vocab_limit=False, minloglevel=2, user_defined_symbols=["<ｘ>"])
    sp = spm.SentencePieceProcessor(model_file=prefix + ".model")
    proto = pb.ModelProto()
    proto.ParseFromString(open(prefix + ".model", "rb").read())
    n = sp.get_piece_size()
    types = [int(TokenType.CONTROL) if sp.is_control(i) else int(TokenType.UNKNOWN) if sp.is_unknown(i) else
             int(TokenType.USER_DEFINED) if proto.pieces[i].type == 4 else int(TokenType.NORMAL) for i in range(n)]
    meta = {"tokenizer.ggml.model": ("t5", {}), "tokenizer.ggml.tokens": ([sp.id_to_piece(i) for i in range(n)], {"etype": ValueType.STRING}),
            "tokenizer.ggml.scores": (np.asarray([sp.get_score(i) for i in range(n)], np.float32), {"etype": ValueType.FLOAT32}),
            "tokenizer.ggml.token_type": (np.asarray(types, np.int32), {"etype": ValueType.INT32}),
            "tokenizer.ggml.unknown_token_id": (2, {"vtype": ValueType.UINT32}), "tokenizer.ggml.eos_token_id": (1, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.padding_token_id": (0, {"vtype": ValueType.UINT32}), "tokenizer.ggml.add_bos_token": (False, {}),
            "tokenizer.ggml.add_eos_token": (False, {}), "tokenizer.ggml.add_space_prefix": (True, {}), "tokenizer.ggml.remove_extra_whitespaces": (True, {}),
            "tokenizer.ggml.precompiled_charsmap": (np.frombuffer(proto.normalizer_spec.precompiled_charsmap, np.uint8), {"etype": ValueType.UINT8})}
    model = _model_with(setup, "ugmcm", meta, n)
    text = ("Ｈｅｌｌｏ ｗｏｒｌｄ ﬁne ① ２３ Ⅷ tests <ｘ> we'll see\u3000Two\u200bspaces ｶﾞ ㌔ CamelCaseWords e-mail@example.com ... end\n") * 24
    text_path = os.path.join(scratch, "ugmcm-text.txt")
    open(text_path, "w", encoding="utf-8").write(text)
    ref_ids = _ref_tokens(setup, model, "ugmcm", text_path)
    tok = Tokenizer(open_gguf(model).metadata)
    assert tok.codec.charsmap is not None
    mine = tok.encode(text)
    assert len(ref_ids) > 200 and mine[:len(ref_ids)] == ref_ids, next((i, mine[max(0, i - 3):i + 3], ref_ids[max(0, i - 3):i + 3]) for i in range(len(ref_ids)) if mine[i] != ref_ids[i])


def test_rwkv_vocab_matches_reference(setup):
    """RWKV: greedy longest match over raw bytes on an escaped vocabulary."""
    words = collections.Counter(w for w in TEXT.split() if w.isascii() and w.isalpha())
    raw = [bytes([b]) for b in range(256)] + [(" " + w).encode() for w, _ in words.most_common(40)] + [b"\n\n", b"  ", "你好".encode(), b"ing", b"'s"]
    raw = list(dict.fromkeys(raw))

    def esc(b: bytes) -> str:
        return "".join(chr(c) if (32 <= c < 127 and c not in (0x5C, 0x27, 0x22)) else f"\\x{c:02x}" for c in b)
    pieces = ["<s>"] + [esc(b) for b in raw]
    types = [int(TokenType.CONTROL)] + [int(TokenType.NORMAL)] * len(raw)
    meta = {"tokenizer.ggml.model": ("rwkv", {}), "tokenizer.ggml.tokens": (pieces, {"etype": ValueType.STRING}),
            "tokenizer.ggml.token_type": (np.asarray(types, np.int32), {"etype": ValueType.INT32}),
            "tokenizer.ggml.bos_token_id": (0, {"vtype": ValueType.UINT32}), "tokenizer.ggml.eos_token_id": (0, {"vtype": ValueType.UINT32}),
            "tokenizer.ggml.unknown_token_id": (0, {"vtype": ValueType.UINT32}), "tokenizer.ggml.add_bos_token": (False, {})}
    model = _model_with(setup, "rwkv", meta, len(pieces))
    ref_ids = _ref_tokens(setup, model, "rwkv")
    mine = Tokenizer(open_gguf(model).metadata).encode(TEXT, add_bos=False)
    assert len(ref_ids) > 200 and mine[:len(ref_ids)] == ref_ids, next((i, mine[max(0, i - 3):i + 3], ref_ids[max(0, i - 3):i + 3]) for i in range(len(ref_ids)) if mine[i] != ref_ids[i])
