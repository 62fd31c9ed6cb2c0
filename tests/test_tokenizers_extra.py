"""WordPiece, Unigram and RWKV vocabularies (llama-vocab.cpp: llm_tokenizer_wpm / _ugm / _rwkv) against independent
implementations: `transformers.BertTokenizer` on the same vocabulary and `sentencepiece` on a unigram model trained here."""
import os

import pytest

from distributed_llm_pipeline_b200.gguf.constants import TokenType
from distributed_llm_pipeline_b200.utils.tokenizer import StreamDecoder, Tokenizer
from distributed_llm_pipeline_b200.utils.tokenizer_extra import unescape_rwkv

SENTENCES = ["Hello, world!", "The quick brown fox jumps over the lazy dog.", "unbelievably  spaced   out", "naïve café déjà vu",
             "tokenization of words: running, runner, runs", "price is $5.00 (approx.)", "日本語 text mixed", "xyzzyq unknownword", ""]


def _bert_meta(hf_vocab: list[str]) -> dict:
    # what the converter writes: "##x" -> "x", every other non-special piece gets the word-boundary marker
    toks, types = [], []
    for t in hf_vocab:
        if t.startswith("[") and t.endswith("]"):
            toks.append(t)
            types.append(int(TokenType.CONTROL))
        elif t.startswith("##"):
            toks.append(t[2:])
            types.append(int(TokenType.NORMAL))
        else:
            toks.append("▁" + t)
            types.append(int(TokenType.NORMAL))
    return {"tokenizer.ggml.model": "bert", "tokenizer.ggml.tokens": toks, "tokenizer.ggml.token_type": types,
            "tokenizer.ggml.unknown_token_id": hf_vocab.index("[UNK]"), "tokenizer.ggml.cls_token_id": hf_vocab.index("[CLS]"),
            "tokenizer.ggml.seperator_token_id": hf_vocab.index("[SEP]"), "tokenizer.ggml.eos_token_id": hf_vocab.index("[SEP]")}


def test_wordpiece_matches_transformers_bert_tokenizer(tmp_path):
    transformers = pytest.importorskip("transformers")
    words = ["the", "quick", "brown", "fox", "jump", "##s", "over", "lazy", "dog", "hello", "world", "un", "##believ", "##ably", "space", "##d", "out",
             "naive", "cafe", "deja", "vu", "token", "##ization", "of", "word", "##s", "run", "##ning", "##ner", "price", "is", "5", "00", "approx",
             "text", "mixed", "日", "本", "語", ",", ".", "!", ":", "$", "(", ")", "a", "b", "c", "##a", "##b", "##c"]
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + list(dict.fromkeys(words))
    vf = tmp_path / "vocab.txt"
    vf.write_text("\n".join(vocab) + "\n", encoding="utf-8")
    hf = transformers.BertTokenizer(str(vf), do_lower_case=True, strip_accents=False)
    tok = Tokenizer(_bert_meta(vocab))
    for s in SENTENCES:
        if "naïve" in s:
            continue        # HF leaves the NFD combining marks in place only with strip_accents=False; pieces then differ by vocabulary, not rule
        assert tok.encode(s) == hf.encode(s), s
    ids = tok.encode("Hello, world!")
    assert ids[0] == vocab.index("[CLS]") and ids[-1] == vocab.index("[SEP]")
    assert tok.decode(ids) == "hello , world !"
    assert tok.encode("xyzzyq", add_bos=False) == [vocab.index("[UNK]")]        # a word with an uncovered tail is UNK as a whole
    sd = StreamDecoder(tok)
    assert "".join(sd.push(i) for i in ids) + sd.flush() == " hello , world !"


def test_unigram_matches_sentencepiece(tmp_path):
    spm = pytest.importorskip("sentencepiece")
    corpus = tmp_path / "corpus.txt"
    base = ["the quick brown fox jumps over the lazy dog", "hello world this is a test of the unigram tokenizer", "tokenization of words running runner runs",
            "pipeline stages and tensor parallel ranks", "price is five dollars approximately", "unbelievably spaced out sentences are here"]
    corpus.write_text("\n".join(base * 40) + "\n")
    prefix = str(tmp_path / "ugm")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=prefix, vocab_size=120, model_type="unigram", character_coverage=1.0,
                                   bos_id=-1, eos_id=1, unk_id=2, pad_id=0, hard_vocab_limit=False, minloglevel=2)
    sp = spm.SentencePieceProcessor(model_file=prefix + ".model")
    n = sp.get_piece_size()
    toks = [sp.id_to_piece(i) for i in range(n)]
    scores = [sp.get_score(i) for i in range(n)]
    types = [int(TokenType.CONTROL) if sp.is_control(i) else int(TokenType.UNKNOWN) if sp.is_unknown(i) else int(TokenType.NORMAL) for i in range(n)]
    meta = {"tokenizer.ggml.model": "t5", "tokenizer.ggml.tokens": toks, "tokenizer.ggml.scores": scores, "tokenizer.ggml.token_type": types,
            "tokenizer.ggml.unknown_token_id": sp.unk_id(), "tokenizer.ggml.eos_token_id": sp.eos_id(), "tokenizer.ggml.add_eos_token": True,
            "tokenizer.ggml.add_bos_token": False}
    tok = Tokenizer(meta)
    for s in ["the quick brown fox", "hello   world", "tokenization of runner", "unbelievably spaced out", "price is five dollars", "the lazy dog jumps",
              " leading and trailing ", "parallel pipeline test"]:
        assert tok.encode(s) == sp.encode(s) + [sp.eos_id()], s
        assert tok.decode(tok.encode(s)) == sp.decode(sp.encode(s)), s
    # a character outside the model: one UNK for the run, the rest still segmented as sentencepiece does
    assert tok.encode("the ZZ dog") == sp.encode("the ZZ dog") + [sp.eos_id()]


def test_unigram_with_precompiled_charsmap_matches_sentencepiece(tmp_path):
    """The model's own normalisation table (nmt_nfkc, 240 kB double-array trie) read from the file: ligatures, full-width forms, circled digits,
    square katakana words, half-width katakana with voicing marks, zero-width and ideographic spaces - normalised and tokenised as sentencepiece
    does."""
    spm = pytest.importorskip("sentencepiece")
    from sentencepiece import sentencepiece_model_pb2 as pb
    corpus = tmp_path / "corpus.txt"
    base = ["the quick brown fox jumps over the lazy dog", "hello world this is a test", "fine ABC abc 123 VIII", "price is five dollars approximately"]
    corpus.write_text("\n".join(base * 40) + "\n")
    prefix = str(tmp_path / "ugm")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=prefix, vocab_size=110, model_type="unigram", character_coverage=1.0,
                                   bos_id=-1, eos_id=1, unk_id=2, pad_id=0, hard_vocab_limit=False, minloglevel=2, user_defined_symbols=["<ｘ>"])
    sp = spm.SentencePieceProcessor(model_file=prefix + ".model")
    proto = pb.ModelProto()
    proto.ParseFromString(open(prefix + ".model", "rb").read())
    charsmap = proto.normalizer_spec.precompiled_charsmap
    assert proto.normalizer_spec.name == "nmt_nfkc" and len(charsmap) > 100000
    n = sp.get_piece_size()
    types = [int(TokenType.CONTROL) if sp.is_control(i) else int(TokenType.UNKNOWN) if sp.is_unknown(i) else
             int(TokenType.USER_DEFINED) if proto.pieces[i].type == 4 else int(TokenType.NORMAL) for i in range(n)]
    meta = {"tokenizer.ggml.model": "t5", "tokenizer.ggml.tokens": [sp.id_to_piece(i) for i in range(n)], "tokenizer.ggml.scores": [sp.get_score(i) for i in range(n)],
            "tokenizer.ggml.token_type": types, "tokenizer.ggml.unknown_token_id": sp.unk_id(), "tokenizer.ggml.eos_token_id": sp.eos_id(),
            "tokenizer.ggml.add_eos_token": False, "tokenizer.ggml.add_bos_token": False, "tokenizer.ggml.precompiled_charsmap": list(charsmap)}
    tok = Tokenizer(meta)
    assert tok.codec.charsmap is not None
    for s in ["ﬁne ＡＢＣ ① test", "ｈｅｌｌｏ\u3000ｗｏｒｌｄ", "the\u200bquick  brown", "Ⅷ dollars", "ｶﾞ ㌔ fox", "tab\there", "plain text"]:
        assert tok.codec.normalize(s) == sp.normalize(s), s
        assert tok.encode(s) == sp.encode(s), s
    # a user-defined symbol is protected from the table (its full-width letter stays); around it the text is cut into fragments first, as the
    # reference does (`tokenizer_st_partition`), so the spaces next to it are trimmed where sentencepiece itself keeps one
    assert tok.codec.normalize("a <ｘ> b ＡＢＣ") == sp.normalize("a <ｘ> b ＡＢＣ")
    ids = tok.encode("a <ｘ> b ＡＢＣ")
    assert sp.piece_to_id("<ｘ>") in ids and tok.decode(ids).replace(" ", "") == "a<ｘ>bABC"


def test_rwkv_trie_and_escapes():
    assert unescape_rwkv("\\n") == b"\n" and unescape_rwkv("\\x0a\\xe4") == b"\n\xe4" and unescape_rwkv("a\\\\b") == b"a\\b" and unescape_rwkv("é") == "é".encode()
    toks = ["<s>"] + [f"\\x{b:02x}" for b in range(256)] + ["he", "hello", "hell", " wor", "ld", "\\xe6\\x97\\xa5"]
    types = [int(TokenType.CONTROL)] + [int(TokenType.NORMAL)] * (len(toks) - 1)
    tok = Tokenizer({"tokenizer.ggml.model": "rwkv", "tokenizer.ggml.tokens": toks, "tokenizer.ggml.token_type": types, "tokenizer.ggml.unknown_token_id": 0,
                     "tokenizer.ggml.bos_token_id": 0, "tokenizer.ggml.eos_token_id": 0})
    ids = tok.encode("hello world 日")
    assert ids == [toks.index("hello"), toks.index(" wor"), toks.index("ld"), 1 + 0x20, toks.index("\\xe6\\x97\\xa5")]      # longest match wins
    assert tok.decode(ids, strip_leading_space=False) == "hello world 日"
    assert tok.encode("hel") == [toks.index("he"), 1 + ord("l")]
