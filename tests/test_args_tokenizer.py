"""Flag table (CLI aliases + environment fall-backs; reference: tests/test-arg-parser.cpp) and tokenizer corner cases
(reference: tests/test-tokenizer-*.cpp): byte fallback, specials, streaming detokenisation of split UTF-8."""
import pytest

from distributed_llm_pipeline_b200.utils.args import COMMON, GENERATION, SAMPLING, SPECULATIVE, build_parser, lora_from_args, sampling_from_args
from distributed_llm_pipeline_b200.utils.tokenizer import StreamDecoder, Tokenizer


def test_flag_aliases_defaults_and_env(monkeypatch):
    ap = build_parser("t", [COMMON, SAMPLING, GENERATION, SPECULATIVE])
    a = ap.parse_args(["-m", "x.gguf", "-c", "512", "--temp", "0", "-n", "7", "-ngl", "99", "-sm", "row", "--top-k", "5"])
    assert (a.model, a.ctx_size, a.temp, a.predict, a.gpu_layers, a.split_mode, a.top_k) == ("x.gguf", 512, 0.0, 7, 99, "row", 5)
    assert a.parallel == 1 and a.batch_size == 2048 and a.ubatch_size == 512 and a.draft_max == 8 and not a.lookup
    assert ap.parse_args(["--n-predict", "3"]).predict == 3 and ap.parse_args(["--n-gpu-layers", "1"]).gpu_layers == 1
    monkeypatch.setenv("DLP_ARG_CTX_SIZE", "4096")
    monkeypatch.setenv("DLP_ARG_VERBOSE", "1")
    monkeypatch.setenv("DLP_ARG_TEMP", "0.25")
    b = build_parser("t", [COMMON, SAMPLING]).parse_args([])
    assert b.ctx_size == 4096 and b.verbose is True and b.temp == 0.25
    assert build_parser("t", [COMMON, SAMPLING]).parse_args(["-c", "64"]).ctx_size == 64        # the command line wins over the env
    with pytest.raises(SystemExit):
        ap.parse_args(["-sm", "diagonal"])
    with pytest.raises(SystemExit):
        ap.parse_args(["--no-such-flag"])
    sp = sampling_from_args(ap.parse_args(["--temp", "0.7", "--mirostat", "2", "--grammar", 'root ::= "a"', "-j", '{"type": "boolean"}']))
    assert sp.temperature == 0.7 and sp.mirostat == 2 and sp.grammar == 'root ::= "a"' and sp.json_schema == {"type": "boolean"} and sp.host_only
    assert lora_from_args(ap.parse_args([])) == []


def test_tokenizer_byte_fallback_specials_and_streaming(tmp_models):
    tok = Tokenizer.from_gguf(tmp_models[("tiny", "F16")])
    text = "héllo ☕ 日本"
    ids = tok.encode(text)
    assert ids[0] == tok.bos_id and tok.decode(ids[1:]).strip() == text          # characters outside the vocabulary go through <0xNN>
    assert tok.encode("", add_bos=False) == []
    assert tok.encode("a</s>b", add_bos=False).count(tok.eos_id) == 1            # special tokens are parsed ...
    assert tok.eos_id not in tok.encode("a</s>b", add_bos=False, parse_special=False)   # ... unless asked not to
    dec = StreamDecoder(tok)
    pieces = [dec.push(t) for t in ids[1:]]
    assert all("�" not in p for p in pieces)                                # no replacement characters from split UTF-8
    assert ("".join(pieces) + dec.flush()).strip() == text
    texts = tok.token_texts()
    assert texts[tok.bos_id] is None and len(texts) == tok.n_vocab
    assert tok.fim_ids() is None
