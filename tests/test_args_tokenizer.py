"""Flag table (CLI aliases + environment fall-backs; reference: tests/test-arg-parser.cpp) and tokenizer corner cases
(reference: tests/test-tokenizer-*.cpp): byte fallback, specials, streaming detokenisation of split UTF-8."""
import os

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


def test_command_lines_written_for_the_reference_parse(tmp_path, monkeypatch):
    """Flags of `llama.cpp/common/arg.cpp` that have no counterpart: placement / affinity settings are accepted with a note, features that do not
    exist are refused with the reason, metadata overrides (`--override-kv`, the rope / YaRN flags) reach the loader, and at most a handful of
    the reference's documented flags is still unknown to the CLI parser."""
    import re
    from distributed_llm_pipeline_b200.utils.args import COMMON, GENERATION, SAMPLING, SPECULATIVE, build_parser, parse_override_kv, sampling_from_args
    monkeypatch.delenv("DLP_OVERRIDE_KV", raising=False)
    notes = []
    ap = build_parser("t", [COMMON, SAMPLING, GENERATION, SPECULATIVE])
    import distributed_llm_pipeline_b200.utils.args as A
    monkeypatch.setattr(A.os, "environ", dict(os.environ))          # keep what finish_compat exports out of this process
    a = ap.parse_args(["-m", "x.gguf", "-ngl", "99", "-tb", "2", "--numa", "distribute", "--no-kv-offload", "--mmap", "-ctk", "f16", "--prio", "2", "--color",
                       "--rope-scaling", "yarn", "--rope-scale", "4", "--yarn-orig-ctx", "2048", "--override-kv", "llama.context_length=int:4096",
                       "-l", "15+1", "-l", "7-inf", "--repeat-last-n", "32", "--dry-multiplier", "0.8", "-r", "User:", "-sys", "be brief", "--samplers", "top_k;top_p"])
    ov = A.os.environ["DLP_OVERRIDE_KV"].split(";")
    assert "llama.context_length=int:4096" in ov and "{arch}.rope.scaling.type=str:yarn" in ov and "{arch}.rope.scaling.factor=float:4.0" in ov
    assert parse_override_kv("{arch}.rope.freq_base=float:5e5", "qwen2") == ("qwen2.rope.freq_base", 500000.0)
    assert parse_override_kv("general.name=str:a:b") == ("general.name", "a:b") and parse_override_kv("x=bool:true") == ("x", True)
    sp = sampling_from_args(a)
    assert sp.logit_bias == {15: 1.0, 7: float("-inf")} and sp.repeat_last_n == 32 and sp.dry_multiplier == 0.8 and a.reverse_prompt == ["User:"]
    for bad in (["--mmproj", "p.gguf"], ["-hf", "org/model"], ["-ctk", "q8_0"], ["--ssl-key-file", "k"]):
        with pytest.raises(SystemExit):
            ap.parse_args(["-m", "x.gguf"] + bad)
    ref = "/root/reference/llama.cpp/common/arg.cpp"
    if os.path.exists(ref):
        names = set(re.findall(r'"(--[a-z0-9][a-z0-9-]*)"', open(ref).read()))
        known = {s for act in ap._actions for s in act.option_strings}
        from distributed_llm_pipeline_b200.serving import orchestrator as orch_mod
        known |= set(re.findall(r'"(--[a-z0-9][a-z0-9-]*)"', open(orch_mod.__file__).read()))
        tools = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools")
        for f in os.listdir(tools):
            if f.endswith(".py"):
                known |= set(re.findall(r'"(--[a-z0-9][a-z0-9-]*)"', open(os.path.join(tools, f)).read()))
        # presets that only name a model to download, and flags of tools that have no counterpart here (tts, diffusion, retrieval chunking ...)
        rest = {n for n in names - known if not re.search(r"-default$|-spec$|^--(diffusion|tts|fim-|hf-|mmproj|embd-|chunk|lookup-cache|hellaswag|winogrande|multiple-choice|kl-divergence|cpu-.*draft|prio.*draft|poll.*draft|threads.*draft|.*-draft$)", n)}
        assert len(rest) < 75, sorted(rest)


def test_override_kv_reaches_the_loader(tmp_models, monkeypatch):
    from distributed_llm_pipeline_b200.gguf import open_gguf
    from distributed_llm_pipeline_b200.models.config import config_from_gguf
    monkeypatch.setenv("DLP_OVERRIDE_KV", "{arch}.rope.scaling.type=str:yarn;{arch}.rope.scaling.factor=float:4.0;{arch}.rope.scaling.original_context_length=int:128;"
                       "{arch}.rope.freq_base=float:50000")
    f = open_gguf(tmp_models[("tiny", "F16")])
    cfg = config_from_gguf(f.metadata, f.tensors)
    assert cfg.yarn_factor == 4.0 and cfg.yarn_orig_ctx == 128 and cfg.rope_theta == 50000.0
    monkeypatch.delenv("DLP_OVERRIDE_KV")
    cfg0 = config_from_gguf(open_gguf(tmp_models[("tiny", "F16")]).metadata)
    assert cfg0.yarn_factor != 4.0 and cfg0.rope_theta == 10000.0
