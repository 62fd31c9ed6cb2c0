"""Grammar (GBNF + JSON schema), the extra samplers and speculative decoding - host-side logic on the tiny CPU model."""
import pytest
import torch

from distributed_llm_pipeline_b200.serving.grammar import Grammar, GrammarConstraint, GrammarError, json_schema_to_gbnf
from distributed_llm_pipeline_b200.serving.sampler import Sampler, SamplingParams
from distributed_llm_pipeline_b200.serving.speculative import ModelDrafter, NgramDrafter, SpecStats, generate_speculative


def test_gbnf_parser_and_matcher():
    g = Grammar('root ::= "yes" | "no"')
    assert g.matches("yes") and g.matches("no") and not g.matches("maybe") and not g.matches("ye")
    g = Grammar('root ::= item+\nitem ::= "- " [a-z]+ "\\n"  # a list\n')
    assert g.matches("- abc\n- de\n") and not g.matches("- Abc\n") and not g.matches("")
    g = Grammar('root ::= [0-9]{2,4} "-" ("a"|"b")? [^x]* .')
    assert g.matches("12-aqqq!") and g.matches("1234-z") and not g.matches("1-a") and not g.matches("12345-a")
    g = Grammar('root ::= "ab" "c"*')
    g.accept("a")
    assert not g.complete
    g.accept("b")
    assert g.complete and g.accepts("cc") and not g.accepts("d")
    with pytest.raises(GrammarError):
        Grammar('root ::= missing')
    with pytest.raises(GrammarError):
        Grammar('start ::= "a"')


def test_json_schema_grammar():
    j = Grammar(json_schema_to_gbnf(None))
    assert j.matches('{"a": [1, 2.5, true, null], "b": {"c": "x\\n"}}') and not j.matches('{"a": }')
    sc = {"type": "object", "properties": {"name": {"type": "string"}, "age": {"type": "integer"},
                                           "tags": {"type": "array", "items": {"enum": ["x", "y"]}}}, "required": ["name", "age"]}
    j = Grammar(json_schema_to_gbnf(sc))
    assert j.matches('{"name": "bob", "age": 3}') and j.matches('{"name":"b","age":-1,"tags":["x","y"]}')
    assert not j.matches('{"age": 3}') and not j.matches('{"name":"b","age":1,"tags":["z"]}')


def test_grammar_constraint_masks_vocabulary_and_eos():
    vocab = ["<s>", "</s>", "a", "b", "ab", "ba", "c", None]
    c = GrammarConstraint('root ::= "ab" "c"?', vocab, eos_ids=(1,))
    assert c.mask() == [False, False, True, False, True, False, False, False]
    c.accept(4)
    assert c.mask() == [False, True, False, False, False, False, True, False]     # complete: EOS or the optional "c"
    c.accept(6)
    assert c.mask() == [False, True, False, False, False, False, False, False]


class _Tok:
    stop_ids = {1}

    def __init__(self, texts):
        self._t = texts

    def token_texts(self):
        return self._t


def test_sampler_grammar_and_extra_samplers():
    texts = ["<s>", "</s>", "a", "b", "ab", "ba", "c", "x"]
    logits = torch.tensor([0.0, 0.1, 0.2, 3.0, 0.5, 2.0, 0.3, 5.0])
    s = Sampler(SamplingParams(temperature=0.0, grammar='root ::= "ab" "c"?'), _Tok(texts))
    assert s.sample(logits) == 4          # "x", "b", "ba" are more likely but break the grammar
    assert s.sample(logits) == 6          # then "c" beats EOS
    assert s.sample(logits) == 1          # then only EOS is left
    # top-n-sigma keeps only logits within n std of the max; with n small only the arg-max survives
    s = Sampler(SamplingParams(temperature=1.0, top_k=0, top_p=1.0, min_p=0.0, top_n_sigma=0.5, seed=1))
    assert all(s.sample(logits) == 7 for _ in range(20))
    # XTC with probability 1 removes every candidate above the threshold except the least likely of them
    s = Sampler(SamplingParams(temperature=1.0, top_k=0, top_p=1.0, min_p=0.0, xtc_probability=1.0, xtc_threshold=0.05, seed=2))
    picks = {s.sample(logits) for _ in range(200)}
    assert 7 not in picks and 3 in picks          # p(x)=0.80 and p(b)=0.11 are above 0.05: x goes, b stays
    # mirostat v2 and dynamic temperature produce valid, seed-reproducible streams
    for kw in (dict(mirostat=2, mirostat_tau=3.0), dict(dynatemp_range=0.5)):
        a = Sampler(SamplingParams(temperature=0.9, seed=7, **kw))
        b = Sampler(SamplingParams(temperature=0.9, seed=7, **kw))
        sa, sb = [a.sample(logits * 0.3) for _ in range(30)], [b.sample(logits * 0.3) for _ in range(30)]
        assert sa == sb and all(0 <= t < 8 for t in sa) and len(set(sa)) > 1
    assert SamplingParams(mirostat=2).host_only and not SamplingParams().host_only


def _engine(path):
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    return LlamaEngine(load_weights(path, "cpu"), n_ctx=256, max_seqs=3, page_size=16, max_batch_tokens=64)


def test_speculative_decoding_equals_greedy(tmp_models):
    eng = _engine(tmp_models[("tiny", "F16")])
    prompt = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3, 250, 77] * 2
    base = eng.generate(prompt, 40)
    st = SpecStats()
    assert generate_speculative(eng, prompt, 40, NgramDrafter(), 8, stats=st) == base
    assert st.n_generated == 40 and st.n_target_passes <= 41
    st = SpecStats()
    draft = _engine(tmp_models[("tiny", "F16")])                      # the model drafting for itself: everything is accepted
    assert generate_speculative(eng, prompt, 40, ModelDrafter(draft), 5, stats=st) == base
    assert st.acceptance > 0.9 and st.tokens_per_pass > 3.5
    assert not eng.seqs and not draft.seqs                            # scratch sequences were released
    st = SpecStats()
    q = _engine(tmp_models[("tiny", "Q8_0")])                         # a different (quantised) draft: still exact output
    assert generate_speculative(eng, prompt, 24, ModelDrafter(q), 4, stats=st) == base[:24]


def test_embed_pooling(tmp_models):
    eng = _engine(tmp_models[("tiny", "F16")])
    ids = [1, 5, 9, 200, 33]
    allv = eng.embed(ids, pooling="none", normalize=False)
    assert allv.shape == (5, eng.cfg.n_embd)
    assert torch.allclose(eng.embed(ids, "mean", False), allv.mean(0), atol=1e-5)
    assert torch.allclose(eng.embed(ids, "last", False), allv[-1], atol=1e-5)
    assert abs(float(eng.embed(ids).norm()) - 1.0) < 1e-4
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel, rms_norm
    ref = ReferenceModel.from_gguf(tmp_models[("tiny", "F16")])
    h = ref.forward(ids, ref.new_cache(16), return_hidden=True)
    want = rms_norm(h, ref.w["output_norm.weight"], ref.cfg.rms_eps)
    assert torch.allclose(allv, want, atol=2e-3, rtol=2e-3)


def test_dry_sampler_penalises_repeated_spans():
    from distributed_llm_pipeline_b200.serving.sampler import apply_dry
    p = SamplingParams(dry_multiplier=1.0, dry_base=2.0, dry_allowed_length=2)
    hist = [5, 6, 7, 8, 1, 5, 6, 7]                 # "... 5 6 7" seen before, followed by 8
    logits = torch.zeros(10)
    out = apply_dry(logits, hist, p)
    assert out[8] == -2.0 and (out[torch.arange(10) != 8] == 0).all()     # match length 3: 1.0 * 2^(3-2)
    assert torch.equal(apply_dry(logits, [1, 2, 3, 4], p), logits)        # no repetition: untouched
    pb = SamplingParams(dry_multiplier=1.0, dry_base=2.0, dry_allowed_length=2, dry_breakers=(6,))
    assert torch.equal(apply_dry(logits, hist, pb), logits)               # the breaker cuts the match to length 1
    s = Sampler(SamplingParams(temperature=0.0, dry_multiplier=5.0))
    s.history = list(hist)
    lg = torch.zeros(10)
    lg[8], lg[3] = 1.0, 0.5
    assert s.sample(lg) == 3


def test_eval_callback_and_memory_breakdown(tmp_models):
    eng = _engine(tmp_models[("tiny", "F16")])
    seen = []
    eng.eval_callback = lambda name, il, t: seen.append((name, il, tuple(t.shape)))
    sid = eng.alloc_seq()
    eng.prefill(sid, [1, 2, 3])
    eng.decode([sid], [4], sample=False)
    assert seen == [("l_out", 0, (3, 256)), ("l_out", 1, (3, 256)), ("l_out", 0, (1, 256)), ("l_out", 1, (1, 256))]
    mb = eng.memory_breakdown()
    assert mb["model"] > 0 and mb["kv_cache"] > 0 and mb["compute"] > 0


def test_mirostat_v1_and_jinja_chat_template():
    logits = torch.tensor([0.0, 0.1, 0.2, 3.0, 0.5, 2.0, 0.3, 5.0]) * 0.5
    a = Sampler(SamplingParams(temperature=1.0, mirostat=1, mirostat_tau=2.0, seed=3))
    b = Sampler(SamplingParams(temperature=1.0, mirostat=1, mirostat_tau=2.0, seed=3))
    sa, sb = [a.sample(logits) for _ in range(40)], [b.sample(logits) for _ in range(40)]
    assert sa == sb and len(set(sa)) > 1 and all(0 <= t < 8 for t in sa) and a.mu != 4.0
    from distributed_llm_pipeline_b200.serving.orchestrator import apply_chat_template
    tmpl = ("{{ bos_token }}{% for m in messages %}<|{{ m['role'] }}|>{{ m['content'] }}<|end|>{% endfor %}"
            "{% if add_generation_prompt %}<|assistant|>{% endif %}")
    msgs = [{"role": "system", "content": "S"}, {"role": "user", "content": "U"}]
    assert apply_chat_template({"tokenizer.chat_template": tmpl, "bos_token": "<s>"}, msgs) == "<s><|system|>S<|end|><|user|>U<|end|><|assistant|>"
    # a template that raises falls back to the marker-based formats
    bad = "{% for m in messages %}{{ raise_exception('no') }}{% endfor %}<|im_start|>"
    assert apply_chat_template({"tokenizer.chat_template": bad}, msgs).endswith("<|im_start|>assistant\n")


def test_infill_sampler_rules():
    """`llama_sampler_init_infill` behaviour on hand-made distributions."""
    import math
    from distributed_llm_pipeline_b200.serving.sampler import apply_infill
    pieces = ["<eos>", "foo", "foobar", "fo", "zed", "qux", "<eot>", "x"]
    eog, eot = {0, 6}, 6

    def lp(probs):
        return torch.log(torch.tensor(probs))

    def alive(out):
        return {i: math.exp(v) for i, v in enumerate(out.tolist()) if v != float("-inf")}

    # 1. end-of-generation mass matters (3 * p_eog * n > p_txt): only EOG tokens survive, renormalised
    out = alive(apply_infill(lp([0.05, 0.3, 0.2, 0.1, 0.2, 0.1, 0.05, 0.0001]), pieces, eog, eot))
    assert set(out) == {0, 6} and abs(out[0] - 0.5) < 1e-4
    # 2. "fo" / "foo" / "foobar" are one continuation: merged into the likeliest ("foo"), which then clears the thresholds
    out = alive(apply_infill(lp([1e-9, 0.35, 0.25, 0.15, 0.15, 0.1 - 2e-9, 1e-9, 1e-9]), pieces, eog, eot))
    text = {i: v for i, v in out.items() if i not in eog}
    assert set(text) == {1} and sum(out.values()) == pytest.approx(1.0, abs=1e-4) and text[1] > 0.99
    # 3. nothing confident: every text token is below 0.2 after merging -> the end-of-turn token is forced
    flat = apply_infill(lp([1e-9, 0.15, 1e-9, 1e-9, 0.15, 0.15, 1e-9, 0.15] + [0.1] * 4), pieces + ["a1", "b2", "c3", "d4"], eog, eot)
    assert alive(flat) == {6: 1.0}
    # through the Sampler: a request flagged infill takes the host path and ends when unsure
    from distributed_llm_pipeline_b200.serving.sampler import Sampler, SamplingParams

    class Tok:
        eot_id, eos_id, stop_ids = 6, 0, {0, 6}

        def token_texts(self):
            return pieces
    sp = SamplingParams(temperature=0.0, infill=True)
    assert sp.host_only
    s = Sampler(sp, Tok())
    assert s.sample(lp([0.05, 0.3, 0.2, 0.1, 0.2, 0.1, 0.05, 0.0001])) in (0, 6)
    assert s.sample(lp([1e-9, 0.35, 0.25, 0.15, 0.15, 0.1 - 2e-9, 1e-9, 1e-9])) == 1


def test_lazy_grammar_is_armed_by_a_trigger():
    """grammar_lazy + trigger words / patterns: free text until the trigger shows up in the output, the grammar (which starts at the
    trigger) from then on (`llama_sampler_init_grammar_lazy_patterns`)."""
    from distributed_llm_pipeline_b200.serving.sampler import Sampler, SamplingParams
    pieces = ["<eos>", "hello ", "world ", "<call>", "{", "}", '"a"', ":", "1", "x", "<ca", "ll>"]

    class Tok:
        eot_id, eos_id, stop_ids = None, 0, {0}

        def token_texts(self):
            return [None] + pieces[1:]
    gr = 'root ::= "<call>" "{" "\\"a\\"" ":" "1" "}"'

    def run(params, script):
        s = Sampler(params, Tok())
        out = []
        for favourite in script:          # logits that prefer `favourite`, then everything else equally
            row = torch.zeros(len(pieces))
            row[favourite] = 5.0
            out.append(s.sample(row))
        return out
    lazy = SamplingParams(temperature=0.0, grammar=gr, grammar_lazy=True, grammar_trigger_words=("<call>",))
    # free text first (a plain grammar would have refused "hello "), then the call is forced token by token whatever the logits prefer
    got = run(lazy, [1, 2, 3, 9, 9, 9, 9, 9, 9])
    assert got[:3] == [1, 2, 3] and "".join(pieces[t] for t in got[3:8]) == '{"a":1}' and got[8] == 0       # ... and EOS once complete
    # a trigger split over two tokens still arms the grammar
    got = run(lazy, [1, 10, 11, 9, 9, 9, 9, 9])
    assert "".join(pieces[t] for t in got[1:8]) == '<call>{"a":1}'
    # regex trigger
    pat = SamplingParams(temperature=0.0, grammar=gr, grammar_lazy=True, grammar_trigger_patterns=(r"<c[a-z]+>",))
    assert "".join(pieces[t] for t in run(pat, [2, 3, 9, 9, 9, 9, 9])[1:]) == '<call>{"a":1}'
    # not lazy: constrained from the first token
    eager = SamplingParams(temperature=0.0, grammar=gr)
    assert run(eager, [1])[0] == 3
