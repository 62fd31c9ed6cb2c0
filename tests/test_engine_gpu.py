"""End-to-end: the CUDA engine (fused kernels, paged KV, CUDA-graph decode) against the fp32 oracle."""
import pytest
import torch

from distributed_llm_pipeline_b200.models.llama import LlamaEngine
from distributed_llm_pipeline_b200.models.reference import ReferenceModel, nmse
from distributed_llm_pipeline_b200.models.weights import load_weights

pytestmark = pytest.mark.gpu
PROMPT = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3, 250, 77]


@pytest.mark.parametrize("key", [("tiny", "F16"), ("tiny", "Q8_0"), ("tiny", "Q4_K_M"), ("tiny128", "F16"), ("tiny-moe", "F16"),
                                 ("tiny-qwen2", "F16"), ("tiny-qwen2", "Q8_0"), ("tiny-qwen3", "F16"), ("tiny-phi3", "F16"),
                                 ("tiny-phi3", "Q8_0"), ("tiny-qwen3moe", "F16"), ("tiny-qwen2moe", "F16"), ("tiny-granite", "F16"), ("tiny-granite", "Q8_0"),
                                 ("tiny-internlm2", "F16"), ("tiny-exaone", "F16"), ("tiny-seed-oss", "F16"), ("tiny-minicpm", "F16"),
                                 ("tiny-granitemoe", "F16"), ("tiny-smollm3", "F16"), ("tiny-smollm3", "Q8_0"),
                                 ("tiny-starcoder2", "F16"), ("tiny-starcoder2", "Q8_0"), ("tiny-falcon", "F16"), ("tiny-olmo", "F16"),
                                 ("tiny-orion", "F16"), ("tiny-command-r", "F16"), ("tiny-gptneox", "F16"), ("tiny-gptneox", "Q8_0"), ("tiny-stablelm", "F16"),
                                 ("tiny-nemotron", "F16"), ("tiny-arcee", "F16"), ("tiny-gpt2", "F16"), ("tiny-starcoder", "F16"),
                                 ("tiny-codeshell", "F16"), ("tiny-qwen", "F16"), ("tiny-olmo2", "F16"), ("tiny-glm4", "F16"), ("tiny-pangu", "F16")])
def test_engine_gpu_matches_oracle(tmp_models, key):
    path = tmp_models[key]
    ref = ReferenceModel.from_gguf(path)
    toks, ref_logits = ref.generate_greedy(PROMPT, 6)
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=4, page_size=16, max_batch_tokens=64)
    sid = eng.alloc_seq()
    assert sid == 0
    lg = eng.prefill(sid, PROMPT)                     # GEMM path (T = 13 > 4): tcgen05 + flash prefill
    assert nmse(ref_logits[0], lg) < 1e-3
    for i, t in enumerate(toks[:-1]):                 # mat-vec path, CUDA graph after the first call
        lg = eng.decode([sid], [t], sample=False)[0]
        assert nmse(ref_logits[i + 1], lg) < 1e-3, i
    assert any(k[:2] == (1, False) for k in eng._graphs)   # every family decodes inside a CUDA graph (MoE routing stays on the device)


def test_graph_decode_feeds_back_device_tokens(tmp_models):
    path = tmp_models[("tiny", "F16")]
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)
    a = eng.generate(PROMPT, 12)
    eng2 = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)
    sid = eng2.alloc_seq()
    lg = eng2.prefill(sid, PROMPT)
    out = [int(lg.argmax())]
    for _ in range(11):                                # eager, host-fed tokens, no graph
        out.append(int(eng2.decode([sid], [out[-1]], sample=False, use_graph=False)[0].argmax()))
    assert a == out


def _eager_greedy(path, prompt, n):
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)
    sid = eng.alloc_seq()
    out = [int(eng.prefill(sid, prompt).argmax())]
    for _ in range(n - 1):
        out.append(int(eng.decode([sid], [out[-1]], sample=False, use_graph=False)[0].argmax()))
    return out


@pytest.mark.parametrize("extra", [[426] * 6, [929] * 5 + [3], [17, 250, 9, 9, 33]])
def test_graph_capture_keeps_the_pending_token(tmp_models, extra):
    """The graph warm-up samples into next_tok; the pending input token has to survive it.  The prompts are chosen so that the
    first generated tokens differ from each other (a dropped token then shows up as a shifted sequence)."""
    path = tmp_models[("tiny", "F16")]
    prompt = PROMPT + extra
    want = _eager_greedy(path, prompt, 12)
    assert len(set(want[:4])) > 1 or len(set(want)) > 1
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)
    assert eng.generate(prompt, 12) == want          # device-fed, graph captured on the first decode
    assert eng.generate(prompt, 12) == want          # replay of the cached graph


def test_device_sampler_is_per_request(tmp_models):
    """Sampling parameters live in per-sequence device records: no graph per seed, and a request's random stream does not depend
    on what the engine did before (same seed + prompt -> same tokens)."""
    path = tmp_models[("tiny", "F16")]
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)
    eng.sampling.update(temp=1.5, top_k=50, top_p=1.0, min_p=0.0, seed=11)
    a = eng.generate(PROMPT, 16)
    n_graphs = len(eng._graphs)
    eng.sampling.update(seed=12)
    b = eng.generate(PROMPT, 16)
    eng.sampling.update(seed=11)
    c = eng.generate(PROMPT, 16)
    assert a == c and a != b
    assert len(eng._graphs) == n_graphs              # new seeds did not capture new graphs
    # two sequences with different parameters share one batched step: the greedy one matches the greedy oracle
    greedy = _eager_greedy(path, PROMPT, 8)
    s0, s1 = eng.alloc_seq(), eng.alloc_seq()
    eng.set_seq_sampling(s0, temp=0.0)
    eng.set_seq_sampling(s1, temp=2.0, top_k=0, top_p=1.0, min_p=0.0, seed=5)
    t0 = int(eng.prefill(s0, PROMPT).argmax())
    eng.prefill(s1, PROMPT)
    assert t0 == greedy[0]
    toks = torch.tensor([t0, t0], dtype=torch.int32)
    got = [t0]
    for _ in range(7):
        toks = eng.decode([s0, s1], toks if len(got) == 1 else None, sample=True).clone()
        got.append(int(toks[0]))
    assert got == greedy


def test_batched_decode_matches_single(tmp_models):
    path = tmp_models[("tiny128", "F16")]
    ref = ReferenceModel.from_gguf(path)
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=8, page_size=16, max_batch_tokens=64)
    prompts = [[1, 2, 3], [9, 8, 7, 6, 5], [100, 200], [4], [5, 6], [7, 7, 7, 7]]
    sids = [eng.alloc_seq() for _ in prompts]
    for s, p in zip(sids, prompts):
        eng.prefill(s, p)
    nxt = [11, 12, 13, 14, 15, 16]
    lg = eng.decode(sids, nxt, sample=False).clone()     # B = 6 > 4 -> tcgen05 GEMM path with BT = 16
    for row, (p, t) in enumerate(zip(prompts, nxt)):
        cache = ref.new_cache(64)
        ref.forward(p, cache)
        ref.advance(cache, len(p))
        assert nmse(ref.forward([t], cache)[-1], lg[row]) < 1e-3


def test_speculative_verify_and_embed_on_gpu(tmp_models):
    """Draft-and-verify on the CUDA engine: the T = k+1 verify forward (GEMM path) must reproduce the mat-vec decode
    path's greedy tokens, except where the oracle's top-2 logits are closer than fp16 noise."""
    from distributed_llm_pipeline_b200.serving.speculative import ModelDrafter, SpecStats, generate_speculative
    path = tmp_models[("tiny128", "F16")]
    mk = lambda: LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=64)  # noqa: E731
    eng, draft = mk(), mk()
    prompt = PROMPT * 2
    base = eng.generate(prompt, 32)
    st = SpecStats()
    out = generate_speculative(eng, prompt, 32, ModelDrafter(draft), 6, stats=st)
    assert st.tokens_per_pass > 2.0 and len(out) == 32
    if out != base:
        i = next(k for k in range(32) if out[k] != base[k])
        ref = ReferenceModel.from_gguf(path)
        cache = ref.new_cache(128)
        lg = ref.forward(prompt + base[:i], cache)[-1]
        assert abs(float(lg[out[i]] - lg[base[i]])) < 2e-2, (i, out[i], base[i])
    ids = [1, 5, 9, 200, 33, 7, 8]
    e = eng.embed(ids, pooling="none", normalize=False)
    from distributed_llm_pipeline_b200.models.reference import rms_norm
    ref = ReferenceModel.from_gguf(path)
    h = ref.forward(ids, ref.new_cache(16), return_hidden=True)
    assert nmse(rms_norm(h, ref.w["output_norm.weight"], ref.cfg.rms_eps), e) < 1e-4


@pytest.mark.parametrize("key", [("tiny", "F16"), ("tiny128", "F16"), ("tiny", "Q8_0"), ("tiny-phi3", "F16")])
def test_engine_onepass_attention_matches_oracle(tmp_models, key):
    """Decode with the one-launch attention step (RoPE + KV write + attention + cluster combine) forced on from the first
    token (normally it takes over at 64 keys), plus the hand-over between the two variants in the middle of a sequence."""
    path = tmp_models[key]
    ref = ReferenceModel.from_gguf(path)
    toks, ref_logits = ref.generate_greedy(PROMPT, 8)
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=4, page_size=16, max_batch_tokens=64)
    assert eng.attn_onepass
    eng.attn_onepass_min = 0
    sid = eng.alloc_seq()
    eng.prefill(sid, PROMPT)
    for i, t in enumerate(toks[:-1]):
        if i == 4:
            eng.attn_onepass_min = 10 ** 6          # back to the two-kernel path: both read / write the same cache
        lg = eng.decode([sid], [t], sample=False)[0]
        assert nmse(ref_logits[i + 1], lg) < 1e-3, i
    assert any(k[:3] == (1, False, True) for k in eng._graphs) and any(k[:3] == (1, False, False) for k in eng._graphs)


@pytest.mark.parametrize("key", [("tiny-moe", "F16"), ("tiny-qwen2moe", "F16"), ("tiny-qwen3moe", "F16")])
@pytest.mark.parametrize("T", [1, 7, 64, 200])
def test_moe_engine_any_token_count_on_device(tmp_models, key, T):
    """Every MoE family, any token count: prefill through the grouped GEMM path and graph-captured decode, vs the fp32 oracle;
    no launch of the step visits the host (the decode graph exists for all families, the shared-expert one included)."""
    path = tmp_models[key]
    ref = ReferenceModel.from_gguf(path)
    g = torch.Generator().manual_seed(T)
    prompt = torch.randint(3, ref.cfg.vocab_size, (T,), generator=g).tolist()
    toks, ref_logits = ref.generate_greedy(prompt, 3)
    eng = LlamaEngine(load_weights(path, "cuda"), n_ctx=256, max_seqs=2, page_size=16, max_batch_tokens=256)
    sid = eng.alloc_seq()
    lg = eng.prefill(sid, prompt)
    assert nmse(ref_logits[0], lg) < 1e-3
    for i, t in enumerate(toks[:-1]):
        lg = eng.decode([sid], [t], sample=False)[0]
        assert nmse(ref_logits[i + 1], lg) < 1e-3, i
    assert any(k[:2] == (1, False) for k in eng._graphs)
