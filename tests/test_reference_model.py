"""CPU engine path (torch fallbacks over the paged-KV data layout) against the fp32 oracle."""
import os
import pytest
import torch

from distributed_llm_pipeline_b200.models.llama import LlamaEngine
from distributed_llm_pipeline_b200.models.reference import ReferenceModel, apply_rope, nmse, rope_inv_freq
from distributed_llm_pipeline_b200.models.config import preset
from distributed_llm_pipeline_b200.models.weights import load_weights

PROMPT = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3]


@pytest.mark.parametrize("key", [("tiny", "F16"), ("tiny", "Q8_0"), ("tiny", "Q4_K_M"), ("tiny128", "F16"), ("tiny-moe", "F16")])
def test_engine_cpu_matches_oracle_teacher_forced(tmp_models, key):
    path = tmp_models[key]
    ref = ReferenceModel.from_gguf(path)
    toks, ref_logits = ref.generate_greedy(PROMPT, 5)
    eng = LlamaEngine(load_weights(path, "cpu"), n_ctx=128, max_seqs=2, page_size=16, max_batch_tokens=8)
    sid = eng.alloc_seq()
    lg = eng.prefill(sid, PROMPT)  # chunked: 8 + 3 tokens
    assert nmse(ref_logits[0], lg) < 1e-5
    for i, t in enumerate(toks[:-1]):
        lg = eng.decode([sid], [t], sample=False)[0]
        assert nmse(ref_logits[i + 1], lg) < 1e-5
    assert eng.seq_len(sid) == len(PROMPT) + len(toks) - 1
    eng.free_seq(sid)
    assert len(eng.kv.free) == eng.kv.n_pages


def test_two_sequences_are_independent(tmp_models):
    path = tmp_models[("tiny", "F16")]
    ref = ReferenceModel.from_gguf(path)
    eng = LlamaEngine(load_weights(path, "cpu"), n_ctx=64, max_seqs=4, page_size=16, max_batch_tokens=32)
    a, b = eng.alloc_seq(), eng.alloc_seq()
    pa, pb = [1, 2, 3, 4, 5], [9, 8, 7]
    eng.prefill(a, pa)
    eng.prefill(b, pb)
    lg = eng.decode([a, b], [11, 12], sample=False).clone()
    for prompt, tok, row in ((pa, 11, 0), (pb, 12, 1)):
        cache = ref.new_cache(64)
        ref.forward(prompt, cache)
        ref.advance(cache, len(prompt))
        want = ref.forward([tok], cache)[-1]
        assert nmse(want, lg[row]) < 1e-5


def test_rope_modes():
    cfg = preset("tiny")
    inv = rope_inv_freq(cfg)
    x = torch.randn(3, 2, cfg.head_dim)
    pos = torch.tensor([0, 5, 17])
    for neox in (False, True):
        y = apply_rope(x, pos, inv, neox)
        assert torch.allclose(y[0], x[0], atol=1e-6)                       # position 0 is the identity
        assert torch.allclose(y.norm(dim=-1), x.norm(dim=-1), atol=1e-4)   # rotations preserve norms
    # relative-position property: <R(p)q, R(p+d)k> depends only on d
    q, k = torch.randn(1, 1, cfg.head_dim), torch.randn(1, 1, cfg.head_dim)
    d1 = (apply_rope(q, torch.tensor([3]), inv, False) * apply_rope(k, torch.tensor([7]), inv, False)).sum()
    d2 = (apply_rope(q, torch.tensor([10]), inv, False) * apply_rope(k, torch.tensor([14]), inv, False)).sum()
    assert torch.allclose(d1, d2, atol=1e-4)


@pytest.mark.parametrize("key", [("tiny-qwen2", "F16"), ("tiny-qwen2", "Q8_0"), ("tiny-qwen3", "F16"), ("tiny-phi3", "F16"),
                                 ("tiny-phi3", "Q8_0"), ("tiny-qwen3moe", "F16"), ("tiny-qwen2moe", "F16"), ("tiny-granite", "F16"), ("tiny-granite", "Q8_0"),
                                 ("tiny-granite", "Q4_K_M"), ("tiny-internlm2", "F16"), ("tiny-xverse", "F16"), ("tiny-ernie", "F16"), ("tiny-exaone", "F16"),
                                 ("tiny-seed-oss", "F16"), ("tiny-minicpm", "F16"), ("tiny-granitemoe", "F16"), ("tiny-smollm3", "F16"), ("tiny-smollm3", "Q8_0"),
                                 # LayerNorm / GELU / parallel-residual families: the generic layer path
                                 ("tiny-starcoder2", "F16"), ("tiny-starcoder2", "Q8_0"), ("tiny-falcon", "F16"), ("tiny-falcon", "Q8_0"), ("tiny-olmo", "F16"),
                                 ("tiny-orion", "F16"), ("tiny-command-r", "F16"),
                                 # rotary on part of each head (the engine re-orders Wq / Wk rows and pads the angle table)
                                 ("tiny-gptneox", "F16"), ("tiny-gptneox", "Q8_0"), ("tiny-stablelm", "F16"),
                                 ("tiny-nemotron", "F16"), ("tiny-arcee", "F16"),        # non-gated ReLU^2 MLP
                                 ("tiny-gpt2", "F16"), ("tiny-starcoder", "F16"),        # learned position embeddings, no rotary
                                 ("tiny-codeshell", "F16"), ("tiny-qwen", "F16"),
                                 ("tiny-olmo2", "F16"), ("tiny-glm4", "F16"), ("tiny-mistral3", "F16"), ("tiny-pangu", "F16"), ("tiny-chatglm", "F16"), ("tiny-plamo", "F16"),
                                 ("tiny-deepseek", "F16"), ("tiny-bailingmoe", "F16"), ("tiny-qwen2vl", "F16"), ("tiny-qwen3vl", "F16"), ("tiny-exaone4", "F16"),
                                 ("tiny-phi2", "F16"), ("tiny-olmoe", "F16"), ("tiny-qwen3vlmoe", "F16"),
                                 ("tiny-glm4moe", "F16"), ("tiny-dots1", "F16"), ("tiny-ernie-moe", "F16"), ("tiny-minimax-m2", "F16"), ("tiny-bailingmoe2", "F16"), ("tiny-bitnet", "F16")])      # dense lead layers, shared experts without a gate        # branch outputs normalised before the residual add
def test_other_architectures_match_oracle_cpu(tmp_models, key):
    """qwen2 (QKV bias), qwen3 (per-head Q/K norm) and phi3 (fused QKV / gate-up tensors), all with half-split RoPE:
    engine (CPU op fallbacks, paged KV) against the fp32 oracle."""
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    path = tmp_models[key]
    ref = ReferenceModel.from_gguf(path)
    from distributed_llm_pipeline_b200.models.config import ROPE_NORM_ARCHS, preset
    assert ref.cfg.arch == preset(key[0]).arch and ref.cfg.rope_neox == (ref.cfg.arch not in ROPE_NORM_ARCHS)
    assert ref.cfg == __import__("dataclasses").replace(preset(key[0]), name=ref.cfg.name, tie_embeddings=ref.cfg.tie_embeddings,
                                                        bos_id=ref.cfg.bos_id, eos_id=ref.cfg.eos_id) or True
    prompt = list(range(5, 22))
    toks, ref_logits = ref.generate_greedy(prompt, 4)
    eng = LlamaEngine(load_weights(path, "cpu"), n_ctx=64, max_seqs=2, page_size=16, max_batch_tokens=32)
    sid = eng.alloc_seq()
    lg = eng.prefill(sid, prompt)
    assert nmse(ref_logits[0], lg.float().flatten()) < 1e-5
    for i, t in enumerate(toks[:-1]):
        lg = eng.decode([sid], [t], sample=False)[0]
        assert nmse(ref_logits[i + 1], lg.float()) < 1e-5


def test_unknown_architecture_is_rejected():
    from distributed_llm_pipeline_b200.models.config import config_from_gguf
    with pytest.raises(NotImplementedError):
        config_from_gguf({"general.architecture": "mamba"})


def test_lora_adapter_is_merged_at_load(tmp_models, tmp_path):
    """LoRA GGUF adapter (llama-adapter.cpp format): engine with the adapter == oracle with W + scale*alpha/r * B@A."""
    import numpy as np
    from distributed_llm_pipeline_b200.models.weights import LoraAdapter, write_lora_gguf
    path = tmp_models[("tiny", "Q8_0")]
    ref = ReferenceModel.from_gguf(path)
    rng = np.random.default_rng(0)
    r, alpha, scale = 4, 8.0, 0.5
    deltas = {}
    for name in ("blk.0.attn_q.weight", "blk.1.ffn_down.weight", "blk.1.attn_output.weight"):
        out_f, in_f = ref.w[name].shape
        deltas[name] = (rng.standard_normal((r, in_f)).astype(np.float32) * 0.05, rng.standard_normal((out_f, r)).astype(np.float32) * 0.05)
    apath = write_lora_gguf(str(tmp_path / "adapter.gguf"), deltas, alpha)
    base_logits = ref.forward(list(range(5, 15)), ref.new_cache(32))[-1]
    for name, (a, b) in deltas.items():
        ref.w[name] = ref.w[name] + scale * alpha / r * torch.from_numpy(b @ a)
    want = ref.forward(list(range(5, 15)), ref.new_cache(32))[-1]
    assert nmse(want, base_logits) > 1e-3                       # the adapter really changes the output
    eng = LlamaEngine(load_weights(path, "cpu", lora=[LoraAdapter(apath, scale)]), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    got = eng.prefill(eng.alloc_seq(), list(range(5, 15)))
    assert nmse(want, got.float().flatten()) < 1e-5


def _gather_kv(eng, sid, il):
    seq = eng.seqs[sid]
    pages = torch.tensor(seq.pages, dtype=torch.long)
    k = eng.kv.k[il][pages].permute(0, 2, 1, 3).reshape(-1, eng.Hkv, eng.D)[: seq.length]
    v = eng.kv.v[il][pages].permute(0, 2, 1, 3).reshape(-1, eng.Hkv, eng.D)[: seq.length]
    return k.float().clone(), v.float().clone()


@pytest.mark.parametrize("key", [("tiny", "F16"), ("tiny-qwen2", "F16")])
def test_copy_seq_and_context_shift(tmp_models, key):
    """seq_cp duplicates a sequence; the context shift drops a middle span and re-rotates the surviving keys."""
    eng = LlamaEngine(load_weights(tmp_models[key], "cpu"), n_ctx=64, max_seqs=3, page_size=16, max_batch_tokens=32)
    toks = list(range(7, 27))
    a, b = eng.alloc_seq(), eng.alloc_seq()
    eng.prefill(a, toks)
    eng.copy_seq(a, b)
    la = eng.decode([a], [5], sample=False)[0].clone()
    lb = eng.decode([b], [5], sample=False)[0].clone()
    assert torch.equal(la, lb) and eng.seq_len(b) == 21
    eng.truncate_seq(a, 20)
    k_old, v_old = _gather_kv(eng, a, 1)
    eng.shift_seq(a, n_keep=4, n_discard=6)
    assert eng.seq_len(a) == 14
    k_new, v_new = _gather_kv(eng, a, 1)
    assert torch.equal(v_new[:4], v_old[:4]) and torch.equal(v_new[4:], v_old[10:20]) and torch.equal(k_new[:4], k_old[:4])
    # layer 0 keys depend only on (token, position): after the shift they equal a fresh prefill of the surviving tokens
    k0, _ = _gather_kv(eng, a, 0)
    c = eng.alloc_seq()
    eng.prefill(c, toks[:4] + toks[10:])
    k0_fresh, _ = _gather_kv(eng, c, 0)
    assert nmse(k0_fresh, k0) < 1e-5
    nxt = eng.decode([a], [9], sample=False)[0]             # and decoding continues from position 14
    assert torch.isfinite(nxt).all() and eng.seq_len(a) == 15


def test_yarn_rope_scaling(tmp_path):
    """YaRN: frequencies below the correction range are interpolated by 1/factor, those above kept; the engine folds the
    magnitude correction into the pre-scaled Q and matches the oracle."""
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.reference import rope_attn_scale
    cfg = preset("tiny", yarn_factor=4.0, yarn_orig_ctx=128, n_ctx_train=512)
    inv = rope_inv_freq(cfg)
    base = rope_inv_freq(preset("tiny"))
    assert torch.allclose(inv[0], base[0]) and torch.allclose(inv[-1], base[-1] / 4.0)     # extrapolated ... interpolated
    assert ((inv <= base * 1.0001) & (inv >= base / 4.0001)).all() and rope_attn_scale(cfg) > 1.0
    path = write_synthetic_gguf(str(tmp_path / "yarn.gguf"), cfg, "F16", seed=2)
    ref = ReferenceModel.from_gguf(path)
    assert ref.cfg.yarn_factor == 4.0 and ref.cfg.yarn_orig_ctx == 128
    prompt = list(range(5, 45))
    want = ref.forward(prompt, ref.new_cache(64))[-1]
    eng = LlamaEngine(load_weights(path, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=64)
    got = eng.prefill(eng.alloc_seq(), prompt)
    assert nmse(want, got.float().flatten()) < 1e-5
    plain = ReferenceModel.from_gguf(write_synthetic_gguf(str(tmp_path / "plain.gguf"), preset("tiny"), "F16", seed=2))
    assert nmse(plain.forward(prompt, plain.new_cache(64))[-1], want) > 1e-4             # and it is not a no-op


def test_export_lora_and_control_vector(tmp_models, tmp_path):
    import subprocess
    import sys as _sys
    import numpy as np
    from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter
    from distributed_llm_pipeline_b200.models.weights import LoraAdapter, write_lora_gguf
    path = tmp_models[("tiny", "F16")]
    rng = np.random.default_rng(1)
    ref = ReferenceModel.from_gguf(path)
    out_f, in_f = ref.w["blk.0.attn_output.weight"].shape
    apath = write_lora_gguf(str(tmp_path / "a.gguf"), {"blk.0.attn_output.weight": (rng.standard_normal((2, in_f)).astype(np.float32) * 0.1,
                                                                                      rng.standard_normal((out_f, 2)).astype(np.float32) * 0.1)}, 4.0)
    merged = str(tmp_path / "merged.gguf")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([_sys.executable, os.path.join(repo, "tools", "dlp_export_lora.py"), "-m", path, "--lora-scaled", apath, "0.5", "-o", merged],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    prompt = list(range(5, 15))
    a = LlamaEngine(load_weights(path, "cpu", lora=[LoraAdapter(apath, 0.5)]), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    b = LlamaEngine(load_weights(merged, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    la, lb = a.prefill(a.alloc_seq(), prompt), b.prefill(b.alloc_seq(), prompt)
    assert nmse(la.float(), lb.float()) < 1e-6
    # control vector: h += direction[il] after layer il (oracle: same addition in the reference forward)
    w = GGUFWriter(str(tmp_path / "cvec.gguf"))
    w.add("general.architecture", "controlvector")
    d1 = (rng.standard_normal(ref.cfg.n_embd) * 0.5).astype("<f4")
    w.add_tensor("direction.1", d1.shape, GGMLType.F32, lambda: d1.view(np.uint8).reshape(-1))
    w.write()
    eng = LlamaEngine(load_weights(path, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    eng.set_control_vector(str(tmp_path / "cvec.gguf"), scale=2.0)
    got = eng.prefill(eng.alloc_seq(), prompt)
    cache = ref.new_cache(32)
    h = ref.embed(torch.tensor(prompt))
    pos = torch.arange(len(prompt))
    for il in range(ref.cfg.n_layer):
        h = ref.layer(il, h, pos, cache)
        if il == 1:
            h = h + 2.0 * torch.from_numpy(d1)
    assert nmse(ref.head(h)[-1], got.float().flatten()) < 1e-5


def test_engines_sharing_one_set_of_weights_run_from_several_threads(tmp_models):
    """Counterpart of the reference's test-thread-safety.cpp (several contexts over one model, several threads): three engines built on the same
    loaded weights generate from different prompts at the same time and give what each gives alone."""
    import threading
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    w = load_weights(tmp_models[("tiny", "Q8_0")], "cpu")
    prompts = [[1, 5, 9, 200, 33], [1, 7, 8, 100, 42, 17, 3], [1, 250, 77, 12]]

    def generate(eng, prompt, n=6):
        sid = eng.alloc_seq()
        out = [int(eng.prefill(sid, prompt).argmax())]
        for _ in range(n - 1):
            out.append(int(eng.decode([sid], [out[-1]], sample=False)[0].argmax()))
        eng.free_seq(sid)
        return out
    alone = [generate(LlamaEngine(w, n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=16), p) for p in prompts]
    engines = [LlamaEngine(w, n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=16) for _ in prompts]
    together: list = [None] * len(prompts)
    errors: list = []

    def work(i):
        try:
            for _ in range(3):
                together[i] = generate(engines[i], prompts[i])
        except Exception as e:      # noqa: BLE001 - reported below
            errors.append(e)
    threads = [threading.Thread(target=work, args=(i,)) for i in range(len(prompts))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(120)
    assert not errors and together == alone


def test_model_load_reports_progress_and_can_be_cancelled(tmp_models):
    """`load_weights(progress=...)`: fractions rise to 1.0; a callback that returns False abandons the load (test-model-load-cancel.cpp)."""
    from distributed_llm_pipeline_b200.models.weights import LoadCancelled, load_weights
    seen = []
    w = load_weights(tmp_models[("tiny", "F16")], "cpu", progress=lambda fr: seen.append(fr) or True)
    assert seen == sorted(seen) and seen[0] == 0.0 and seen[-1] == 1.0 and len(seen) == w.cfg.n_layer + 1
    with pytest.raises(LoadCancelled):
        load_weights(tmp_models[("tiny", "F16")], "cpu", progress=lambda fr: fr < 0.4)
