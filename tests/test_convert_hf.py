"""HF -> GGUF converter, cross-validated against an INDEPENDENT implementation: random-init Hugging Face models (transformers,
CPU, fp32) are converted and the oracle / engine logits must match the HF forward pass - this pins the graph conventions
(RoPE pair layout and the Q/K row permutation, bias, per-head Q/K norm, fused tensors, GQA) to the upstream definition."""
import importlib
import os
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "tools"))

transformers = pytest.importorskip("transformers")
pytest.importorskip("safetensors")

COMMON = dict(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
              num_key_value_heads=2, max_position_embeddings=512, rms_norm_eps=1e-5, tie_word_embeddings=False,
              bos_token_id=1, eos_token_id=2, pad_token_id=0)
CASES = {
    "llama": ("LlamaConfig", "LlamaForCausalLM", {}),
    "llama31": ("LlamaConfig", "LlamaForCausalLM", {"rope_parameters": {"rope_type": "llama3", "rope_theta": 10000.0, "factor": 8.0,
                                                                          "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                                                                          "original_max_position_embeddings": 64}}),
    "mistral": ("MistralConfig", "MistralForCausalLM", {"sliding_window": None}),
    "qwen2": ("Qwen2Config", "Qwen2ForCausalLM", {}),
    "qwen2-yarn": ("Qwen2Config", "Qwen2ForCausalLM", {"rope_parameters": {"rope_type": "yarn", "rope_theta": 10000.0, "factor": 4.0,
                                                                             "original_max_position_embeddings": 128}}),
    "qwen3": ("Qwen3Config", "Qwen3ForCausalLM", {"head_dim": 128}),
    "granite": ("GraniteConfig", "GraniteForCausalLM", {"embedding_multiplier": 12.0, "residual_multiplier": 0.22, "logits_scaling": 8.0,
                                                        "attention_multiplier": 0.015625}),
    "phi3": ("Phi3Config", "Phi3ForCausalLM", {"num_key_value_heads": 4}),
    "mixtral": ("MixtralConfig", "MixtralForCausalLM", {"intermediate_size": 256, "num_local_experts": 4, "num_experts_per_tok": 2}),
    "qwen2moe": ("Qwen2MoeConfig", "Qwen2MoeForCausalLM", {"moe_intermediate_size": 128, "shared_expert_intermediate_size": 256,
                                                             "num_experts": 4, "num_experts_per_tok": 2, "norm_topk_prob": False,
                                                             "decoder_sparse_step": 1, "mlp_only_layers": []}),
    "smollm3": ("SmolLM3Config", "SmolLM3ForCausalLM", {"num_hidden_layers": 4, "no_rope_layer_interval": 4, "use_sliding_window": False}),
    "seed_oss": ("SeedOssConfig", "SeedOssForCausalLM", {"head_dim": 64, "attention_bias": True, "attention_out_bias": False}),
    "ernie4_5": ("Ernie4_5Config", "Ernie4_5ForCausalLM", {"use_bias": False, "head_dim": 64}),
    "granitemoe": ("GraniteMoeConfig", "GraniteMoeForCausalLM", {"intermediate_size": 256, "num_local_experts": 4, "num_experts_per_tok": 2,
                                                                 "embedding_multiplier": 12.0, "residual_multiplier": 0.22, "logits_scaling": 8.0,
                                                                 "attention_multiplier": 0.015625}),
    "starcoder2": ("Starcoder2Config", "Starcoder2ForCausalLM", {"norm_epsilon": 1e-5, "use_bias": True, "hidden_act": "gelu_pytorch_tanh", "sliding_window": None,
                                                                 "tie_word_embeddings": True}),
    "falcon": ("FalconConfig", "FalconForCausalLM", {"multi_query": True, "new_decoder_architecture": False, "parallel_attn": True, "bias": False, "alibi": False,
                                                     "num_kv_heads": None, "activation": "gelu_pytorch_tanh", "tie_word_embeddings": False, "ffn_hidden_size": 1024}),
    "falcon-new-arch": ("FalconConfig", "FalconForCausalLM", {"multi_query": False, "new_decoder_architecture": True, "parallel_attn": True, "bias": False, "alibi": False,
                                                              "num_kv_heads": 2, "activation": "gelu_pytorch_tanh", "tie_word_embeddings": False, "ffn_hidden_size": 1024}),
    "gptneox": ("GPTNeoXConfig", "GPTNeoXForCausalLM", {"rotary_pct": 0.25, "use_parallel_residual": True, "hidden_act": "gelu_pytorch_tanh", "layer_norm_eps": 1e-5,
                                                      "attention_bias": True, "intermediate_size": 1024}),
    "gptneox-seq": ("GPTNeoXConfig", "GPTNeoXForCausalLM", {"rotary_pct": 0.5, "use_parallel_residual": False, "hidden_act": "gelu_pytorch_tanh", "layer_norm_eps": 1e-5,
                                                          "attention_bias": True, "intermediate_size": 1024}),
    "stablelm": ("StableLmConfig", "StableLmForCausalLM", {"partial_rotary_factor": 0.25, "layer_norm_eps": 1e-5, "use_qkv_bias": False}),
    "nemotron": ("NemotronConfig", "NemotronForCausalLM", {"partial_rotary_factor": 0.5, "norm_eps": 1e-5, "hidden_act": "relu2", "intermediate_size": 1024,
                                                         "attention_bias": False, "mlp_bias": False, "head_dim": 64}),
    "arcee": ("ArceeConfig", "ArceeForCausalLM", {"hidden_act": "relu2", "intermediate_size": 1024, "head_dim": 64, "attention_bias": False, "mlp_bias": False}),
    "gpt2": ("GPT2Config", "GPT2LMHeadModel", {"n_embd": 256, "n_layer": 2, "n_head": 4, "n_positions": 512, "n_inner": 1024, "activation_function": "gelu_pytorch_tanh",
                                               "layer_norm_epsilon": 1e-5, "tie_word_embeddings": True}),
    "starcoder": ("GPTBigCodeConfig", "GPTBigCodeForCausalLM", {"n_embd": 256, "n_layer": 2, "n_head": 4, "n_positions": 512, "n_inner": 1024, "multi_query": True,
                                                                "activation_function": "gelu_pytorch_tanh", "layer_norm_epsilon": 1e-5, "tie_word_embeddings": True}),
    "olmo2": ("Olmo2Config", "Olmo2ForCausalLM", {"num_key_value_heads": 2}),
    "glm4": ("Glm4Config", "Glm4ForCausalLM", {"partial_rotary_factor": 0.5, "head_dim": 64, "attention_bias": True}),
    "olmo": ("OlmoConfig", "OlmoForCausalLM", {"clip_qkv": None, "num_key_value_heads": 4}),
    "olmoe": ("OlmoeConfig", "OlmoeForCausalLM", {"intermediate_size": 128, "num_experts": 4, "num_experts_per_tok": 2, "norm_topk_prob": False,
                                                  "num_key_value_heads": 4, "clip_qkv": None}),
    "glm4moe": ("Glm4MoeConfig", "Glm4MoeForCausalLM", {"num_hidden_layers": 3, "head_dim": 64, "moe_intermediate_size": 128, "num_experts_per_tok": 2,
                                                        "n_shared_experts": 1, "n_routed_experts": 4, "routed_scaling_factor": 1.5, "first_k_dense_replace": 1,
                                                        "use_qk_norm": True, "attention_bias": True, "norm_topk_prob": True}),
    "dots1": ("Dots1Config", "Dots1ForCausalLM", {"num_hidden_layers": 3, "num_key_value_heads": 4, "head_dim": 64, "moe_intermediate_size": 128,
                                                  "num_experts_per_tok": 2, "n_shared_experts": 1, "n_routed_experts": 4, "routed_scaling_factor": 2.5,
                                                  "first_k_dense_replace": 1, "norm_topk_prob": True, "sliding_window": None,
                                                  "layer_types": ["full_attention"] * 3}),
    "ernie4_5-moe": ("Ernie4_5_MoeConfig", "Ernie4_5_MoeForCausalLM", {"num_hidden_layers": 3, "moe_intermediate_size": 128, "moe_k": 2, "moe_num_experts": 4,
                                                                        "moe_num_shared_experts": 1, "moe_layer_start_index": 1, "use_bias": False, "head_dim": 64}),
    "minimax-m2": ("MiniMaxM2Config", "MiniMaxM2ForCausalLM", {"intermediate_size": 128, "head_dim": 64, "num_local_experts": 4, "num_experts_per_tok": 2,
                                                               "rope_parameters": {"rope_type": "default", "rope_theta": 10000.0, "partial_rotary_factor": 0.5}}),
    "exaone4": ("Exaone4Config", "Exaone4ForCausalLM", {"head_dim": 64, "sliding_window": None, "sliding_window_pattern": None,
                                                        "layer_types": ["full_attention", "full_attention"]}),
    "phi2": ("PhiConfig", "PhiForCausalLM", {"partial_rotary_factor": 0.5, "hidden_act": "gelu_new", "layer_norm_eps": 1e-5, "intermediate_size": 1024,
                                             "qk_layernorm": False, "tie_word_embeddings": False}),
    "command-r": ("CohereConfig", "CohereForCausalLM", {"logit_scale": 0.0625, "layer_norm_eps": 1e-5, "use_qk_norm": False, "tie_word_embeddings": True}),
    "qwen3moe": ("Qwen3MoeConfig", "Qwen3MoeForCausalLM", {"head_dim": 64, "moe_intermediate_size": 128, "num_experts": 4,
                                                             "num_experts_per_tok": 2, "norm_topk_prob": True}),
}


@pytest.mark.parametrize("family", list(CASES))
def test_converted_model_matches_hf_forward(tmp_path, family):
    cfg_cls, model_cls, extra = CASES[family]
    if not hasattr(transformers, cfg_cls):
        pytest.skip(f"transformers has no {cfg_cls}")
    torch.manual_seed(0)
    cfg = getattr(transformers, cfg_cls)(**{**COMMON, **extra})
    model = getattr(transformers, model_cls)(cfg).eval().float()
    with torch.no_grad():                      # make biases / norm weights non-trivial
        for n, p in model.named_parameters():
            if n.endswith("bias"):
                p.normal_(0, 0.05)
            elif "norm" in n:
                p.copy_(1 + 0.1 * torch.randn_like(p))
    d = tmp_path / family
    model.save_pretrained(str(d), safe_serialization=True)
    conv = importlib.import_module("convert_hf_to_gguf")
    out = conv.convert(str(d), str(tmp_path / f"{family}.gguf"), "F32")
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel, nmse
    from distributed_llm_pipeline_b200.models.weights import load_weights
    ids = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3, 250] * (4 if "yarn" in family else 1)
    with torch.no_grad():
        want = model(torch.tensor([ids])).logits[0]
    ref = ReferenceModel.from_gguf(out)
    assert ref.cfg.arch == ("llama" if family in ("llama", "llama31", "mistral", "mixtral") else family if family in ("command-r", "ernie4_5-moe", "minimax-m2") else family.split("-")[0])
    if family == "qwen2-yarn":
        assert ref.cfg.yarn_factor == 4.0 and ref.cfg.yarn_orig_ctx == 128
    if family == "llama31":
        assert "rope_freqs.weight" in ref.w and float(ref.w["rope_freqs.weight"].max()) == 8.0
    got = ref.forward(ids, ref.new_cache(64))
    assert nmse(want, got) < 1e-9, family
    eng = LlamaEngine(load_weights(out, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    lg = eng.prefill(eng.alloc_seq(), ids)
    assert nmse(want[-1], lg.float().flatten()) < 1e-5


@pytest.mark.parametrize("family", ["llama", "qwen2"])
def test_converted_lora_adapter_matches_merged_hf_model(tmp_path, family):
    """convert_lora_to_gguf.py: base model and a PEFT-layout adapter converted separately; engine(base + adapter) == the HF model with
    W + (alpha / r) B A merged in (for llama this pins the Q/K row permutation of the B factors)."""
    import json
    from safetensors.torch import save_file
    cfg_cls, model_cls, extra = CASES[family]
    torch.manual_seed(1)
    cfg = getattr(transformers, cfg_cls)(**{**COMMON, **extra})
    model = getattr(transformers, model_cls)(cfg).eval().float()
    base_dir = tmp_path / "base"
    model.save_pretrained(str(base_dir), safe_serialization=True)
    r, alpha = 4, 8.0
    targets = ["self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj", "mlp.down_proj"]
    lora = {}
    with torch.no_grad():
        for il, layer in enumerate(model.model.layers):
            for tname in targets:
                mod = layer.get_submodule(tname)
                A, B = torch.randn(r, mod.in_features) * 0.05, torch.randn(mod.out_features, r) * 0.05
                lora[f"base_model.model.model.layers.{il}.{tname}.lora_A.weight"] = A
                lora[f"base_model.model.model.layers.{il}.{tname}.lora_B.weight"] = B
                mod.weight += (alpha / r) * (B @ A)             # the HF model now IS base + adapter
    ad_dir = tmp_path / "adapter"
    ad_dir.mkdir()
    save_file(lora, str(ad_dir / "adapter_model.safetensors"))
    json.dump({"peft_type": "LORA", "r": r, "lora_alpha": alpha, "target_modules": [t.split(".")[-1] for t in targets], "bias": "none"},
              open(ad_dir / "adapter_config.json", "w"))
    conv = importlib.import_module("convert_hf_to_gguf")
    base_gguf = conv.convert(str(base_dir), str(tmp_path / "base.gguf"), "F32")
    lconv = importlib.import_module("convert_lora_to_gguf")
    ad_gguf = lconv.convert(str(ad_dir), str(base_dir), str(tmp_path / "adapter.gguf"))
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.reference import nmse
    from distributed_llm_pipeline_b200.models.weights import LoraAdapter, load_weights
    ids = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3, 250]
    with torch.no_grad():
        want = model(torch.tensor([ids])).logits[0, -1]
    eng = LlamaEngine(load_weights(base_gguf, "cpu", lora=[LoraAdapter(ad_gguf)]), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    got = eng.prefill(eng.alloc_seq(), ids).float().flatten()
    assert nmse(want, got) < 1e-5
    plain = LlamaEngine(load_weights(base_gguf, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    assert nmse(want, plain.prefill(plain.alloc_seq(), ids).float().flatten()) > 1e-3      # the adapter matters
