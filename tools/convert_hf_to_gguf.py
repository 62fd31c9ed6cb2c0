#!/usr/bin/env python
"""convert-hf-to-gguf: Hugging Face checkpoint directory (config.json + *.safetensors) -> GGUF.

Parity: `llama.cpp/convert_hf_to_gguf.py` for the architecture families this engine runs: LlamaForCausalLM / MistralForCausalLM
/ MixtralForCausalLM (arch `llama`, Q/K rows permuted from the HF half-split layout to ggml's adjacent-pair rotation; experts
stacked into `ffn_*_exps`), Qwen2ForCausalLM (`qwen2`, biases), Qwen2MoeForCausalLM (`qwen2moe`: shared expert), Qwen3ForCausalLM / Qwen3MoeForCausalLM (`qwen3`, `qwen3moe`,
per-head Q/K norms), Phi3ForCausalLM (`phi3`, fused qkv / gate_up kept fused), GraniteForCausalLM (`granite`: llama tensors plus the
embedding / residual / attention / logit multipliers).
Tokenizer: `tokenizer.json` (byte-level BPE -> `gpt2` model with merges) or `tokenizer.model` (SentencePiece -> `llama` model)
when present, else a synthetic vocabulary (random-init test checkpoints).  Output types: F32 / F16 / BF16 / Q8_0.
"""
import argparse
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, quantize  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import FTYPE, TokenType, ValueType  # noqa: E402
from distributed_llm_pipeline_b200.gguf.synth import synthetic_vocab  # noqa: E402

ARCHS = {"LlamaForCausalLM": "llama", "MistralForCausalLM": "llama", "MixtralForCausalLM": "llama", "Qwen2ForCausalLM": "qwen2",
         "Qwen2MoeForCausalLM": "qwen2moe", "Qwen3ForCausalLM": "qwen3", "Qwen3MoeForCausalLM": "qwen3moe", "Phi3ForCausalLM": "phi3", "GraniteForCausalLM": "granite",
         "SmolLM3ForCausalLM": "smollm3", "SeedOssForCausalLM": "seed_oss", "Ernie4_5ForCausalLM": "ernie4_5", "GraniteMoeForCausalLM": "granitemoe",
         "Starcoder2ForCausalLM": "starcoder2", "FalconForCausalLM": "falcon", "OlmoForCausalLM": "olmo", "CohereForCausalLM": "command-r",
         "GPTNeoXForCausalLM": "gptneox", "StableLmForCausalLM": "stablelm", "NemotronForCausalLM": "nemotron", "ArceeForCausalLM": "arcee",
         "GPT2LMHeadModel": "gpt2", "GPTBigCodeForCausalLM": "starcoder", "Olmo2ForCausalLM": "olmo2", "Glm4ForCausalLM": "glm4", "PhiForCausalLM": "phi2", "OlmoeForCausalLM": "olmoe", "Glm4MoeForCausalLM": "glm4moe", "Dots1ForCausalLM": "dots1",
         "Ernie4_5_MoeForCausalLM": "ernie4_5-moe", "MiniMaxM2ForCausalLM": "minimax-m2", "Exaone4ForCausalLM": "exaone4"}
# families whose GGUF graph rotates adjacent pairs: HF's half-split Q/K rows are permuted on the way in (ernie4_5 already stores
# adjacent pairs upstream, so it is not in the list)
PERMUTE_QK = ("llama", "granite", "granitemoe", "smollm3", "olmo", "arcee")
LAYERNORM = ("starcoder2", "falcon", "olmo", "command-r", "gptneox", "stablelm", "nemotron", "gpt2", "starcoder", "phi2")
LAYER_MAP = {
    "input_layernorm.weight": "attn_norm.weight", "post_attention_layernorm.weight": "ffn_norm.weight",
    "self_attn.q_proj.weight": "attn_q.weight", "self_attn.k_proj.weight": "attn_k.weight", "self_attn.v_proj.weight": "attn_v.weight",
    "self_attn.q_proj.bias": "attn_q.bias", "self_attn.k_proj.bias": "attn_k.bias", "self_attn.v_proj.bias": "attn_v.bias",
    "self_attn.o_proj.weight": "attn_output.weight", "self_attn.q_norm.weight": "attn_q_norm.weight",
    "self_attn.k_norm.weight": "attn_k_norm.weight", "self_attn.qkv_proj.weight": "attn_qkv.weight",
    "mlp.gate_proj.weight": "ffn_gate.weight", "mlp.up_proj.weight": "ffn_up.weight", "mlp.down_proj.weight": "ffn_down.weight",
    "mlp.gate_up_proj.weight": "ffn_up.weight",
    # olmo2 / glm4: norms behind the branches
    "post_feedforward_layernorm.weight": "post_ffw_norm.weight", "post_self_attn_layernorm.weight": "post_attention_norm.weight",
    "post_mlp_layernorm.weight": "post_ffw_norm.weight",
    # LayerNorm families: norm and projection biases, starcoder2's non-gated MLP
    "input_layernorm.bias": "attn_norm.bias", "post_attention_layernorm.bias": "ffn_norm.bias", "self_attn.o_proj.bias": "attn_output.bias",
    "mlp.c_fc.weight": "ffn_up.weight", "mlp.c_fc.bias": "ffn_up.bias", "mlp.c_proj.weight": "ffn_down.weight", "mlp.c_proj.bias": "ffn_down.bias",
    # phi (phi-1 / phi-1.5 / phi-2)
    "self_attn.dense.weight": "attn_output.weight", "self_attn.dense.bias": "attn_output.bias",
    "mlp.fc1.weight": "ffn_up.weight", "mlp.fc1.bias": "ffn_up.bias", "mlp.fc2.weight": "ffn_down.weight", "mlp.fc2.bias": "ffn_down.bias",
}
# falcon keeps its own module names (transformer.h.N....); QKV is regrouped separately
FALCON_MAP = {"input_layernorm.weight": "attn_norm.weight", "input_layernorm.bias": "attn_norm.bias",
              "ln_mlp.weight": "attn_norm.weight", "ln_mlp.bias": "attn_norm.bias", "ln_attn.weight": "attn_norm_2.weight", "ln_attn.bias": "attn_norm_2.bias",
              "self_attention.dense.weight": "attn_output.weight", "mlp.dense_h_to_4h.weight": "ffn_up.weight", "mlp.dense_4h_to_h.weight": "ffn_down.weight",
              "self_attention.query_key_value.weight": "attn_qkv.weight"}
GPT2_MAP = {"ln_1.weight": "attn_norm.weight", "ln_1.bias": "attn_norm.bias", "ln_2.weight": "ffn_norm.weight", "ln_2.bias": "ffn_norm.bias",
            "attn.c_attn.weight": "attn_qkv.weight", "attn.c_attn.bias": "attn_qkv.bias", "attn.c_proj.weight": "attn_output.weight",
            "attn.c_proj.bias": "attn_output.bias", "mlp.c_fc.weight": "ffn_up.weight", "mlp.c_fc.bias": "ffn_up.bias",
            "mlp.c_proj.weight": "ffn_down.weight", "mlp.c_proj.bias": "ffn_down.bias"}
NEOX_MAP = {"input_layernorm.weight": "attn_norm.weight", "input_layernorm.bias": "attn_norm.bias", "post_attention_layernorm.weight": "ffn_norm.weight",
            "post_attention_layernorm.bias": "ffn_norm.bias", "attention.query_key_value.weight": "attn_qkv.weight", "attention.query_key_value.bias": "attn_qkv.bias",
            "attention.dense.weight": "attn_output.weight", "attention.dense.bias": "attn_output.bias", "mlp.dense_h_to_4h.weight": "ffn_up.weight",
            "mlp.dense_h_to_4h.bias": "ffn_up.bias", "mlp.dense_4h_to_h.weight": "ffn_down.weight", "mlp.dense_4h_to_h.bias": "ffn_down.bias"}
TOP_MAP = {"model.embed_tokens.weight": "token_embd.weight", "model.norm.weight": "output_norm.weight", "lm_head.weight": "output.weight",
           "model.norm.bias": "output_norm.bias", "transformer.word_embeddings.weight": "token_embd.weight",
           "transformer.ln_f.weight": "output_norm.weight", "transformer.ln_f.bias": "output_norm.bias",
           "transformer.wte.weight": "token_embd.weight", "transformer.wpe.weight": "position_embd.weight",
           "gpt_neox.embed_in.weight": "token_embd.weight", "gpt_neox.final_layer_norm.weight": "output_norm.weight",
           "gpt_neox.final_layer_norm.bias": "output_norm.bias", "embed_out.weight": "output.weight",
           "model.final_layernorm.weight": "output_norm.weight", "model.final_layernorm.bias": "output_norm.bias", "lm_head.bias": "output.bias"}


def permute_qk(w: np.ndarray, n_head: int) -> np.ndarray:
    """HF stores Q/K rows for rotate_half (first half | second half of every head); ggml's llama graph rotates adjacent pairs."""
    return w.reshape(n_head, 2, w.shape[0] // n_head // 2, *w.shape[1:]).swapaxes(1, 2).reshape(w.shape)


def load_safetensors(d: str) -> dict:
    from safetensors import safe_open
    out = {}
    for fn in sorted(os.listdir(d)):
        if fn.endswith(".safetensors"):
            with safe_open(os.path.join(d, fn), framework="pt") as f:
                for k in f.keys():
                    out[k] = f.get_tensor(k)
    return out


def add_tokenizer(w: GGUFWriter, d: str, n_vocab: int, cfg: dict):
    tj, tm = os.path.join(d, "tokenizer.json"), os.path.join(d, "tokenizer.model")
    bos, eos = cfg.get("bos_token_id"), cfg.get("eos_token_id")
    eos = eos[0] if isinstance(eos, list) else eos
    if os.path.exists(tj):
        t = json.load(open(tj))
        model = t.get("model", {})
        if model.get("type") == "BPE":
            vocab = model["vocab"]
            toks = [f"[PAD{i}]" for i in range(n_vocab)]
            types = [int(TokenType.UNUSED)] * n_vocab
            for s, i in vocab.items():
                if i < n_vocab:
                    toks[i], types[i] = s, int(TokenType.NORMAL)
            for a in t.get("added_tokens", []):
                if a["id"] < n_vocab:
                    toks[a["id"]], types[a["id"]] = a["content"], int(TokenType.CONTROL if a.get("special") else TokenType.USER_DEFINED)
            merges = [m if isinstance(m, str) else " ".join(m) for m in model.get("merges", [])]
            w.add("tokenizer.ggml.model", "gpt2")
            w.add("tokenizer.ggml.pre", "llama-bpe" if "llama" in str(cfg.get("model_type", "")) else "default")
            w.add("tokenizer.ggml.tokens", toks, etype=ValueType.STRING)
            w.add("tokenizer.ggml.token_type", np.asarray(types, np.int32), etype=ValueType.INT32)
            w.add("tokenizer.ggml.merges", merges, etype=ValueType.STRING)
            _ids(w, bos, eos)
            return "bpe"
    if os.path.exists(tm):
        import sentencepiece as spm
        sp = spm.SentencePieceProcessor(model_file=tm)
        toks, scores, types = [], [], []
        for i in range(n_vocab):
            if i < sp.get_piece_size():
                toks.append(sp.id_to_piece(i))
                scores.append(sp.get_score(i))
                types.append(int(TokenType.CONTROL if sp.is_control(i) else TokenType.UNKNOWN if sp.is_unknown(i)
                                 else TokenType.BYTE if sp.is_byte(i) else TokenType.NORMAL))
            else:
                toks.append(f"[PAD{i}]"); scores.append(-1000.0); types.append(int(TokenType.UNUSED))
        w.add("tokenizer.ggml.model", "llama")
        w.add("tokenizer.ggml.tokens", toks, etype=ValueType.STRING)
        w.add("tokenizer.ggml.scores", np.asarray(scores, np.float32), etype=ValueType.FLOAT32)
        w.add("tokenizer.ggml.token_type", np.asarray(types, np.int32), etype=ValueType.INT32)
        _ids(w, bos, eos)
        return "spm"
    toks, scores, types = synthetic_vocab(n_vocab)
    w.add("tokenizer.ggml.model", "llama")
    w.add("tokenizer.ggml.tokens", toks, etype=ValueType.STRING)
    w.add("tokenizer.ggml.scores", scores, etype=ValueType.FLOAT32)
    w.add("tokenizer.ggml.token_type", types, etype=ValueType.INT32)
    _ids(w, 1, 2)
    return "synthetic"


def _ids(w, bos, eos):
    if bos is not None:
        w.add("tokenizer.ggml.bos_token_id", int(bos), ValueType.UINT32)
    if eos is not None:
        w.add("tokenizer.ggml.eos_token_id", int(eos), ValueType.UINT32)


def convert(src: str, dst: str, outtype: str = "F16") -> str:
    cfg = json.load(open(os.path.join(src, "config.json")))
    hf_arch = (cfg.get("architectures") or ["?"])[0]
    if hf_arch not in ARCHS:
        raise SystemExit(f"unsupported architecture {hf_arch}; supported: {sorted(ARCHS)}")
    a = ARCHS[hf_arch]
    if a in ("gpt2", "starcoder"):       # the GPT-2 config spells its sizes differently
        cfg = {**cfg, "hidden_size": cfg["n_embd"], "num_hidden_layers": cfg["n_layer"], "num_attention_heads": cfg["n_head"],
               "intermediate_size": cfg.get("n_inner") or 4 * cfg["n_embd"], "max_position_embeddings": cfg.get("n_positions", 1024),
               "num_key_value_heads": (1 if cfg.get("multi_query", True) else cfg["n_head"]) if a == "starcoder" else cfg["n_head"],
               "tie_word_embeddings": cfg.get("tie_word_embeddings", True)}
    E, L = cfg["hidden_size"], cfg["num_hidden_layers"]
    H = cfg["num_attention_heads"]
    Hkv = cfg.get("num_key_value_heads") or H
    if a == "gptneox":      # always multi-head
        Hkv = H
    if a == "falcon":       # multi_query: one K/V head; new_decoder_architecture: num_kv_heads groups
        Hkv = int(cfg.get("num_kv_heads") or H) if cfg.get("new_decoder_architecture") else (1 if cfg.get("multi_query", True) else H)
    D = cfg.get("head_dim") or E // H
    rope = cfg.get("rope_parameters") or {}
    theta = float(cfg.get("rope_theta") or rope.get("rope_theta") or 10000.0)
    w = GGUFWriter(dst)
    w.add("general.architecture", a)
    w.add("general.name", os.path.basename(os.path.normpath(src)))
    w.add("general.file_type", FTYPE[outtype], ValueType.UINT32)
    w.add(f"{a}.vocab_size", cfg["vocab_size"], ValueType.UINT32)
    w.add(f"{a}.context_length", cfg.get("max_position_embeddings", 2048), ValueType.UINT32)
    w.add(f"{a}.embedding_length", E, ValueType.UINT32)
    w.add(f"{a}.block_count", L, ValueType.UINT32)
    w.add(f"{a}.feed_forward_length", (cfg.get("ffn_hidden_size") or 4 * E) if a == "falcon" else cfg["intermediate_size"], ValueType.UINT32)
    if a in ("granite", "granitemoe"):
        w.add(f"{a}.embedding_scale", float(cfg.get("embedding_multiplier", 1.0)), ValueType.FLOAT32)
        w.add(f"{a}.residual_scale", float(cfg.get("residual_multiplier", 1.0)), ValueType.FLOAT32)
        w.add(f"{a}.logit_scale", float(cfg.get("logits_scaling", 1.0)), ValueType.FLOAT32)
        w.add(f"{a}.attention.scale", float(cfg.get("attention_multiplier", 1.0)), ValueType.FLOAT32)
    n_exp = int(cfg.get("num_local_experts") or cfg.get("num_experts") or cfg.get("n_routed_experts") or cfg.get("moe_num_experts") or 0)
    if n_exp:
        w.add(f"{a}.expert_count", n_exp, ValueType.UINT32)
        w.add(f"{a}.expert_used_count", int(cfg.get("num_experts_per_tok") or cfg.get("moe_k") or 2), ValueType.UINT32)
        if cfg.get("moe_intermediate_size"):
            w.add(f"{a}.expert_feed_forward_length", int(cfg["moe_intermediate_size"]), ValueType.UINT32)
        if a == "ernie4_5-moe":
            if cfg.get("moe_layer_interval", 1) != 1 or cfg.get("moe_layer_end_index", L - 1) not in (-1, L - 1):
                raise SystemExit("ernie4_5-moe: only checkpoints whose layers all route behind the dense lead layers are supported (as in the reference loader)")
            w.add(f"{a}.leading_dense_block_count", int(cfg.get("moe_layer_start_index", 0)), ValueType.UINT32)
            w.add(f"{a}.interleave_moe_layer_step", 1, ValueType.UINT32)
            if cfg.get("moe_num_shared_experts"):
                w.add(f"{a}.expert_shared_feed_forward_length", int(cfg["moe_num_shared_experts"]) * int(cfg["moe_intermediate_size"]), ValueType.UINT32)
        if a == "minimax-m2":
            w.add(f"{a}.expert_gating_func", 2, ValueType.UINT32)      # sigmoid scores, selection bias, renormalised weights
            w.add(f"{a}.expert_feed_forward_length", int(cfg["intermediate_size"]), ValueType.UINT32)
        if a in ("glm4moe", "dots1"):      # DeepSeek-V3 style routing: sigmoid scores, selection bias, scaled weights, dense lead layers, ungated shared experts
            w.add(f"{a}.expert_shared_count", int(cfg.get("n_shared_experts", 1)), ValueType.UINT32)
            w.add(f"{a}.leading_dense_block_count", int(cfg.get("first_k_dense_replace", 0)), ValueType.UINT32)
            w.add(f"{a}.expert_weights_scale", float(cfg.get("routed_scaling_factor", 1.0)), ValueType.FLOAT32)
            w.add(f"{a}.expert_weights_norm", bool(cfg.get("norm_topk_prob", True)))
            w.add(f"{a}.expert_gating_func", 2, ValueType.UINT32)
        if a == "qwen2moe" and cfg.get("shared_expert_intermediate_size"):
            w.add(f"{a}.expert_shared_feed_forward_length", int(cfg["shared_expert_intermediate_size"]), ValueType.UINT32)
    w.add(f"{a}.attention.head_count", H, ValueType.UINT32)
    w.add(f"{a}.attention.head_count_kv", Hkv, ValueType.UINT32)
    if a in LAYERNORM:
        w.add(f"{a}.attention.layer_norm_epsilon", float(cfg.get("norm_epsilon") or cfg.get("layer_norm_epsilon") or cfg.get("layer_norm_eps") or cfg.get("norm_eps") or 1e-5), ValueType.FLOAT32)
        if a == "command-r":
            w.add(f"{a}.logit_scale", float(cfg.get("logit_scale", 1.0)), ValueType.FLOAT32)
        if a == "olmo" and cfg.get("clip_qkv"):
            raise SystemExit("olmo checkpoints with clip_qkv are not supported (the fused RoPE / cache-write kernel does not clamp)")
    else:
        w.add(f"{a}.attention.layer_norm_rms_epsilon", float(cfg.get("rms_norm_eps", 1e-5)), ValueType.FLOAT32)
    w.add(f"{a}.rope.freq_base", theta, ValueType.FLOAT32)
    rot = (cfg.get("rotary_pct") or cfg.get("partial_rotary_factor") or rope.get("partial_rotary_factor")) if a in ("gptneox", "stablelm", "nemotron", "glm4", "phi2", "glm4moe", "minimax-m2") else None
    w.add(f"{a}.rope.dimension_count", int(D * float(rot)) if rot else int(cfg["rotary_dim"]) if (a == "minimax-m2" and cfg.get("rotary_dim")) else D, ValueType.UINT32)
    if a == "gptneox":
        w.add(f"{a}.use_parallel_residual", bool(cfg.get("use_parallel_residual", True)))
    if D * H != E:
        w.add(f"{a}.attention.key_length", D, ValueType.UINT32)
        w.add(f"{a}.attention.value_length", D, ValueType.UINT32)
    sc = cfg.get("rope_scaling") or (rope if rope.get("rope_type") not in (None, "default") else None)
    if sc:
        kind = sc.get("rope_type") or sc.get("type")
        if kind in ("linear", "yarn"):
            w.add(f"{a}.rope.scaling.type", kind)
            w.add(f"{a}.rope.scaling.factor", float(sc.get("factor", 1.0)), ValueType.FLOAT32)
            if kind == "yarn" and sc.get("original_max_position_embeddings"):
                w.add(f"{a}.rope.scaling.original_context_length", int(sc["original_max_position_embeddings"]), ValueType.UINT32)
    rope_factors = None
    if sc and (sc.get("rope_type") or sc.get("type")) == "llama3":
        # Llama-3.1 / 3.2 frequency-dependent scaling becomes a per-frequency divisor tensor (`rope_freqs.weight`), exactly as
        # the reference converter does: long wavelengths are divided by `factor`, short ones kept, a smooth ramp in between
        import math
        factor, lo, hi = float(sc.get("factor", 8.0)), float(sc.get("low_freq_factor", 1.0)), float(sc.get("high_freq_factor", 4.0))
        old_ctx = float(sc.get("original_max_position_embeddings", 8192))
        freqs = theta ** (-np.arange(0, D, 2, dtype=np.float64) / D)
        out = []
        for f in freqs:
            wavelen = 2 * math.pi / f
            if wavelen < old_ctx / hi:
                out.append(1.0)
            elif wavelen > old_ctx / lo:
                out.append(factor)
            else:
                smooth = (old_ctx / wavelen - lo) / (hi - lo)
                out.append(1.0 / ((1.0 - smooth) / factor + smooth))
        rope_factors = np.asarray(out, dtype="<f4")
    tok_kind = add_tokenizer(w, src, cfg["vocab_size"], cfg)
    tensors = load_safetensors(src)
    qtype = {"F32": GGMLType.F32, "F16": GGMLType.F16, "BF16": GGMLType.BF16, "Q8_0": GGMLType.Q8_0}[outtype]
    n = 0
    # routed experts: either fused 3-D tensors (transformers >= 5: experts.gate_up_proj [n, 2F, E], experts.down_proj [n, E, F]) or
    # one matrix per expert (experts.<e>.w1/w3/w2 or .gate_proj/.up_proj/.down_proj) -> stacked ffn_{gate,up,down}_exps
    moe: dict[str, dict[str, dict[int, np.ndarray]]] = {}
    for name in list(tensors):
        m = re.match(r"model\.layers\.(\d+)\.(?:mlp|block_sparse_moe)\.(.+)", name)
        if not m or not n_exp:
            continue
        il, rest = m.group(1), m.group(2)
        t = tensors[name]
        if rest.startswith("shared_experts."):       # glm4moe: always-on experts, no gate
            tensors[f"__moe__.blk.{il}.ffn_{rest.split('.')[1].split('_')[0]}_shexp.weight"] = tensors.pop(name)
        elif rest in ("gate.e_score_correction_bias", "gate.moe_statics.e_score_correction_bias", "moe_statics.e_score_correction_bias", "e_score_correction_bias"):
            tensors[f"__moe__.blk.{il}.exp_probs_b.bias"] = tensors.pop(name).reshape(-1)
        elif rest.startswith("shared_expert"):
            sh = {"shared_expert_gate.weight": "ffn_gate_inp_shexp.weight", "shared_expert.gate_proj.weight": "ffn_gate_shexp.weight",
                  "shared_expert.up_proj.weight": "ffn_up_shexp.weight", "shared_expert.down_proj.weight": "ffn_down_shexp.weight"}[rest]
            t2 = tensors.pop(name)
            tensors[f"__moe__.blk.{il}.{sh}"] = t2.reshape(-1) if sh == "ffn_gate_inp_shexp.weight" else t2
        elif rest in ("gate.weight", "router.layer.weight"):
            tensors[f"__moe__.blk.{il}.ffn_gate_inp.weight"] = tensors.pop(name)
        elif rest == "input_linear.weight":        # granitemoe: [n_expert, 2F, E] = gate rows then up rows
            g, u = t.float().chunk(2, dim=1)
            tensors[f"__moe__.blk.{il}.ffn_gate_exps.weight"], tensors[f"__moe__.blk.{il}.ffn_up_exps.weight"] = g.contiguous(), u.contiguous()
            tensors.pop(name)
        elif rest == "output_linear.weight":
            tensors[f"__moe__.blk.{il}.ffn_down_exps.weight"] = tensors.pop(name)
        elif rest == "experts.gate_up_proj":
            g, u = t.float().chunk(2, dim=1)
            tensors[f"__moe__.blk.{il}.ffn_gate_exps.weight"], tensors[f"__moe__.blk.{il}.ffn_up_exps.weight"] = g.contiguous(), u.contiguous()
            tensors.pop(name)
        elif rest == "experts.down_proj":
            tensors[f"__moe__.blk.{il}.ffn_down_exps.weight"] = tensors.pop(name)
        else:
            me = re.match(r"experts\.(\d+)\.(w1|w2|w3|gate_proj|up_proj|down_proj)\.weight", rest)
            if me:
                kind = {"w1": "gate", "gate_proj": "gate", "w3": "up", "up_proj": "up", "w2": "down", "down_proj": "down"}[me.group(2)]
                moe.setdefault(il, {}).setdefault(kind, {})[int(me.group(1))] = tensors.pop(name).float().numpy()
    for il, kinds in moe.items():
        for kind, per in kinds.items():
            import torch
            tensors[f"__moe__.blk.{il}.ffn_{kind}_exps.weight"] = torch.from_numpy(np.stack([per[e] for e in range(n_exp)]))
    for name in sorted(tensors, key=lambda k: (0 if k.startswith("model.embed") else 1, [int(x) if x.isdigit() else x for x in re.split(r"(\d+)", k)])):
        if name.endswith("rotary_emb.inv_freq"):
            continue
        if name.startswith("__moe__."):
            out = name[len("__moe__."):]
        elif name in TOP_MAP:
            out = TOP_MAP[name]
        elif a in ("gpt2", "starcoder"):
            m = re.match(r"transformer\.h\.(\d+)\.(.+)", name)
            if m and m.group(2) in ("attn.bias", "attn.masked_bias"):
                continue
            if not m or m.group(2) not in GPT2_MAP:
                raise SystemExit(f"unmapped tensor {name}")
            out = f"blk.{m.group(1)}.{GPT2_MAP[m.group(2)]}"
        elif a == "gptneox":
            m = re.match(r"gpt_neox\.layers\.(\d+)\.(.+)", name)
            if m and m.group(2) in ("attention.bias", "attention.masked_bias", "attention.rotary_emb.inv_freq"):
                continue
            if not m or m.group(2) not in NEOX_MAP:
                raise SystemExit(f"unmapped tensor {name}")
            out = f"blk.{m.group(1)}.{NEOX_MAP[m.group(2)]}"
        elif a == "falcon":
            m = re.match(r"transformer\.h\.(\d+)\.(.+)", name)
            if not m or m.group(2) not in FALCON_MAP:
                raise SystemExit(f"unmapped tensor {name}")
            out = f"blk.{m.group(1)}.{FALCON_MAP[m.group(2)]}"
        else:
            m = re.match(r"model\.layers\.(\d+)\.(.+)", name)
            if not m or m.group(2) not in LAYER_MAP:
                raise SystemExit(f"unmapped tensor {name}")
            out = f"blk.{m.group(1)}.{LAYER_MAP[m.group(2)]}"
        if name == "lm_head.weight" and cfg.get("tie_word_embeddings"):
            continue
        t = tensors[name].float().numpy()
        if a == "gpt2" and t.ndim == 2 and out.startswith("blk."):
            t = t.T             # GPT-2 keeps its projections as Conv1D: [in, out]
        if a == "falcon" and out.endswith("attn_qkv.weight"):
            # HF groups the rows per K/V head as [that group's query heads..., its key head, its value head]; GGUF wants [all Q | all K | all V]
            g = t.reshape(Hkv, H // Hkv + 2, D, E)
            t = np.concatenate([g[:, :-2].reshape(H * D, E), g[:, -2].reshape(Hkv * D, E), g[:, -1].reshape(Hkv * D, E)], 0)
        if a == "gptneox" and ".attn_qkv." in out:
            # HF interleaves per head [q_h | k_h | v_h]; GGUF wants [all Q | all K | all V] (weight rows and bias entries alike)
            g = t.reshape(H, 3, D, *t.shape[1:])
            t = np.concatenate([g[:, 0].reshape(H * D, *t.shape[1:]), g[:, 1].reshape(H * D, *t.shape[1:]), g[:, 2].reshape(H * D, *t.shape[1:])], 0)
        if a in ("olmo2", "exaone4"):        # no pre-norms: HF's two layer norms are the post-branch ones
            out = out.replace("ffn_norm.weight", "post_attention_norm.weight")
        if a == "nemotron" and out.endswith("_norm.weight"):
            t = t + 1.0          # LayerNorm1p: the checkpoint stores weight - 1
        if a in ("seed_oss", "glm4moe") and out.endswith("ffn_norm.weight"):
            out = out.replace("ffn_norm.weight", "post_attention_norm.weight")
        if a in PERMUTE_QK and out.endswith("attn_q.weight"):
            t = permute_qk(t, H)
        if a in PERMUTE_QK and out.endswith("attn_k.weight"):
            t = permute_qk(t, Hkv)
        ttype = qtype if t.ndim >= 2 and out != "position_embd.weight" and not out.endswith("_norm.weight") and not out.endswith("ffn_gate_inp.weight") and not out.endswith("ffn_gate_inp_shexp.weight") else GGMLType.F32
        if ttype == GGMLType.Q8_0 and t.shape[-1] % 32:
            ttype = GGMLType.F16
        arr = np.ascontiguousarray(t)
        w.add_tensor(out, arr.shape, ttype, lambda arr=arr, ttype=ttype: quantize(arr.reshape(-1, arr.shape[-1]) if arr.ndim > 1 else arr, ttype))
        n += 1
    if rope_factors is not None:
        w.add_tensor("rope_freqs.weight", rope_factors.shape, GGMLType.F32, lambda: rope_factors.view(np.uint8).reshape(-1))
    w.write()
    print(f"{src} ({hf_arch}) -> {dst}: arch {a}, {n} tensors, {outtype}, tokenizer {tok_kind}")
    return dst


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("model_dir")
    ap.add_argument("--outfile", required=True)
    ap.add_argument("--outtype", default="F16", choices=["F32", "F16", "BF16", "Q8_0"])
    a = ap.parse_args(argv)
    convert(a.model_dir, a.outfile, a.outtype)


if __name__ == "__main__":
    main()
