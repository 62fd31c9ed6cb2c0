"""Random-init GGUF checkpoint synthesis.

Writes loadable llama/mixtral-architecture GGUF files in F16 / BF16 / Q8_0 / Q4_K_M with the tensor and
KV names the reference expects (`llama.cpp/src/llama-arch.cpp:128-337`), so that the *same file* can be
served by this engine and by the reference's `llama-bench` for A/B runs.  The Q4_K_M type mix follows
the public recipe (`llama.cpp/src/llama-quant.cpp:302-405`): mostly Q4_K, `output.weight` Q6_K, and
"more bits" (Q6_K) for attn_v / ffn_down in the first/last eighth and every third layer in between.

Two payload modes:
  * ``exact=True``  - draw float weights, then quantise (small models; oracle tests need real values);
  * ``exact=False`` - synthesise valid random blocks directly (multi-GB models; no float tensors exist).
"""
from __future__ import annotations

import os

import numpy as np

from ..models.config import ModelConfig
from .constants import BLOCK_GEOMETRY, FTYPE, GGMLType, TokenType, ValueType
from .quants import quantize, random_blocks
from .writer import GGUFWriter


def use_more_bits(i_layer: int, n_layer: int) -> bool:
    return i_layer < n_layer // 8 or i_layer >= 7 * n_layer // 8 or (i_layer - n_layer // 8) % 3 == 2


def tensor_type_for(name: str, ftype: str, i_layer: int, cfg: ModelConfig) -> GGMLType:
    """Per-tensor storage type for a given file type."""
    if name.endswith("_norm.weight") or name in ("rope_freqs.weight", "position_embd.weight") or name.endswith("ffn_gate_inp.weight") \
            or name.endswith("ffn_gate_inp_shexp.weight") \
            or name.endswith(".bias"):
        return GGMLType.F32
    if ftype == "F32":
        return GGMLType.F32
    if ftype == "F16":
        return GGMLType.F16
    if ftype == "BF16":
        return GGMLType.BF16
    if ftype == "Q8_0":
        return GGMLType.Q8_0
    if ftype == "Q6_K":
        return GGMLType.Q6_K
    if ftype in ("Q4_0", "Q4_1", "Q5_0", "Q5_1"):      # legacy block types: uniform (decoded to fp16 at load by the engine)
        return GGMLType[ftype]
    if ftype in ("Q2_K", "Q3_K"):                      # low-bit K types keep the output matrix at Q6_K like the reference
        return GGMLType.Q6_K if name == "output.weight" else GGMLType[ftype]
    if ftype in ("Q4_K_M", "Q5_K_M"):
        base = GGMLType.Q4_K if ftype == "Q4_K_M" else GGMLType.Q5_K
        if name == "output.weight":
            return GGMLType.Q6_K
        if name == "token_embd.weight":
            return base
        if ".attn_v." in name:
            if use_more_bits(i_layer, cfg.n_layer):
                return GGMLType.Q6_K
            if cfg.n_layer >= 80 and base == GGMLType.Q4_K:  # 70B: shared KV heads get Q5_K
                return GGMLType.Q5_K
            return base
        if ".ffn_down" in name and use_more_bits(i_layer, cfg.n_layer):
            return GGMLType.Q6_K
        return base
    if ftype in ("IQ4_NL", "IQ4_XS", "TQ1_0", "TQ2_0"):     # decoded to fp16 at load; embeddings / output keep more bits as in the reference
        if name == "output.weight":
            return GGMLType.Q6_K
        if name == "token_embd.weight":
            return GGMLType.Q4_K if ftype.startswith("TQ") else GGMLType[ftype]
        if ftype == "IQ4_XS" and ".attn_v." in name and cfg.n_expert == 0 and use_more_bits(i_layer, cfg.n_layer):
            return GGMLType.Q5_K
        return GGMLType[ftype]
    if ftype == "MXFP4_MOE":                                # expert matrices in MXFP4, everything else Q8_0 (llama-quant.cpp)
        return GGMLType.MXFP4 if "_exps." in name else GGMLType.Q8_0
    raise ValueError(f"unknown file type {ftype!r}")


def synthetic_vocab(n_vocab: int) -> tuple[list[str], np.ndarray, np.ndarray]:
    """A self-consistent SentencePiece-style vocabulary: specials, 256 byte tokens, then word pieces."""
    tokens = ["<unk>", "<s>", "</s>"]
    types = [TokenType.UNKNOWN, TokenType.CONTROL, TokenType.CONTROL]
    scores = [0.0, 0.0, 0.0]
    for b in range(256):
        tokens.append(f"<0x{b:02X}>")
        types.append(TokenType.BYTE)
        scores.append(0.0)
    syll = ["a", "e", "i", "o", "u", "ka", "ri", "to", "me", "su", "no", "la", "ve", "di", "po", "an", "er", "is",
            "th", "in", "on", "st", "ch", "qu", "ly", "ng", "ed", "es", "or", "al"]
    singles = [chr(c) for c in range(32, 127)]
    idx = 0
    while len(tokens) < n_vocab:
        if idx < len(singles):
            piece = singles[idx]
            piece = "▁" if piece == " " else piece
        else:
            k = idx - len(singles)
            parts = []
            n = k
            for _ in range(1 + (k % 3)):
                parts.append(syll[n % len(syll)])
                n //= len(syll)
            piece = ("▁" if k % 2 == 0 else "") + "".join(parts) + (str(k) if k >= len(syll) ** 2 else "")
        idx += 1
        if piece in ("<unk>", "<s>", "</s>"):
            continue
        tokens.append(piece)
        types.append(TokenType.NORMAL)
        scores.append(-float(len(tokens)))
    # de-duplicate while preserving count
    seen: dict[str, int] = {}
    for i, t in enumerate(tokens):
        if t in seen:
            tokens[i] = f"{t}#{i}"
        seen[tokens[i]] = i
    return tokens[:n_vocab], np.asarray(scores[:n_vocab], np.float32), np.asarray(types[:n_vocab], np.int32)


def norm_layout(cfg: ModelConfig) -> dict:
    """Which norm tensors a family's files carry: {"attn": "wb" | "w" | "", "ffn": ..., "out": ...} ("" = the tensor is absent)."""
    if cfg.norm_type != "layer":
        return {"attn": "w", "ffn": "w", "out": "w"}
    return {"starcoder2": {"attn": "wb", "ffn": "wb", "out": "wb"}, "orion": {"attn": "wb", "ffn": "wb", "out": "wb"},
            "gptneox": {"attn": "wb", "ffn": "wb", "out": "wb"}, "stablelm": {"attn": "wb", "ffn": "wb", "out": "wb"},
            "nemotron": {"attn": "wb", "ffn": "wb", "out": "wb"}, "gpt2": {"attn": "wb", "ffn": "wb", "out": "wb"},
            "starcoder": {"attn": "wb", "ffn": "wb", "out": "wb"}, "codeshell": {"attn": "wb", "ffn": "wb", "out": "wb"},
            "falcon": {"attn": "wb", "ffn": "", "out": "wb"}, "command-r": {"attn": "w", "ffn": "", "out": "w"},
            "olmo": {"attn": "", "ffn": "", "out": ""}}.get(cfg.arch, {"attn": "wb", "ffn": "" if cfg.parallel_residual else "wb", "out": "wb"})


def _tensor_plan(cfg: ModelConfig, rope_freqs: bool):
    """(name, shape, layer, init_std) for every tensor of the architecture; shapes row-major [out, in]."""
    E, F = cfg.n_embd, cfg.n_ff
    nl = norm_layout(cfg)
    all_bias = cfg.arch in ("starcoder2", "gptneox", "gpt2", "starcoder", "codeshell", "phi2")       # biases on the attention output and the MLP as well

    def norm(name, il, kind):
        return ([(name + ".weight", (E,), il, 0.0)] if "w" in kind else []) + ([(name + ".bias", (E,), il, 0.05)] if "b" in kind else [])
    plan = [("token_embd.weight", (cfg.vocab_size, E), -1, 0.02)]
    if cfg.learned_pos:
        plan.append(("position_embd.weight", (cfg.n_ctx_train, E), -1, 0.02))
    if rope_freqs:
        plan.append(("rope_freqs.weight", (cfg.rope_dim // 2,), -1, 0.0))
    for il in range(cfg.n_layer):
        p = f"blk.{il}."
        if cfg.pre_norm:
            plan += norm(p + "attn_norm", il, nl["attn"])
        if cfg.fused_qkv:
            plan.append((p + "attn_qkv.weight", (cfg.q_dim + 2 * cfg.kv_dim, E), il, 0.02))
            if cfg.qkv_bias:
                plan.append((p + "attn_qkv.bias", (cfg.q_dim + 2 * cfg.kv_dim,), il, 0.05))
        else:
            plan += [(p + "attn_q.weight", (cfg.q_dim, E), il, 0.02),
                     (p + "attn_k.weight", (cfg.kv_dim, E), il, 0.02),
                     (p + "attn_v.weight", (cfg.kv_dim, E), il, 0.02)]
        if cfg.qkv_bias and not cfg.fused_qkv:
            plan += [(p + "attn_q.bias", (cfg.q_dim,), il, 0.05), (p + "attn_k.bias", (cfg.kv_dim,), il, 0.05),
                     (p + "attn_v.bias", (cfg.kv_dim,), il, 0.05)]
        if cfg.qk_norm:
            plan += [(p + "attn_q_norm.weight", (cfg.head_dim,), il, 0.0), (p + "attn_k_norm.weight", (cfg.head_dim,), il, 0.0)]
        if cfg.qk_norm_full:
            plan += [(p + "attn_q_norm.weight", (cfg.q_dim,), il, 0.0), (p + "attn_k_norm.weight", (cfg.kv_dim,), il, 0.0)]
        if cfg.sub_norm:
            plan.append((p + "attn_sub_norm.weight", (cfg.q_dim,), il, 0.0))
        plan.append((p + "attn_output.weight", (E, cfg.q_dim), il, 0.02))
        if all_bias or cfg.attn_out_bias:
            plan.append((p + "attn_output.bias", (E,), il, 0.05))
        if cfg.post_norm:
            plan.append((p + "post_attention_norm.weight", (E,), il, 0.0))
        if cfg.norm_type == "layer":
            plan += norm(p + "ffn_norm", il, nl["ffn"])
        elif cfg.pre_norm and not (cfg.parallel_residual and cfg.arch == "plamo"):
            plan.append((p + cfg.ffn_norm_name, (E,), il, 0.0))
        if cfg.ffn_act in ("gelu", "relu2"):       # non-gated MLP
            plan += [(p + "ffn_up.weight", (F, E), il, 0.02), (p + "ffn_down.weight", (E, F), il, 0.02)]
            if all_bias:
                plan += [(p + "ffn_up.bias", (F,), il, 0.05), (p + "ffn_down.bias", (E,), il, 0.05)]
        elif cfg.fused_gate_up:
            plan += [(p + "ffn_up.weight", (2 * F, E), il, 0.02), (p + "ffn_down.weight", (E, F), il, 0.02)]
        elif cfg.is_moe and not cfg.layer_is_moe(il):      # deepseek / bailingmoe / glm4moe: dense lead layers with their own width (ernie4_5-moe: interleaved too)
            Fd = cfg.n_ff_dense or F
            plan += [(p + "ffn_gate.weight", (Fd, E), il, 0.02), (p + "ffn_down.weight", (E, Fd), il, 0.02), (p + "ffn_up.weight", (Fd, E), il, 0.02)]
        elif cfg.is_moe:
            plan += [
                (p + "ffn_gate_inp.weight", (cfg.n_expert, E), il, 0.02),
                *([(p + "exp_probs_b.bias", (cfg.n_expert,), il, 0.05)] if cfg.moe_sel_bias else []),
                (p + "ffn_gate_exps.weight", (cfg.n_expert, F, E), il, 0.02),
                (p + "ffn_down_exps.weight", (cfg.n_expert, E, F), il, 0.02),
                (p + "ffn_up_exps.weight", (cfg.n_expert, F, E), il, 0.02),
            ]
            if cfg.n_ff_shared:     # qwen2moe shared expert + its scalar gate
                Fs = cfg.n_ff_shared
                if cfg.arch == "qwen2moe":
                    plan.append((p + "ffn_gate_inp_shexp.weight", (E,), il, 0.0))
                plan += [(p + "ffn_gate_shexp.weight", (Fs, E), il, 0.02), (p + "ffn_up_shexp.weight", (Fs, E), il, 0.02),
                         (p + "ffn_down_shexp.weight", (E, Fs), il, 0.02)]
        else:
            plan += [
                (p + "ffn_gate.weight", (F, E), il, 0.02),
                (p + "ffn_down.weight", (E, F), il, 0.02),
                (p + "ffn_up.weight", (F, E), il, 0.02),
            ]
            if cfg.sub_norm:
                plan.append((p + "ffn_sub_norm.weight", (F,), il, 0.0))
        if cfg.post_norm:
            plan.append((p + "post_ffw_norm.weight", (E,), il, 0.0))
    plan += norm("output_norm", -1, nl["out"])
    if not cfg.tie_embeddings:
        plan.append(("output.weight", (cfg.vocab_size, E), -1, 0.02))
    if cfg.arch == "phi2":
        plan.append(("output.bias", (cfg.vocab_size,), -1, 0.05))
    return plan


def write_synthetic_gguf(path: str, cfg: ModelConfig, ftype: str = "F16", seed: int = 0, exact: bool | None = None,
                         rope_freqs: bool = False, vocab: bool = True) -> str:
    """Write a random-init checkpoint for `cfg` at `path` and return the path."""
    if exact is None:
        exact = cfg.n_params() < 300_000_000
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    w = GGUFWriter(path)
    a = cfg.arch
    w.add("general.architecture", a)
    w.add("general.name", f"{cfg.name}-random-init")
    w.add("general.file_type", FTYPE[ftype], ValueType.UINT32)
    w.add("general.quantization_version", 2, ValueType.UINT32)
    w.add(f"{a}.vocab_size", cfg.vocab_size, ValueType.UINT32)
    w.add(f"{a}.context_length", cfg.n_ctx_train, ValueType.UINT32)
    w.add(f"{a}.embedding_length", cfg.n_embd, ValueType.UINT32)
    w.add(f"{a}.block_count", cfg.n_layer, ValueType.UINT32)
    w.add(f"{a}.feed_forward_length", (cfg.n_ff_dense or cfg.n_ff) * (2 if a == "qwen" else 1), ValueType.UINT32)
    w.add(f"{a}.attention.head_count", cfg.n_head, ValueType.UINT32)
    w.add(f"{a}.attention.head_count_kv", cfg.n_head_kv, ValueType.UINT32)
    w.add(f"{a}.attention.layer_norm_epsilon" if cfg.norm_type == "layer" else f"{a}.attention.layer_norm_rms_epsilon", float(cfg.rms_eps), ValueType.FLOAT32)
    w.add(f"{a}.rope.freq_base", float(cfg.rope_theta), ValueType.FLOAT32)
    w.add(f"{a}.rope.dimension_count", cfg.rope_dim, ValueType.UINT32)
    if cfg.head_dim * cfg.n_head != cfg.n_embd:
        w.add(f"{a}.attention.key_length", cfg.head_dim, ValueType.UINT32)
        w.add(f"{a}.attention.value_length", cfg.head_dim, ValueType.UINT32)
    if cfg.embedding_scale != 1.0:
        w.add(f"{a}.embedding_scale", float(cfg.embedding_scale), ValueType.FLOAT32)
    if cfg.residual_scale != 1.0:
        w.add(f"{a}.residual_scale", float(cfg.residual_scale), ValueType.FLOAT32)
    if cfg.logit_scale != 1.0:      # command-r multiplies the logits by its key, the granite family divides by it
        w.add(f"{a}.logit_scale", float(1.0 / cfg.logit_scale if a == "command-r" else cfg.logit_scale), ValueType.FLOAT32)
    if cfg.attention_scale > 0.0:
        w.add(f"{a}.attention.scale", float(cfg.attention_scale), ValueType.FLOAT32)
    if cfg.yarn_factor > 1.0:
        w.add(f"{a}.rope.scaling.type", "yarn")
        w.add(f"{a}.rope.scaling.factor", float(cfg.yarn_factor), ValueType.FLOAT32)
        w.add(f"{a}.rope.scaling.original_context_length", int(cfg.yarn_orig_ctx), ValueType.UINT32)
    elif cfg.rope_scale_linear != 1.0:
        w.add(f"{a}.rope.scaling.type", "linear")
        w.add(f"{a}.rope.scaling.factor", float(1.0 / cfg.rope_scale_linear), ValueType.FLOAT32)
    if a == "gptneox":
        w.add(f"{a}.use_parallel_residual", bool(cfg.parallel_residual))
    if a in ("qwen3vl", "qwen3vlmoe"):
        w.add(f"{a}.n_deepstack_layers", 0, ValueType.UINT32)
    if a in ("qwen2vl", "qwen3vl", "qwen3vlmoe"):      # temporal / height / width / extra shares of the rotary pairs (text tokens use one position for all of them)
        q = cfg.rope_dim // 8
        w.add(f"{a}.rope.dimension_sections", np.asarray([q, (cfg.rope_dim // 2 - q) // 2, cfg.rope_dim // 2 - q - (cfg.rope_dim // 2 - q) // 2, 0], np.int32), etype=ValueType.INT32)
    if cfg.is_moe:
        if a != "llama":
            w.add(f"{a}.expert_feed_forward_length", cfg.n_ff, ValueType.UINT32)
        w.add(f"{a}.expert_count", cfg.n_expert, ValueType.UINT32)
        w.add(f"{a}.expert_used_count", cfg.n_expert_used, ValueType.UINT32)
        if cfg.n_expert_groups > 1:
            w.add(f"{a}.expert_group_count", cfg.n_expert_groups, ValueType.UINT32)
            w.add(f"{a}.expert_group_used_count", cfg.n_group_used, ValueType.UINT32)
        if cfg.n_ff_shared and a not in ("deepseek", "bailingmoe", "glm4moe", "dots1"):
            w.add(f"{a}.expert_shared_feed_forward_length", cfg.n_ff_shared, ValueType.UINT32)
        if a == "ernie4_5-moe":
            w.add(f"{a}.leading_dense_block_count", cfg.n_layer_dense_lead, ValueType.UINT32)
            w.add(f"{a}.interleave_moe_layer_step", cfg.moe_layer_step, ValueType.UINT32)
        if a in ("deepseek", "bailingmoe", "glm4moe", "dots1", "bailingmoe2"):
            w.add(f"{a}.leading_dense_block_count", cfg.n_layer_dense_lead, ValueType.UINT32)
            w.add(f"{a}.expert_shared_count", cfg.n_ff_shared // cfg.n_ff, ValueType.UINT32)
            w.add(f"{a}.expert_weights_scale", float(cfg.expert_weights_scale), ValueType.FLOAT32)
            if a in ("bailingmoe", "glm4moe", "dots1", "bailingmoe2"):
                w.add(f"{a}.expert_weights_norm", bool(cfg.moe_norm_topk))
        if a in ("glm4moe", "dots1", "minimax-m2", "bailingmoe2"):
            w.add(f"{a}.expert_gating_func", 2 if cfg.moe_gating == "sigmoid" else 1, ValueType.UINT32)
    if vocab:
        toks, scores, types = synthetic_vocab(cfg.vocab_size)
        w.add("tokenizer.ggml.model", "llama")
        w.add("tokenizer.ggml.pre", "default")
        w.add("tokenizer.ggml.tokens", toks, etype=ValueType.STRING)
        w.add("tokenizer.ggml.scores", scores, etype=ValueType.FLOAT32)
        w.add("tokenizer.ggml.token_type", types, etype=ValueType.INT32)
        w.add("tokenizer.ggml.bos_token_id", 1, ValueType.UINT32)
        w.add("tokenizer.ggml.eos_token_id", 2, ValueType.UINT32)
        w.add("tokenizer.ggml.unknown_token_id", 0, ValueType.UINT32)
        w.add("tokenizer.ggml.add_bos_token", True)
        w.add("tokenizer.ggml.add_eos_token", False)

    pools: dict = {}

    def make_producer(name, shape, il, std, ttype, tseed):
        def produce():
            rng = np.random.default_rng(tseed)
            n = int(np.prod(shape))
            if std == 0.0:
                if name == "rope_freqs.weight":
                    return np.ones(shape, np.float32).view(np.uint8).reshape(-1)
                return (1.0 + 0.1 * rng.standard_normal(n, dtype=np.float32)).astype("<f4").view(np.uint8)
            if name.endswith(".bias"):
                return (rng.standard_normal(n, dtype=np.float32) * std).astype("<f4").view(np.uint8)
            if exact:
                x = rng.standard_normal(n, dtype=np.float32) * std
                return quantize(x.reshape(-1, shape[-1]), ttype)
            # multi-GB checkpoints: draw one pool of valid random blocks per type and fill each tensor
            # with a rotated tiling of it (bandwidth-bound instead of RNG-bound; still random-init)
            per, nbytes = BLOCK_GEOMETRY[ttype]
            pool_elems = 1 << 24
            if n <= pool_elems:
                return random_blocks(ttype, n, rng, std)
            key = (ttype, std)
            if key not in pools:
                pools[key] = random_blocks(ttype, pool_elems, np.random.default_rng(seed + 77), std
                                           ).reshape(-1, nbytes)
            pool = pools[key]
            shift = int(rng.integers(0, pool.shape[0]))
            rolled = np.concatenate([pool[shift:], pool[:shift]])
            return np.resize(rolled, (n // per, nbytes)).reshape(-1)
        return produce

    for i, (name, shape, il, std) in enumerate(_tensor_plan(cfg, rope_freqs)):
        ttype = tensor_type_for(name, ftype, max(il, 0), cfg)
        w.add_tensor(name, shape, ttype, make_producer(name, shape, il, std, ttype, seed * 100003 + i))
    return w.write()
