"""Model hyper-parameters and presets.

Parity: hparams are read from the same GGUF KV names the reference uses
(`llama.cpp/src/llama-arch.cpp:128-268`); presets are public architecture facts (TinyLlama-1.1B,
Llama-3-8B/70B, Mixtral-8x7B) for the five BASELINE.json configurations.
"""
from __future__ import annotations

import math
from dataclasses import asdict, dataclass, field, replace
from typing import Any


@dataclass(frozen=True)
class ModelConfig:
    name: str = "llama"
    arch: str = "llama"
    vocab_size: int = 32000
    n_embd: int = 4096
    n_layer: int = 32
    n_head: int = 32
    n_head_kv: int = 8
    head_dim: int = 128
    n_ff: int = 14336
    n_ctx_train: int = 8192
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_dim: int = 128
    rope_neox: bool = False  # llama arch = adjacent-pair ("norm") rotation
    rope_scale_linear: float = 1.0
    # YaRN context extension (`llama.cpp/src/llama-context.cpp:82-132`, kernel `ggml-cuda/rope.cu:15-40`): factor > 1 enables it
    yarn_factor: float = 1.0
    yarn_orig_ctx: int = 0
    yarn_ext_factor: float = 1.0
    yarn_attn_factor: float = 1.0
    yarn_beta_fast: float = 32.0
    yarn_beta_slow: float = 1.0
    n_expert: int = 0
    n_expert_used: int = 0
    bos_id: int = 1
    eos_id: int = 2
    tie_embeddings: bool = False
    qkv_bias: bool = False     # qwen2: bias on Wq/Wk/Wv
    qk_norm: bool = False      # qwen3: per-head RMSNorm on Q and K before RoPE
    fused_qkv: bool = False    # phi3: one attn_qkv tensor and one [2*n_ff] ffn_up (gate | up) tensor in the file
    # granite: four scalar multipliers on top of the llama graph (`llama.cpp/src/models/granite.cpp`, keys `llama-arch.cpp`)
    embedding_scale: float = 1.0   # h = embed(tok) * embedding_scale
    residual_scale: float = 1.0    # h += residual_scale * block(h)
    logit_scale: float = 1.0       # logits /= logit_scale
    attention_scale: float = 0.0   # replaces 1/sqrt(head_dim) when > 0
    # qwen2moe: routed experts without re-normalised top-k weights plus an always-on shared expert behind a sigmoid gate
    moe_norm_topk: bool = True     # renormalise the selected experts' weights (Mixtral, qwen3moe); qwen2moe: False
    n_ff_shared: int = 0           # > 0: shared-expert FFN width (`ffn_*_shexp`; qwen2moe also has the gate vector `ffn_gate_inp_shexp`)
    # deepseek (v1), bailingmoe: the first layers are dense (their own FFN width), routed weights are multiplied by a constant, and the
    # shared experts are added without a gate (`llama.cpp/src/models/deepseek.cpp:86-120`)
    n_layer_dense_lead: int = 0
    moe_layer_step: int = 1        # ernie4_5-moe: behind the lead layers only every step-th layer routes, the others are dense
    n_ff_dense: int = 0            # FFN width of the dense lead layers (0: n_ff)
    expert_weights_scale: float = 1.0      # read and written back; applied only with moe_scale_w (the deepseek-v1 / bailingmoe builders pass scale_w = false)
    # glm4moe (GLM-4.5), dots1: DeepSeek-V3 style routing - sigmoid scores, a per-expert bias added for the SELECTION only, weights = the
    # unbiased scores of the selected experts (renormalised when the file says so) times expert_weights_scale (llama-graph.cpp:1037-1143)
    sub_norm: bool = False         # bitnet: one more RMSNorm in front of the attention output matrix and of the MLP's down matrix
    moe_gating: str = "softmax"    # "softmax" | "sigmoid"
    moe_sel_bias: bool = False     # `blk.N.exp_probs_b.bias`
    moe_scale_w: bool = False
    n_expert_groups: int = 0       # > 1: experts are picked from the n_group_used best groups only (group score = its two best experts; bailingmoe2)
    n_group_used: int = 0
    # LayerNorm / GELU / parallel-residual families (starcoder2, falcon, olmo, orion, command-r): the generic layer path
    norm_type: str = "rms"         # "rms" | "layer" (mean-centred; weight and bias optional per file: olmo has neither)
    ffn_act: str = "swiglu"        # "swiglu" (gated SiLU) | "gelu" | "relu2" (non-gated: ffn_up -> activation -> ffn_down, biases optional)
    parallel_residual: bool = False  # h + attn(n(h)) + ffn(n'(h)) instead of the sequential form (falcon, command-r)
    fused_gate_up: bool = False    # phi3, glm4: ffn_up holds [gate | up]
    # olmo2, glm4: each branch output is normalised before it joins the residual stream (`post_attention_norm`, `post_ffw_norm`);
    # olmo2 has no norm in front of the branches at all and normalises Q and K over their full width
    attn_out_bias: bool = False    # a bias on the attention output projection (pangu-embedded and the LayerNorm families that carry one)
    post_norm: bool = False
    pre_norm: bool = True
    qk_norm_full: bool = False

    @property
    def generic(self) -> bool:
        """True for families that run the unfused layer path (models/llama.py `_layer_generic`)."""
        return (self.norm_type != "rms" or self.ffn_act != "swiglu" or self.parallel_residual or self.post_norm or not self.pre_norm or self.attn_out_bias
                or self.qk_norm_full or self.sub_norm)

    # smollm3: every `no_rope_step`-th layer attends without rotary position encoding (`llama.cpp/src/models/smollm3.cpp:26`)
    no_rope_step: int = 0
    learned_pos: bool = False      # gpt2, starcoder: h = tok_embd[id] + position_embd[pos], no rotary anywhere

    @property
    def ffn_norm_name(self) -> str:
        """Tensor holding the pre-FFN norm weight: seed_oss files call it post_attention_norm (`llama.cpp/src/models/seed-oss.cpp:86`)."""
        return "post_attention_norm.weight" if self.arch in ("seed_oss", "glm4moe") else "ffn_norm.weight"

    def layer_is_moe(self, il: int) -> bool:
        return self.n_expert > 0 and il >= self.n_layer_dense_lead and (il + 1) % max(1, self.moe_layer_step) == 0

    def layer_uses_rope(self, il: int) -> bool:
        return self.no_rope_step <= 0 or (il + 1) % self.no_rope_step != 0

    @property
    def n_gqa(self) -> int:
        return self.n_head // self.n_head_kv

    @property
    def q_dim(self) -> int:
        return self.n_head * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_head_kv * self.head_dim

    @property
    def is_moe(self) -> bool:
        return self.n_expert > 0

    def n_params(self) -> int:
        attn = self.n_embd * (self.q_dim + 2 * self.kv_dim) + self.q_dim * self.n_embd
        ffn = 3 * self.n_embd * self.n_ff * max(1, self.n_expert) + (self.n_embd * self.n_expert if self.is_moe else 0)
        emb = self.vocab_size * self.n_embd * (1 if self.tie_embeddings else 2)
        return self.n_layer * (attn + ffn + 2 * self.n_embd) + emb + self.n_embd

    def to_dict(self) -> dict[str, Any]:
        return asdict(self)


PRESETS: dict[str, ModelConfig] = {
    # tiny shapes for CPU tests / smoke
    "tiny": ModelConfig(name="tiny", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64,
                        n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),
    "tiny128": ModelConfig(name="tiny128", vocab_size=1024, n_embd=512, n_layer=2, n_head=4, n_head_kv=2,
                           head_dim=128, n_ff=1024, n_ctx_train=1024, rope_theta=10000.0, rope_dim=128),
    # enough heads / width for every tensor-parallel degree up to 8 (per-rank K of the row-parallel mat-vecs stays a multiple of 256)
    "tiny-tp8": ModelConfig(name="tiny-tp8", vocab_size=2048, n_embd=2048, n_layer=2, n_head=16, n_head_kv=8,
                            head_dim=128, n_ff=2048, n_ctx_train=1024, rope_theta=10000.0, rope_dim=128),
    "tiny-moe-tp": ModelConfig(name="tiny-moe-tp", vocab_size=512, n_embd=512, n_layer=2, n_head=8, n_head_kv=4, head_dim=64, n_ff=512,
                               n_ctx_train=512, rope_theta=10000.0, rope_dim=64, n_expert=4, n_expert_used=2),
    "tiny-moe": ModelConfig(name="tiny-moe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                            head_dim=64, n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, n_expert=4,
                            n_expert_used=2),
    # one tiny config per extra architecture family (llama.cpp/src/models/{qwen2,qwen3,phi3}.cpp)
    "tiny-qwen2": ModelConfig(name="tiny-qwen2", arch="qwen2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                              head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True,
                              qkv_bias=True),
    "tiny-qwen3": ModelConfig(name="tiny-qwen3", arch="qwen3", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                              head_dim=128, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=128, rope_neox=True,
                              qk_norm=True),
    "tiny-qwen2moe": ModelConfig(name="tiny-qwen2moe", arch="qwen2moe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                 head_dim=64, n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qkv_bias=True,
                                 n_expert=4, n_expert_used=2, moe_norm_topk=False, n_ff_shared=512),
    "tiny-qwen3moe": ModelConfig(name="tiny-qwen3moe", arch="qwen3moe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                 head_dim=64, n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm=True,
                                 n_expert=4, n_expert_used=2),
    "qwen3-30b-a3b": ModelConfig(name="qwen3-30b-a3b", arch="qwen3moe", vocab_size=151936, n_embd=2048, n_layer=48, n_head=32, n_head_kv=4,
                                 head_dim=128, n_ff=768, n_ctx_train=40960, rms_eps=1e-6, rope_theta=1000000.0, rope_dim=128,
                                 rope_neox=True, qk_norm=True, n_expert=128, n_expert_used=8, bos_id=151643, eos_id=151645),
    "tiny-granite": ModelConfig(name="tiny-granite", arch="granite", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, embedding_scale=12.0,
                                residual_scale=0.22, logit_scale=8.0, attention_scale=0.015625),
    "granite-3-8b": ModelConfig(name="granite-3-8b", arch="granite", vocab_size=49155, n_embd=4096, n_layer=40, n_head=32, n_head_kv=8,
                                head_dim=128, n_ff=12800, n_ctx_train=4096, rope_theta=10000.0, rope_dim=128, embedding_scale=12.0,
                                residual_scale=0.22, logit_scale=16.0, attention_scale=0.0078125),
    # the llama graph under other architecture names (one tiny config each; see SUPPORTED_ARCHS)
    "tiny-internlm2": ModelConfig(name="tiny-internlm2", arch="internlm2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                  head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),
    "tiny-xverse": ModelConfig(name="tiny-xverse", arch="xverse", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                               head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),
    "tiny-ernie": ModelConfig(name="tiny-ernie", arch="ernie4_5", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                              head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),
    "tiny-exaone": ModelConfig(name="tiny-exaone", arch="exaone", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                               head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True),
    "tiny-seed-oss": ModelConfig(name="tiny-seed-oss", arch="seed_oss", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                 head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qkv_bias=True),
    "tiny-minicpm": ModelConfig(name="tiny-minicpm", arch="minicpm", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, embedding_scale=12.0,
                                residual_scale=0.7, logit_scale=2.0),
    "tiny-granitemoe": ModelConfig(name="tiny-granitemoe", arch="granitemoe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                   head_dim=64, n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, embedding_scale=12.0,
                                   residual_scale=0.22, logit_scale=8.0, attention_scale=0.015625, n_expert=4, n_expert_used=2),
    "tiny-smollm3": ModelConfig(name="tiny-smollm3", arch="smollm3", vocab_size=512, n_embd=256, n_layer=4, n_head=4, n_head_kv=2,
                                head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, no_rope_step=4),
    "tiny-phi3": ModelConfig(name="tiny-phi3", arch="phi3", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                             head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True,
                             fused_qkv=True, fused_gate_up=True),
    # LayerNorm families
    "tiny-starcoder2": ModelConfig(name="tiny-starcoder2", arch="starcoder2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                   head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qkv_bias=True,
                                   norm_type="layer", ffn_act="gelu"),
    "tiny-falcon": ModelConfig(name="tiny-falcon", arch="falcon", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=1,
                               head_dim=64, n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, fused_qkv=True,
                               norm_type="layer", ffn_act="gelu", parallel_residual=True),
    "tiny-olmo": ModelConfig(name="tiny-olmo", arch="olmo", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                             head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, norm_type="layer"),
    "tiny-orion": ModelConfig(name="tiny-orion", arch="orion", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                              head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, norm_type="layer"),
    "tiny-gptneox": ModelConfig(name="tiny-gptneox", arch="gptneox", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                                head_dim=64, n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=16, rope_neox=True, fused_qkv=True, qkv_bias=True,
                                norm_type="layer", ffn_act="gelu", parallel_residual=True),
    "tiny-phi2": ModelConfig(name="tiny-phi2", arch="phi2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4,
                             head_dim=64, n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, fused_qkv=True, qkv_bias=True,
                             norm_type="layer", ffn_act="gelu", parallel_residual=True),
    "tiny-stablelm": ModelConfig(name="tiny-stablelm", arch="stablelm", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                 head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, norm_type="layer"),
    "tiny-nemotron": ModelConfig(name="tiny-nemotron", arch="nemotron", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                 head_dim=64, n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, norm_type="layer", ffn_act="relu2"),
    "tiny-arcee": ModelConfig(name="tiny-arcee", arch="arcee", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                              head_dim=64, n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, ffn_act="relu2"),
    "tiny-gpt2": ModelConfig(name="tiny-gpt2", arch="gpt2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4, head_dim=64, n_ff=1024,
                             n_ctx_train=512, rope_dim=64, rope_neox=True, fused_qkv=True, qkv_bias=True, norm_type="layer", ffn_act="gelu", no_rope_step=1,
                             learned_pos=True, tie_embeddings=True),
    "tiny-starcoder": ModelConfig(name="tiny-starcoder", arch="starcoder", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=1, head_dim=64,
                                  n_ff=1024, n_ctx_train=512, rope_dim=64, rope_neox=True, fused_qkv=True, qkv_bias=True, norm_type="layer", ffn_act="gelu",
                                  no_rope_step=1, learned_pos=True, tie_embeddings=True),
    "tiny-codeshell": ModelConfig(name="tiny-codeshell", arch="codeshell", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64,
                                  n_ff=1024, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, fused_qkv=True, qkv_bias=True,
                                  norm_type="layer", ffn_act="gelu"),
    "tiny-qwen": ModelConfig(name="tiny-qwen", arch="qwen", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4, head_dim=64, n_ff=512,
                             n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, fused_qkv=True, qkv_bias=True),
    "tiny-baichuan": ModelConfig(name="tiny-baichuan", arch="baichuan", vocab_size=512, n_embd=128, n_layer=32, n_head=2, n_head_kv=2, head_dim=64,
                                 n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),     # 32 layers: the reference keys the 7B graph on that
    "tiny-olmo2": ModelConfig(name="tiny-olmo2", arch="olmo2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                              n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, post_norm=True, pre_norm=False, qk_norm_full=True),
    "tiny-glm4": ModelConfig(name="tiny-glm4", arch="glm4", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                             n_ctx_train=512, rope_theta=10000.0, rope_dim=32, qkv_bias=True, fused_gate_up=True, post_norm=True),
    "tiny-mistral3": ModelConfig(name="tiny-mistral3", arch="mistral3", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64,
                                 n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64),
    "tiny-pangu": ModelConfig(name="tiny-pangu", arch="pangu-embedded", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64,
                              n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qkv_bias=True, attn_out_bias=True),
    "tiny-chatglm": ModelConfig(name="tiny-chatglm", arch="chatglm", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                                n_ctx_train=512, rope_theta=10000.0, rope_dim=32, fused_qkv=True, qkv_bias=True, fused_gate_up=True),
    "tiny-plamo": ModelConfig(name="tiny-plamo", arch="plamo", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4, head_dim=64, n_ff=512,
                              n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, parallel_residual=True),
    "tiny-deepseek": ModelConfig(name="tiny-deepseek", arch="deepseek", vocab_size=512, n_embd=256, n_layer=3, n_head=4, n_head_kv=4, head_dim=64, n_ff=128,
                                 n_ctx_train=512, rope_theta=10000.0, rope_dim=64, n_expert=4, n_expert_used=2, moe_norm_topk=False, n_ff_shared=256,
                                 n_layer_dense_lead=1, n_ff_dense=512, expert_weights_scale=1.5),
    "tiny-bailingmoe": ModelConfig(name="tiny-bailingmoe", arch="bailingmoe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=128,
                                   n_ctx_train=512, rope_theta=10000.0, rope_dim=64, n_expert=4, n_expert_used=2, moe_norm_topk=True, n_ff_shared=128,
                                   expert_weights_scale=2.0),
    # qwen2vl, text only: multi-section rotary with the same position in every section is the plain half-split rotation
    "tiny-qwen2vl": ModelConfig(name="tiny-qwen2vl", arch="qwen2vl", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                                n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qkv_bias=True),
    "tiny-qwen3vl": ModelConfig(name="tiny-qwen3vl", arch="qwen3vl", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                                n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm=True),
    "tiny-qwen3vlmoe": ModelConfig(name="tiny-qwen3vlmoe", arch="qwen3vlmoe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64,
                                   n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm=True, n_expert=4, n_expert_used=2),
    "tiny-olmoe": ModelConfig(name="tiny-olmoe", arch="olmoe", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=4, head_dim=64, n_ff=256,
                              n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm_full=True, n_expert=4, n_expert_used=2,
                              moe_norm_topk=False),
    "tiny-glm4moe": ModelConfig(name="tiny-glm4moe", arch="glm4moe", vocab_size=512, n_embd=256, n_layer=3, n_head=4, n_head_kv=2, head_dim=64, n_ff=256,
                                n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, qkv_bias=True, qk_norm=True, n_expert=4, n_expert_used=2,
                                moe_norm_topk=True, n_ff_shared=256, n_layer_dense_lead=1, n_ff_dense=512, expert_weights_scale=1.5,
                                moe_gating="sigmoid", moe_sel_bias=True, moe_scale_w=True),
    "tiny-dots1": ModelConfig(name="tiny-dots1", arch="dots1", vocab_size=512, n_embd=256, n_layer=3, n_head=4, n_head_kv=4, head_dim=64, n_ff=256,
                              n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm=True, n_expert=4, n_expert_used=2,
                              moe_norm_topk=True, n_ff_shared=256, n_layer_dense_lead=1, n_ff_dense=512, expert_weights_scale=2.5,
                              moe_gating="sigmoid", moe_sel_bias=True, moe_scale_w=True),
    "tiny-ernie-moe": ModelConfig(name="tiny-ernie-moe", arch="ernie4_5-moe", vocab_size=512, n_embd=256, n_layer=4, n_head=4, n_head_kv=2, head_dim=64, n_ff=256,
                                  n_ctx_train=512, rope_theta=10000.0, rope_dim=64, n_expert=4, n_expert_used=2, moe_norm_topk=True, n_ff_shared=256,
                                  n_layer_dense_lead=1, n_ff_dense=512, moe_sel_bias=True),
    "tiny-minimax-m2": ModelConfig(name="tiny-minimax-m2", arch="minimax-m2", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=256,
                                   n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, qk_norm_full=True, n_expert=4, n_expert_used=2,
                                   moe_norm_topk=True, moe_gating="sigmoid", moe_sel_bias=True),
    "tiny-bailingmoe2": ModelConfig(name="tiny-bailingmoe2", arch="bailingmoe2", vocab_size=512, n_embd=256, n_layer=3, n_head=4, n_head_kv=2, head_dim=64,
                                    n_ff=256, n_ctx_train=512, rope_theta=10000.0, rope_dim=32, rope_neox=True, fused_qkv=True, qk_norm=True, n_expert=8,
                                    n_expert_used=2, moe_norm_topk=True, n_ff_shared=256, n_layer_dense_lead=1, n_ff_dense=512, expert_weights_scale=2.5,
                                    moe_gating="sigmoid", moe_sel_bias=True, moe_scale_w=True, n_expert_groups=4, n_group_used=2),
    "tiny-bitnet": ModelConfig(name="tiny-bitnet", arch="bitnet", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                               n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, sub_norm=True, tie_embeddings=True),
    "tiny-exaone4": ModelConfig(name="tiny-exaone4", arch="exaone4", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2, head_dim=64, n_ff=512,
                                n_ctx_train=512, rope_theta=10000.0, rope_dim=64, rope_neox=True, qk_norm=True, post_norm=True, pre_norm=False),
    "tiny-command-r": ModelConfig(name="tiny-command-r", arch="command-r", vocab_size=512, n_embd=256, n_layer=2, n_head=4, n_head_kv=2,
                                  head_dim=64, n_ff=512, n_ctx_train=512, rope_theta=10000.0, rope_dim=64, norm_type="layer",
                                  parallel_residual=True, logit_scale=16.0, tie_embeddings=True),
    "qwen2.5-7b": ModelConfig(name="qwen2.5-7b", arch="qwen2", vocab_size=152064, n_embd=3584, n_layer=28, n_head=28,
                              n_head_kv=4, head_dim=128, n_ff=18944, n_ctx_train=32768, rms_eps=1e-6, rope_theta=1000000.0,
                              rope_dim=128, rope_neox=True, qkv_bias=True, bos_id=151643, eos_id=151645),
    "qwen3-8b": ModelConfig(name="qwen3-8b", arch="qwen3", vocab_size=151936, n_embd=4096, n_layer=36, n_head=32,
                            n_head_kv=8, head_dim=128, n_ff=12288, n_ctx_train=40960, rms_eps=1e-6, rope_theta=1000000.0,
                            rope_dim=128, rope_neox=True, qk_norm=True, bos_id=151643, eos_id=151645),
    "mistral-7b": ModelConfig(name="mistral-7b", vocab_size=32000, n_embd=4096, n_layer=32, n_head=32, n_head_kv=8,
                              head_dim=128, n_ff=14336, n_ctx_train=32768, rope_theta=10000.0, rope_dim=128),
    "phi-3-medium": ModelConfig(name="phi-3-medium", arch="phi3", vocab_size=32064, n_embd=5120, n_layer=40, n_head=40,
                                n_head_kv=10, head_dim=128, n_ff=17920, n_ctx_train=4096, rope_theta=10000.0, rope_dim=128,
                                rope_neox=True, fused_qkv=True),
    "tinyllama-1.1b": ModelConfig(name="tinyllama-1.1b", vocab_size=32000, n_embd=2048, n_layer=22, n_head=32,
                                  n_head_kv=4, head_dim=64, n_ff=5632, n_ctx_train=2048, rope_theta=10000.0,
                                  rope_dim=64),
    "llama-3-8b": ModelConfig(name="llama-3-8b", vocab_size=128256, n_embd=4096, n_layer=32, n_head=32, n_head_kv=8,
                              head_dim=128, n_ff=14336, n_ctx_train=8192, rope_theta=500000.0, rope_dim=128,
                              bos_id=128000, eos_id=128001),
    "llama-3-70b": ModelConfig(name="llama-3-70b", vocab_size=128256, n_embd=8192, n_layer=80, n_head=64,
                               n_head_kv=8, head_dim=128, n_ff=28672, n_ctx_train=8192, rope_theta=500000.0,
                               rope_dim=128, bos_id=128000, eos_id=128001),
    "mixtral-8x7b": ModelConfig(name="mixtral-8x7b", vocab_size=32000, n_embd=4096, n_layer=32, n_head=32,
                                n_head_kv=8, head_dim=128, n_ff=14336, n_ctx_train=32768, rope_theta=1000000.0,
                                rope_dim=128, n_expert=8, n_expert_used=2),
}


# GGUF `general.architecture` values this engine runs.  llama covers LLaMA 1-3, Mistral and Mixtral (MoE tensors present);
# qwen2 adds Q/K/V biases, qwen3 per-head Q/K RMSNorm (qwen3moe: the same attention with Mixtral-style routed experts of
# `expert_feed_forward_length`), phi3 fused QKV / gate-up tensors; all but llama rotate half-split
# ("neox") pairs (`llama.cpp/src/llama-model.cpp` llama_model_rope_type).
# The llama graph under other names (reference builders differ from llama.cpp/src/models/llama.cpp only in what is listed):
#   internlm2, xverse, ernie4_5 - none (adjacent-pair rotation);   exaone, seed_oss - half-split rotation (seed_oss: optional QKV bias);
#   minicpm, granitemoe - the granite builder (four scalars; granitemoe with routed experts);   smollm3 - no rotation on every 4th layer.
SUPPORTED_ARCHS = {"llama", "qwen2", "qwen2moe", "qwen3", "qwen3moe", "phi3", "granite",
                   "internlm2", "xverse", "ernie4_5", "exaone", "seed_oss", "minicpm", "granitemoe", "smollm3",
                   "starcoder2", "falcon", "olmo", "orion", "command-r", "gptneox", "stablelm", "nemotron", "arcee",
                   "gpt2", "starcoder", "codeshell", "qwen", "baichuan", "olmo2", "glm4", "mistral3", "pangu-embedded", "chatglm", "plamo", "deepseek", "bailingmoe", "qwen2vl", "qwen3vl", "exaone4", "phi2", "olmoe", "qwen3vlmoe", "glm4moe", "dots1", "ernie4_5-moe", "minimax-m2", "bailingmoe2", "bitnet"}
ROPE_NORM_ARCHS = {"llama", "granite", "granitemoe", "minicpm", "internlm2", "xverse", "ernie4_5", "ernie4_5-moe", "smollm3", "olmo", "command-r", "arcee", "baichuan", "glm4", "mistral3", "chatglm", "deepseek", "bailingmoe"}     # llama_model_rope_type
# LayerNorm families and what differs beyond the norm (llama.cpp/src/models/{starcoder2,falcon,olmo,orion,command-r}.cpp):
#   starcoder2: biases on every matrix, non-gated GELU MLP;  falcon: fused QKV, GELU MLP, attention and MLP in parallel from the same
#   (or a second) norm;  olmo: norms without parameters;  orion: llama with LayerNorm;  command-r: parallel, logits * logit_scale
#   gptneox (pythia): fused QKV with bias, rotary on part of each head, GELU MLP with biases, parallel blocks with a second norm for the MLP
#   stablelm: rotary on part of each head, gated SiLU MLP
#   nemotron: rotary on part of each head, non-gated ReLU^2 MLP;   arcee keeps RMSNorm but has the same non-gated ReLU^2 MLP
#   gpt2, starcoder (v1): learned position embeddings instead of rotary (starcoder: multi-query), fused QKV, biases everywhere, GELU MLP
#   codeshell: the same block with rotary.   (qwen v1 and baichuan-7B are RMSNorm / SwiGLU graphs and run the fused path.)
#   phi2: one LayerNorm feeds attention and MLP in parallel, fused QKV, biases everywhere including the output matrix, rotary on part of each head
LAYERNORM_ARCHS = {"starcoder2", "falcon", "olmo", "orion", "command-r", "gptneox", "stablelm", "nemotron", "gpt2", "starcoder", "codeshell", "phi2"}
LEARNED_POS_ARCHS = {"gpt2", "starcoder"}


def preset(name: str, **overrides) -> ModelConfig:
    if name not in PRESETS:
        raise KeyError(f"unknown model preset {name!r}; known: {sorted(PRESETS)}")
    return replace(PRESETS[name], **overrides)


def config_from_gguf(meta: dict[str, Any], tensors: dict | None = None) -> ModelConfig:
    """Build a ModelConfig from GGUF metadata (keys as in `llama-arch.cpp:128-268`)."""
    arch = meta.get("general.architecture", "llama")
    if arch not in SUPPORTED_ARCHS:
        raise NotImplementedError(f"architecture {arch!r} not supported (supported: {sorted(SUPPORTED_ARCHS)})")

    def g(key, default=None):
        return meta.get(f"{arch}.{key}", default)

    n_embd = int(g("embedding_length"))
    n_head = int(g("attention.head_count"))
    n_head_kv = int(g("attention.head_count_kv", n_head))
    head_dim = int(g("attention.key_length", n_embd // n_head))
    vocab = g("vocab_size")
    if vocab is None:
        toks = meta.get("tokenizer.ggml.tokens")
        if toks is not None:
            vocab = len(toks)
        elif tensors is not None and "token_embd.weight" in tensors:
            vocab = tensors["token_embd.weight"].shape[0]
        else:
            raise ValueError("cannot determine vocab size")
    scale = 1.0
    if meta.get(f"{arch}.rope.scaling.type", "none") == "linear":
        scale = 1.0 / float(meta.get(f"{arch}.rope.scaling.factor", 1.0))
    yarn = {}
    if meta.get(f"{arch}.rope.scaling.type", "none") == "yarn":
        yarn = dict(yarn_factor=float(g("rope.scaling.factor", 1.0)),
                    yarn_orig_ctx=int(g("rope.scaling.original_context_length", g("context_length", 2048))),
                    yarn_ext_factor=float(g("rope.scaling.yarn_ext_factor", 1.0)), yarn_attn_factor=float(g("rope.scaling.attn_factor", 1.0)),
                    yarn_beta_fast=float(g("rope.scaling.yarn_beta_fast", 32.0)), yarn_beta_slow=float(g("rope.scaling.yarn_beta_slow", 1.0)))
    tie = tensors is not None and "output.weight" not in tensors
    n_layer = int(g("block_count")) - int(g("nextn_predict_layers", 0) or 0)      # trailing multi-token-prediction layers are not part of the forward pass (glm4-moe.cpp:24)
    if arch == "exaone4" and (n_layer == 64 or int(g("attention.sliding_window", 0) or 0) > 0):
        raise NotImplementedError("exaone4 checkpoints with sliding-window layers (the 32B model) are not supported")
    if arch == "bitnet" and tensors is not None and any(k.endswith(".scale") for k in tensors):
        raise NotImplementedError("bitnet files with separate per-matrix scale tensors (I2_S exports) are not supported: use a TQ1_0 / TQ2_0 / F16 file")
    if arch == "mistral3" and float(g("attention.temperature_scale", 0.0) or 0.0) != 0.0:
        raise NotImplementedError("mistral3 checkpoints with a position-dependent attention temperature are not supported")
    if tensors is not None and "blk.0.ffn_down.weight" in tensors and int(g("feed_forward_length")) // (2 if arch == "qwen" else 1) != tensors["blk.0.ffn_down.weight"].shape[-1]:
        raise ValueError(f"{arch}.feed_forward_length = {g('feed_forward_length')} but blk.0.ffn_down.weight has {tensors['blk.0.ffn_down.weight'].shape[-1]} "
                         "input columns: the file's metadata does not describe its tensors")
    # minicpm: defaults of older exports (llama-model.cpp:756-770), overridden by the keys when present
    dflt = dict(embedding_scale=12.0, residual_scale=1.4 / math.sqrt(n_layer), logit_scale=256.0 / n_embd) if arch == "minicpm" else \
        dict(embedding_scale=1.0, residual_scale=1.0, logit_scale=1.0)
    return ModelConfig(
        name=str(meta.get("general.name", arch)), arch=arch, vocab_size=int(vocab), n_embd=n_embd,
        n_layer=n_layer, n_head=n_head, n_head_kv=n_head_kv, head_dim=head_dim,
        n_ff=int(g("expert_feed_forward_length") or g("feed_forward_length")) if int(g("expert_count", 0) or 0) > 0
        else int(g("feed_forward_length")) // (2 if arch == "qwen" else 1),      # qwen v1 files store twice the MLP width (llama-model.cpp: n_ff/2)
        n_ctx_train=int(g("context_length", 2048)),
        rms_eps=float(g("attention.layer_norm_epsilon", 1e-5) if arch in LAYERNORM_ARCHS else g("attention.layer_norm_rms_epsilon", 1e-5)),
        norm_type="layer" if arch in LAYERNORM_ARCHS else "rms", ffn_act="gelu" if arch in ("starcoder2", "falcon", "gptneox", "gpt2", "starcoder", "codeshell", "phi2") else "relu2" if arch in ("nemotron", "arcee") else "swiglu",
        attn_out_bias=tensors is not None and "blk.0.attn_output.bias" in tensors and arch not in LAYERNORM_ARCHS,
        post_norm=arch in ("olmo2", "glm4", "exaone4"), pre_norm=arch not in ("olmo2", "exaone4"), qk_norm_full=arch in ("olmo2", "olmoe", "minimax-m2"),
        parallel_residual=arch in ("falcon", "command-r", "plamo", "phi2") or (arch == "gptneox" and bool(g("use_parallel_residual", True)))
        or (arch == "stablelm" and tensors is not None and "blk.0.ffn_norm.weight" not in tensors), fused_gate_up=arch in ("phi3", "glm4", "chatglm"), rope_theta=float(g("rope.freq_base", 10000.0)),
        rope_dim=int(g("rope.dimension_count", head_dim)), rope_neox=arch not in ROPE_NORM_ARCHS, rope_scale_linear=scale,
        embedding_scale=float(g("embedding_scale") or dflt["embedding_scale"]), residual_scale=float(g("residual_scale") or dflt["residual_scale"]),
        logit_scale=(1.0 / float(g("logit_scale") or 1.0)) if arch == "command-r" else float(g("logit_scale") or dflt["logit_scale"]), attention_scale=float(g("attention.scale", 0.0) or 0.0),
        no_rope_step=4 if arch == "smollm3" else 1 if arch in LEARNED_POS_ARCHS else 0, learned_pos=arch in LEARNED_POS_ARCHS,
        moe_gating="sigmoid" if ((arch == "glm4moe" and int(g("expert_gating_func", 0) or 0) in (0, 2)) or (arch in ("dots1", "minimax-m2", "bailingmoe2") and int(g("expert_gating_func", 0) or 0) == 2)) else "softmax",
        moe_sel_bias=tensors is not None and any(k.endswith(".exp_probs_b.bias") for k in tensors), moe_scale_w=arch in ("glm4moe", "dots1", "bailingmoe2"),
        n_expert_groups=int(g("expert_group_count", 0) or 0), n_group_used=int(g("expert_group_used_count", 0) or 0),
        moe_layer_step=int(g("interleave_moe_layer_step", 1) or 1), sub_norm=arch == "bitnet",
        moe_norm_topk=bool(g("expert_weights_norm", False)) if arch in ("bailingmoe", "glm4moe", "dots1", "bailingmoe2") else arch not in ("qwen2moe", "deepseek", "olmoe"),
        n_ff_shared=(int(g("expert_shared_feed_forward_length", 0) or g("expert_feed_forward_length")) * int(g("expert_shared_count", 1) or 1)) if arch == "bailingmoe2" else
        int(g("expert_shared_feed_forward_length", 0) or 0) or (int(g("expert_shared_count", 0) or 0) * int(g("expert_feed_forward_length", 0) or 0)),
        n_layer_dense_lead=int(g("leading_dense_block_count", 0) or 0), expert_weights_scale=float(g("expert_weights_scale", 1.0) or 1.0),
        n_ff_dense=int(g("feed_forward_length")) if (int(g("leading_dense_block_count", 0) or 0) > 0 or int(g("interleave_moe_layer_step", 1) or 1) > 1) else 0,
        qkv_bias=arch in ("qwen2", "qwen2moe", "qwen2vl") or (tensors is not None and ("blk.0.attn_q.bias" in tensors or "blk.0.attn_qkv.bias" in tensors)),
        qk_norm=arch in ("qwen3", "qwen3moe", "qwen3vl", "qwen3vlmoe") or (arch not in ("olmo2", "olmoe", "minimax-m2") and tensors is not None and "blk.0.attn_q_norm.weight" in tensors),
        fused_qkv=arch == "phi3" or (tensors is not None and "blk.0.attn_qkv.weight" in tensors),
        n_expert=int(g("expert_count", 0) or 0), n_expert_used=int(g("expert_used_count", 0) or 0),
        bos_id=int(meta.get("tokenizer.ggml.bos_token_id", 1)), eos_id=int(meta.get("tokenizer.ggml.eos_token_id", 2)),
        tie_embeddings=tie, **yarn)
