"""FP32 PyTorch oracle for the llama / mixtral forward pass.

This is the numerical reference every CUDA kernel and every parallel layout is tested against
(NMSE metric and thresholds as in `llama.cpp/tests/test-backend-ops.cpp:273-287,3716-3718`), and the
timed CPU baseline for config 1.  The graph it evaluates is the one the reference builds per ubatch
(`llama.cpp/src/models/llama.cpp:31-164`): RMSNorm -> Wq/Wk/Wv -> RoPE(Q,K) -> KV write -> causal
GQA attention -> Wo -> +res -> RMSNorm -> SwiGLU FFN | top-k MoE -> +res; final norm; lm_head.
MoE routing matches `build_moe_ffn` (`llama.cpp/src/llama-graph.cpp:1039-1139`): softmax over all
experts, top-k, re-normalised weights.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch

from ..gguf import GGUFModelFiles, open_gguf
from .config import ModelConfig, config_from_gguf


def nmse(a: torch.Tensor, b: torch.Tensor) -> float:
    """Normalised mean squared error of `b` against reference `a`."""
    a = a.detach().double().flatten().cpu()
    b = b.detach().double().flatten().cpu()
    return float(((a - b) ** 2).sum() / (a ** 2).sum().clamp_min(1e-30))


def rope_inv_freq(cfg: ModelConfig, freq_factors: torch.Tensor | None = None) -> torch.Tensor:
    i = torch.arange(0, cfg.rope_dim // 2, dtype=torch.float64)
    inv = cfg.rope_theta ** (-2.0 * i / cfg.rope_dim)
    if freq_factors is not None:
        inv = inv / freq_factors.double()
    if cfg.yarn_factor > 1.0:
        # YaRN: interpolate low frequencies by 1/factor, keep high frequencies, blend over the correction range
        # (`rope_yarn` / `ggml_rope_yarn_corr_dims`, ggml-cuda/rope.cu:15-40, ggml.c)
        def corr_dim(n_rot):
            return cfg.rope_dim * math.log(cfg.yarn_orig_ctx / (n_rot * 2 * math.pi)) / (2 * math.log(cfg.rope_theta))
        low = max(0.0, math.floor(corr_dim(cfg.yarn_beta_fast)))
        high = min(cfg.rope_dim - 1.0, math.ceil(corr_dim(cfg.yarn_beta_slow)))
        ramp = (1.0 - ((i - low) / max(0.001, high - low)).clamp(0, 1)) * cfg.yarn_ext_factor
        inv = inv / cfg.yarn_factor * (1 - ramp) + inv * ramp
    else:
        inv = inv * cfg.rope_scale_linear
    return inv.float()


def rope_attn_scale(cfg: ModelConfig) -> float:
    """Multiplier on the attention logits from YaRN's magnitude correction (cos/sin are scaled by
    mscale = attn_factor * (1 + 0.1 ln(factor)) on both Q and K, i.e. the scores by mscale^2)."""
    base = cfg.attention_scale * math.sqrt(cfg.head_dim) if cfg.attention_scale > 0.0 else 1.0   # granite: explicit score scale
    if cfg.yarn_factor <= 1.0 or cfg.yarn_ext_factor == 0.0:
        return base
    return base * (cfg.yarn_attn_factor * (1.0 + 0.1 * math.log(cfg.yarn_factor))) ** 2


def apply_rope(x: torch.Tensor, pos: torch.Tensor, inv_freq: torch.Tensor, neox: bool) -> torch.Tensor:
    """x: [T, H, D]; pos: [T]. Rotates the first 2*len(inv_freq) dims of every head."""
    n_rot = inv_freq.numel() * 2
    ang = pos.to(torch.float32)[:, None] * inv_freq[None, :]
    cos = ang.cos()[:, None, :]
    sin = ang.sin()[:, None, :]
    xr = x[..., :n_rot]
    if neox:
        x0, x1 = xr[..., : n_rot // 2], xr[..., n_rot // 2:]
        out = torch.cat([x0 * cos - x1 * sin, x0 * sin + x1 * cos], dim=-1)
    else:
        x0, x1 = xr[..., 0::2], xr[..., 1::2]
        out = torch.stack([x0 * cos - x1 * sin, x0 * sin + x1 * cos], dim=-1).flatten(-2)
    return torch.cat([out, x[..., n_rot:]], dim=-1)


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w


@dataclass
class RefKVCache:
    k: list[torch.Tensor] = field(default_factory=list)  # per layer [n_ctx, Hkv, D]
    v: list[torch.Tensor] = field(default_factory=list)
    length: int = 0


def limit_to_groups(sel: torch.Tensor, n_groups: int, n_used: int) -> torch.Tensor:
    """Group-limited expert selection (DeepSeek-V3 / bailingmoe2; llama-graph.cpp:1075-1098): the experts form `n_groups` equal groups, a
    group scores the sum of its two best selection scores, and only experts of the `n_used` best groups stay eligible."""
    if n_groups <= 1:
        return sel
    g = sel.reshape(*sel.shape[:-1], n_groups, -1)
    keep = g.topk(2, dim=-1).values.sum(-1).topk(n_used, dim=-1).indices
    mask = torch.zeros(g.shape[:-1], dtype=torch.bool, device=sel.device).scatter_(-1, keep, True)
    return g.masked_fill(~mask[..., None], float("-inf")).reshape(sel.shape)


class ReferenceModel:
    """Dense fp32 weights + eager forward.  Small models only (it dequantises everything)."""

    def __init__(self, cfg: ModelConfig, weights: dict[str, torch.Tensor]):
        self.cfg = cfg
        self.w = weights
        ff = weights.get("rope_freqs.weight")
        self.inv_freq = rope_inv_freq(cfg, ff)
        self.inv_freq_off = torch.zeros_like(self.inv_freq)      # zero angles = identity rotation (layers without RoPE)

    lora: dict = {}     # weight name -> (A [r, K], B [N, r], scale)
    observer = None     # fn(weight_name, x [T, K], expert | None): sees the input of every weight mat-mul (imatrix collection)

    def mm(self, x: torch.Tensor, name: str, expert: int | None = None) -> torch.Tensor:
        w = self.w[name] if expert is None else self.w[name][expert]
        if self.observer is not None:
            self.observer(name, x, expert)
        y = x @ w.T
        ad = self.lora.get(name) if (self.lora and expert is None) else None
        if ad is not None:          # trainable low-rank update: W + scale * B A  (tools/dlp_finetune.py --lora)
            a, b, scale = ad
            y = y + scale * ((x @ a.T) @ b.T)
        return y

    def norm(self, x: torch.Tensor, name: str) -> torch.Tensor:
        """RMSNorm, or for the LayerNorm families the mean-centred norm with whatever of `<name>.weight` / `.bias` the file has."""
        c, w = self.cfg, self.w
        if c.norm_type == "layer":
            return torch.nn.functional.layer_norm(x, (x.shape[-1],), w.get(name + ".weight"), w.get(name + ".bias"), c.rms_eps)
        return rms_norm(x, w[name + ".weight"], c.rms_eps)

    def layer_inv_freq(self, il: int) -> torch.Tensor:
        return self.inv_freq if self.cfg.layer_uses_rope(il) else self.inv_freq_off

    @classmethod
    def from_gguf(cls, path_or_files: str | GGUFModelFiles) -> "ReferenceModel":
        f = open_gguf(path_or_files) if isinstance(path_or_files, str) else path_or_files
        cfg = config_from_gguf(f.metadata, f.tensors)
        w = {name: torch.from_numpy(np.array(f.array(name), dtype=np.float32, copy=True)) for name in f.names()}
        return cls(cfg, w)

    def new_cache(self, n_ctx: int) -> RefKVCache:
        c = self.cfg
        return RefKVCache([torch.zeros(n_ctx, c.n_head_kv, c.head_dim) for _ in range(c.n_layer)],
                          [torch.zeros(n_ctx, c.n_head_kv, c.head_dim) for _ in range(c.n_layer)], 0)

    # ---- sub-blocks (exposed for per-op tests) --------------------------------------------
    def attention(self, il: int, x: torch.Tensor, pos: torch.Tensor, cache: RefKVCache) -> torch.Tensor:
        c, w = self.cfg, self.w
        p = f"blk.{il}."
        T = x.shape[0]
        if c.fused_qkv:   # phi3: one [q | k | v] matrix (llama.cpp/src/models/phi3.cpp)
            qkv = self.mm(x, p + "attn_qkv.weight")
            if p + "attn_qkv.bias" in w:      # gptneox
                qkv = qkv + w[p + "attn_qkv.bias"]
            q, k, v = qkv.split([c.q_dim, c.kv_dim, c.kv_dim], dim=-1)
        else:
            q, k, v = (self.mm(x, p + f"attn_{n}.weight") for n in "qkv")
        if c.qkv_bias and not c.fused_qkv:    # qwen2 (llama.cpp/src/models/qwen2.cpp)
            q, k, v = (t + w[p + f"attn_{n}.bias"] for t, n in zip((q, k, v), "qkv"))
        if c.qk_norm_full:    # olmo2: RMSNorm over the whole projected width, before the heads are split
            q = rms_norm(q, w[p + "attn_q_norm.weight"], c.rms_eps)
            k = rms_norm(k, w[p + "attn_k_norm.weight"], c.rms_eps)
        q = q.reshape(T, c.n_head, c.head_dim)
        k = k.reshape(T, c.n_head_kv, c.head_dim)
        v = v.reshape(T, c.n_head_kv, c.head_dim)
        if c.qk_norm:     # qwen3: RMSNorm over head_dim, before RoPE (llama.cpp/src/models/qwen3.cpp)
            q = rms_norm(q, w[p + "attn_q_norm.weight"], c.rms_eps)
            k = rms_norm(k, w[p + "attn_k_norm.weight"], c.rms_eps)
        q = apply_rope(q, pos, self.layer_inv_freq(il), c.rope_neox)
        k = apply_rope(k, pos, self.layer_inv_freq(il), c.rope_neox)
        start = cache.length
        cache.k[il][start:start + T] = k
        cache.v[il][start:start + T] = v
        K = cache.k[il][: start + T]
        V = cache.v[il][: start + T]
        Kx = K.repeat_interleave(c.n_gqa, dim=1)
        Vx = V.repeat_interleave(c.n_gqa, dim=1)
        scores = torch.einsum("thd,shd->hts", q, Kx) * (rope_attn_scale(c) / math.sqrt(c.head_dim))
        kpos = torch.arange(start + T)
        mask = kpos[None, :] > pos[:, None]
        scores = scores.masked_fill(mask[None], float("-inf"))
        att = torch.softmax(scores, dim=-1)
        o = torch.einsum("hts,shd->thd", att, Vx).reshape(T, c.q_dim)
        if p + "attn_sub_norm.weight" in w:       # bitnet: the attention result is normalised before the output matrix
            o = rms_norm(o, w[p + "attn_sub_norm.weight"], c.rms_eps)
        out = self.mm(o, p + "attn_output.weight")
        return out + w[p + "attn_output.bias"] if p + "attn_output.bias" in w else out

    def ffn(self, il: int, x: torch.Tensor) -> torch.Tensor:
        c, w = self.cfg, self.w
        p = f"blk.{il}."
        if c.ffn_act in ("gelu", "relu2"):   # non-gated MLP with optional biases; GELU in its tanh form, as ggml computes it
            up = self.mm(x, p + "ffn_up.weight")
            if p + "ffn_up.bias" in w:
                up = up + w[p + "ffn_up.bias"]
            a = torch.nn.functional.gelu(up, approximate="tanh") if c.ffn_act == "gelu" else torch.relu(up) ** 2
            dn = self.mm(a, p + "ffn_down.weight")
            return dn + w[p + "ffn_down.bias"] if p + "ffn_down.bias" in w else dn
        if c.fused_gate_up:   # phi3: ffn_up holds [gate | up]; swiglu = silu(first half) * second half
            g, u = self.mm(x, p + "ffn_up.weight").split([c.n_ff, c.n_ff], dim=-1)
            return self.mm(torch.nn.functional.silu(g) * u, p + "ffn_down.weight")
        if not c.is_moe or p + "ffn_gate_inp.weight" not in w:      # dense model, or one of the dense lead layers of deepseek / bailingmoe
            g = self.mm(x, p + "ffn_gate.weight")
            u = self.mm(x, p + "ffn_up.weight")
            a = torch.nn.functional.silu(g) * u
            if p + "ffn_sub_norm.weight" in w:    # bitnet
                a = rms_norm(a, w[p + "ffn_sub_norm.weight"], c.rms_eps)
            return self.mm(a, p + "ffn_down.weight")
        logits = self.mm(x, p + "ffn_gate_inp.weight")
        probs = torch.sigmoid(logits) if c.moe_gating == "sigmoid" else torch.softmax(logits, dim=-1)
        sel = probs + w[p + "exp_probs_b.bias"] if p + "exp_probs_b.bias" in w else probs     # the bias steers which experts are picked ...
        sel = limit_to_groups(sel, c.n_expert_groups, c.n_group_used)
        topi = sel.topk(c.n_expert_used, dim=-1).indices
        topw = probs.gather(-1, topi)                                                          # ... their weights are the unbiased scores
        if c.moe_norm_topk:
            topw = topw / topw.sum(-1, keepdim=True).clamp_min(6.103515625e-5)
        if c.moe_scale_w:
            topw = topw * c.expert_weights_scale
        out = torch.zeros_like(x)
        if c.n_ff_shared:   # qwen2moe: shared expert, scaled by sigmoid(x . gate_vector)  (llama.cpp/src/models/qwen2moe.cpp)
            g = self.mm(x, p + "ffn_gate_shexp.weight")
            u = self.mm(x, p + "ffn_up_shexp.weight")
            sh = self.mm(torch.nn.functional.silu(g) * u, p + "ffn_down_shexp.weight")
            gate = torch.sigmoid(x @ w[p + "ffn_gate_inp_shexp.weight"].reshape(-1, 1)) if p + "ffn_gate_inp_shexp.weight" in w else 1.0
            out = out + gate * sh
        for e in range(c.n_expert):
            sel = (topi == e)
            rows = sel.any(-1).nonzero().flatten()
            if rows.numel() == 0:
                continue
            xe = x[rows]
            g = self.mm(xe, p + "ffn_gate_exps.weight", e)
            u = self.mm(xe, p + "ffn_up_exps.weight", e)
            y = self.mm(torch.nn.functional.silu(g) * u, p + "ffn_down_exps.weight", e)
            we = (topw[rows] * sel[rows]).sum(-1, keepdim=True)
            out[rows] += we * y
        return out

    def layer(self, il: int, h: torch.Tensor, pos: torch.Tensor, cache: RefKVCache) -> torch.Tensor:
        c, w = self.cfg, self.w
        p = f"blk.{il}."
        if c.parallel_residual:     # falcon, command-r: both branches read the block input (falcon-40b: through two different norms)
            xn = self.norm(h, p + "attn_norm")
            if p + "ffn_norm.weight" in w:      # gptneox: attention reads attn_norm(h), the MLP ffn_norm(h)
                return h + self.attention(il, xn, pos, cache) + self.ffn(il, self.norm(h, p + "ffn_norm"))
            xa = self.norm(h, p + "attn_norm_2") if p + "attn_norm_2.weight" in w else xn
            return h + self.attention(il, xa, pos, cache) + self.ffn(il, xn)
        if c.post_norm:             # olmo2 (no pre-norm) / glm4 (both): the branch output is normalised before the residual add
            a = self.attention(il, self.norm(h, p + "attn_norm") if c.pre_norm else h, pos, cache)
            h = h + self.norm(a, p + "post_attention_norm")
            f = self.ffn(il, self.norm(h, p + "ffn_norm") if c.pre_norm else h)
            return h + self.norm(f, p + "post_ffw_norm")
        h = h + c.residual_scale * self.attention(il, self.norm(h, p + "attn_norm"), pos, cache)
        h = h + c.residual_scale * self.ffn(il, self.norm(h, p + c.ffn_norm_name[: -len(".weight")]))
        return h

    def embed(self, tokens: torch.Tensor, pos: torch.Tensor | None = None) -> torch.Tensor:
        h = self.w["token_embd.weight"][tokens] * self.cfg.embedding_scale
        if self.cfg.learned_pos:        # gpt2, starcoder: learned absolute positions (llama.cpp/src/models/gpt2.cpp:20-25)
            h = h + self.w["position_embd.weight"][torch.arange(tokens.numel()) if pos is None else pos]
        return h

    def head(self, h: torch.Tensor) -> torch.Tensor:
        c, w = self.cfg, self.w
        out_w = w.get("output.weight", w["token_embd.weight"])
        xn = self.norm(h, "output_norm")
        if self.observer is not None:
            self.observer("output.weight" if "output.weight" in w else "token_embd.weight", xn, None)
        lg = xn @ out_w.T / c.logit_scale
        return lg + w["output.bias"] if "output.bias" in w else lg      # phi2

    # ---- full forward ---------------------------------------------------------------------
    def forward(self, tokens, cache: RefKVCache, layers: range | None = None, hidden: torch.Tensor | None = None,
                return_hidden: bool = False) -> torch.Tensor:
        with torch.no_grad():
            return self.forward_grad(tokens, cache, layers, hidden, return_hidden)

    def forward_grad(self, tokens, cache: RefKVCache, layers: range | None = None, hidden: torch.Tensor | None = None,
                     return_hidden: bool = False) -> torch.Tensor:
        """Run `tokens` (appended at cache.length) through `layers` (default: all) and return logits
        [T, vocab] - or the hidden state when this is a non-final pipeline stage."""
        c = self.cfg
        tokens = torch.as_tensor(tokens, dtype=torch.long)
        T = tokens.numel()
        pos = torch.arange(cache.length, cache.length + T)
        layers = range(c.n_layer) if layers is None else layers
        h = self.embed(tokens, pos) if hidden is None else hidden
        for il in layers:
            h = self.layer(il, h, pos, cache)
        last_stage = len(layers) == 0 or layers[-1] == c.n_layer - 1
        if return_hidden or not last_stage:
            return h
        return self.head(h)

    def advance(self, cache: RefKVCache, n: int):
        cache.length += n

    @torch.no_grad()
    def generate_greedy(self, prompt: list[int], n_predict: int, n_ctx: int = 512) -> tuple[list[int], torch.Tensor]:
        cache = self.new_cache(n_ctx)
        logits = self.forward(prompt, cache)
        self.advance(cache, len(prompt))
        out = []
        last = logits[-1]
        all_logits = [last]
        for _ in range(n_predict):
            tok = int(last.argmax())
            out.append(tok)
            last = self.forward([tok], cache)[-1]
            self.advance(cache, 1)
            all_logits.append(last)
        return out, torch.stack(all_logits)
