"""Shard-aware GGUF -> device weight loading for llama / mixtral checkpoints.

Parity: replaces `llama_model::load_tensors` + `load_all_data` (`llama.cpp/src/llama-model.cpp:2465-2550`,
`llama-model-loader.cpp:957-1197`).  The reference's client streams every tensor to its workers (RPC SET_TENSOR)
or splits rows per device inside one process; here every rank opens the same mmapped file and uploads only
  * the layers of its pipeline stage (layer ranges follow the reference's cumulative-split rule,
    `llama-model.cpp:2492-2537`, see parallel/grid.py), and inside the stage
  * its tensor-parallel slice: Wq/Wk/Wv/gate/up by output rows (column-parallel, whole heads per rank),
    Wo/down by K range (row-parallel), lm_head by vocab rows.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from ..gguf import GGMLType, GGUFModelFiles, open_gguf
from ..ops.weights import DeviceWeight, scale_raw, slice_raw, stack_rows
from .config import ModelConfig, config_from_gguf


@dataclass
class LayerWeights:
    attn_norm: torch.Tensor
    wqkv: DeviceWeight                                 # [q | k | v] rows - or [q | k] when `wv` is set
    wo: DeviceWeight
    ffn_norm: torch.Tensor
    qkv_bias: torch.Tensor | None = None               # fp32 [local q | k | v rows]  (qwen2)
    wv: DeviceWeight | None = None                     # V rows kept in their own block type (Q4_K_M / Q5_K_M mixes, tp == 1)
    q_norm: torch.Tensor | None = None                 # fp32 [head_dim]              (qwen3)
    k_norm: torch.Tensor | None = None
    w_gate: DeviceWeight | None = None
    w_up: DeviceWeight | None = None
    w_down: DeviceWeight | None = None
    # MoE
    router: torch.Tensor | None = None                 # fp32 [n_expert, E]
    shared_gate: torch.Tensor | None = None            # fp32 [E]: sigmoid gate of the shared expert (qwen2moe; w_gate/w_up/w_down hold it)
    attn_sub_norm: torch.Tensor | None = None          # bitnet: fp32 [q_dim] / [n_ff] RMSNorm weights in front of wo / w_down
    ffn_sub_norm: torch.Tensor | None = None
    sel_bias: torch.Tensor | None = None               # fp32 [n_expert]: added to the router scores for the expert selection only (glm4moe)
    e_gate: list[DeviceWeight] = field(default_factory=list)
    e_up: list[DeviceWeight] = field(default_factory=list)
    e_down: list[DeviceWeight] = field(default_factory=list)
    # stacked [n_expert * rows, K] views of the same memory + fp16 router (device-side MoE decode)
    e_gate_s: DeviceWeight | None = None
    e_up_s: DeviceWeight | None = None
    e_down_s: DeviceWeight | None = None
    router_w: DeviceWeight | None = None
    router_wp: DeviceWeight | None = None              # router_w zero-padded to a multiple of 128 rows
    # LayerNorm / biased families (generic layer path): every one optional
    attn_norm_b: torch.Tensor | None = None
    ffn_norm_b: torch.Tensor | None = None
    attn_norm2: torch.Tensor | None = None             # falcon-40b: the attention branch has its own norm
    attn_norm2_b: torch.Tensor | None = None
    bo: torch.Tensor | None = None                     # fp32 [E] attention output bias (added once, after the TP reduction)
    b_up: torch.Tensor | None = None                   # fp32 [local F]
    b_down: torch.Tensor | None = None                 # fp32 [E]
    post_attn_norm: torch.Tensor | None = None         # olmo2, glm4: norm of each branch output before the residual add
    post_ffn_norm: torch.Tensor | None = None
    q_norm_full: torch.Tensor | None = None            # olmo2: RMSNorm weights over the whole Q / K width
    k_norm_full: torch.Tensor | None = None


@dataclass
class ModelWeights:
    cfg: ModelConfig
    layers: dict[int, LayerWeights]
    layer_range: range
    tp_rank: int
    tp_size: int
    tok_embd: DeviceWeight | None
    out_norm: torch.Tensor | None
    lm_head: DeviceWeight | None          # vocab rows [v0, v1) of the output matrix
    vocab_range: tuple[int, int]
    rope_freqs: torch.Tensor | None
    n_head_local: int = 0
    n_head_kv_local: int = 0
    n_ff_local: int = 0
    out_norm_b: torch.Tensor | None = None
    pos_embd: torch.Tensor | None = None      # fp32 [n_ctx_train, E] learned positions (gpt2, starcoder; first stage only)
    out_b: torch.Tensor | None = None         # fp32 bias of the output matrix, rows [v0, v1) (phi2)

    def nbytes(self) -> int:
        tot = 0
        for lw in self.layers.values():
            for w in [lw.wqkv, lw.wv, lw.wo, lw.w_gate, lw.w_up, lw.w_down, *lw.e_gate, *lw.e_up, *lw.e_down]:
                if w is not None:
                    tot += w.nbytes
        for w in (self.tok_embd, self.lm_head):
            if w is not None:
                tot += w.nbytes
        return tot


def _split(n: int, parts: int, rank: int, align: int = 1) -> tuple[int, int]:
    if n % (parts * align):
        raise ValueError(f"cannot split {n} into {parts} parts aligned to {align}")
    per = n // parts
    return rank * per, (rank + 1) * per


import os as _os
SPLIT_V = _os.environ.get("DLP_SPLIT_V", "0") == "1"


class LoraAdapter:
    """A GGUF LoRA adapter (`general.type = adapter`, `adapter.type = lora`, `adapter.lora.alpha`, tensor pairs
    `<base tensor>.lora_a` [r, in] / `.lora_b` [out, r]) as written by the reference's `convert_lora_to_gguf.py` and read by
    `llama_adapter_lora_init` (`llama.cpp/src/llama-adapter.cpp`).  The reference keeps adapters separate and adds
    `scale * B(A x)` per mat-mul (`build_lora_mm`); this engine folds `W += scale * alpha / r * B @ A` into the weights at
    load time (decode streams the merged fp16 matrix once; no extra launches)."""

    def __init__(self, path: str, scale: float = 1.0):
        self.f = open_gguf(path)
        md = self.f.metadata
        if md.get("general.type", "adapter") != "adapter" or md.get("adapter.type", "lora") != "lora":
            raise ValueError(f"{path}: not a LoRA adapter GGUF")
        self.alpha = float(md.get("adapter.lora.alpha", 0.0))
        self.scale = scale
        self.targets = {n[: -len(".lora_a")] for n in self.f.tensors if n.endswith(".lora_a")}

    def delta(self, name: str) -> np.ndarray | None:
        if name not in self.targets:
            return None
        a = np.asarray(self.f.array(name + ".lora_a"), dtype=np.float32)      # [r, in]
        b = np.asarray(self.f.array(name + ".lora_b"), dtype=np.float32)      # [out, r]
        r = a.shape[0]
        return (self.scale * (self.alpha / r if self.alpha else 1.0)) * (b @ a)


def write_lora_gguf(path: str, deltas: dict[str, tuple[np.ndarray, np.ndarray]], alpha: float, arch: str = "llama") -> str:
    """Write a LoRA adapter file ({tensor name: (A [r, in], B [out, r])}) - the test / tooling counterpart of the loader."""
    from ..gguf import GGUFWriter
    w = GGUFWriter(path)
    w.add("general.architecture", arch)
    w.add("general.type", "adapter")
    w.add("adapter.type", "lora")
    w.add("adapter.lora.alpha", float(alpha))
    for name, (a, b) in deltas.items():
        for suffix, m in ((".lora_a", a), (".lora_b", b)):
            m = np.ascontiguousarray(m, dtype="<f4")
            w.add_tensor(name + suffix, m.shape, GGMLType.F32, lambda m=m: m.view(np.uint8).reshape(-1))
    return w.write()


class LoadCancelled(RuntimeError):
    """Raised by `load_weights` when its progress callback returns False."""


def load_weights(src: str | GGUFModelFiles, device, layer_range: range | None = None, tp_rank: int = 0, tp_size: int = 1,
                 first_stage: bool = True, last_stage: bool = True, shard_vocab: bool = True,
                 lora: list[LoraAdapter] | None = None, progress=None) -> ModelWeights:
    """`progress(fraction) -> bool` is called before every layer and once at the end (1.0); returning False abandons the load with
    `LoadCancelled` (`llama_model_params.progress_callback`, the reference's test-model-load-cancel.cpp)."""
    f = open_gguf(src) if isinstance(src, str) else src
    lora = lora or []
    cfg = config_from_gguf(f.metadata, f.tensors)
    layer_range = range(cfg.n_layer) if layer_range is None else layer_range
    dev = torch.device(device)
    if cfg.n_head % tp_size or cfg.n_ff % tp_size:
        raise ValueError(f"tp={tp_size} must divide n_head={cfg.n_head} and n_ff={cfg.n_ff}")
    # KV heads: shard when possible, otherwise replicate groups (tp > n_head_kv)
    kv_rep = max(1, tp_size // cfg.n_head_kv)
    n_kv_local = max(1, cfg.n_head_kv // tp_size)
    kv_rank = tp_rank // kv_rep
    hq0, hq1 = _split(cfg.n_head, tp_size, tp_rank)
    D = cfg.head_dim

    def f32(name):
        return torch.from_numpy(np.array(f.array(name), dtype=np.float32, copy=True)).to(dev)

    def mat(name, rows=None, cols=None, sub=None, scale: float = 1.0, perm=None) -> DeviceWeight:
        """`perm`: gather of the (sliced) output rows - the partial-rotary head re-ordering below."""
        t = f.tensors[name]
        raw = f.raw(name)
        shape = t.shape
        ttype = t.ggml_type
        if sub is not None:  # one expert of a stacked [n_expert, N, K] tensor
            raw = raw[sub]
            shape = shape[1:]
        deltas = [d for d in (ad.delta(name) for ad in lora) if d is not None] if sub is None else []
        if scale != 1.0 and not deltas:       # scalar folded into the matrix (granite residual multiplier): block scales only
            raw, ttype = scale_raw(raw, ttype, shape, scale)
            scale = 1.0
        try:
            if deltas:
                raise ValueError("LoRA target: merge densely")
            raw2, shp = slice_raw(raw, ttype, shape, rows, cols)
        except ValueError:
            # K range not aligned to the quant block (tiny models / odd TP sizes): decode once, slice densely, keep fp16
            from ..gguf.quants import dequantize, quantize
            dense = dequantize(raw, ttype, shape)
            for d in deltas:
                dense = dense + d.reshape(dense.shape)
            if scale != 1.0:
                dense = dense * np.float32(scale)
            if rows is not None:
                dense = dense[rows[0]:rows[1]]
            if cols is not None:
                dense = dense[:, cols[0]:cols[1]]
            if perm is not None:
                dense = dense[perm]
            dense = np.ascontiguousarray(dense)
            return DeviceWeight.from_raw(quantize(dense, GGMLType.F16), GGMLType.F16, dense.shape, dev)
        if perm is not None:
            raw2 = np.ascontiguousarray(raw2.reshape(shp[0], -1)[perm])
        return DeviceWeight.from_raw(raw2, ttype, shp, dev)

    # Rotary on only the first n_rot dims of each head with half-split pairs (i, i + n_rot/2) - gptneox, stablelm: the kernels rotate
    # pairs (j, j + D/2) over the whole head.  Re-ordering the rows of Wq / Wk inside every head so that the rotated dims land on
    # [0, n_rot/2) and [D/2, D/2 + n_rot/2) - the rest gets a zero angle - gives the same scores, since Q.K is invariant under one
    # permutation applied to both.  V and everything downstream are untouched.
    def head_perm(n_heads: int):
        r = cfg.rope_dim // 2
        if cfg.rope_dim >= D or not cfg.rope_neox:
            return None
        one = list(range(r)) + list(range(cfg.rope_dim, cfg.rope_dim + D // 2 - r)) + list(range(r, cfg.rope_dim)) + \
            list(range(cfg.rope_dim + D // 2 - r, D))
        return np.concatenate([np.asarray(one) + h * D for h in range(n_heads)])
    perm_q, perm_k = head_perm(hq1 - hq0), head_perm(n_kv_local)

    layers: dict[int, LayerWeights] = {}
    for il in layer_range:
        if progress is not None and progress((il - layer_range.start) / max(1, len(layer_range))) is False:
            raise LoadCancelled(f"load of {getattr(f, 'path', 'model')} cancelled at layer {il}")
        p = f"blk.{il}."
        qr = (hq0 * D, hq1 * D)
        kr = (kv_rank * n_kv_local * D, (kv_rank + 1) * n_kv_local * D)
        if cfg.fused_qkv:   # phi3: rows [q | k | v] of one tensor
            ko, vo = cfg.q_dim, cfg.q_dim + cfg.kv_dim
            wq = mat(p + "attn_qkv.weight", rows=qr, perm=perm_q)
            wk = mat(p + "attn_qkv.weight", rows=(ko + kr[0], ko + kr[1]), perm=perm_k)
            wv = mat(p + "attn_qkv.weight", rows=(vo + kr[0], vo + kr[1]))
        else:
            wq = mat(p + "attn_q.weight", rows=qr, perm=perm_q)
            wk = mat(p + "attn_k.weight", rows=kr, perm=perm_k)
            wv = mat(p + "attn_v.weight", rows=kr)
        wv_sep = None
        if wq.wtype == wk.wtype == wv.wtype:
            wqkv = stack_rows([wq, wk, wv])
        elif SPLIT_V and wq.wtype == wk.wtype and tp_size == 1 and dev.type == "cuda" and wq.dense is None and wv.dense is None:
            # Q4_K_M / Q5_K_M store attn_v at more bits in some layers: keep every matrix in its block type and run V as a
            # second (small) mat-vec / GEMM into the tail columns of the QKV buffer instead of inflating all three to fp16.
            # Measured on B200 (Llama-3-8B Q4_K_M): 343.7 vs 349.6 tok/s decode, 20.8 k vs 23.3 k tok/s prefill - the extra
            # launch and the small-N block-quant kernels cost more than the 36 MB/layer they save, so this is opt-in.
            wqkv, wv_sep = stack_rows([wq, wk]), wv
        else:  # mixed-type files (Q4_K_M stores attn_v at higher precision): decode the odd ones to fp16
            dt = [w.to_dense(torch.float16) if w.dense is None else w.dense for w in (wq, wk, wv)]
            dense = torch.cat([d.float() for d in dt], 0)
            wqkv = (DeviceWeight(GGMLType.F16, dense.shape[0], dense.shape[1], [], dense) if dev.type == "cpu"
                    else DeviceWeight(GGMLType.F16, dense.shape[0], dense.shape[1], [dense.to(torch.float16).contiguous()]))
        def norm_w(name):      # LayerNorm files may carry no weight (olmo) or no such norm at all (parallel blocks): ones / None
            if name in f:
                return f32(name)
            return torch.ones(cfg.n_embd, dtype=torch.float32, device=dev) if cfg.norm_type == "layer" else None

        def opt(name):
            return f32(name) if name in f else None
        lw = LayerWeights(attn_norm=norm_w(p + "attn_norm.weight") if cfg.pre_norm else None, wqkv=wqkv, wv=wv_sep,
                          wo=mat(p + "attn_output.weight", cols=(hq0 * D, hq1 * D), scale=cfg.residual_scale),
                          ffn_norm=None if ((cfg.parallel_residual and p + cfg.ffn_norm_name not in f) or not cfg.pre_norm) else norm_w(p + cfg.ffn_norm_name))
        if cfg.generic:
            lw.attn_norm_b, lw.ffn_norm_b = opt(p + "attn_norm.bias"), opt(p + "ffn_norm.bias")
            lw.attn_norm2, lw.attn_norm2_b = opt(p + "attn_norm_2.weight"), opt(p + "attn_norm_2.bias")
            lw.bo, lw.b_down = opt(p + "attn_output.bias"), opt(p + "ffn_down.bias")
            lw.post_attn_norm, lw.post_ffn_norm = opt(p + "post_attention_norm.weight"), opt(p + "post_ffw_norm.weight")
            if cfg.sub_norm:
                if tp_size > 1:
                    raise NotImplementedError("bitnet's sub-norms span the full attention / MLP width: not split across tensor-parallel ranks")
                lw.attn_sub_norm, lw.ffn_sub_norm = f32(p + "attn_sub_norm.weight"), f32(p + "ffn_sub_norm.weight")
            if cfg.qk_norm_full:
                if tp_size > 1:
                    raise NotImplementedError("Q/K norms over the full projected width (olmo2) are not split across tensor-parallel ranks")
                lw.q_norm_full, lw.k_norm_full = f32(p + "attn_q_norm.weight"), f32(p + "attn_k_norm.weight")
                if perm_q is not None:      # partial rotary (minimax-m2): the norm weights follow the re-ordered projection rows
                    lw.q_norm_full = lw.q_norm_full[torch.as_tensor(perm_q, device=dev)].contiguous()
                    lw.k_norm_full = lw.k_norm_full[torch.as_tensor(perm_k, device=dev)].contiguous()
        if cfg.qkv_bias:
            if cfg.fused_qkv:       # gptneox: one bias vector laid out like the fused matrix
                fb = f32(p + "attn_qkv.bias")
                bq, bk, bv = fb[qr[0]:qr[1]], fb[cfg.q_dim + kr[0]:cfg.q_dim + kr[1]], fb[cfg.q_dim + cfg.kv_dim + kr[0]:cfg.q_dim + cfg.kv_dim + kr[1]]
            else:
                bq, bk, bv = f32(p + "attn_q.bias")[qr[0]:qr[1]], f32(p + "attn_k.bias")[kr[0]:kr[1]], f32(p + "attn_v.bias")[kr[0]:kr[1]]
            if perm_q is not None:
                bq, bk = bq[torch.as_tensor(perm_q, device=bq.device)], bk[torch.as_tensor(perm_k, device=bk.device)]
            lw.qkv_bias = torch.cat([bq, bk, bv]).contiguous()
        if cfg.qk_norm:
            lw.q_norm, lw.k_norm = f32(p + "attn_q_norm.weight"), f32(p + "attn_k_norm.weight")
            if perm_q is not None:      # partial rotary (glm4moe): the per-head norm weights follow the re-ordered rows of their head
                one = torch.as_tensor(perm_q[:D], device=lw.q_norm.device)
                lw.q_norm, lw.k_norm = lw.q_norm[one].contiguous(), lw.k_norm[one].contiguous()
        f0, f1 = _split(cfg.n_ff, tp_size, tp_rank, 256 if tp_size > 1 else 1)
        if cfg.ffn_act in ("gelu", "relu2"):     # non-gated MLP: up -> activation -> down
            lw.w_up = mat(p + "ffn_up.weight", rows=(f0, f1))
            lw.w_down = mat(p + "ffn_down.weight", cols=(f0, f1), scale=cfg.residual_scale)
            if p + "ffn_up.bias" in f:
                lw.b_up = f32(p + "ffn_up.bias")[f0:f1].contiguous()
        elif cfg.fused_gate_up:   # phi3: ffn_up rows [gate | up]
            lw.w_gate = mat(p + "ffn_up.weight", rows=(f0, f1))
            lw.w_up = mat(p + "ffn_up.weight", rows=(cfg.n_ff + f0, cfg.n_ff + f1))
            lw.w_down = mat(p + "ffn_down.weight", cols=(f0, f1), scale=cfg.residual_scale)
        elif cfg.is_moe and p + "ffn_gate_inp.weight" in f:
            lw.router = f32(p + "ffn_gate_inp.weight")
            lw.sel_bias = f32(p + "exp_probs_b.bias") if p + "exp_probs_b.bias" in f else None
            w_scale = cfg.expert_weights_scale if cfg.moe_scale_w else 1.0     # a constant on the routed weights: folded into the down projections
            for e in range(cfg.n_expert):
                lw.e_gate.append(mat(p + "ffn_gate_exps.weight", rows=(f0, f1), sub=e))
                lw.e_up.append(mat(p + "ffn_up_exps.weight", rows=(f0, f1), sub=e))
                lw.e_down.append(mat(p + "ffn_down_exps.weight", cols=(f0, f1), sub=e, scale=cfg.residual_scale * w_scale))
            if dev.type == "cuda" and all(w.wtype == lw.e_gate[0].wtype for w in lw.e_gate + lw.e_up) and \
                    all(w.wtype == lw.e_down[0].wtype for w in lw.e_down):
                def restack(ws):
                    st = stack_rows(ws)
                    n = ws[0].N
                    views = [DeviceWeight(st.wtype, n, st.K, [pl[e * n:(e + 1) * n] for pl in st.planes]) for e in range(len(ws))]
                    return st, views
                lw.e_gate_s, lw.e_gate = restack(lw.e_gate)
                lw.e_up_s, lw.e_up = restack(lw.e_up)
                lw.e_down_s, lw.e_down = restack(lw.e_down)
                lw.router_w = DeviceWeight(GGMLType.F16, cfg.n_expert, cfg.n_embd, [lw.router.to(torch.float16).contiguous()])
                # the same router padded to one full 128-row GEMM tile (token counts above the mat-vec range use the tcgen05 GEMM)
                rp = torch.zeros(max(128, (cfg.n_expert + 127) // 128 * 128), cfg.n_embd, dtype=torch.float16, device=lw.router.device)
                rp[:cfg.n_expert] = lw.router.to(torch.float16)
                lw.router_wp = DeviceWeight(GGMLType.F16, rp.shape[0], cfg.n_embd, [rp])
            if cfg.n_ff_shared:   # qwen2moe shared expert (split across TP ranks like a dense FFN) + its gate vector
                s0, s1 = _split(cfg.n_ff_shared, tp_size, tp_rank, 256 if tp_size > 1 else 1)
                lw.w_gate = mat(p + "ffn_gate_shexp.weight", rows=(s0, s1))
                lw.w_up = mat(p + "ffn_up_shexp.weight", rows=(s0, s1))
                lw.w_down = mat(p + "ffn_down_shexp.weight", cols=(s0, s1), scale=cfg.residual_scale)
                if p + "ffn_gate_inp_shexp.weight" in f:      # qwen2moe gates it; deepseek / bailingmoe add it as it is
                    lw.shared_gate = f32(p + "ffn_gate_inp_shexp.weight").reshape(-1)
        else:
            if cfg.is_moe:        # a dense lead layer of a mixture-of-experts model: its own FFN width
                f0, f1 = _split(cfg.n_ff_dense or cfg.n_ff, tp_size, tp_rank, 256 if tp_size > 1 else 1)
            lw.w_gate = mat(p + "ffn_gate.weight", rows=(f0, f1))
            lw.w_up = mat(p + "ffn_up.weight", rows=(f0, f1))
            lw.w_down = mat(p + "ffn_down.weight", cols=(f0, f1), scale=cfg.residual_scale)
        layers[il] = lw

    tok = mat("token_embd.weight") if first_stage else None
    out_norm = out_norm_b = lm = None
    v0, v1 = 0, cfg.vocab_size
    if last_stage:
        # granite: logits / logit_scale, folded into the final norm (weight and, for LayerNorm files, bias)
        out_norm = (f32("output_norm.weight") if "output_norm.weight" in f else torch.ones(cfg.n_embd, dtype=torch.float32, device=dev)) / cfg.logit_scale
        out_norm_b = f32("output_norm.bias") / cfg.logit_scale if "output_norm.bias" in f else None
        name = "output.weight" if "output.weight" in f else "token_embd.weight"
        if shard_vocab and tp_size > 1 and cfg.vocab_size % (2 * tp_size) == 0:
            v0, v1 = _split(cfg.vocab_size, tp_size, tp_rank)
        lm = mat(name, rows=(v0, v1))
    freqs = f32("rope_freqs.weight") if "rope_freqs.weight" in f else None
    if progress is not None and progress(1.0) is False:
        raise LoadCancelled("load cancelled")
    return ModelWeights(cfg, layers, layer_range, tp_rank, tp_size, tok, out_norm, lm, (v0, v1), freqs,
                        n_head_local=hq1 - hq0, n_head_kv_local=n_kv_local, n_ff_local=cfg.n_ff // tp_size, out_norm_b=out_norm_b,
                        pos_embd=f32("position_embd.weight") if (cfg.learned_pos and first_stage) else None,
                        out_b=f32("output.bias")[v0:v1].contiguous() if (last_stage and "output.bias" in f) else None)
