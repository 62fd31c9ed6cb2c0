"""Llama / Mixtral executor: paged KV cache, prefill + decode step functions, CUDA-graph decode.

Parity: this is the `llama_context` decode loop (`llama.cpp/src/llama-context.cpp:1042-1104,1375-1724`) plus the
graph `llm_build_llama` builds per ubatch (`llama.cpp/src/models/llama.cpp:31-164`), restated B200-first:
there is no graph builder / scheduler / allocator at run time - the step is a fixed sequence of fused kernels
over pre-allocated buffers (captured once into a CUDA graph per decode batch size), the KV cache is paged
instead of a contiguous ring (`llama-kv-cache.cpp:100-150`), positions / slots advance on the device, and
sampling happens on the device so only token ids cross PCIe.

One engine instance = one (pipeline stage, tensor-parallel rank).  `comm` supplies the TP reduction and the
PP activation hand-off (see parallel/).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import os

import torch

from ..ops import api as ops
from .config import ModelConfig
from .reference import limit_to_groups, rope_attn_scale, rope_inv_freq
from .weights import ModelWeights


@dataclass
class Sequence:
    sid: int
    length: int = 0
    pages: list[int] = field(default_factory=list)


class PagedKVCache:
    """Fixed-size pages [n_pages, Hkv, page, D] per layer + a host-side free list and block table."""

    def __init__(self, cfg: ModelConfig, layers, n_head_kv: int, n_pages: int, page_size: int, max_seqs: int,
                 max_pages_per_seq: int, device):
        self.page_size = page_size
        self.n_pages = n_pages
        self.max_pages = max_pages_per_seq
        shape = (n_pages, n_head_kv, page_size, cfg.head_dim)
        self.k = {il: torch.zeros(shape, dtype=torch.float16, device=device) for il in layers}
        self.v = {il: torch.zeros(shape, dtype=torch.float16, device=device) for il in layers}
        self.free = list(range(n_pages - 1, -1, -1))
        self.block_table = torch.zeros(max_seqs, max_pages_per_seq, dtype=torch.int32, device=device)
        self.block_table_host = torch.zeros(max_seqs, max_pages_per_seq, dtype=torch.int32)
        if torch.device(device).type == "cuda":
            self.block_table_host = self.block_table_host.pin_memory()

    def ensure(self, seq: Sequence, new_len: int):
        need = (new_len + self.page_size - 1) // self.page_size
        if need > self.max_pages:
            raise RuntimeError(f"sequence {seq.sid}: context {new_len} exceeds n_ctx")
        grew = False
        while len(seq.pages) < need:
            if not self.free:
                raise RuntimeError("KV cache exhausted (no free pages)")
            pg = self.free.pop()
            self.block_table_host[seq.sid, len(seq.pages)] = pg
            seq.pages.append(pg)
            grew = True
        if grew:
            self.block_table[seq.sid].copy_(self.block_table_host[seq.sid], non_blocking=True)

    def release(self, seq: Sequence):
        self.free.extend(reversed(seq.pages))
        seq.pages.clear()
        seq.length = 0

    def bytes(self) -> int:
        return sum(t.numel() * 2 for t in self.k.values()) * 2


class NullComm:
    """Single-rank communicator (tp = pp = 1)."""
    tp_size = 1
    tp_rank = 0

    def allreduce_delta(self, delta, il, which):
        return delta

    def allgather_logits(self, logits, vocab_range, vocab_size):
        return logits

    fused = False


class LlamaEngine:
    def __init__(self, weights: ModelWeights, n_ctx: int = 4096, max_seqs: int = 8, page_size: int = 64,
                 max_batch_tokens: int = 2048, n_pages: int | None = None, comm=None, first_stage: bool = True,
                 last_stage: bool = True, attn_splits: int | None = None):
        self.w = weights
        self.cfg = cfg = weights.cfg
        dev = (weights.tok_embd or weights.lm_head or next(iter(weights.layers.values())).wqkv).device
        self.device = dev
        self.comm = comm or NullComm()
        self.first_stage, self.last_stage = first_stage, last_stage
        self.n_ctx, self.max_seqs, self.page_size = n_ctx, max_seqs, page_size
        self.max_tokens = max(max_batch_tokens, max_seqs)
        self.Hq, self.Hkv, self.D = weights.n_head_local, weights.n_head_kv_local, cfg.head_dim
        self.E, self.F = cfg.n_embd, weights.n_ff_local
        if dev.type == "cuda" and self.D not in (64, 128):
            raise NotImplementedError(f"head dimension {self.D}: the attention kernels are built for 64 and 128 (phi-2 has 80, gemma 256)")
        max_pages = (n_ctx + page_size - 1) // page_size
        if n_pages is None:
            n_pages = max_pages * max_seqs
        self.kv = PagedKVCache(cfg, weights.layer_range, self.Hkv, n_pages, page_size, max_seqs, max_pages, dev)
        self.inv_freq = rope_inv_freq(cfg, weights.rope_freqs.cpu() if weights.rope_freqs is not None else None).to(dev)
        if cfg.rope_dim < cfg.head_dim:     # partial rotary: zero angles for the dims left alone (Wq / Wk rows were re-ordered at load, models/weights.py)
            self.inv_freq = torch.cat([self.inv_freq, torch.zeros(cfg.head_dim // 2 - cfg.rope_dim // 2, dtype=self.inv_freq.dtype, device=dev)])
        # layers without rotary encoding (smollm3: every 4th) run the same kernels on a zero-angle table: the rotation is the identity
        self.inv_freq_off = torch.zeros_like(self.inv_freq) if cfg.no_rope_step > 0 else None
        self.attn_scale = rope_attn_scale(cfg)
        self.seqs: dict[int, Sequence] = {}
        self._free_sids = list(range(max_seqs - 1, -1, -1))
        M = self.max_tokens + 8   # head-room: the sequence-parallel path pads T up to a multiple of tp
        f16, f32, i32 = torch.float16, torch.float32, torch.int32
        qkv_dim = (self.Hq + 2 * self.Hkv) * self.D
        self.h = torch.zeros(M, self.E, dtype=f32, device=dev)
        self.fused_tp = bool(getattr(self.comm, "fused", False)) and self.comm.tp_size > 1 and not cfg.is_moe
        if cfg.generic and self.fused_tp:
            raise NotImplementedError(f"the {cfg.arch} family runs the generic layer path: tensor parallelism through --tp-impl nccl")
        if cfg.generic and cfg.is_moe and (cfg.norm_type != "rms" or cfg.parallel_residual or cfg.post_norm or not cfg.pre_norm):
            raise NotImplementedError("routed experts on the generic layer path need the sequential pre-RMSNorm block form")
        self.sp_ok = False
        self._greedy_now = False
        if self.fused_tp:
            if not self.comm.ready:
                self.comm.setup(self.E, 2 * len(weights.layer_range))
            self.hb = [self.h, torch.zeros(4, self.E, dtype=f32, device=dev)]   # ping-pong residual (decode rows only)
            self.cur, self.pending = 0, None
            tp = self.comm.tp_size
            self.sp_ok = not getattr(self.comm, "no_sp", False)
            if self.sp_ok:
                if not getattr(self.comm, "sp_ready", False):
                    self.comm.sp_setup(self.E, M)
                self.h_sh = torch.zeros(M // tp + 2, self.E, dtype=f32, device=dev)   # token-sharded residual (SP prefill)
                self.G = [self.comm.sp_tensor(g) for g in (0, 1)]                            # gather buffers as fp16 [Tmax, E] views
        self.delta = torch.zeros(M, self.E, dtype=f32, device=dev)
        self.xn = torch.zeros(M, self.E, dtype=f16, device=dev)
        self.qkv32 = torch.zeros(min(M, 4), qkv_dim, dtype=f32, device=dev)
        self.qkv16 = torch.zeros(M, qkv_dim, dtype=f16, device=dev)
        self.q = torch.zeros(M, self.Hq * self.D, dtype=f16, device=dev)
        self.att = torch.zeros(M, self.Hq * self.D, dtype=f16, device=dev)
        f_buf = max(self.F, (cfg.n_ff_dense // max(1, self.comm.tp_size)) if cfg.n_layer_dense_lead else 0)     # dense lead layers may be wider than the experts
        self.act = torch.zeros(M, f_buf, dtype=f16, device=dev)
        # generic layer path: fp32 MLP pre-activation (bias + GELU run between the two mat-muls), second norm output (falcon-40b)
        self.up32 = torch.zeros(M, self.F, dtype=f32, device=dev) if cfg.ffn_act != "swiglu" else None
        self.xn2 = torch.zeros(M, self.E, dtype=f16, device=dev) if cfg.post_norm or any(
            lw.attn_norm2 is not None or (cfg.parallel_residual and lw.ffn_norm is not None) for lw in weights.layers.values()) else None
        self.tokens = torch.zeros(M, dtype=i32, device=dev)
        self.pos = torch.zeros(M, dtype=i32, device=dev)
        self.slot = torch.zeros(M, dtype=i32, device=dev)
        self.kv_len = torch.zeros(M, dtype=i32, device=dev)
        self.active = torch.zeros(max_seqs, dtype=i32, device=dev)   # sequence slot of each decode batch row
        self._active_host: list[int] = []
        self.seq_len_dev = torch.zeros(max_seqs, dtype=i32, device=dev)
        self.next_tok = torch.zeros(max_seqs, dtype=i32, device=dev)
        self.step_ctr = torch.zeros(1, dtype=i32, device=dev)
        # per-sequence sampler records (dlp_sample_params: top_k, top_p, min_p, temp, seed, draw counter) read by the device sampler
        self.samp = torch.zeros(max_seqs, ops.SAMPLE_REC_WORDS, dtype=i32, device=dev)
        self._samp_host: dict[int, tuple] = {}      # what each record currently holds
        self._seq_sampling: dict[int, tuple] = {}   # per-sequence overrides of self.sampling (set_seq_sampling)
        self.max_graphs = int(os.environ.get("DLP_MAX_GRAPHS", str(max(24, 4 * max_seqs + 8))))   # (batch sizes x attention variants x greedy)
        v0, v1 = weights.vocab_range
        self.logits = torch.zeros(max(max_seqs, 1), v1 - v0, dtype=f32, device=dev) if last_stage else None
        if last_stage and self.comm.tp_size > 1 and hasattr(self.comm, "reserve_logits") and (v1 - v0) != cfg.vocab_size:
            self.comm.reserve_logits(max_seqs, v1 - v0, f32, dev)     # before any graph is captured
        self.attn_splits = attn_splits if attn_splits is not None else (16 if dev.type == "cuda" else 1)
        ns = self.attn_splits
        self.attn_ws = (torch.zeros(max_seqs * self.Hq * ns * self.D, dtype=f32, device=dev),
                        torch.zeros(max_seqs * self.Hq * ns * 2, dtype=f32, device=dev),
                        torch.zeros(max_seqs * self.Hq, dtype=i32, device=dev)) if ns > 1 else None
        self.scratch = None  # fp16 decode buffer for block-quantised weights on the GEMM path (allocated lazily)
        self._graphs: dict[tuple, torch.cuda.CUDAGraph] = {}
        self.sampling = dict(top_k=40, top_p=0.95, min_p=0.05, temp=0.0, seed=0)
        # decode attention in one launch (RoPE + KV write + split-KV + cluster combine); plain-llama layers with full-width RoPE
        self.attn_onepass = (dev.type == "cuda" and os.environ.get("DLP_ATTN_ONEPASS", "1") == "1" and cfg.rope_dim == self.D
                             and not cfg.qkv_bias and not cfg.qk_norm)
        self.attn_cluster = int(os.environ.get("DLP_ATTN_CLUSTER", "16"))
        # the cluster launch only pays once the context spans more than one 64-key split; below this many keys the two-kernel
        # path (RoPE kernel + attention) is used (profiles/r1_attn_onepass_ab.txt).  The host knows every sequence length, so the choice is part of the graph key.
        self.attn_onepass_min = int(os.environ.get("DLP_ATTN_ONEPASS_MIN", "64"))
        self._onepass_now = False
        self._graph_launches: dict[tuple, int] = {}    # kernels inside each captured graph
        self.launches_replayed = 0                      # kernels launched through graph replays so far
        self.rope_cs = (torch.zeros(max_seqs, self.D // 2, 2, dtype=f32, device=dev)
                        if self.attn_onepass and os.environ.get("DLP_ROPE_TABLE", "1") == "1" else None)
        self.cvec: dict[int, torch.Tensor] = {}     # layer -> fp32 [E] added to the residual stream after that layer
        self.eval_callback = None      # fn(name, layer, tensor) called after every layer (eager paths only; disables graphs)

    # ---- sequence management ------------------------------------------------------------------
    def alloc_seq(self) -> int:
        if not self._free_sids:
            raise RuntimeError("no free sequence slots")
        sid = self._free_sids.pop()
        self.seqs[sid] = Sequence(sid)
        return sid

    def free_seq(self, sid: int):
        self.kv.release(self.seqs.pop(sid))
        self._free_sids.append(sid)
        self._seq_sampling.pop(sid, None)
        self._samp_host.pop(sid, None)

    def set_seq_sampling(self, sid: int, top_k: int = 40, top_p: float = 0.95, min_p: float = 0.05, temp: float = 0.0, seed: int = 0):
        """Sampler parameters of one sequence for the device sampler; also restarts that sequence's random stream, so a request
        with a given seed draws the same tokens whatever ran before it (the reference keeps one seeded RNG per slot,
        `common/sampling.cpp:179-311`).  Sequences without their own record follow `self.sampling`."""
        self._seq_sampling[sid] = (int(top_k), float(top_p), float(min_p), float(temp), int(seed))
        self._samp_host.pop(sid, None)

    def _sync_sampling(self, sids):
        for sid in sids:
            s = self.sampling
            want = self._seq_sampling.get(sid) or (s["top_k"], s["top_p"], s["min_p"], s["temp"], s["seed"])
            if self._samp_host.get(sid) != want:
                self.samp[sid].copy_(ops.pack_sample_record(*want))
                self._samp_host[sid] = want

    def seq_len(self, sid: int) -> int:
        return self.seqs[sid].length

    def truncate_seq(self, sid: int, new_len: int):
        """Keep only the first `new_len` cached tokens of a sequence (prompt-prefix reuse, rollback after errors).
        Parity: `llama_memory_seq_rm(seq, p0=new_len, p1=-1)` (`llama-kv-cache.cpp:224-320`); pages past the cut
        return to the free list."""
        seq = self.seqs[sid]
        new_len = max(0, min(new_len, seq.length))
        keep_pages = (new_len + self.page_size - 1) // self.page_size
        while len(seq.pages) > keep_pages:
            self.kv.free.append(seq.pages.pop())
        seq.length = new_len
        self.seq_len_dev[sid] = new_len

    def copy_seq(self, src: int, dst: int, n: int | None = None):
        """Copy the first `n` cached tokens (default: all) of sequence `src` into `dst` (which is overwritten).
        Parity: `llama_memory_seq_cp` (`llama-kv-cache.cpp:322-392`) - used for prompt sharing between slots / beams."""
        s, d = self.seqs[src], self.seqs[dst]
        n = s.length if n is None else min(n, s.length)
        self.truncate_seq(dst, 0)
        self.kv.ensure(d, n)
        npg = (n + self.page_size - 1) // self.page_size
        if npg:
            sp = torch.tensor(s.pages[:npg], dtype=torch.long, device=self.device)
            dp = torch.tensor(d.pages[:npg], dtype=torch.long, device=self.device)
            for il in self.w.layer_range:
                self.kv.k[il][dp] = self.kv.k[il][sp]
                self.kv.v[il][dp] = self.kv.v[il][sp]
        d.length = n
        self.seq_len_dev[dst] = n

    def shift_seq(self, sid: int, n_keep: int, n_discard: int):
        """Context shift: drop cached tokens [n_keep, n_keep + n_discard) and slide the rest down, re-rotating their keys by
        -n_discard positions so that attention sees contiguous positions again.  Parity: `llama_memory_seq_rm` +
        `llama_memory_seq_add(-n_discard)` and the K-shift graph (`llama-kv-cache.cpp:1379-1484`), as driven by the server's
        context shift (`server-context.cpp:1944-2003`).  Rare, latency-insensitive: gathers / scatters pages with torch ops."""
        from .reference import apply_rope
        seq = self.seqs[sid]
        n, ps = seq.length, self.page_size
        n_keep = max(0, min(n_keep, n))
        n_discard = max(0, min(n_discard, n - n_keep))
        if n_discard == 0:
            return
        pages = torch.tensor(seq.pages[: (n + ps - 1) // ps], dtype=torch.long, device=self.device)
        tail = n - n_keep - n_discard
        delta = torch.full((tail,), -n_discard, dtype=torch.int32, device=self.device)
        for il in self.w.layer_range:
            for which, cache in (("k", self.kv.k[il]), ("v", self.kv.v[il])):
                flat = cache[pages].permute(0, 2, 1, 3).reshape(-1, self.Hkv, self.D)          # [tokens, Hkv, D]
                moved = flat[n_keep + n_discard:n].clone()
                if which == "k" and tail:
                    # R(p - d) k = R(-d) R(p) k: rotate the stored (already rotated) keys by the position delta
                    moved = apply_rope(moved.float(), delta, self._inv_freq(il), self.cfg.rope_neox).to(flat.dtype)
                flat[n_keep:n_keep + tail] = moved
                cache[pages] = flat.view(len(pages), ps, self.Hkv, self.D).permute(0, 2, 1, 3)
        self.truncate_seq(sid, n_keep + tail)

    def _scratch_for(self, *ws):
        if not ops.TWO_PASS_DEQUANT:   # block-quant decode is fused into the GEMM: no scratch
            return None
        need = 0
        for w in ws:
            if w is not None and w.dense is None and w.wtype.name.startswith("Q"):
                need += w.N * w.K * 2
        if need == 0:
            return None
        if self.scratch is None or self.scratch.numel() < need:
            self.scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self.scratch

    def _inv_freq(self, il: int) -> torch.Tensor:
        return self.inv_freq if self.inv_freq_off is None or self.cfg.layer_uses_rope(il) else self.inv_freq_off

    def _rope(self, il: int, qkv: torch.Tensor, T: int) -> torch.Tensor:
        """(+bias, per-head Q/K norm,) RoPE, paged K/V write: one launch for every architecture family."""
        lw = self.w.layers[il]
        return ops.rope_kv(qkv, self.pos[:T], self.slot[:T], self._inv_freq(il), self.kv.k[il], self.kv.v[il], self.Hq, self.Hkv,
                           self.D, self.cfg.rope_neox, q_out=self.q[:T], bias=lw.qkv_bias, q_norm=lw.q_norm, k_norm=lw.k_norm,
                           eps=self.cfg.rms_eps, prefetch=lw.wo if T <= 4 else None, attn_scale=self.attn_scale)

    # ---- one transformer layer over T tokens held in the static buffers ------------------------
    def _layer(self, il: int, T: int, attn_fn):
        lw, cfg, comm = self.w.layers[il], self.cfg, self.comm
        h = self.h[:T]
        tp = comm.tp_size > 1
        small = T <= 4 and not tp  # fused mat-vec path (norm in prologue, residual add in epilogue)
        # --- attention block ---
        if small:
            qkv = self.qkv32[:T]
            if lw.wv is None:
                ops.linear(lw.wqkv, h, "f32", norm_w=lw.attn_norm, eps=cfg.rms_eps, out=qkv)
            else:   # V in its own block type: second mat-vec into the tail columns
                ops.linear(lw.wqkv, h, "f32", norm_w=lw.attn_norm, eps=cfg.rms_eps, out=qkv[:, :lw.wqkv.N])
                ops.linear(lw.wv, h, "f32", norm_w=lw.attn_norm, eps=cfg.rms_eps, out=qkv[:, lw.wqkv.N:])
        else:
            xn = ops.rmsnorm(h, lw.attn_norm, cfg.rms_eps, out=self.xn[:T])
            qkv = self.qkv16[:T]
            if lw.wv is None:
                ops.linear(lw.wqkv, xn, "f16", out=qkv, scratch=self._scratch_for(lw.wqkv))
            else:
                ops.linear(lw.wqkv, xn, "f16", out=qkv[:, :lw.wqkv.N], scratch=self._scratch_for(lw.wqkv))
                ops.linear(lw.wv, xn, "f16", out=qkv[:, lw.wqkv.N:], scratch=self._scratch_for(lw.wv))
        if small and self._onepass_now and attn_fn == self._attn_decode:
            att = self._attn_decode_onepass(il, qkv, T)
        else:
            q = self._rope(il, qkv, T)
            att = attn_fn(il, q, T)
        if tp:
            d = ops.linear(lw.wo, att, "f32", out=self.delta[:T], scratch=self._scratch_for(lw.wo))
            d = comm.allreduce_delta(d, il, 0)
            xn = ops.rmsnorm(h, lw.ffn_norm, cfg.rms_eps, delta=d, out=self.xn[:T])
        elif small:
            ops.linear(lw.wo, att, "add", out=h)
        else:
            ops.linear(lw.wo, att, "add", out=h, scratch=self._scratch_for(lw.wo))
            xn = ops.rmsnorm(h, lw.ffn_norm, cfg.rms_eps, out=self.xn[:T])
        # --- feed-forward block ---
        if lw.router is not None:       # (a mixture-of-experts model may still have dense layers: deepseek's lead layers)
            self._moe(lw, T, None if small else xn)
            return
        if small:
            act = ops.linear(lw.w_gate, h, "swiglu", w_up=lw.w_up, norm_w=lw.ffn_norm, eps=cfg.rms_eps, out=self.act[:T])
            ops.linear(lw.w_down, act, "add", out=h)
            return
        act = ops.linear(lw.w_gate, xn, "swiglu", w_up=lw.w_up, out=self.act[:T], scratch=self._scratch_for(lw.w_gate, lw.w_up))
        if tp:
            d = ops.linear(lw.w_down, act, "f32", out=self.delta[:T], scratch=self._scratch_for(lw.w_down))
            d = comm.allreduce_delta(d, il, 1)
            h.add_(d)
        else:
            ops.linear(lw.w_down, act, "add", out=h, scratch=self._scratch_for(lw.w_down))

    def _norm(self, h, w, b, delta=None, out=None):
        if self.cfg.norm_type == "layer":
            return ops.layernorm(h, w, b, self.cfg.rms_eps, delta=delta, out=out)
        return ops.rmsnorm(h, w, self.cfg.rms_eps, delta=delta, out=out)

    def _layer_generic(self, il: int, T: int, attn_fn):
        """LayerNorm / GELU / parallel-residual families (starcoder2, falcon, olmo, orion, command-r; `build_norm` LLM_NORM,
        `build_ffn` LLM_FFN_GELU + LLM_FFN_SEQ in llama-graph.cpp:785-990): the same mat-vec / GEMM / attention kernels as `_layer`
        with the norm, the bias + activation and the residual adds as separate launches instead of fused prologues / epilogues."""
        lw, cfg, comm = self.w.layers[il], self.cfg, self.comm
        h = self.h[:T]
        tp = comm.tp_size > 1
        if cfg.pre_norm:
            xn = self._norm(h, lw.attn_norm, lw.attn_norm_b, out=self.xn[:T])
        else:                                                     # olmo2: the branches read the residual stream as it is
            xn = self.xn[:T].copy_(h)
        xa, xf = xn, xn
        if lw.attn_norm2 is not None:                         # falcon-40b: the attention branch has its own norm of the block input
            xa = self._norm(h, lw.attn_norm2, lw.attn_norm2_b, out=self.xn2[:T])
        elif cfg.parallel_residual and lw.ffn_norm is not None:   # gptneox: the MLP branch has its own norm of the block input
            xf = self._norm(h, lw.ffn_norm, lw.ffn_norm_b, out=self.xn2[:T])
        qkv = self.qkv16[:T]
        if lw.wv is None:
            ops.linear(lw.wqkv, xa, "f16", out=qkv, scratch=self._scratch_for(lw.wqkv))
        else:
            ops.linear(lw.wqkv, xa, "f16", out=qkv[:, :lw.wqkv.N], scratch=self._scratch_for(lw.wqkv))
            ops.linear(lw.wv, xa, "f16", out=qkv[:, lw.wqkv.N:], scratch=self._scratch_for(lw.wv))
        if lw.q_norm_full is not None:                            # olmo2: RMSNorm over the whole Q and K width, before the rotation
            qd, kd = self.Hq * self.D, self.Hkv * self.D
            for sl, wn in ((slice(0, qd), lw.q_norm_full), (slice(qd, qd + kd), lw.k_norm_full)):
                v = qkv[:, sl].float()
                qkv[:, sl] = (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + cfg.rms_eps) * wn).to(qkv.dtype)
        q = self._rope(il, qkv, T)
        att = attn_fn(il, q, T)
        if lw.attn_sub_norm is not None:                           # bitnet: RMSNorm over the attention result, in front of the output matrix
            v = att.float()
            att = (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + cfg.rms_eps) * lw.attn_sub_norm).to(att.dtype)
        d = ops.linear(lw.wo, att, "f32", out=self.delta[:T], scratch=self._scratch_for(lw.wo))
        if tp:
            d = comm.allreduce_delta(d, il, 0)
        if lw.bo is not None:
            d.add_(lw.bo)
        if cfg.post_norm:               # the attention output is normalised, then joins the residual stream
            h.add_(self._norm(d, lw.post_attn_norm, None, out=self.xn2[:T]))
            ffn_in = self._norm(h, lw.ffn_norm, lw.ffn_norm_b, out=self.xn[:T]) if cfg.pre_norm else self.xn[:T].copy_(h)
        elif cfg.parallel_residual:       # the MLP reads the norm of the block INPUT: add the attention branch to h only now
            h.add_(d)
            ffn_in = xf
        else:
            ffn_in = self._norm(h, lw.ffn_norm, lw.ffn_norm_b, delta=d, out=self.xn[:T])       # h += d inside
        if lw.router is not None:         # olmoe: routed experts behind the same pre-norm (adds its output to h)
            self._moe(lw, T, ffn_in)
            return
        if cfg.ffn_act in ("gelu", "relu2"):
            up = ops.linear(lw.w_up, ffn_in, "f32", out=self.up32[:T], scratch=self._scratch_for(lw.w_up))
            act = ops.bias_act(up, lw.b_up, cfg.ffn_act, out=self.act[:T])
        else:
            act = ops.linear(lw.w_gate, ffn_in, "swiglu", w_up=lw.w_up, out=self.act[:T], scratch=self._scratch_for(lw.w_gate, lw.w_up))
        if lw.ffn_sub_norm is not None:                            # bitnet: RMSNorm over the gated activation, in front of the down matrix
            v = act.float()
            act = (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + cfg.rms_eps) * lw.ffn_sub_norm).to(act.dtype)
        d = ops.linear(lw.w_down, act, "f32", out=self.delta[:T], scratch=self._scratch_for(lw.w_down))
        if tp:
            d = comm.allreduce_delta(d, il, 1)
        if lw.b_down is not None:
            d.add_(lw.b_down)
        if cfg.post_norm:
            h.add_(self._norm(d, lw.post_ffn_norm, None, out=self.xn2[:T]))
        else:
            h.add_(d)

    def _layer_fused(self, il: int, T: int, attn_fn):
        """Decode layer with the tensor-parallel all-reduces fused into the mat-vecs (parallel/fused.py):
        row-parallel GEMVs push partial rows into every peer's reduce buffer, the following GEMV's prologue
        sums them into the residual stream while it computes RMSNorm.  No NCCL call, no extra launch."""
        lw, cfg, c = self.w.layers[il], self.cfg, self.comm
        li = il - self.w.layer_range.start
        hin = self.hb[self.cur]
        if self.pending is not None:
            c.gemv(lw.wqkv, hin, ops.OUT_F32, T, out=self.qkv32[:T], norm_w=lw.attn_norm, eps=cfg.rms_eps,
                   ar_in=c.ar_in(self.pending, self.hb[1 - self.cur]))
            self.cur ^= 1
        else:
            c.gemv(lw.wqkv, hin, ops.OUT_F32, T, out=self.qkv32[:T], norm_w=lw.attn_norm, eps=cfg.rms_eps)
        if self._onepass_now and attn_fn == self._attn_decode:
            att = self._attn_decode_onepass(il, self.qkv32[:T], T)
        else:
            q = self._rope(il, self.qkv32[:T], T)
            att = attn_fn(il, q, T)
        c.gemv(lw.wo, att, 4, T, ar_out=c.ar_out(2 * li))
        act = c.gemv(lw.w_gate, self.hb[self.cur], ops.OUT_SWIGLU_F16, T, out=self.act[:T], w_up=lw.w_up, norm_w=lw.ffn_norm,
                     eps=cfg.rms_eps, ar_in=c.ar_in(2 * li, self.hb[1 - self.cur]))
        self.cur ^= 1
        c.gemv(lw.w_down, act, 4, T, ar_out=c.ar_out(2 * li + 1))
        self.pending = 2 * li + 1

    def _moe(self, lw, T: int, xn):
        """Top-k routed experts (softmax over all experts, top-k, renormalise: llama-graph.cpp:1039-1139)."""
        cfg, comm = self.cfg, self.comm
        h = self.h[:T]
        if (T <= 4 and lw.e_gate_s is not None and cfg.n_expert <= 32 and cfg.n_embd % 256 == 0 and self.F % 256 == 0
                and lw.router_w.N % 2 == 0 and cfg.moe_norm_topk and not cfg.n_ff_shared and self.device.type == "cuda"
                and cfg.moe_gating == "softmax" and lw.sel_bias is None and cfg.n_expert_groups <= 1
                and os.environ.get("DLP_MOE_GEMV", "1") == "1"):
            return self._moe_device(lw, T)
        if xn is None:
            xn = ops.rmsnorm(h, lw.ffn_norm, cfg.rms_eps, out=self.xn[:T])
        if self._moe_grouped_ok(lw):
            return self._moe_grouped(lw, T, xn)
        # host-routed fallback (CPU engine, mixed expert types): not graph-capturable
        logits = xn.float() @ lw.router.T
        probs = torch.sigmoid(logits) if cfg.moe_gating == "sigmoid" else torch.softmax(logits, -1)
        sel = probs + lw.sel_bias if lw.sel_bias is not None else probs      # DeepSeek-V3 style: the bias picks the experts, the unbiased scores weigh them
        topi = limit_to_groups(sel, cfg.n_expert_groups, cfg.n_group_used).topk(cfg.n_expert_used, -1).indices
        topw = probs.gather(-1, topi)
        if cfg.moe_norm_topk:
            topw = topw / topw.sum(-1, keepdim=True).clamp_min(6.103515625e-5)
        out = torch.zeros(T, self.E, dtype=torch.float32, device=h.device)
        if lw.w_gate is not None:        # always-on shared expert: behind a sigmoid gate (qwen2moe) or added as it is (deepseek, bailingmoe);
            act = ops.linear(lw.w_gate, xn, "swiglu", w_up=lw.w_up, scratch=self._scratch_for(lw.w_gate, lw.w_up))   # partial sums under TP, the gate is replicated
            ys = ops.linear(lw.w_down, act, "f32", scratch=self._scratch_for(lw.w_down))
            out += (torch.sigmoid(xn.float() @ lw.shared_gate)[:, None] * ys) if lw.shared_gate is not None else ys
        topi_h = topi.cpu()
        for e in torch.unique(topi_h).tolist():
            rows = (topi_h == e).any(-1).nonzero().flatten().to(h.device)
            xe = xn[rows].contiguous()
            act = ops.linear(lw.e_gate[e], xe, "swiglu", w_up=lw.e_up[e], scratch=self._scratch_for(lw.e_gate[e], lw.e_up[e]))
            ye = ops.linear(lw.e_down[e], act, "f32", scratch=self._scratch_for(lw.e_down[e]))
            we = (topw[rows] * (topi[rows] == e)).sum(-1, keepdim=True)
            out.index_add_(0, rows, we * ye)
        if comm.tp_size > 1:
            out = comm.allreduce_delta(out, -1, 1)
        h.add_(out)

    def _moe_grouped_ok(self, lw) -> bool:
        cfg = self.cfg
        return (self.device.type == "cuda" and lw.e_gate_s is not None and lw.router_wp is not None and cfg.n_expert <= 256
                and cfg.moe_gating == "softmax" and lw.sel_bias is None and cfg.n_expert_groups <= 1     # (the device router implements softmax top-k only)
                and cfg.n_expert_used <= 32 and not ops.TWO_PASS_DEQUANT and os.environ.get("DLP_MOE_GROUPED", "1") == "1")

    def _moe_grouped(self, lw, T: int, xn):
        """Any token count, any MoE family, tp >= 1, entirely on the device (graph-capturable): router GEMM -> route -> plan
        (expert-sorted order + token-tile table) -> gather -> ONE grouped tcgen05 GEMM per projection over all experts ->
        weighted combine in slot order.  Under tensor parallelism every expert's n_ff is sharded like a dense FFN, so the
        combine produces a partial sum that takes the usual all-reduce.  ref: mmid.cu + mul_mat_q ids path, topk-moe.cu."""
        cfg, comm = self.cfg, self.comm
        k, nE, E, F = cfg.n_expert_used, cfg.n_expert, self.E, self.F
        A = T * k
        if not hasattr(self, "moe_g"):
            dev, M = self.device, self.max_tokens + 8
            self.moe_g = dict(logits=torch.zeros(M, lw.router_wp.N, dtype=torch.float32, device=dev),
                              ids=torch.zeros(M * k, dtype=torch.int32, device=dev), w=torch.zeros(M * k, dtype=torch.float32, device=dev),
                              plan=ops.MoePlan(M * k, nE, dev), xs=torch.zeros(M * k, E, dtype=torch.float16, device=dev),
                              act=torch.zeros(M * k, F, dtype=torch.float16, device=dev), y=torch.zeros(M * k, E, dtype=torch.float32, device=dev))
        b = self.moe_g
        h = self.h[:T]
        ops.linear(lw.router_wp, xn, "f32", out=b["logits"][:T])     # router (padded to a full 128-row tile for the GEMM sizes)
        ops.moe_route_ex(b["logits"][:T], nE, k, cfg.moe_norm_topk, b["ids"][:A], b["w"][:A])
        per = max(1, A // nE)
        bt = 16 if per <= 16 else 32 if per <= 32 else 64 if per <= 96 else 128
        ops.moe_plan(b["ids"][:A], nE, bt, b["plan"])
        ops.moe_gather(xn, b["plan"], k, A, b["xs"][:A])
        ops.linear_grouped(lw.e_gate_s, b["xs"][:A], b["act"][:A], "swiglu", nE, F, bt, b["plan"], w_up=lw.e_up_s)
        ops.linear_grouped(lw.e_down_s, b["act"][:A], b["y"][:A], "f32", nE, E, bt, b["plan"])
        tp = comm.tp_size > 1
        tgt = h
        if tp:
            tgt = self.delta[:T]
            tgt.zero_()
        ops.moe_combine(b["y"][:A], b["plan"], b["w"][:A], T, k, tgt)
        if lw.w_gate is not None:        # always-on shared expert, gated (qwen2moe) or not (deepseek, bailingmoe): device ops, no sync
            act = ops.linear(lw.w_gate, xn, "swiglu", w_up=lw.w_up)
            ys = ops.linear(lw.w_down, act, "f32")
            if lw.shared_gate is not None:
                tgt.addcmul_(torch.sigmoid(xn.float() @ lw.shared_gate)[:, None], ys)
            else:
                tgt.add_(ys)
        if tp:
            h.add_(comm.allreduce_delta(tgt, -1, 1))

    def _moe_device(self, lw, T: int):
        """Decode-path MoE entirely on the device (graph-capturable): fused-norm router mat-vec -> dlp_moe_route ->
        per selected expert a fused-norm gate/up+SwiGLU mat-vec and a weighted down mat-vec accumulating into h."""
        cfg = self.cfg
        k, F = cfg.n_expert_used, self.F
        if not hasattr(self, "moe_buf"):
            dev = self.device
            self.moe_buf = dict(logits=torch.zeros(4, cfg.n_expert, dtype=torch.float32, device=dev),
                                ids=torch.zeros(4 * k, dtype=torch.int32, device=dev),
                                w=torch.zeros(4 * k, dtype=torch.float32, device=dev),
                                act=torch.zeros(4 * k, F, dtype=torch.float16, device=dev))
        b = self.moe_buf
        h = self.h[:T]
        # Under tensor parallelism every expert's n_ff is sharded: the down mat-vecs accumulate this rank's partial sums into a
        # zeroed delta, which takes one all-reduce per layer (the bandwidth-bound mat-vec kernels use every SM for a single token,
        # where the grouped GEMM has two token tiles to offer: config 5 ran 0.63x the reference at one request through that path)
        tp = self.comm.tp_size > 1
        tgt = self.delta[:T] if tp else h
        if tp:
            tgt.zero_()
        for t in range(T):   # router per token (T <= 4): the mat-vec path supports T in {1,2,4}, the router is tiny
            ops.linear(lw.router_w, h[t:t + 1], "f32", norm_w=lw.ffn_norm, eps=cfg.rms_eps, out=b["logits"][t:t + 1])
        ops.moe_route(b["logits"][:T], k, b["ids"][:T * k], b["w"][:T * k])
        for t in range(T):
            for j in range(k):
                job = t * k + j
                ops.gemv_moe(lw.e_gate_s, h[t:t + 1], b["act"][job:job + 1], "swiglu", b["ids"][job:job + 1], F, w_up=lw.e_up_s,
                             norm_w=lw.ffn_norm, eps=cfg.rms_eps)
            for j in range(k):   # all gate/up mat-vecs read h[t] before any down mat-vec updates it
                job = t * k + j
                ops.gemv_moe(lw.e_down_s, b["act"][job:job + 1], tgt[t:t + 1], "add", b["ids"][job:job + 1], self.E,
                             out_scale=b["w"][job:job + 1])
        if tp:
            h.add_(self.comm.allreduce_delta(tgt, -1, 1))

    # ---- attention flavours ---------------------------------------------------------------------
    def _attn_decode(self, il, q, T):
        return ops.attn_paged(q, self.kv.k[il], self.kv.v[il], self.kv.block_table, self.kv_len[:T], self.Hq, self.Hkv,
                              self.D, bt_row=self.active[:T], n_splits=self.attn_splits, workspace=self.attn_ws,
                              out=self.att[:T])

    def _attn_decode_onepass(self, il, qkv, T):
        rope_on = self.inv_freq_off is None or self.cfg.layer_uses_rope(il)     # off: no (cos, sin) table, angles from the zero table
        return ops.attn_decode_rope(qkv, self._inv_freq(il), self.kv.k[il], self.kv.v[il], self.kv.block_table, self.kv_len[:T],
                                    self.Hq, self.Hkv, self.D, bt_row=self.active[:T], n_splits=self.attn_cluster,
                                    neox=self.cfg.rope_neox, attn_scale=self.attn_scale, out=self.att[:T],
                                    rope_cs=self.rope_cs[:T] if (self.rope_cs is not None and rope_on) else None)

    def _attn_prefill_for(self, sid: int, pos0: int):
        def fn(il, q, T):
            return ops.attn_prefill(q, self.kv.k[il], self.kv.v[il], self.kv.block_table[sid], pos0, self.Hq, self.Hkv,
                                    self.D, out=self.att[:T])
        return fn

    # ---- stage-level forward ------------------------------------------------------------------
    def _embed(self, T: int, ids: torch.Tensor | None = None):
        ops.dequant_rows(self.w.tok_embd, self.tokens[:T] if ids is None else ids, out=self.h[:T])
        if self.cfg.embedding_scale != 1.0:
            self.h[:T].mul_(self.cfg.embedding_scale)
        if self.w.pos_embd is not None:      # learned absolute positions; self.pos is the device-side position of every batch row
            self.h[:T].add_(self.w.pos_embd.index_select(0, self.pos[:T].long()))

    def _run_layers(self, T: int, attn_fn, fused: bool = False):
        if fused:
            if self.cvec:
                raise NotImplementedError("control vectors are not supported on the fused tensor-parallel decode path")
            self.cur, self.pending = 0, None
            self.comm.begin_step()
            for il in self.w.layer_range:
                self._layer_fused(il, T, attn_fn)
            return
        layer = self._layer_generic if self.cfg.generic else self._layer
        for il in self.w.layer_range:
            layer(il, T, attn_fn)
            if il in self.cvec:                     # control vector: h += direction[il] (llama-adapter.cpp apply_to)
                self.h[:T].add_(self.cvec[il])
            if self.eval_callback is not None:      # tensor inspection hook (ggml-backend.cpp:1584-1616, examples/eval-callback)
                self.eval_callback("l_out", il, self.h[:T])

    def set_control_vector(self, src, scale: float = 1.0, layer_range: tuple[int, int] | None = None):
        """Load a control vector (GGUF with `direction.<layer>` tensors [n_embd], as written by the reference's
        cvector-generator; `llama_adapter_cvec`, `llama.cpp/src/llama-adapter.cpp`) or a {layer: tensor} dict; `None` clears.
        Cached decode graphs are dropped because the set of launches changes."""
        self.cvec = {}
        self._graphs.clear()
        if src is None:
            return
        if isinstance(src, str):
            from ..gguf import open_gguf
            import numpy as np
            f = open_gguf(src)
            src = {int(n.split(".")[1]): torch.from_numpy(np.array(f.array(n), dtype=np.float32, copy=True))
                   for n in f.tensors if n.startswith("direction.")}
        lo, hi = layer_range or (0, self.cfg.n_layer)
        for il, v in src.items():
            if lo <= il < hi and il in self.w.layers:
                if v.numel() != self.E:
                    raise ValueError(f"control vector for layer {il} has {v.numel()} elements, model has n_embd={self.E}")
                self.cvec[il] = (v.float() * scale).to(self.device)

    def memory_breakdown(self) -> dict:
        """Bytes held on this rank's device by category (`llama_memory_breakdown_print`, llama.h:1501)."""
        def nb(*ts):
            return sum(t.numel() * t.element_size() for t in ts if t is not None)
        act = nb(self.h, self.delta, self.xn, self.qkv32, self.qkv16, self.q, self.att, self.act, self.logits, *(self.attn_ws or ()))
        return {"model": self.w.nbytes(), "kv_cache": self.kv.bytes(), "compute": act,
                "graphs": len(self._graphs), "device": str(self.device)}

    def _forward_sp(self, T: int, attn_fn, logits_row: int | None, local_logits: bool = False):
        """Tensor-parallel forward for T > 4 tokens in the sequence-parallel layout (parallel/fused.py): the residual
        stream is sharded by token rows; every row-parallel GEMM pushes its partial tiles to the owner ranks from its
        epilogue (GEMM -> reduce-scatter), the owners fold + RMSNorm + push fp16 rows to everyone (all-gather), and the
        next column-parallel GEMM's TMA producer consumes the gathered rows as their flags arrive.  No NCCL call."""
        c, cfg, w = self.comm, self.cfg, self.w
        tp, rank = c.tp_size, c.tp_rank
        c.begin_sp_step()
        Tp = (T + tp - 1) // tp * tp
        rows = Tp // tp
        r0 = rank * rows
        nv = max(0, min(T, r0 + rows) - r0)          # valid (non-padding) rows of this rank's shard
        hs = self.h_sh[:rows]
        if nv < rows:
            hs[nv:].zero_()
        if nv > 0:
            if self.first_stage:
                ops.dequant_rows(w.tok_embd, self.tokens[r0:r0 + nv], out=hs[:nv])
                if cfg.embedding_scale != 1.0:
                    hs[:nv].mul_(cfg.embedding_scale)
            else:
                hs[:nv].copy_(self.h[r0:r0 + nv])
        if Tp > T:
            self.att[T:Tp].zero_()
        E, qd, Fl = self.E, self.Hq * self.D, self.F
        pending = None
        for il in w.layer_range:
            lw = w.layers[il]
            c.sp_norm_ag(hs, rows, lw.attn_norm, cfg.rms_eps, g=0, red=pending)
            c.sp_gemm(lw.wqkv, c.sp_gather_view(0, Tp), E, Tp, rows, 0, y=self.qkv16[:Tp], x_from_g=0)
            q = self._rope(il, self.qkv16[:T], T)
            attn_fn(il, q, T)
            c.sp_gemm(lw.wo, self.att[:Tp], qd, Tp, rows, 5, rs_to=0)
            c.sp_norm_ag(hs, rows, lw.ffn_norm, cfg.rms_eps, g=1, red=0)
            c.sp_gemm(lw.w_gate, c.sp_gather_view(1, Tp), E, Tp, rows, 3, y=self.act[:Tp], w_up=lw.w_up, x_from_g=1)
            c.sp_gemm(lw.w_down, self.act[:Tp], Fl, Tp, rows, 5, rs_to=1)
            pending = 1
        if self.last_stage and logits_row is not None:
            c.sp_norm_ag(hs, rows, w.out_norm, cfg.rms_eps, g=0, red=pending)
            owner = logits_row // rows
            lg = c.sp_gemm(w.lm_head, self.G[0][logits_row:logits_row + 1], E, 1, Tp, 1, y=self.logits[:1], x_from_g=0, x_owner=owner)
            return c.allgather_logits(lg, w.vocab_range, cfg.vocab_size)
        if self.last_stage:   # logits for every row (batched decode): all rows of the gathered, normalised stream
            c.sp_norm_ag(hs, rows, w.out_norm, cfg.rms_eps, g=0, red=pending)
            lg = c.sp_gemm(w.lm_head, c.sp_gather_view(0, Tp), E, T, rows, 1, y=self.logits[:T], x_from_g=0)
            if local_logits:      # the caller finishes with the sharded arg-max: no gather
                return lg
            return c.allgather_logits(lg, w.vocab_range, cfg.vocab_size)
        # pipeline hand-off: fold the last reduction, then rebuild the full [T, E] stream for the next stage
        c.sp_norm_ag(hs, rows, None, cfg.rms_eps, g=None, red=pending)
        full = torch.empty(tp, rows, E, dtype=torch.float32, device=self.device)
        torch.distributed.all_gather_into_tensor(full.view(-1), hs.reshape(-1), group=c.group)
        self.h[:T].copy_(full.view(Tp, E)[:T])
        return None

    def _finish_fused(self, T: int, greedy: bool = False):
        """Fold the last pending all-reduce: into the lm_head prologue (last stage) or a stand-alone consumer.  Greedy steps
        finish with the sharded arg-max (8-byte candidates exchanged through symmetric memory) instead of a logits all-gather."""
        c, cfg = self.comm, self.cfg
        hin, hout = self.hb[self.cur], self.hb[1 - self.cur]
        if self.last_stage:
            c.gemv(self.w.lm_head, hin, ops.OUT_F32, T, out=self.logits[:T], norm_w=self.w.out_norm, eps=cfg.rms_eps,
                   ar_in=c.ar_in(self.pending, hout))
            if greedy:
                c.argmax_sharded(self.logits[:T], self.w.vocab_range[0], self.next_tok[:T])
                return None
            return c.allgather_logits(self.logits[:T], self.w.vocab_range, cfg.vocab_size)
        c.consume(hin, hout, self.pending, T)
        if self.cur == 0:           # hand-off buffer for the next stage is always self.h
            self.h[:T].copy_(hout[:T])
        return None

    def _head(self, rows: torch.Tensor | slice, n: int) -> torch.Tensor:
        """logits for `n` rows of h (already gathered to the front of the xn buffer) -> self.logits[:n]."""
        cfg = self.cfg
        hsel = self.h[rows]
        if n <= 4 and n != 3 and cfg.norm_type == "rms":
            lg = ops.linear(self.w.lm_head, hsel.contiguous(), "f32", norm_w=self.w.out_norm, eps=cfg.rms_eps, out=self.logits[:n])
        else:
            xn = self._norm(hsel.contiguous(), self.w.out_norm, self.w.out_norm_b)
            lg = ops.linear(self.w.lm_head, xn, "f32", out=self.logits[:n], scratch=self._scratch_for(self.w.lm_head))
        if self.w.out_b is not None:        # phi2: the output matrix has a bias
            lg.add_(self.w.out_b)
        return self.comm.allgather_logits(lg, self.w.vocab_range, cfg.vocab_size)

    # ---- public: prefill ------------------------------------------------------------------------
    @torch.no_grad()
    def prefill(self, sid: int, tokens, hidden_in: torch.Tensor | None = None, chunk: int | None = None,
                all_logits: bool = False):
        """Append `tokens` to sequence `sid`.  Returns logits [vocab] of the last token (last stage) or the
        hidden states [T, E] to hand to the next pipeline stage.  `all_logits=True` (single chunk, tp == 1 or the
        NCCL TP path) returns logits [T, vocab] for every appended token (speculative verification, perplexity)."""
        seq = self.seqs[sid]
        tokens = torch.as_tensor(tokens, dtype=torch.int32)
        n = tokens.numel()
        chunk = min(chunk or self.max_tokens, self.max_tokens)
        if all_logits and self.sp_ok and n > 4:     # refuse before anything is appended to the cache
            raise NotImplementedError("all_logits is not available on the sequence-parallel prefill path")
        outs = []
        for c0 in range(0, n, chunk):
            T = min(chunk, n - c0)
            pos0 = seq.length
            self.kv.ensure(seq, pos0 + T)
            pos = torch.arange(pos0, pos0 + T, dtype=torch.int32)
            pages = torch.tensor(seq.pages, dtype=torch.int32)
            slot = pages[(pos // self.page_size).long()] * self.page_size + pos % self.page_size
            self.pos[:T].copy_(pos, non_blocking=True)
            self.slot[:T].copy_(slot.to(torch.int32), non_blocking=True)
            sp = self.sp_ok and T > 4
            if self.first_stage:
                self.tokens[:T].copy_(tokens[c0:c0 + T], non_blocking=True)
                if not sp:
                    self._embed(T)
            else:
                self.h[:T].copy_(hidden_in[c0:c0 + T])
            if sp:
                sp_logits = self._forward_sp(T, self._attn_prefill_for(sid, pos0), T - 1)
            else:
                self._run_layers(T, self._attn_prefill_for(sid, pos0))
            seq.length = pos0 + T
            if not self.last_stage:
                outs.append(self.h[:T].clone())
        self.seq_len_dev[sid] = seq.length
        if not self.last_stage:
            return torch.cat(outs, 0)
        if sp:
            if all_logits:
                raise NotImplementedError("all_logits is not available on the sequence-parallel prefill path")
            return sp_logits[0]
        if all_logits:
            if n > chunk:
                raise ValueError("all_logits needs the tokens to fit one chunk")
            xn = self._norm(self.h[:T], self.w.out_norm, self.w.out_norm_b)
            lg = ops.linear(self.w.lm_head, xn, "f32", scratch=self._scratch_for(self.w.lm_head))
            if self.w.out_b is not None:
                lg.add_(self.w.out_b)
            return self.comm.allgather_logits(lg, self.w.vocab_range, self.cfg.vocab_size)
        return self._head(slice(T - 1, T), 1)[0]

    @torch.no_grad()
    def embed(self, tokens, pooling: str = "mean", normalize: bool = True) -> torch.Tensor:
        """Sentence embedding of `tokens`: final-norm hidden states pooled over the sequence (`mean` | `last` | `cls` |
        `none` -> [T, E]) and L2-normalised.  Parity: `llama_get_embeddings_seq` with `--pooling`
        (`llama.cpp/src/llama-context.cpp` embeddings output, `examples/embedding`); runs in a scratch sequence slot."""
        if self.fused_tp or not (self.first_stage and self.last_stage):
            raise NotImplementedError("embed() runs on a single-stage engine (tp == 1 or the NCCL TP path)")
        tokens = list(tokens)
        if len(tokens) > self.max_tokens:
            raise ValueError(f"{len(tokens)} tokens exceed the batch limit {self.max_tokens}")
        sid = self.alloc_seq()
        try:
            seq = self.seqs[sid]
            T = len(tokens)
            self.kv.ensure(seq, T)
            pos = torch.arange(T, dtype=torch.int32)
            pages = torch.tensor(seq.pages, dtype=torch.int32)
            self.pos[:T].copy_(pos)
            self.slot[:T].copy_((pages[(pos // self.page_size).long()] * self.page_size + pos % self.page_size).to(torch.int32))
            self.tokens[:T].copy_(torch.as_tensor(tokens, dtype=torch.int32))
            self._embed(T)
            self._run_layers(T, self._attn_prefill_for(sid, 0))
            xn = self._norm(self.h[:T], self.w.out_norm, self.w.out_norm_b).float()
        finally:
            self.free_seq(sid)
        if pooling == "mean":
            e = xn.mean(0)
        elif pooling == "last":
            e = xn[-1]
        elif pooling == "cls":
            e = xn[0]
        elif pooling == "none":
            e = xn
        else:
            raise ValueError(f"unknown pooling {pooling!r}")
        return torch.nn.functional.normalize(e, dim=-1) if normalize else e

    # ---- public: decode -------------------------------------------------------------------------
    def _decode_body(self, B: int, sample: bool):
        """One decode step for the sequence slots listed in self.active[:B]; everything stays on the device."""
        ops.advance(self.seq_len_dev, self.active, self.pos, self.slot, self.kv_len, self.kv.block_table, self.page_size, B,
                    inv_freq=self.inv_freq, rope_cs=self.rope_cs)
        fused = self.fused_tp and B in (1, 2, 4)
        sp = self.sp_ok and B > 4
        if sp:   # batched decode in the sequence-parallel layout (embedding happens per token shard inside)
            self.tokens[:B].copy_(self.next_tok[:B])
            greedy = sample and self._greedy_now and self.last_stage
            lg = self._forward_sp(B, self._attn_decode, None, local_logits=greedy)
            if greedy:
                self.comm.argmax_sharded(lg, self.w.vocab_range[0], self.next_tok[:B])
                return
        else:
            if self.first_stage:
                self._embed(B, self.next_tok[:B])  # input ids = previously sampled (or host-provided) tokens
            self._run_layers(B, self._attn_decode, fused)
        if fused:
            lg = self._finish_fused(B, greedy=sample and self._greedy_now)
            if sample and self._greedy_now:
                return      # sharded arg-max already wrote next_tok on every rank
        if self.last_stage:
            if not fused and not sp:
                lg = self._head(slice(0, B), B)
            self.logits_full = lg  # full-vocab logits (gathered across TP ranks); static across graph replays
            if sample:   # records are indexed by sequence slot through the row -> slot map: nothing of a request is baked in
                ops.sample_dev(lg, self.samp, self.active[:B], self.next_tok[:B])

    @torch.no_grad()
    def decode(self, sids: list[int], tokens=None, hidden_in: torch.Tensor | None = None, sample: bool = True,
               use_graph: bool = True):
        """One token for each sequence in `sids` (any subset of slots; batch row i belongs to sids[i]).
        tokens: host ints / tensor, or None to feed back the previously sampled tokens (device resident).
        Returns next-token ids (device int32 [B]) when sampling, else logits [B, vocab]; or hidden states."""
        B = len(sids)
        for sid in sids:
            seq = self.seqs[sid]
            self.kv.ensure(seq, seq.length + 1)
        if sids != self._active_host:
            self.active[:B].copy_(torch.tensor(sids, dtype=torch.int32), non_blocking=False)
            self._active_host = list(sids)
        if tokens is not None and self.first_stage:
            self.next_tok[:B].copy_(torch.as_tensor(tokens, dtype=torch.int32), non_blocking=True)
        if hidden_in is not None:
            self.h[:B].copy_(hidden_in)
        moe_lw = next((lw for lw in self.w.layers.values() if lw.router is not None), None)
        moe_host = moe_lw is not None and not self._moe_grouped_ok(moe_lw)   # routing through the host: no graph
        graphable = (use_graph and self.eval_callback is None and self.device.type == "cuda" and not moe_host
                     and not getattr(self.comm, "no_graph", False)    # (the sequence-parallel path is capturable too: its epochs are
                     and not (self.sp_ok and B > 4 and os.environ.get("DLP_SP_GRAPH", "1") == "0"))   # positions inside the pass on top of a device-side counter; DLP_SP_GRAPH=0: A/B)
        self._onepass_now = self.attn_onepass and max(self.seqs[sid].length for sid in sids) + 1 >= self.attn_onepass_min
        if sample and self.last_stage:
            self._sync_sampling(sids)
        # all rows greedy on the fused tensor-parallel path: sharded arg-max instead of gathering the logits (host-side fact, so
        # it is part of the graph key; at most one extra graph per batch size)
        self._greedy_now = bool(sample and self.last_stage and self.fused_tp and (B in (1, 2, 4) or (self.sp_ok and 4 < B <= 64)) and
                                os.environ.get("DLP_TP_ARGMAX", "1") == "1" and all(self._samp_host[sid][3] <= 0 for sid in sids))
        if graphable:
            # only the batch size, the output kind and the attention variant shape the graph (sampler parameters are device data)
            key = (B, sample, self._onepass_now, self._greedy_now)
            g = self._graphs.pop(key, None)
            if g is None:
                while len(self._graphs) + 2 > self.max_graphs:    # least recently used graphs go first
                    old = next(iter(self._graphs))
                    del self._graphs[old]
                    self._graph_launches.pop(old, None)
                if self.attn_onepass:   # capture the other attention variant now as well: the switch at 64 keys must not stall
                    want = self._onepass_now
                    self._onepass_now = not want
                    self._capture(B, sample, key[:2] + (not want,) + key[3:])
                    self._onepass_now = want
                g = self._capture(B, sample, key)
                self._graphs.pop(key)
            self._graphs[key] = g      # most recently used last
            g.replay()
            self.launches_replayed += self._graph_launches.get(key, 0)
        else:
            self._decode_body(B, sample)
        for sid in sids:
            self.seqs[sid].length += 1
        if not self.last_stage:
            return self.h[:B]
        return self.next_tok[:B] if sample else self.logits_full[:B]

    def _capture(self, B: int, sample: bool, key=None):
        # warm-up (also lets every kernel set its attributes), then roll back the bookkeeping it advanced
        saved = self.seq_len_dev.clone()
        saved_ctr = self.step_ctr.clone()
        saved_h = self.h[:B].clone()   # non-first stages consume their input hidden state in place: the warm-up must not eat it
        saved_tok = self.next_tok.clone()   # the warm-up samples into next_tok: the pending input token must survive it
        saved_samp = self.samp.clone()      # ... and advances the draw counters
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._decode_body(B, sample)
        torch.cuda.current_stream().wait_stream(s)
        self.seq_len_dev.copy_(saved)
        self.step_ctr.copy_(saved_ctr)
        self.h[:B].copy_(saved_h)
        self.next_tok.copy_(saved_tok)
        self.samp.copy_(saved_samp)
        g = torch.cuda.CUDAGraph()
        n0 = ops.launch_count
        with torch.cuda.graph(g):
            self._decode_body(B, sample)
        self.launches_per_graph = getattr(self, "launches_per_graph", {})
        self.launches_per_graph[(B, sample)] = ops.launch_count - n0
        self._graph_launches[key or (B, sample)] = ops.launch_count - n0
        self.seq_len_dev.copy_(saved)
        self.step_ctr.copy_(saved_ctr)
        self._graphs[key or (B, sample)] = g
        return g

    # ---- convenience ------------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, prompt: list[int], n_predict: int, sid: int | None = None, stop_ids: tuple[int, ...] = ()):
        """Greedy / sampled generation for one sequence (slot 0 path); returns the generated ids."""
        own = sid is None
        if own:
            sid = self.alloc_seq()
        logits = self.prefill(sid, prompt)
        s = self.sampling
        if s["temp"] <= 0:
            tok = ops.argmax(logits[None])
        else:   # first token from the prefill logits: same per-sequence record (and random stream) as the decode steps
            self._sync_sampling([sid])
            tok = ops.sample_dev(logits[None], self.samp, torch.tensor([sid], dtype=torch.int32, device=self.device),
                                 torch.empty(1, dtype=torch.int32, device=self.device))
        self.next_tok[:1].copy_(tok)
        out = [int(tok[0])]
        for _ in range(n_predict - 1):
            if out[-1] in stop_ids:
                break
            nt = self.decode([sid])
            out.append(int(nt[0]))
        if own:
            self.free_seq(sid)
        return out
