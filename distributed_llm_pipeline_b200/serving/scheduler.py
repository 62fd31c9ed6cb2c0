"""Serving core: request queue, slots, continuous batching with chunked prefill, stop conditions, timings, metrics.

Parity: the slot engine of `llama-server` / `llama-cli` (`llama.cpp/tools/server/server-context.cpp`): slot state
machine (`:34-41`), `update_slots` (`:1917-2860`) = add one token per generating slot, then prompt chunks up to
`n_batch`, then decode and sample per slot; prompt-prefix reuse by longest common prefix (`:960-1060`); deferred
requests when all slots are busy; per-request timings (`timings.prompt_per_second`, `predicted_per_second`) and the
Prometheus counters of `/metrics` (`:3225-3300`).  Differences by design: the KV cache is paged so prefix reuse is a
length truncation (no cell bookkeeping), decode runs as one CUDA-graph replay per step for whatever subset of slots is
generating, and the default sampler chain runs on the device.
"""
from __future__ import annotations

import itertools
import threading
import time
from collections import deque
from dataclasses import dataclass, field
from enum import Enum
from typing import Callable

import torch

from ..utils.tokenizer import StreamDecoder, Tokenizer
from .sampler import Sampler, SamplingParams


class SlotState(Enum):
    IDLE = 0
    PROMPT = 1
    GENERATING = 2


@dataclass
class Request:
    prompt: str | list[int]
    n_predict: int = 200
    sampling: SamplingParams = field(default_factory=SamplingParams)
    stop: list[str] = field(default_factory=list)
    on_event: Callable[[str, str], None] | None = None  # ("token" | "log" | "done" | "error", content)
    ignore_eos: bool = False
    cache_prompt: bool = True
    n_keep: int = 0             # tokens of the prompt that survive a context shift (-1 = the whole prompt)
    n_probs: int = 0            # > 0: record the log-probability of every generated token and of the n_probs most likely ones
    post_sampling_probs: bool = False   # record the probabilities the token was drawn from (after the sampler chain) instead
    n_draft: int = 0            # > 0: speculative decoding with prompt-lookup drafts of up to this many tokens (greedy requests)
    id: int = 0
    # filled in by the scheduler
    prompt_ids: list[int] = field(default_factory=list)
    output_ids: list[int] = field(default_factory=list)
    probs: list[dict] = field(default_factory=list)       # per generated token: {"id", "token", "logprob", "top_logprobs": [...]}
    text: str = ""
    finish_reason: str | None = None
    stop_type: str = "none"     # llama-server's name for it: "none" | "eos" | "limit" | "word"
    stopping_word: str = ""
    truncated: bool = False     # generation ended because the context was full
    id_slot: int = -1
    tokens_cached: int = 0      # tokens of this request resident in the slot's cache when it ended
    t_submit: float = 0.0
    t_start: float = 0.0
    t_first_token: float = 0.0
    t_done: float = 0.0
    n_prompt_processed: int = 0
    n_prompt_cached: int = 0
    done: threading.Event = field(default_factory=threading.Event)
    cancelled: bool = False

    def timings(self) -> dict:
        p_ms = max(1e-6, (self.t_first_token - self.t_start) * 1e3) if self.t_first_token else 0.0
        g_ms = max(1e-6, (self.t_done - self.t_first_token) * 1e3) if self.t_done and self.t_first_token else 0.0
        n_gen = max(0, len(self.output_ids) - 1)
        return {
            "prompt_n": self.n_prompt_processed, "cache_n": self.n_prompt_cached, "prompt_ms": round(p_ms, 3),
            "prompt_per_token_ms": round(p_ms / max(1, self.n_prompt_processed), 4),
            "prompt_per_second": round(self.n_prompt_processed / p_ms * 1e3, 2) if p_ms else 0.0,
            "predicted_n": len(self.output_ids), "predicted_ms": round(g_ms, 3), "predicted_per_token_ms": round(g_ms / max(1, n_gen), 4),
            "predicted_per_second": round(n_gen / g_ms * 1e3, 2) if g_ms and n_gen else 0.0,
            "ttft_ms": round((self.t_first_token - self.t_submit) * 1e3, 3) if self.t_first_token else None,
        }


@dataclass
class Slot:
    idx: int
    sid: int | None = None
    state: SlotState = SlotState.IDLE
    req: Request | None = None
    cache_tokens: list[int] = field(default_factory=list)  # tokens whose KV is resident (prompt cache)
    n_past: int = 0
    sampler: Sampler | None = None
    decoder: StreamDecoder | None = None
    next_token: int = 0
    t_last_used: float = 0.0
    last_req: Request | None = None     # the slot's latest request, kept after it ends (GET /slots shows its parameters)


class Metrics:
    """Counters exported at /metrics (names follow the reference's Prometheus endpoint)."""

    def __init__(self):
        self.lock = threading.Lock()
        self.c = dict(prompt_tokens_total=0, prompt_seconds_total=0.0, tokens_predicted_total=0,
                      tokens_predicted_seconds_total=0.0, n_decode_total=0, n_busy_slots_total=0, n_tokens_max=0,
                      requests_total=0, requests_failed_total=0)
        self.g = dict(requests_processing=0, requests_deferred=0, kv_pages_free=0, kv_pages_total=0)

    def add(self, **kw):
        with self.lock:
            for k, v in kw.items():
                self.c[k] = self.c.get(k, 0) + v

    def prometheus(self) -> str:
        with self.lock:
            c, g = dict(self.c), dict(self.g)
        lines = []

        def emit(name, kind, helptext, value):
            lines.append(f"# HELP llamacpp:{name} {helptext}")
            lines.append(f"# TYPE llamacpp:{name} {kind}")
            lines.append(f"llamacpp:{name} {value}")

        emit("prompt_tokens_total", "counter", "Number of prompt tokens processed.", c["prompt_tokens_total"])
        emit("prompt_seconds_total", "counter", "Prompt process time", round(c["prompt_seconds_total"], 6))
        emit("tokens_predicted_total", "counter", "Number of generation tokens processed.", c["tokens_predicted_total"])
        emit("tokens_predicted_seconds_total", "counter", "Predict process time", round(c["tokens_predicted_seconds_total"], 6))
        emit("n_decode_total", "counter", "Total number of decode steps", c["n_decode_total"])
        emit("n_tokens_max", "counter", "Largest observed n_tokens.", c["n_tokens_max"])
        emit("n_busy_slots_per_decode", "counter", "Average number of busy slots per decode step",
             round(c["n_busy_slots_total"] / max(1, c["n_decode_total"]), 4))
        emit("prompt_tokens_seconds", "gauge", "Average prompt throughput in tokens/s.",
             round(c["prompt_tokens_total"] / c["prompt_seconds_total"], 3) if c["prompt_seconds_total"] else 0)
        emit("predicted_tokens_seconds", "gauge", "Average generation throughput in tokens/s.",
             round(c["tokens_predicted_total"] / c["tokens_predicted_seconds_total"], 3) if c["tokens_predicted_seconds_total"] else 0)
        emit("requests_processing", "gauge", "Number of requests processing.", g["requests_processing"])
        emit("requests_deferred", "gauge", "Number of requests deferred.", g["requests_deferred"])
        emit("kv_cache_pages_free", "gauge", "Free pages of the paged KV cache.", g["kv_pages_free"])
        emit("kv_cache_pages_total", "gauge", "Pages of the paged KV cache.", g["kv_pages_total"])
        return "\n".join(lines) + "\n"


class Scheduler:
    """Owns the engine; `step()` advances every active slot by one scheduling quantum."""

    def __init__(self, engine, tokenizer: Tokenizer, n_batch: int | None = None, log: Callable[[str], None] | None = None,
                 n_slots: int | None = None, draft_engine=None, n_draft_default: int = 0):
        self.eng, self.tok = engine, tokenizer
        # `-md / --model-draft`: a smaller engine with the same vocabulary proposes tokens, the target verifies a whole block in
        # one forward (server-context.cpp:2029-2068 + common/speculative.cpp).  One draft sequence per slot, kept in step with it.
        self.draft_eng = draft_engine
        self.n_draft_default = n_draft_default or (16 if draft_engine is not None else 0)
        self._drafters: dict[int, object] = {}
        self.n_slots = min(n_slots or engine.max_seqs, engine.max_seqs)
        self.n_batch = min(n_batch or engine.max_tokens, engine.max_tokens)
        self.slots = [Slot(i) for i in range(self.n_slots)]
        self.queue: deque[Request] = deque()
        self.lock = threading.Lock()
        self.wake = threading.Condition(self.lock)
        self.metrics = Metrics()
        self.metrics.g["kv_pages_total"] = engine.kv.n_pages
        self._ids = itertools.count(1)
        self._stop = False
        self._thread: threading.Thread | None = None
        self._jobs: list = []
        self.context_shift = False          # --context-shift: slide the window instead of stopping at n_ctx
        self.log = log or (lambda s: None)

    def warmup(self) -> float:
        """Run one tiny prefill and one decode step per batch size before traffic arrives, so that kernel set-up and CUDA-graph
        capture (one graph per batch size, both attention variants) are paid at start-up and not by the first requests - the
        reference server warms its context up the same way (`common_init_from_params` warm-up run, common/common.cpp).
        Works through any engine (the distributed one replays the calls on every rank).  Returns the seconds spent."""
        eng = self.eng
        if eng.device.type != "cuda":
            return 0.0
        t0 = time.time()
        n = min(self.n_slots, eng.max_seqs)
        sids = [eng.alloc_seq() for _ in range(n)]
        try:
            for s in sids:
                eng.prefill(s, [self.tok.bos_id if 0 <= self.tok.bos_id < self.tok.n_vocab else 1, 5, 6, 7, 8, 9])
            for B in range(1, n + 1):
                eng.decode(sids[:B], torch.full((B,), 5, dtype=torch.int32), sample=True)
            torch.cuda.synchronize()
        finally:
            for s in sids:
                eng.free_seq(s)
        return time.time() - t0

    # ---- submission -------------------------------------------------------------------------------
    def submit(self, req: Request) -> Request:
        req.id = next(self._ids)
        req.t_submit = time.time()
        if isinstance(req.prompt, str):
            req.prompt_ids = self.tok.encode(req.prompt)
        else:
            req.prompt_ids = list(req.prompt)
        if not req.prompt_ids:
            req.prompt_ids = [self.tok.bos_id]
        limit = self.eng.n_ctx - 1
        if len(req.prompt_ids) > limit:  # keep the tail, like the reference's prompt truncation
            req.prompt_ids = req.prompt_ids[:1] + req.prompt_ids[-(limit - 1):]
        with self.wake:
            self.queue.append(req)
            self.metrics.add(requests_total=1)
            self.metrics.g["requests_deferred"] = len(self.queue)
            self.wake.notify_all()
        return req

    def cancel(self, req: Request):
        req.cancelled = True

    # ---- slot selection with prompt-prefix reuse --------------------------------------------------------
    @staticmethod
    def _lcp(a: list[int], b: list[int]) -> int:
        n = 0
        for x, y in zip(a, b):
            if x != y:
                break
            n += 1
        return n

    def _pick_slot(self, req: Request) -> Slot | None:
        want = getattr(req, "want_slot", -1)
        if 0 <= want < len(self.slots):       # `id_slot`: the request waits for that slot (server-context.cpp get_slot_by_id)
            return self.slots[want] if self.slots[want].state == SlotState.IDLE else None
        idle = [s for s in self.slots if s.state == SlotState.IDLE]
        if not idle:
            return None
        if req.cache_prompt:
            best = max(idle, key=lambda s: (self._lcp(s.cache_tokens, req.prompt_ids), -s.t_last_used))
            if self._lcp(best.cache_tokens, req.prompt_ids) > 0:
                return best
        return min(idle, key=lambda s: (len(s.cache_tokens) > 0, s.t_last_used))  # prefer empty, then least recently used

    def _launch(self, slot: Slot, req: Request):
        eng = self.eng
        if slot.sid is None:
            slot.sid = eng.alloc_seq()
        keep = self._lcp(slot.cache_tokens, req.prompt_ids) if req.cache_prompt else 0
        keep = min(keep, len(req.prompt_ids) - 1)  # always evaluate at least the last prompt token
        eng.truncate_seq(slot.sid, keep)
        slot.cache_tokens = slot.cache_tokens[:keep]
        slot.n_past = keep
        slot.req, slot.last_req, slot.state = req, req, SlotState.PROMPT
        slot.sampler = Sampler(req.sampling, self.tok)
        slot.sampler.history = list(req.prompt_ids)      # penalties and DRY look back over the prompt too (server-context.cpp: prompt tokens
        #                                                  are accepted into the sampler without advancing the grammar)
        if hasattr(eng, "set_seq_sampling"):    # the device sampler's record of this sequence; restarts its random stream
            p = req.sampling
            eng.set_seq_sampling(slot.sid, p.top_k, p.top_p, p.min_p, p.temperature, p.seed)
        slot.decoder = StreamDecoder(self.tok)
        req.n_prompt_cached = keep
        req.t_start = time.time()
        self._emit(req, "log", f"slot {slot.idx}: request {req.id} | prompt {len(req.prompt_ids)} tokens, "
                               f"{keep} reused from cache, n_predict {req.n_predict}\n")

    def _emit(self, req: Request, kind: str, content: str):
        if req.on_event is not None:
            try:
                req.on_event(kind, content)
            except Exception:  # a broken consumer must not take the engine down
                req.cancelled = True

    # ---- one scheduling quantum ----------------------------------------------------------------------
    @torch.no_grad()
    def step(self) -> bool:
        """Returns False when there was nothing to do."""
        with self.lock:
            jobs, self._jobs = self._jobs, []
        for fn, fut in jobs:            # engine-side calls from other threads (embeddings), between batches
            try:
                fut["value"] = fn(self.eng)
            except Exception as e:       # noqa: BLE001 - handed back to the caller
                fut["error"] = e
            fut["done"].set()
        with self.lock:
            while self.queue:
                slot = self._pick_slot(self.queue[0])
                if slot is None:
                    break
                self._launch(slot, self.queue.popleft())
            self.metrics.g["requests_deferred"] = len(self.queue)
        active = [s for s in self.slots if s.state != SlotState.IDLE]
        self.metrics.g["requests_processing"] = len(active)
        self.metrics.g["kv_pages_free"] = len(self.eng.kv.free)
        if not active:
            return False
        # 1) prompt chunks, sharing the n_batch budget
        budget = self.n_batch
        for s in active:
            if s.state != SlotState.PROMPT or budget <= 0:
                continue
            req = s.req
            if req.cancelled:
                self._finish(s, "cancelled")
                continue
            remaining = req.prompt_ids[s.n_past:]
            chunk = remaining[:budget]
            t0 = time.time()
            try:
                logits = self.eng.prefill(s.sid, chunk)
            except RuntimeError as e:
                self._fail(s, str(e))
                continue
            budget -= len(chunk)
            s.n_past += len(chunk)
            s.cache_tokens.extend(chunk)
            req.n_prompt_processed += len(chunk)
            # counters as server-context.cpp:523-547 keeps them: every forward is a decode, n_tokens_max is the longest slot context,
            # and the token sampled from the prompt's last logits is a predicted token
            self.metrics.add(prompt_tokens_total=len(chunk), n_decode_total=1, n_busy_slots_total=1)
            self.metrics.c["n_tokens_max"] = max(self.metrics.c["n_tokens_max"], s.n_past)
            if s.n_past == len(req.prompt_ids):  # prompt complete: sample the first token from its logits
                tok = self._sample_rows(logits[None], [s])[0]
                self.metrics.add(prompt_seconds_total=time.time() - req.t_start, tokens_predicted_total=1)
                req.t_first_token = time.time()
                s.state = SlotState.GENERATING
                self._accept(s, tok)
            else:
                self.metrics.add(prompt_seconds_total=0.0)
            _ = t0
        # 2) one decode token for every generating slot (single batched step)
        gen = [s for s in self.slots if s.state == SlotState.GENERATING]
        if gen:
            for s in list(gen):
                if s.req.cancelled:
                    self._finish(s, "cancelled")
                    gen.remove(s)
            for s in list(gen):          # speculative slots: draft from the sequence's own history, verify in one forward
                if self._speculate(s):
                    gen.remove(s)
            gen = [s for s in gen if s.state == SlotState.GENERATING]
            if gen:
                t0 = time.time()
                sids = [s.sid for s in gen]
                toks = torch.tensor([s.next_token for s in gen], dtype=torch.int32)
                try:
                    # every sequence carries its own sampler record on the device: mixed parameters share one step
                    device_sampling = all(self._device_samplable(s) for s in gen)
                    if device_sampling:
                        out = self.eng.decode(sids, toks, sample=True)
                        new = out.cpu().tolist()
                        for s in gen:
                            s.sampler.accept(s.next_token)
                    else:
                        logits = self.eng.decode(sids, toks, sample=False)
                        new = self._sample_rows(logits, gen)
                except RuntimeError as e:
                    for s in gen:
                        self._fail(s, str(e))
                    return True
                dt = time.time() - t0
                self.metrics.add(n_decode_total=1, n_busy_slots_total=len(gen), tokens_predicted_total=len(gen),
                                 tokens_predicted_seconds_total=dt)
                for s, t in zip(gen, new):
                    s.cache_tokens.append(s.next_token)
                    s.n_past += 1
                    self.metrics.c["n_tokens_max"] = max(self.metrics.c["n_tokens_max"], s.n_past)
                    self._accept(s, int(t))
        return True

    def _speculate(self, s: Slot) -> bool:
        """Draft-and-verify step for one slot (`server-context.cpp` speculative path with the n-gram "lookup" drafter of
        `common/ngram-cache.cpp`): the pending token plus up to n_draft drafted tokens go through ONE forward that returns logits
        for every position; the longest prefix the model agrees with is accepted.  Returns False when the slot should take the
        normal batched decode step instead (no draft found, sampling is not greedy, engine cannot return all logits)."""
        req = s.req
        n_draft = req.n_draft if req.n_draft > 0 else (self.n_draft_default if req.n_draft == 0 else 0)     # n_draft < 0: off for this request
        if n_draft <= 0 or req.n_probs > 0 or not req.sampling.greedy or req.sampling.host_only or not hasattr(self.eng, "copy_seq"):
            return False
        from .speculative import ModelDrafter, NgramDrafter
        history = s.cache_tokens + [s.next_token]
        room = min(n_draft, req.n_predict - len(req.output_ids), self.eng.n_ctx - len(history) - 2, self.eng.max_tokens - 1)
        draft: list[int] = []
        if room > 0 and self.draft_eng is not None and len(history) + room < self.draft_eng.n_ctx:
            d = self._drafters.get(s.idx)
            if d is None:
                d = self._drafters[s.idx] = ModelDrafter(self.draft_eng)
                d.start(history)
            try:
                draft = d.draft(history, room)
            except (RuntimeError, NotImplementedError) as e:     # the draft model must never take the request down with it
                self.log(f"draft model failed ({e}); continuing without it for slot {s.idx}")
                draft = []
        elif room > 0:
            draft = NgramDrafter().draft(history, room)
        if not draft:
            return False
        t0 = time.time()
        block = [s.next_token] + draft
        try:
            pred = self.eng.prefill(s.sid, block, all_logits=True).argmax(-1).tolist()
        except (RuntimeError, NotImplementedError):
            return False
        n_ok = 0
        while n_ok < len(draft) and pred[n_ok] == draft[n_ok]:
            n_ok += 1
        self.eng.truncate_seq(s.sid, len(history) + n_ok)              # drop the rejected tail of the block from the cache
        self.metrics.add(n_decode_total=1, n_busy_slots_total=1, tokens_predicted_total=n_ok + 1, tokens_predicted_seconds_total=time.time() - t0,
                         draft_tokens_total=len(draft), draft_tokens_accepted_total=n_ok)
        s.cache_tokens.append(s.next_token)                            # the pending token is in the cache now
        s.n_past += 1
        for t in draft[:n_ok]:                                         # accepted drafts: emitted, and already cached
            self._accept(s, int(t))
            if s.state != SlotState.GENERATING:
                break
            s.cache_tokens.append(int(t))
            s.n_past += 1
        else:
            self._accept(s, int(pred[n_ok]))                           # the model's own next token becomes the pending one
        if s.sid is not None and self.eng.seq_len(s.sid) != len(s.cache_tokens):
            self.eng.truncate_seq(s.sid, len(s.cache_tokens))          # finished inside the block: cache mirrors cache_tokens again
        d = self._drafters.get(s.idx)
        if d is not None:
            d.accept(len(history) + n_ok)                              # the draft cache keeps the agreed prefix only
        return True

    def _device_samplable(self, s: Slot) -> bool:
        p = s.req.sampling
        return not p.host_only and s.req.n_probs <= 0 and self.eng.device.type == "cuda"

    def _sample_rows(self, logits: torch.Tensor, slots: list[Slot]) -> list[int]:
        rows = logits.float().cpu()
        out = []
        for i, s in enumerate(slots):
            tok = s.sampler.sample(rows[i])
            if s.req.n_probs > 0:       # `n_probs` / OpenAI `logprobs`: probabilities of the full softmax, before the sampler chain
                lp = torch.log_softmax(rows[i], -1)      # (server-context.cpp populate_token_probs, post_sampling_probs = false)
                top = lp.topk(min(s.req.n_probs, lp.numel()))
                txt = self.tok.token_texts()

                def entry(j, **val):
                    return {"id": int(j), "token": txt[int(j)] or "", "bytes": list(self.tok.token_bytes(int(j))), **val}
                dist = getattr(s.sampler, "last_dist", None)
                if s.req.post_sampling_probs and dist is not None:
                    # the distribution the token was drawn from; padded with the likeliest other tokens at probability 0
                    ids, pr = dist
                    order = torch.argsort(pr, descending=True, stable=True)[:s.req.n_probs]
                    cand = [(int(ids[k]), float(pr[k])) for k in order]
                    seen = {j for j, _ in cand}
                    cand += [(int(j), 0.0) for j in lp.topk(min(2 * s.req.n_probs + 1, lp.numel())).indices if int(j) not in seen][:s.req.n_probs - len(cand)]
                    chosen = next((q for j, q in zip(ids.tolist(), pr.tolist()) if j == tok), 0.0)
                    s.req.probs.append({**entry(tok, prob=chosen, logprob=float(lp[tok])), "top_probs": [entry(j, prob=q) for j, q in cand]})
                else:
                    s.req.probs.append({**entry(tok, logprob=float(lp[tok])),
                                        "top_logprobs": [entry(j, logprob=float(v)) for v, j in zip(top.values, top.indices)]})
            out.append(tok)
        return out

    # ---- token acceptance / completion -------------------------------------------------------------------
    def _accept(self, s: Slot, tok: int):
        req = s.req
        req.output_ids.append(tok)
        s.next_token = tok
        is_stop = tok in self.tok.stop_ids and not req.ignore_eos
        if not is_stop:
            piece = s.decoder.push(tok)
            if piece:
                req.text += piece
                hit = next((st for st in req.stop if st and st in req.text), None)
                if hit is not None:
                    cut = req.text.index(hit)
                    emitted = len(req.text) - len(piece)
                    if cut > emitted:
                        self._emit(req, "token", req.text[emitted:cut])
                    req.text = req.text[:cut]
                    req.stop_type, req.stopping_word = "word", hit
                    self._finish(s, "stop")
                    return
                self._emit(req, "token", piece)
        if is_stop:
            req.output_ids.pop()
            req.stop_type = "eos"
            self._finish(s, "stop")
        elif len(req.output_ids) >= req.n_predict:
            req.stop_type = "limit"
            self._finish(s, "length")
        elif s.n_past + 1 >= self.eng.n_ctx:
            if self.context_shift and hasattr(self.eng, "shift_seq"):
                # keep the first n_keep tokens, drop half of the rest (server-context.cpp:1944-2003)
                n_keep = (req.n_keep if req.n_keep >= 0 else len(req.prompt_ids)) + (1 if self.tok.add_bos else 0)     # n_keep does not count BOS
                n_keep = min(n_keep, self.eng.n_ctx - 4)
                n_discard = (s.n_past - n_keep) // 2
                self.eng.shift_seq(s.sid, n_keep, n_discard)
                del s.cache_tokens[n_keep:n_keep + n_discard]
                s.n_past -= n_discard
                self.metrics.add(context_shifts_total=1)
                self._emit(req, "log", f"slot {s.idx}: context shift, kept {n_keep}, discarded {n_discard} tokens\n")
            else:
                req.stop_type, req.truncated = "limit", True
                self._finish(s, "length")  # context full (the reference stops as well unless --context-shift)

    def _finish(self, s: Slot, reason: str):
        req = s.req
        tail = s.decoder.flush() if s.decoder else ""
        if tail and reason != "cancelled":
            req.text += tail
            self._emit(req, "token", tail)
        req.finish_reason = reason
        req.t_done = time.time()
        req.id_slot, req.tokens_cached = s.idx, s.n_past
        t = req.timings()
        self._emit(req, "log", f"slot {s.idx}: request {req.id} done ({reason}) | prompt eval {t['prompt_n']} tok "
                               f"{t['prompt_per_second']} tok/s | eval {t['predicted_n']} tok {t['predicted_per_second']} tok/s\n")
        self._emit(req, "done", reason)
        s.state, s.req, s.t_last_used = SlotState.IDLE, None, time.time()
        self._drop_drafter(s)
        if not req.cache_prompt:
            self.eng.truncate_seq(s.sid, 0)
            s.cache_tokens, s.n_past = [], 0
        req.done.set()

    def _fail(self, s: Slot, msg: str):
        req = s.req
        self.metrics.add(requests_failed_total=1)
        req.finish_reason = "error"
        req.t_done = time.time()
        self._emit(req, "error", msg)
        self._emit(req, "done", "error")
        try:
            self.eng.truncate_seq(s.sid, 0)
        except Exception:
            pass
        s.state, s.req, s.cache_tokens, s.n_past = SlotState.IDLE, None, [], 0
        self._drop_drafter(s)
        req.done.set()

    def _drop_drafter(self, s: Slot):
        d = self._drafters.pop(s.idx, None)
        if d is not None:
            try:
                d.close()
            except Exception:
                pass

    # ---- background loop --------------------------------------------------------------------------------
    def start(self):
        self._thread = threading.Thread(target=self._loop, name="dlp-scheduler", daemon=True)
        self._thread.start()

    def stop(self):
        self._stop = True
        with self.wake:
            self.wake.notify_all()
        if self._thread:
            self._thread.join(timeout=5)

    def _loop(self):
        if self.eng.device.type == "cuda":
            torch.cuda.set_device(self.eng.device)
        while not self._stop:
            try:
                idle = not self.step()
            except Exception as e:  # noqa: BLE001 - an engine fault must not leave clients waiting on a dead thread
                self._engine_fault(e)
                continue
            if idle:
                with self.wake:
                    if not self.queue and not self._jobs and not self._stop:
                        self.wake.wait(timeout=0.05)

    def _engine_fault(self, e: Exception):
        """Fail every request in flight with the error.  A CUDA fault poisons the context for good: say so and leave, so that a
        supervisor (dlp-orchestrator --spawn, systemd, ...) can restart the service instead of clients hanging forever."""
        import traceback
        msg = f"engine fault: {type(e).__name__}: {e}".split("\n")[0]
        self.log(msg + "\n")
        traceback.print_exc()
        for s in self.slots:
            if s.req is not None:
                try:
                    self._fail(s, msg)
                except Exception:  # noqa: BLE001
                    s.req.done.set()
                    s.state, s.req = SlotState.IDLE, None
        with self.lock:
            pending, self.queue = list(self.queue), deque()
        for req in pending:
            req.finish_reason = "error"
            self._emit(req, "error", msg)
            self._emit(req, "done", "error")
            req.done.set()
        if "CUDA" in str(e) or "cuda" in type(e).__name__.lower() or "Accelerator" in type(e).__name__:
            import os
            import sys
            sys.stderr.write("fatal device error: the engine service exits\n")
            sys.stderr.flush()
            os._exit(3)

    # ---- blocking convenience ------------------------------------------------------------------------------
    def generate(self, prompt, **kw) -> Request:
        req = self.submit(Request(prompt=prompt, **kw))
        if self._thread is None:
            while not req.done.is_set():
                self.step()
        else:
            req.done.wait()
        return req

    def run_on_engine(self, fn, timeout: float = 120.0):
        """Run `fn(engine)` on the scheduler thread (the only thread that touches the engine) and return its value."""
        fut = {"done": threading.Event()}
        with self.wake:
            self._jobs.append((fn, fut))
            self.wake.notify()
        if self._thread is None:
            self.step()
        if not fut["done"].wait(timeout):
            raise TimeoutError("engine job timed out")
        if "error" in fut:
            raise fut["error"]
        return fut["value"]

    def embed(self, text_or_ids, pooling: str = "mean", normalize: bool = True) -> list[float]:
        """Pooled sentence embedding through the engine (llama-server `/embedding`, `server-context.cpp` embedding
        tasks).  Needs one engine sequence slot beyond the serving slots (build_service reserves it)."""
        ids = self.tok.encode(text_or_ids) if isinstance(text_or_ids, str) else list(text_or_ids)
        ids = ids[: self.eng.max_tokens]
        self.metrics.add(prompt_tokens_total=len(ids))

        def job(eng):
            if not eng._free_sids:       # every engine sequence is parked in a slot: give up one idle slot's cache
                idle = next((s for s in self.slots if s.state == SlotState.IDLE and s.sid is not None), None)
                if idle is None:
                    raise RuntimeError("no free sequence slot for an embedding request")
                eng.free_seq(idle.sid)
                idle.sid, idle.cache_tokens, idle.n_past = None, [], 0
            return eng.embed(ids, pooling, normalize).cpu().tolist()
        return self.run_on_engine(job)

    def slots_info(self) -> list[dict]:
        """GET /slots.  The reference's fields (`server_slot::to_json`, server-context.cpp:360-420: id, n_ctx, speculative, is_processing and,
        once the slot has served a task, id_task / params / next_token) plus this scheduler's own (state, n_past, n_cached, request)."""
        from dataclasses import asdict
        out = []
        for s in self.slots:
            d = {"id": s.idx, "n_ctx": self.eng.n_ctx, "speculative": self.draft_eng is not None, "is_processing": s.state != SlotState.IDLE,
                 "state": s.state.name.lower(), "n_past": s.n_past, "n_cached": len(s.cache_tokens), "request": s.req.id if s.req else None}
            r = s.last_req
            if r is not None:
                sp = {k: v for k, v in asdict(r.sampling).items() if k not in ("logit_bias", "dry_breakers", "json_schema", "grammar_trigger_words",
                                                                                "grammar_trigger_patterns")}
                sp.update(max_tokens=r.n_predict, n_predict=r.n_predict, n_keep=r.n_keep, ignore_eos=r.ignore_eos, n_probs=r.n_probs,
                          stream=False, lora=[])
                sp["speculative.n_max"] = r.n_draft or self.n_draft_default
                d.update(id_task=r.id, params=sp,
                         next_token=[{"has_next_token": r.finish_reason is None, "has_new_line": "\n" in r.text,
                                      "n_remain": max(0, r.n_predict - len(r.output_ids)), "n_decoded": len(r.output_ids)}])
            out.append(d)
        return out
