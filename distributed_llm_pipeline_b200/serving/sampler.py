"""Host-side sampler chain (oracle for the device sampler and the CPU-worker path).

Parity: default chain order of the reference (`llama.cpp/common/common.h:201-210`, `common/sampling.cpp:179-311`):
penalties -> top-k -> top-p -> min-p -> temperature -> dist.  Implemented samplers: repetition/frequency/presence
penalties (`llama-sampling.cpp:2777`), top-k (`:1348`), top-p (`:1543`), min-p (`:1704`), typical (`:1814`),
DRY (`:3216`), temperature (`:1918`), dynamic temperature (`temp_ext :2115`), XTC (`:2222`), mirostat v1 / v2 (`:2341`, `:2445`), top-n-sigma (`:2876`),
grammar (`:2625`, see grammar.py), greedy (`:1011`), seeded multinomial (`:1255`), logit bias (`:3455`).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch


@dataclass
class SamplingParams:
    temperature: float = 0.8
    top_k: int = 40
    top_p: float = 0.95
    min_p: float = 0.05
    typical_p: float = 1.0
    repeat_penalty: float = 1.0
    frequency_penalty: float = 0.0
    presence_penalty: float = 0.0
    repeat_last_n: int = 64
    seed: int = 0
    logit_bias: dict[int, float] = field(default_factory=dict)
    top_n_sigma: float = -1.0          # keep logits within n standard deviations of the maximum (<= 0: off)
    dynatemp_range: float = 0.0        # entropy-scaled temperature in [temp - range, temp + range]
    dynatemp_exponent: float = 1.0
    xtc_probability: float = 0.0       # "exclude top choices": with this probability drop all but the least likely
    xtc_threshold: float = 0.1         # candidate whose probability is >= threshold
    mirostat: int = 0                  # 1 = mirostat v1, 2 = mirostat v2 (both replace top-k/top-p/min-p)
    mirostat_tau: float = 5.0
    mirostat_eta: float = 0.1
    dry_multiplier: float = 0.0        # DRY ("don't repeat yourself"): penalty = multiplier * base^(match_len - allowed_length)
    dry_base: float = 1.75
    dry_allowed_length: int = 2
    dry_penalty_last_n: int = -1       # -1 = whole history
    dry_breakers: tuple = ()           # token ids that end a repeated span (the reference's sequence breakers, tokenised)
    grammar: str = ""                  # GBNF text; "" = unconstrained
    json_schema: dict | None = None    # converted to GBNF when `grammar` is empty
    # lazy grammar (`llama_sampler_init_grammar_lazy_patterns`, llama-sampling.cpp:2625-2775): the grammar stays off until the
    # generated text contains one of the trigger words / matches one of the trigger regexes; from the start of that match on, the
    # output must follow the grammar (tool-call syntaxes: free text first, then a strictly formed call)
    grammar_lazy: bool = False
    grammar_trigger_words: tuple = ()
    grammar_trigger_patterns: tuple = ()
    infill: bool = False               # fill-in-the-middle filter: prefer ending over low-confidence text, merge prefix-sharing tokens

    @property
    def host_only(self) -> bool:
        """True when the chain needs a sampler the device kernel does not implement."""
        return bool(self.repeat_penalty != 1.0 or self.frequency_penalty or self.presence_penalty or self.logit_bias
                    or self.typical_p < 1.0 or self.top_n_sigma > 0 or self.dynatemp_range > 0 or self.xtc_probability > 0
                    or self.mirostat or self.grammar or self.json_schema is not None or self.dry_multiplier > 0 or self.infill)

    @property
    def greedy(self) -> bool:
        return self.temperature <= 0.0


def apply_penalties(logits: torch.Tensor, history: list[int], p: SamplingParams) -> torch.Tensor:
    # repeat_last_n: 0 disables the penalties, -1 covers the whole history (llama-sampling.cpp:2777 `penalty_last_n`)
    if not history or p.repeat_last_n == 0 or (p.repeat_penalty == 1.0 and p.frequency_penalty == 0.0 and p.presence_penalty == 0.0):
        return logits
    logits = logits.clone()
    recent = history[-p.repeat_last_n:] if p.repeat_last_n > 0 else history
    ids, counts = torch.tensor(recent).unique(return_counts=True)
    vals = logits[ids]
    vals = torch.where(vals > 0, vals / p.repeat_penalty, vals * p.repeat_penalty)
    vals = vals - counts.float() * p.frequency_penalty - p.presence_penalty
    logits[ids] = vals
    return logits


def apply_dry(logits: torch.Tensor, history: list[int], p: SamplingParams) -> torch.Tensor:
    """DRY sampler (`llama-sampling.cpp:3216`): a token that would extend a span already present in the history by a match of
    length n >= allowed_length is penalised by multiplier * base^(n - allowed_length)."""
    if p.dry_multiplier <= 0 or len(history) < 2:
        return logits
    h = history if p.dry_penalty_last_n < 0 else history[-p.dry_penalty_last_n:]
    last = h[-1]
    if last in p.dry_breakers:
        return logits
    best: dict[int, int] = {}
    L = len(h)
    for i in range(L - 2, -1, -1):
        if h[i] != last:
            continue
        n = 1                                   # length of the common suffix of h[:i+1] and h
        while i - n >= 0 and n < 50 and h[i - n] == h[L - 1 - n] and h[i - n] not in p.dry_breakers:
            n += 1
        nxt = h[i + 1]
        if n > best.get(nxt, 0):
            best[nxt] = n
    out = None
    for tok, n in best.items():
        if n >= p.dry_allowed_length:
            if out is None:
                out = logits.clone()
            out[tok] -= p.dry_multiplier * p.dry_base ** (n - p.dry_allowed_length)
    return logits if out is None else out


def filter_candidates(logits: torch.Tensor, top_k: int, top_p: float, min_p: float, typical_p: float = 1.0):
    """Returns (sorted candidate logits, candidate ids) after top-k / typical / top-p / min-p."""
    V = logits.numel()
    k = V if top_k <= 0 else min(top_k, V)
    vals, ids = torch.sort(logits.float(), descending=True, stable=True)
    vals, ids = vals[:k], ids[:k]
    if typical_p < 1.0:
        probs = torch.softmax(vals, -1)
        ent = -(probs * torch.log(probs.clamp_min(1e-30))).sum()
        dev = (-torch.log(probs.clamp_min(1e-30)) - ent).abs()
        order = torch.argsort(dev, stable=True)
        cum = probs[order].cumsum(0)
        keep = int((cum < typical_p).sum()) + 1
        sel = order[:keep].sort().values
        vals, ids = vals[sel], ids[sel]
    if top_p < 1.0:
        probs = torch.softmax(vals, -1)
        cum = probs.cumsum(0)
        keep = int((cum < top_p).sum()) + 1
        vals, ids = vals[:keep], ids[:keep]
    if min_p > 0.0:
        rel = torch.exp(vals - vals[0])
        keep = max(1, int((rel >= min_p).sum()))
        vals, ids = vals[:keep], ids[:keep]
    return vals, ids


def sample_reference(logits: torch.Tensor, top_k: int, top_p: float, min_p: float, temp: float, seed: int) -> torch.Tensor:
    """Row-wise chain on [R, V] logits; returns int32 [R]."""
    out = []
    for r in range(logits.shape[0]):
        row = logits[r].float().cpu()
        if temp <= 0:
            out.append(int(row.argmax()))
            continue
        vals, ids = filter_candidates(row, top_k, top_p, min_p)
        probs = torch.softmax(vals / temp, -1)
        g = torch.Generator().manual_seed(seed * 1000003 + r)
        out.append(int(ids[torch.multinomial(probs, 1, generator=g)]))
    return torch.tensor(out, dtype=torch.int32)


def apply_infill(row: torch.Tensor, pieces: list, eog_ids, eot_id: int, n_cand: int = 64) -> torch.Tensor:
    """Fill-in-the-middle candidate filter (`llama_sampler_init_infill`, llama-sampling.cpp:3495-3670), on the `n_cand` most likely
    tokens (the reference runs it after top-k):

    1. when the end-of-generation mass is not negligible against the text mass (3 * p_eog * n > p_txt) only the EOG tokens stay:
       an infill should stop rather than ramble;
    2. a token whose text is a prefix of another candidate's text describes the same continuation: each such pair is merged into
       its more likely member (its probability gains the other's);
    3. text tokens below 0.2 are dropped; nothing left -> end of text; after renormalising, those below 1 / (n_text + 1) go too.

    Returns logits (log-probabilities of the survivors, -inf elsewhere)."""
    V = row.numel()
    probs = torch.softmax(row.float(), -1)
    vals, ids = probs.topk(min(n_cand, V))
    ids_l, p = ids.tolist(), vals.clone()
    p = p / p.sum()
    is_eog = torch.tensor([i in eog_ids for i in ids_l])
    out = torch.full_like(row, float("-inf"), dtype=torch.float32)
    p_eog, p_txt = float(p[is_eog].sum()), float(p[~is_eog].sum())
    if 3.0 * p_eog * len(ids_l) > p_txt:
        keep = ids[is_eog]
        out[keep] = torch.log(p[is_eog] / p_eog)
        return out
    text = [pieces[i] if (i < len(pieces) and pieces[i]) else "" for i in ids_l]
    alive = [True] * len(ids_l)
    pl = p.tolist()
    for i0 in range(len(ids_l)):
        for i1 in range(len(ids_l)):
            if not alive[i0]:
                break
            if i0 == i1 or not alive[i1]:
                continue
            if text[i0] and text[i1].startswith(text[i0]):
                dst, src = (i1, i0) if pl[i1] > pl[i0] else (i0, i1)
                pl[dst] += pl[src]
                pl[src], alive[src] = 0.0, False
    eog_l = is_eog.tolist()
    kept = [i for i in range(len(ids_l)) if alive[i] and (eog_l[i] or pl[i] >= 0.2)]
    n_txt = sum(1 for i in kept if not eog_l[i])
    if n_txt == 0:
        out[eot_id] = 0.0
        return out
    tot = sum(pl[i] for i in kept)
    thold = 1.0 / (n_txt + 1)
    kept = [i for i in kept if eog_l[i] or pl[i] / tot >= thold]
    tot2 = sum(pl[i] for i in kept)
    for i in kept:
        out[ids_l[i]] = float(torch.log(torch.tensor(pl[i] / tot2))) if pl[i] > 0 else float("-inf")
    return out


class Sampler:
    """Stateful per-request sampler (history for penalties, private RNG stream)."""

    def __init__(self, params: SamplingParams, tokenizer=None):
        self.p = params
        self.history: list[int] = []
        self.gen = torch.Generator().manual_seed(params.seed)
        self.mu = 2.0 * params.mirostat_tau
        self.constraint = None
        self._infill = None
        if params.infill:
            if tokenizer is None:
                raise ValueError("the infill sampler needs the tokenizer")
            eot = int(tokenizer.eot_id) if tokenizer.eot_id is not None else int(tokenizer.eos_id)
            self._infill = (tokenizer.token_texts(), set(int(i) for i in tokenizer.stop_ids), eot)
        gtext = params.grammar
        if not gtext and params.json_schema is not None:
            from .grammar import json_schema_to_gbnf
            gtext = json_schema_to_gbnf(params.json_schema)
        if gtext:
            if tokenizer is None:
                raise ValueError("grammar-constrained sampling needs the tokenizer")
            from .grammar import GrammarConstraint
            self.constraint = GrammarConstraint(gtext, tokenizer.token_texts(), tuple(tokenizer.stop_ids))
            if params.grammar_lazy and (params.grammar_trigger_words or params.grammar_trigger_patterns):
                import re
                self._pending = self.constraint          # armed by a trigger
                self.constraint = None
                self._trig = [re.compile(re.escape(w)) for w in params.grammar_trigger_words] + [re.compile(p) for p in params.grammar_trigger_patterns]
                self._texts = tokenizer.token_texts()
                self._seen = ""

    _pending = None

    def accept(self, token: int):
        self.history.append(token)
        if self.constraint is not None:
            self.constraint.accept(token)
        elif self._pending is not None:
            self._seen += self._texts[token] or ""
            hit = min((m.start() for m in (t.search(self._seen) for t in self._trig) if m), default=None)
            if hit is not None:
                tail = self._seen[hit:]
                if self._pending.g.accepts(tail):        # the grammar starts at the trigger and consumes what is already out
                    self._pending.g.accept(tail)
                    self.constraint, self._pending = self._pending, None
                else:                                    # the triggering token ran past what the grammar allows: stay unconstrained
                    self._seen = self._seen[hit + 1:]

    def _pick(self, row: torch.Tensor) -> int:
        p = self.p
        self.last_dist = None          # (token ids, probabilities) the token was finally drawn from (`post_sampling_probs`)
        if p.greedy:
            tok = int(row.argmax())
            self.last_dist = (torch.tensor([tok]), torch.tensor([1.0]))
            return tok
        if p.top_n_sigma > 0:
            row = row.masked_fill(row < row.max() - p.top_n_sigma * row.std(), float("-inf"))
        if p.mirostat == 1:
            # mirostat v1 (`llama-sampling.cpp:2341`): estimate the Zipf exponent s from the top candidates, derive the k that
            # gives the target surprise, sample from the top-k, then move mu towards tau
            vals, ids = torch.sort(row.float(), descending=True, stable=True)
            probs = torch.softmax(vals / p.temperature, -1)
            m = min(100, probs.numel() - 1)
            i = torch.arange(m, dtype=torch.float32)
            t = torch.log((i + 2) / (i + 1))
            b = torch.log(probs[:m].clamp_min(1e-30) / probs[1:m + 1].clamp_min(1e-30))
            s_hat = float((t * b).sum() / (t * t).sum().clamp_min(1e-30))
            eps = s_hat - 1.0
            n = float(probs.numel())
            k = ((eps * (2.0 ** self.mu)) / (1.0 - n ** (-eps))) ** (1.0 / s_hat) if abs(eps) > 1e-6 else n
            k = int(min(max(round(k), 1), probs.numel()))
            pk = probs[:k] / probs[:k].sum()
            j = int(torch.multinomial(pk, 1, generator=self.gen))
            self.mu -= p.mirostat_eta * (float(-torch.log2(pk[j])) - p.mirostat_tau)
            return int(ids[j])
        if p.mirostat == 2:
            probs = torch.softmax(row / p.temperature, -1)
            surprise = -torch.log2(probs.clamp_min(1e-30))
            keep = surprise <= self.mu
            if not bool(keep.any()):
                keep = surprise == surprise.min()
            probs = torch.where(keep, probs, torch.zeros_like(probs))
            probs = probs / probs.sum()
            tok = int(torch.multinomial(probs, 1, generator=self.gen))
            self.mu -= p.mirostat_eta * (float(-torch.log2(probs[tok])) - p.mirostat_tau)
            self.last_dist = (torch.arange(probs.numel()), probs)
            return tok
        vals, ids = filter_candidates(row, p.top_k, p.top_p, p.min_p, p.typical_p)
        ok = torch.isfinite(vals)
        vals, ids = vals[ok], ids[ok]
        temp = p.temperature
        if p.dynatemp_range > 0 and vals.numel() > 1:
            pr = torch.softmax(vals, -1)
            ent = float(-(pr * torch.log(pr.clamp_min(1e-30))).sum())
            max_ent = float(torch.log(torch.tensor(float(vals.numel()))))
            lo, hi = max(0.0, temp - p.dynatemp_range), temp + p.dynatemp_range
            temp = lo + (hi - lo) * (ent / max_ent) ** p.dynatemp_exponent
        probs = torch.softmax(vals / max(temp, 1e-6), -1)
        if p.xtc_probability > 0 and float(torch.rand(1, generator=self.gen)) < p.xtc_probability:
            above = (probs >= p.xtc_threshold).nonzero().flatten()
            if above.numel() >= 2:                         # candidates are sorted: drop all but the last one above
                keep = torch.ones_like(probs, dtype=torch.bool)
                keep[above[:-1]] = False
                probs = torch.where(keep, probs, torch.zeros_like(probs))
                probs = probs / probs.sum()
        self.last_dist = (ids, probs)
        return int(ids[torch.multinomial(probs, 1, generator=self.gen)])

    def sample(self, logits: torch.Tensor) -> int:
        p = self.p
        row = logits.float().cpu()
        if p.logit_bias:
            row = row.clone()        # an fp32 CPU input comes back as the caller's own tensor: never write into it
            for tid, bias in p.logit_bias.items():
                row[tid] += bias
        row = apply_penalties(row, self.history, p)
        row = apply_dry(row, self.history, p)
        if self._infill is not None:
            row = apply_infill(row, *self._infill, n_cand=max(p.top_k, 1) if p.top_k > 0 else 64)
        tok = self._pick(row)
        if self.constraint is not None and not self.constraint.allowed(tok):
            # the unconstrained pick breaks the grammar: mask the whole vocabulary and pick again (the reference checks
            # the sampled token first for the same reason: the full mask is the expensive part)
            allowed = torch.tensor(self.constraint.mask(), dtype=torch.bool)
            if not bool(allowed.any()):
                raise RuntimeError("grammar: no token can continue the output")
            tok = self._pick(row.masked_fill(~allowed, float("-inf")))
        self.accept(tok)
        return tok
