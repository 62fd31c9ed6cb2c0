"""Host-side sampler chain (oracle for the device sampler and the CPU-worker path).

Parity: default chain order of the reference (`llama.cpp/common/common.h:201-210`, `common/sampling.cpp:179-311`):
penalties -> top-k -> top-p -> min-p -> temperature -> dist.  Implemented samplers: repetition/frequency/presence
penalties (`llama-sampling.cpp:2777`), top-k (`:1348`), top-p (`:1543`), min-p (`:1704`), typical (`:1814`),
temperature (`:1918`), greedy (`:1011`), seeded multinomial (`:1255`), logit bias (`:3455`).
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

    @property
    def greedy(self) -> bool:
        return self.temperature <= 0.0


def apply_penalties(logits: torch.Tensor, history: list[int], p: SamplingParams) -> torch.Tensor:
    if not history or (p.repeat_penalty == 1.0 and p.frequency_penalty == 0.0 and p.presence_penalty == 0.0):
        return logits
    logits = logits.clone()
    recent = history[-p.repeat_last_n:] if p.repeat_last_n > 0 else history
    ids, counts = torch.tensor(recent).unique(return_counts=True)
    vals = logits[ids]
    vals = torch.where(vals > 0, vals / p.repeat_penalty, vals * p.repeat_penalty)
    vals = vals - counts.float() * p.frequency_penalty - p.presence_penalty
    logits[ids] = vals
    return logits


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


class Sampler:
    """Stateful per-request sampler (history for penalties, private RNG stream)."""

    def __init__(self, params: SamplingParams):
        self.p = params
        self.history: list[int] = []
        self.gen = torch.Generator().manual_seed(params.seed)

    def accept(self, token: int):
        self.history.append(token)

    def sample(self, logits: torch.Tensor) -> int:
        p = self.p
        row = logits.float().cpu()
        for tid, bias in p.logit_bias.items():
            row[tid] += bias
        row = apply_penalties(row, self.history, p)
        if p.greedy:
            tok = int(row.argmax())
        else:
            vals, ids = filter_candidates(row, p.top_k, p.top_p, p.min_p, p.typical_p)
            probs = torch.softmax(vals / p.temperature, -1)
            tok = int(ids[torch.multinomial(probs, 1, generator=self.gen)])
        self.accept(tok)
        return tok
