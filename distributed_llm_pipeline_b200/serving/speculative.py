"""Speculative decoding: draft k tokens cheaply, verify them with ONE batched forward of the target model.

Parity: `llama.cpp/common/speculative.cpp` + `examples/speculative` (draft model), `common/ngram-cache.cpp` +
`examples/lookup` (prompt-lookup drafts from the sequence's own history), and the server's `--model-draft` path.
Verification is the greedy rule of `common_sampler_sample_and_accept_n`: position i of the drafted block is accepted
while the target's arg-max after position i-1 equals draft[i]; the first mismatch is replaced by the target's own token
(so the output is exactly the target's greedy continuation), the KV caches are rolled back with `truncate_seq`.

On B200 the verify step is a T = k+1 forward: for T <= 4 it stays on the mat-vec path, above that it is one tcgen05 GEMM
pass over the weights - the weights are streamed once for up to k+1 tokens instead of once per token.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch


@dataclass
class SpecStats:
    n_drafted: int = 0
    n_accepted: int = 0
    n_target_passes: int = 0
    n_generated: int = 0

    @property
    def acceptance(self) -> float:
        return self.n_accepted / max(1, self.n_drafted)

    @property
    def tokens_per_pass(self) -> float:
        return self.n_generated / max(1, self.n_target_passes)


class NgramDrafter:
    """Prompt-lookup drafts: find the most recent earlier occurrence of the last n tokens (n = n_max .. n_min) and propose
    what followed it (ngram-cache.cpp's "context" cache, without the static/dynamic cache files)."""

    def __init__(self, n_min: int = 1, n_max: int = 3):
        self.n_min, self.n_max = n_min, n_max

    def start(self, prompt: list[int]):
        pass

    def draft(self, history: list[int], k: int) -> list[int]:
        L = len(history)
        for n in range(min(self.n_max, L - 1), self.n_min - 1, -1):
            tail = history[L - n:]
            for s in range(L - n - 1, -1, -1):
                if history[s:s + n] == tail:
                    cont = history[s + n:s + n + k]
                    if cont:
                        return cont
        return []

    def accept(self, history_len: int):
        pass

    def close(self):
        pass


class ModelDrafter:
    """Draft with a smaller engine that shares the tokenizer (examples/speculative).  Keeps its own KV sequence in step
    with the accepted history."""

    def __init__(self, engine):
        self.eng = engine
        self.sid = None
        self.n_fed = 0          # tokens of the accepted history that are in the draft model's cache

    def start(self, prompt: list[int]):
        self.sid = self.eng.alloc_seq()
        self.n_fed = 0

    def draft(self, history: list[int], k: int) -> list[int]:
        eng, sid = self.eng, self.sid
        eng.truncate_seq(sid, min(self.n_fed, len(history) - 1))
        have = eng.seq_len(sid)
        logits = eng.prefill(sid, history[have:])
        out = [int(logits.argmax())]
        for _ in range(k - 1):
            out.append(int(eng.decode([sid], [out[-1]], sample=False)[0].argmax()))
        self._hist_len = len(history)
        return out

    def accept(self, history_len: int):
        # drafted tokens beyond the accepted prefix are dropped at the next draft() by truncate_seq
        self.n_fed = min(history_len, self.eng.seq_len(self.sid))

    def close(self):
        if self.sid is not None:
            self.eng.free_seq(self.sid)
            self.sid = None


@torch.no_grad()
def generate_speculative(engine, prompt: list[int], n_predict: int, drafter=None, n_draft: int = 8,
                         stop_ids: tuple[int, ...] = (), stats: SpecStats | None = None) -> list[int]:
    """Greedy generation of `n_predict` tokens with draft-and-verify.  Returns the generated ids (identical to plain greedy
    decoding of `engine` up to floating-point ties between the GEMM and mat-vec paths)."""
    drafter = drafter or NgramDrafter()
    stats = stats if stats is not None else SpecStats()
    sid = engine.alloc_seq()
    drafter.start(prompt)
    try:
        history = list(prompt)
        logits = engine.prefill(sid, prompt)
        stats.n_target_passes += 1
        nxt = int(logits.argmax())          # target's token for position len(history)
        out: list[int] = []
        while len(out) < n_predict:
            out.append(nxt)
            history.append(nxt)
            if nxt in stop_ids or len(out) >= n_predict:
                break
            room = min(n_draft, n_predict - len(out), engine.n_ctx - len(history) - 1, engine.max_tokens - 1)
            draft = drafter.draft(history, room) if room > 0 else []
            block = [nxt] + draft                                   # cached length is len(history) - 1 here
            lg = engine.prefill(sid, block, all_logits=True)        # [len(block), vocab]
            stats.n_target_passes += 1
            pred = lg.argmax(-1).tolist()                           # pred[i] = target token after block[:i+1]
            n_ok = 0
            while n_ok < len(draft) and pred[n_ok] == draft[n_ok] and draft[n_ok] not in stop_ids \
                    and len(out) + n_ok < n_predict - 1:
                n_ok += 1
            stats.n_drafted += len(draft)
            stats.n_accepted += n_ok
            out.extend(draft[:n_ok])
            history.extend(draft[:n_ok])
            engine.truncate_seq(sid, len(history))                  # drop the rejected tail of the block
            drafter.accept(len(history))
            nxt = pred[n_ok]
        stats.n_generated += len(out)
        return out
    finally:
        drafter.close()
        engine.free_seq(sid)
