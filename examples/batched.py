#!/usr/bin/env python
"""batched: one prompt, N continuations decoded together - the prompt is evaluated once and its KV pages are shared by reference
(`copy_seq`), each sequence samples with its own seed.   (llama.cpp/examples/batched)"""
import time

import torch
from common import load, parser

ap = parser(__doc__)
ap.add_argument("-p", "--prompt", default="Hello my name is")
ap.add_argument("-n", "--predict", type=int, default=24)
ap.add_argument("-np", "--parallel", type=int, default=4)
ap.add_argument("--temp", type=float, default=0.8)
a = ap.parse_args()
eng, tok = load(a, max_seqs=a.parallel)
ids = tok.encode(a.prompt)
sids = [eng.alloc_seq() for _ in range(a.parallel)]
logits = eng.prefill(sids[0], ids)
for s in sids[1:]:
    eng.copy_seq(sids[0], s)                     # shares the prompt's pages; a page is copied only when a sequence writes into it
outs = [[] for _ in sids]
gens = [torch.Generator().manual_seed(100 + i) for i in range(a.parallel)]
probs = torch.softmax(logits.float().cpu() / max(a.temp, 1e-3), -1)
cur = [int(torch.multinomial(probs, 1, generator=g)) for g in gens]
t0 = time.time()
for step in range(a.predict):
    for o, t in zip(outs, cur):
        o.append(t)
    lg = eng.decode(sids, cur, sample=False).float().cpu()          # one batched step for all sequences
    pr = torch.softmax(lg / max(a.temp, 1e-3), -1)
    cur = [int(torch.multinomial(pr[i], 1, generator=gens[i])) for i in range(a.parallel)]
dt = time.time() - t0
for i, o in enumerate(outs):
    print(f"--- sequence {i}\n{a.prompt}{tok.decode(o, strip_leading_space=False)}")
print(f"\n[{a.parallel} x {a.predict} tokens in {dt:.2f} s = {a.parallel * a.predict / dt:.1f} tok/s aggregate]")
