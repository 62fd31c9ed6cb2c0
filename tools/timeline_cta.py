#!/usr/bin/env python
"""Per-CTA timeline of the decode mat-vecs (dlp_set_trace, all-CTA mode): spread of CTA begin / main-loop-done / end times per
kernel - shows how much of each kernel boundary is tail (slowest CTA) and how much is launch + dependency wait."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from bench import MODEL_DIR  # noqa: E402
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402
from distributed_llm_pipeline_b200.ops import lib as _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--ftype", default="F16")
a = ap.parse_args()
path = os.path.join(MODEL_DIR, f"llama-3-8b-{a.layers}L-{a.ftype}.gguf")
if not os.path.exists(path):
    write_synthetic_gguf(path, preset("llama-3-8b", n_layer=a.layers), a.ftype, seed=0)
w = load_weights(path, "cuda:0")
CAP = 1 << 16
eng = LlamaEngine(w, n_ctx=2048, max_seqs=1, page_size=64, max_batch_tokens=512)
buf = torch.zeros(CAP + 2, dtype=torch.int64, device="cuda")
buf[1] = CAP | (1 << 63) - (1 << 64) if False else CAP      # plain capacity first
sid = eng.alloc_seq()
g = torch.Generator().manual_seed(0)
eng.prefill(sid, torch.randint(3, eng.cfg.vocab_size, (512,), generator=g, dtype=torch.int32))
_lib.check(_lib.load().dlp_set_trace(buf.data_ptr()), "set_trace")
for _ in range(6):
    eng.decode([sid])
torch.cuda.synchronize()
buf[1] = torch.tensor(CAP - (1 << 63), dtype=torch.int64)   # sign bit = bit 63 set: all-CTA mode; low 32 bits = CAP
buf[0] = 0
eng.decode([sid])
torch.cuda.synchronize()
_lib.check(_lib.load().dlp_set_trace(None), "set_trace")
n = int(buf[0])
rec = buf[2:2 + min(n, CAP)].cpu().numpy().astype(np.uint64)
t = (rec & np.uint64(0xFFFFFFFFFFFF)).astype(np.int64)
ident = (rec >> np.uint64(48)).astype(np.int64)
allc = (ident & 0xC000) == 0xC000
cta = (ident >> 4) & 0x3FF
ev = ident & 0xF
t0 = t.min()
print(f"{n} records, {int(allc.sum())} per-CTA")
# split per-CTA records into kernels: a kernel = run of records between successive global minima of 'begin' of CTA 0
order = np.argsort(t)
t, cta, ev, allc = t[order], cta[order], ev[order], allc[order]
b0 = np.where(allc & (cta == 0) & (ev == 0))[0]
bounds = list(t[b0]) + [t.max() + 1]
for k in range(len(b0)):
    lo, hi = bounds[k], bounds[k + 1]
    # events of kernel k: begin within [lo, hi) belong to it; main_done/end are matched by order per CTA (first after begin)
    sel = allc & (t >= lo - 3000) & (t < hi + 30000)
    out = {}
    for e in (0, 1, 2, 3):
        m = sel & (ev == e)
        # take for each CTA the first event of type e at or after that CTA's begin in this kernel
        vals = []
        for c in np.unique(cta[m]):
            tb = t[sel & (ev == 0) & (cta == c) & (t >= lo - 3000) & (t < hi)]
            if tb.size == 0:
                continue
            te = t[m & (cta == c) & (t >= tb[0])]
            if te.size:
                vals.append(te[0])
        out[e] = np.array(vals)
    if out[0].size == 0:
        continue
    base = out[0].min()
    f = lambda x: f"{(x.min() - base) / 1e3:6.2f}..{(np.median(x) - base) / 1e3:6.2f}..{(x.max() - base) / 1e3:6.2f}" if x.size else "-"
    print(f"kernel {k:2d} @{(base - t0) / 1e3:8.2f} us  n_cta {out[0].size:3d} | begin {f(out[0])} | waited {f(out[1])} | main_done {f(out[2])} | end {f(out[3])}")
