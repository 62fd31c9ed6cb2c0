#!/usr/bin/env python
"""Device-side timeline of one decode step (graph replay, PDL on): %globaltimer stamps written by CTA (0,0,0) of the mat-vec,
RoPE and attention kernels (dlp_set_trace).  Shows where the time between the big mat-vecs goes."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import MODEL_DIR  # noqa: E402
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402
from distributed_llm_pipeline_b200.ops import lib as _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=6)
ap.add_argument("--prefill", type=int, default=512)
ap.add_argument("--ftype", default="F16")
a = ap.parse_args()
path = os.path.join(MODEL_DIR, f"llama-3-8b-{a.layers}L-{a.ftype}.gguf")
if not os.path.exists(path):
    write_synthetic_gguf(path, preset("llama-3-8b", n_layer=a.layers), a.ftype, seed=0)
w = load_weights(path, "cuda:0")
NAMES = {1: "gemv", 2: "rope", 3: "attn"}
EV = {0: "begin", 1: "waited", 2: "main_done", 3: "end"}
CAP = 4096
for onepass in (0, 1):
    eng = LlamaEngine(w, n_ctx=2048, max_seqs=1, page_size=64, max_batch_tokens=max(a.prefill, 16))
    eng.attn_onepass = bool(onepass)
    buf = torch.zeros(CAP + 2, dtype=torch.int64, device="cuda")
    buf[1] = CAP
    _lib.check(_lib.load().dlp_set_trace(buf.data_ptr()), "set_trace")
    sid = eng.alloc_seq()
    g = torch.Generator().manual_seed(0)
    eng.prefill(sid, torch.randint(3, eng.cfg.vocab_size, (a.prefill,), generator=g, dtype=torch.int32))
    for _ in range(6):                      # capture + warm replays
        eng.decode([sid])
    torch.cuda.synchronize()
    buf[0] = 0
    eng.decode([sid])
    torch.cuda.synchronize()
    _lib.check(_lib.load().dlp_set_trace(None), "set_trace")
    n = int(buf[0])
    rec = buf[2:2 + min(n, CAP)].cpu().tolist()
    ev = sorted(((r & 0xFFFFFFFFFFFF), (r >> 48) >> 8, (r >> 48) & 0xFF) for r in rec)
    t0 = ev[0][0]
    print(f"==== onepass={onepass}: {n} records, step span {(ev[-1][0] - t0) / 1e3:.1f} us")
    # print layers 2..3 in full (steady state), relative to the first record of that window
    gemv_begins = [i for i, e in enumerate(ev) if e[1] == 1 and e[2] == 0]
    per_layer = 4
    lo = gemv_begins[1 + 2 * per_layer] if len(gemv_begins) > 3 * per_layer else 0
    hi = gemv_begins[1 + 3 * per_layer] if len(gemv_begins) > 1 + 3 * per_layer else len(ev)
    base = ev[lo][0]
    for t, k, e in ev[lo:hi + 4]:
        print(f"   {(t - base) / 1e3:8.2f} us  {NAMES.get(k, k):5s} {EV.get(e, e)}")
    print(f"   layer period: {(ev[hi][0] - base) / 1e3:.2f} us")
