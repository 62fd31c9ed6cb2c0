#!/usr/bin/env python
"""Short deterministic workload for ncu: one pp<N> prefill + a few eager decode steps of the bench model."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import ensure_model  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-3-8b")
ap.add_argument("--ftype", default="F16")
ap.add_argument("--prefill", type=int, default=512)
ap.add_argument("--decode", type=int, default=3)
ap.add_argument("--layers", type=int, default=0, help="profile only the first N layers (0 = all)")
ap.add_argument("--emulate-tp", type=int, default=0, help="rank 0 of an N-way tensor-parallel step on this one GPU (parallel/fused.py make_emulated)")
a = ap.parse_args()
if a.layers:      # same layer shapes, fewer layers: a 4-layer file is written in seconds instead of a minute
    from bench import MODEL_DIR
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.config import preset
    path = os.path.join(MODEL_DIR, f"{a.model}-{a.layers}L-{a.ftype}.gguf")
    if not os.path.exists(path):
        write_synthetic_gguf(path, preset(a.model, n_layer=a.layers), a.ftype, seed=0)
else:
    path = ensure_model(a.model, a.ftype, 0, lambda: None)
comm = None
if a.emulate_tp > 1:
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    comm = FusedComm.make_emulated(a.emulate_tp)
w = load_weights(path, "cuda:0", tp_rank=0, tp_size=max(1, a.emulate_tp))
eng = LlamaEngine(w, n_ctx=2048, max_seqs=1, page_size=64, max_batch_tokens=max(a.prefill, 16), comm=comm)
sid = eng.alloc_seq()
g = torch.Generator().manual_seed(0)
if a.prefill:
    eng.prefill(sid, torch.randint(3, eng.cfg.vocab_size, (a.prefill,), generator=g, dtype=torch.int32))
else:
    eng.prefill(sid, [1])
torch.cuda.synchronize()
for _ in range(a.decode):
    eng.decode([sid], [5], use_graph=False, sample=a.emulate_tp > 1)   # (emulated TP: ends with the sharded arg-max)
torch.cuda.synchronize()
print("prof_step done")
