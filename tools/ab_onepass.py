#!/usr/bin/env python
"""Decode ms/step with the one-pass attention kernel on / off, for a short and the full layer stack (same process, no trace)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import MODEL_DIR, ensure_model  # noqa: E402
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402

layers = 6
path = os.path.join(MODEL_DIR, f"llama-3-8b-{layers}L-F16.gguf")
if not os.path.exists(path):
    write_synthetic_gguf(path, preset("llama-3-8b", n_layer=layers), "F16", seed=0)
w = load_weights(path, "cuda:0")
for ctx in (1, 32, 64, 512, 1900):
    row = []
    for onepass, cluster in ((0, 8), (1, 8), (1, 16)):
        eng = LlamaEngine(w, n_ctx=2048, max_seqs=1, page_size=64, max_batch_tokens=2048)
        eng.attn_onepass = bool(onepass)
        eng.attn_onepass_min = 0
        eng.attn_cluster = cluster
        sid = eng.alloc_seq()
        g = torch.Generator().manual_seed(0)
        eng.prefill(sid, torch.randint(3, eng.cfg.vocab_size, (ctx,), generator=g, dtype=torch.int32))
        for _ in range(6):
            eng.decode([sid])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(32):
            eng.decode([sid])
        e1.record()
        torch.cuda.synchronize()
        row.append(e0.elapsed_time(e1) / 32 * 1e3)
        del eng
    print(f"ctx {ctx:5d}: old {row[0]:7.1f}  onepass-8 {row[1]:7.1f}  onepass-16 {row[2]:7.1f} us/step (6 layers)")
