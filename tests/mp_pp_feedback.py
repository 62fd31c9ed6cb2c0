"""2-stage pipeline, greedy generation with device-to-device token feedback; must equal the single-process engine."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_llm_pipeline_b200.models.llama import LlamaEngine
from distributed_llm_pipeline_b200.models.weights import load_weights
from distributed_llm_pipeline_b200.parallel.grid import init_grid, shutdown, stage_layer_ranges
from distributed_llm_pipeline_b200.parallel.pp import PipelineStage

path = sys.argv[1]
grid = init_grid(pp=2, device="cpu", timeout_s=90)
full = LlamaEngine(load_weights(path, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=16)
want = full.generate([1, 5, 9, 3], 6)
cfg = full.cfg
ranges = stage_layer_ranges(cfg.n_layer, 2)
w = load_weights(path, "cpu", layer_range=ranges[grid.stage], first_stage=grid.first_stage, last_stage=grid.last_stage)
eng = LlamaEngine(w, n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=16, first_stage=grid.first_stage, last_stage=grid.last_stage)
st = PipelineStage(eng, grid)
sid = eng.alloc_seq()
lg = st.prefill(sid, [1, 5, 9, 3])
got = []
if grid.last_stage:
    first = int(lg.argmax())
    got.append(first)
    t = torch.tensor([first], dtype=torch.int32)
else:
    t = torch.zeros(1, dtype=torch.int32)
torch.distributed.broadcast(t, src=grid.world - 1)      # seed token for round 0 comes from the host side
n = 5
for i in range(n):
    out = st.decode_round([[sid]], [t] if i == 0 else None, sample=True, feedback=i < n - 1)[0]
    if grid.last_stage:
        got.append(int(out[0]))
if grid.last_stage:
    assert got == want, (got, want)
    print("FEEDBACK_OK", got)
shutdown(0)
