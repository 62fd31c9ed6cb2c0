"""Multi-process worker used by test_parallel_*.py (launched under torchrun, gloo on CPU or nccl on GPUs).

Checks that a tensor-/pipeline-parallel engine reproduces the single-process fp32 oracle, teacher-forced."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from distributed_llm_pipeline_b200.models.llama import LlamaEngine
from distributed_llm_pipeline_b200.models.reference import ReferenceModel, nmse
from distributed_llm_pipeline_b200.models.weights import load_weights
from distributed_llm_pipeline_b200.parallel.comm import NcclComm
from distributed_llm_pipeline_b200.parallel.grid import init_grid, stage_layer_ranges


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", required=True)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--impl", default="nccl")
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--tol", type=float, default=1e-3)
    ap.add_argument("--greedy", type=int, default=0, help="also generate this many tokens device-fed and compare with the teacher-forced arg-max")
    a = ap.parse_args()
    grid = init_grid(pp=a.pp, device=a.device, timeout_s=90)
    comm = None
    if grid.tp > 1:
        if a.impl == "fused":
            from distributed_llm_pipeline_b200.parallel.fused import FusedComm
            comm = FusedComm(grid)
        else:
            comm = NcclComm(grid.tp_group, grid.tp, grid.tp_rank)
    ref = ReferenceModel.from_gguf(a.model)
    cfg = ref.cfg
    ranges = stage_layer_ranges(cfg.n_layer, grid.pp)
    w = load_weights(a.model, grid.device, layer_range=ranges[grid.stage], tp_rank=grid.tp_rank, tp_size=grid.tp,
                     first_stage=grid.first_stage, last_stage=grid.last_stage)
    eng = LlamaEngine(w, n_ctx=128, max_seqs=4, page_size=16, max_batch_tokens=32, comm=comm,
                      first_stage=grid.first_stage, last_stage=grid.last_stage)
    if grid.pp > 1:
        from distributed_llm_pipeline_b200.parallel.pp import PipelineStage
        stage = PipelineStage(eng, grid)
    prompts = [[1, 5, 9, 200, 33, 7, 8], [4, 4, 2], [9, 1], [3, 6, 9, 12]][: a.batch]
    nxt = [[11, 12, 13, 14, 15], [21, 22, 23, 24, 25], [31, 32, 33, 34, 35], [41, 42, 43, 44, 45]][: a.batch]
    worst = 0.0
    sids = [eng.alloc_seq() for _ in prompts]
    caches = []
    for sid, p in zip(sids, prompts):
        cache = ref.new_cache(128)
        want = ref.forward(p, cache)[-1]
        ref.advance(cache, len(p))
        caches.append(cache)
        got = stage.prefill(sid, p) if grid.pp > 1 else eng.prefill(sid, p)
        if grid.last_stage:
            worst = max(worst, nmse(want, got))
    for step in range(5):
        toks = [n[step] for n in nxt]
        got = stage.decode(sids, toks, sample=False) if grid.pp > 1 else eng.decode(sids, toks, sample=False)
        for row, (cache, t) in enumerate(zip(caches, toks)):
            want = ref.forward([t], cache)[-1]
            ref.advance(cache, 1)
            if grid.last_stage:
                worst = max(worst, nmse(want, got[row]))
    greedy_ok = True
    if a.greedy and grid.pp == 1:
        # device-fed generation (sampled on the device: sharded arg-max on the fused path) vs arg-max of teacher-forced logits
        p = [1, 5, 9, 200, 33, 7, 8, 426, 17]
        sid = eng.alloc_seq()
        want = [int(eng.prefill(sid, p).argmax())]
        for _ in range(a.greedy - 1):
            want.append(int(eng.decode([sid], [want[-1]], sample=False)[0].argmax()))
        eng.free_seq(sid)
        got = eng.generate(p, a.greedy)
        greedy_ok = got == want
        if not greedy_ok:
            print(f"rank {grid.rank}: greedy mismatch {got} != {want}")
    if comm is not None and getattr(comm, "fused", False):
        comm.check_error()
        if grid.rank == 0:
            print(f"symmetric memory: {getattr(comm, 'symm_backend', None)}, multicast {'on' if getattr(comm, 'mc_base', 0) else 'off'} {getattr(comm, 'symm_note', '')}")
    if not greedy_ok:
        worst = 1.0
    res = torch.tensor([worst], dtype=torch.float64)
    if grid.world > 1:
        if grid.device.type == "cuda":
            res = res.to(grid.device)
        dist.all_reduce(res, op=dist.ReduceOp.MAX)
    if grid.rank == 0:
        print(f"MP_RESULT worst_nmse={float(res):.3e} world={grid.world} pp={grid.pp} tp={grid.tp} impl={a.impl}")
    ok = float(res) < a.tol
    from distributed_llm_pipeline_b200.parallel.grid import shutdown
    shutdown(0 if ok else 1)


if __name__ == "__main__":
    main()
