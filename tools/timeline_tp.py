#!/usr/bin/env python
"""Device-side timeline of one tensor-parallel decode step (CUDA-graph replay, PDL on).

CTA (0,0,0) of the decode kernels stamps %globaltimer at: begin, dependency wait returned, all-reduce values consumed
(mat-vecs that fold an all-reduce into their prologue), main loop done, end (dlp_set_trace).  Run it

  * under torchrun on N GPUs (real fused TP: every rank records, rank 0 prints its own view), or
  * on one GPU with --emulate-tp N (rank 0 of an N-way step with its stores landing in its own buffers: no NVLink hop).

For every kernel of a steady-state layer it prints the phase durations; the summary splits the layer period into
launch/dependency wait, all-reduce wait + fold, prologue + main loop, epilogue (stores to the peers) and the gap to the next kernel.
"""
import argparse
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import MODEL_DIR  # noqa: E402
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402
from distributed_llm_pipeline_b200.ops import lib as _lib  # noqa: E402
from distributed_llm_pipeline_b200.parallel.grid import init_grid, shutdown  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=8)
ap.add_argument("--ftype", default="F16")
ap.add_argument("--emulate-tp", type=int, default=0)
ap.add_argument("--ctx", type=int, default=64, help="keys in the cache when the step is recorded")
ap.add_argument("--out", default="")
a = ap.parse_args()

grid = init_grid()
world, rank = grid.world, grid.rank
path = os.path.join(MODEL_DIR, f"llama-3-8b-{a.layers}L-{a.ftype}.gguf")
if rank == 0 and not os.path.exists(path):
    os.makedirs(MODEL_DIR, exist_ok=True)
    write_synthetic_gguf(path, preset("llama-3-8b", n_layer=a.layers), a.ftype, seed=0)
if world > 1:
    torch.distributed.barrier()
from distributed_llm_pipeline_b200.parallel.fused import FusedComm  # noqa: E402
tp = a.emulate_tp if world == 1 and a.emulate_tp > 1 else world
comm = None
if world > 1:
    comm = FusedComm(grid)
elif tp > 1:
    comm = FusedComm.make_emulated(tp)
w = load_weights(path, grid.device, tp_rank=grid.tp_rank if world > 1 else 0, tp_size=tp)
eng = LlamaEngine(w, n_ctx=2048, max_seqs=1, page_size=64, max_batch_tokens=16, comm=comm)
NAMES = {1: "gemv", 2: "rope", 3: "attn"}
EV = {0: "begin", 1: "waited", 2: "main_done", 3: "end", 4: "ar_done"}
CAP = 8192
buf = torch.zeros(CAP + 2, dtype=torch.int64, device=grid.device)
buf[1] = CAP
_lib.check(_lib.load().dlp_set_trace(buf.data_ptr()), "set_trace")
sid = eng.alloc_seq()
eng.prefill(sid, [1])
eng.next_tok[:1].fill_(5)
for _ in range(max(8, a.ctx)):      # capture + warm replays; the context grows to --ctx keys
    eng.decode([sid])
torch.cuda.synchronize()
if world > 1:
    torch.distributed.barrier()
steps = []
for rep in range(5):
    buf[0] = 0
    eng.decode([sid])
    torch.cuda.synchronize()
    n = int(buf[0])
    rec = buf[2:2 + min(n, CAP)].cpu().tolist()
    steps.append(sorted(((r & 0xFFFFFFFFFFFF), (r >> 48) >> 8, (r >> 48) & 0xFF) for r in rec))
_lib.check(_lib.load().dlp_set_trace(None), "set_trace")
if comm is not None:
    comm.check_error()

# group the records of one step into kernels: a kernel starts at its `begin` record
def kernels_of(ev):
    ks, open_k = [], {}
    for t, kind, e in ev:
        if e == 0:
            k = {"kind": NAMES.get(kind, str(kind)), "begin": t}
            ks.append(k)
            open_k[kind] = k
        elif kind in open_k:
            open_k[kind][EV.get(e, str(e))] = t
    return ks


spans = [(s[-1][0] - s[0][0]) / 1e3 for s in steps]
ks = kernels_of(steps[len(steps) // 2])
per_layer = 5 if eng._onepass_now else 6
summary = {"tp": tp, "world": world, "emulated": world == 1 and tp > 1, "layers": a.layers, "ftype": a.ftype, "step_span_us": spans,
           "symm": getattr(comm, "symm_backend", None), "nvls_multicast": bool(getattr(comm, "mc_base", 0))}
if rank == 0:
    print(f"tp={tp} world={world} ({'emulated rank 0' if summary['emulated'] else 'real'}) layers={a.layers} symmetric memory: {summary['symm']}, multicast {summary['nvls_multicast']}")
    print(f"step span over 5 replays: {', '.join(f'{x:.1f}' for x in spans)} us  ({len(ks)} traced kernels)")
    # steady-state layers: skip the first layer's kernels, print two layers in full
    gemv_idx = [i for i, k in enumerate(ks) if k["kind"] == "gemv"]
    n_g = 4
    lo = gemv_idx[n_g * 2] if len(gemv_idx) > n_g * 4 else 0
    hi = gemv_idx[n_g * 4] if len(gemv_idx) > n_g * 4 else len(ks) - 1
    base = ks[lo]["begin"]
    print(f"{'kernel':6s} {'begin':>8s} {'waited':>8s} {'ar_done':>8s} {'main':>8s} {'end':>8s}   (us, relative)")
    agg = {"launch_wait": 0.0, "ar_wait_fold": 0.0, "prologue_main": 0.0, "epilogue": 0.0}
    for k in ks[lo:hi]:
        f = lambda key: f"{(k[key] - base) / 1e3:8.2f}" if key in k else "       -"
        print(f"{k['kind']:6s} {f('begin')} {f('waited')} {f('ar_done')} {f('main_done')} {f('end')}")
    # per-layer phase sums over the steady-state window (serial chain: every kernel waits for its predecessor)
    n_layers_win = 2
    for k in ks[lo:hi]:
        wt = k.get("waited", k["begin"])
        ar = k.get("ar_done", wt)
        md = k.get("main_done", ar)
        en = k.get("end", md)
        agg["ar_wait_fold"] += (ar - wt) / 1e3
        agg["prologue_main"] += (md - ar) / 1e3
        agg["epilogue"] += (en - md) / 1e3
    period = (ks[hi]["begin"] - base) / 1e3 / n_layers_win
    busy = sum(agg.values()) / n_layers_win
    agg = {k: round(v / n_layers_win, 2) for k, v in agg.items()}
    agg["launch_wait"] = round(period - busy, 2)     # time not inside any traced kernel's post-wait body
    summary.update({"layer_period_us": round(period, 2), "per_layer_us": agg})
    print(f"layer period {period:.2f} us: " + ", ".join(f"{k} {v:.2f}" for k, v in agg.items()))
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(summary, open(a.out, "w"))
if world > 1:
    shutdown(0)
