#!/usr/bin/env python
"""Headline benchmark: single-stream decode tokens/s of Llama-3-8B (random-init GGUF), tensor-parallel over N GPUs.

Contract (see task statement): `python bench.py --gpus N --steps K --warmup W`; for N > 1 the driver launches it
under torchrun with one rank per GPU.  One *step* = one decoded token (the reference instrument is llama-bench
`tg`, which also times single-stream generation: llama.cpp/tools/llama-bench/llama-bench.cpp:1991-2010).
Rank 0 prints ONE JSON line.  `--impl reference` runs the unmodified reference (`baseline/_ref/bin/llama-bench`,
built offline from /root/reference) on the same GGUF file and prints the same line with "impl": "reference".

Timing: W untimed warm-up tokens, then exactly K tokens between barrier + synchronize, CUDA events on the launching
stream, max over ranks.  The weights (16 GB fp16) are far larger than the 126 MB L2, so every step streams them from
HBM ("inputs larger than L2").  SM clocks / throttle reasons are sampled with nvidia-smi during the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

MODEL_DIR = os.environ.get("DLP_MODEL_DIR", "/tmp/dlp_models")
BASELINE_PUBLISHED = None  # the reference publishes no number for this metric/config


def model_path(model: str, ftype: str) -> str:
    return os.path.join(MODEL_DIR, f"{model}-{ftype}.gguf")


def ensure_model(model: str, ftype: str, rank: int, barrier) -> str:
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.config import preset
    path = model_path(model, ftype)
    if rank == 0 and not os.path.exists(path):
        tmp = path + ".tmp"
        write_synthetic_gguf(tmp, preset(model), ftype, seed=0)
        os.replace(tmp, path)
    barrier()
    return path


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (profiling recipe's clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.rows: list[list[str]] = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.ops import api as ops
    from distributed_llm_pipeline_b200.parallel.grid import init_grid

    grid = init_grid(pp=args.pp)
    world, rank = grid.world, grid.rank
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    dev = grid.device

    def barrier():
        if world > 1:
            dist.barrier()

    path = ensure_model(args.model, args.ftype, rank, barrier)
    comm = None
    emu = args.emulate_tp if world == 1 else 0
    if emu > 1:
        # per-rank view of an emu-way tensor-parallel step on ONE GPU: rank 0's shard, kernels and launch count exactly as under
        # real TP, its stores landing in its own buffers (no NVLink hop, nobody to wait for).  A profiling aid, not a headline.
        from distributed_llm_pipeline_b200.parallel.fused import FusedComm
        comm = FusedComm.make_emulated(emu)
    if grid.tp > 1:
        fused_ok = False
        if args.tp_impl == "fused":
            try:
                from distributed_llm_pipeline_b200.parallel.fused import FusedComm
                comm = FusedComm(grid)
                fused_ok = True
            except ImportError:
                args.tp_impl = "nccl"
        if not fused_ok:
            from distributed_llm_pipeline_b200.parallel.comm import NcclComm
            comm = NcclComm(grid.tp_group, grid.tp, grid.tp_rank)
    if grid.pp > 1:
        raise SystemExit("bench.py times the TP family (config 3); the PP path is exercised by tools/dlp_bench.py")
    t_load = time.time()
    weights = load_weights(path, dev, tp_rank=grid.tp_rank, tp_size=emu if emu > 1 else grid.tp)
    eng = LlamaEngine(weights, n_ctx=args.n_ctx, max_seqs=1, page_size=64, max_batch_tokens=max(args.prefill, 16), comm=comm)
    torch.cuda.synchronize()
    load_s = time.time() - t_load
    sid = eng.alloc_seq()

    # ---- prefill (pp<args.prefill>) : timed separately, reported as a secondary number ----
    gen = torch.Generator().manual_seed(1234)
    prompt = torch.randint(3, eng.cfg.vocab_size, (args.prefill,), generator=gen, dtype=torch.int32)
    prefill_ms = []
    for it in range(3):
        if it:
            eng.free_seq(sid)
            sid = eng.alloc_seq()
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        logits = eng.prefill(sid, prompt)
        e1.record()
        torch.cuda.synchronize()
        prefill_ms.append(e0.elapsed_time(e1))
    t_pre = torch.tensor([min(prefill_ms[1:])], device=dev)
    if world > 1:
        dist.all_reduce(t_pre, op=dist.ReduceOp.MAX)
    prefill_tok_s = args.prefill / (t_pre.item() / 1e3)

    # ---- tensor-parallel sanity: the fused path must produce the tokens of the NCCL baseline path ----
    tp_check = None
    if grid.tp > 1 and args.tp_impl == "fused" and not args.no_tp_check:
        from distributed_llm_pipeline_b200.parallel.comm import NcclComm
        ref_eng = LlamaEngine(weights, n_ctx=256, max_seqs=1, page_size=64, max_batch_tokens=16,
                              comm=NcclComm(grid.tp_group, grid.tp, grid.tp_rank))
        probe = [1, 17, 4242, 99, 1234, 7, 31000, 5]
        want = ref_eng.generate(probe, 8)
        eng.free_seq(sid)
        got = eng.generate(probe, 8)
        sid = eng.alloc_seq()
        tp_check = {"fused_equals_nccl": got == want, "tokens": got[:8]}
        if got != want:
            tp_check["nccl_tokens"] = want[:8]
        del ref_eng

    # ---- decode: device-resident loop (CUDA graph replay, token fed back on the device) ----
    eng.free_seq(sid)
    sid = eng.alloc_seq()
    eng.prefill(sid, [1])  # llama-bench tg starts from an empty context
    eng.next_tok[:1].fill_(5)
    for _ in range(args.warmup):
        eng.decode([sid])
    clocks = ClockSampler(dev.index or 0)
    barrier()
    torch.cuda.synchronize()
    n0 = ops.launch_count
    r0 = getattr(eng, "launches_replayed", 0)
    if rank == 0:
        clocks.start()
    # The timed region is a block of exactly K steps between CUDA events.  Under tensor parallelism the ranks spin on each
    # other inside the kernels, so one host hiccup on any rank lands in everybody's block; the block is therefore repeated
    # (ranks re-synchronised immediately before each e0) and the MEDIAN block is reported, every block listed in `blocks_ms`.
    n_blocks = max(1, args.blocks)
    blocks_ms = []
    launches = 0
    for b in range(n_blocks):
        if eng.seq_len(sid) + args.steps + 2 > args.n_ctx:      # stay inside the context window
            eng.free_seq(sid)
            sid = eng.alloc_seq()
            eng.prefill(sid, [1])
            eng.next_tok[:1].fill_(5)
            eng.decode([sid])
        barrier()
        torch.cuda.synchronize()
        n0 = ops.launch_count
        r0 = getattr(eng, "launches_replayed", 0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            eng.decode([sid])
        e1.record()
        barrier()
        torch.cuda.synchronize()
        t_blk = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t_blk, op=dist.ReduceOp.MAX)
        blocks_ms.append(t_blk.item())
        launches = (getattr(eng, "launches_replayed", 0) - r0) + (ops.launch_count - n0)   # inside graph replays + eager
    clk = clocks.stop() if rank == 0 else {}
    if hasattr(comm, "check_error"):
        comm.check_error()      # a peer that never signalled makes the kernels fall through: that run has no number
    t_dec = torch.tensor([statistics.median(blocks_ms)], device=dev)
    ms_per_step = t_dec.item() / args.steps

    # ---- end to end through the public API: pinned H2D token in, D2H token out, every step ----
    tok_host = torch.zeros(1, dtype=torch.int32).pin_memory()
    out_host = torch.zeros(1, dtype=torch.int32).pin_memory()
    tok_host[0] = 7
    if eng.seq_len(sid) + args.steps + 8 > args.n_ctx:
        eng.free_seq(sid)
        sid = eng.alloc_seq()
        eng.prefill(sid, [1])
    for _ in range(3):
        out_host.copy_(eng.decode([sid], tok_host), non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        nt = eng.decode([sid], tok_host)          # H2D copy of this step's input from pinned memory
        out_host.copy_(nt, non_blocking=True)     # D2H read of this step's result
        torch.cuda.current_stream().synchronize()
        tok_host[0] = out_host[0]
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    t_e2e = torch.tensor([max(e0.elapsed_time(e1) / 1e3, wall)], device=dev)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_tok_s = args.steps / t_e2e.item()

    value = 1e3 / ms_per_step
    res = {
        "metric": "decode_tok_s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": (value / BASELINE_PUBLISHED) if BASELINE_PUBLISHED else None, "dtype": "fp16" if args.ftype == "F16" else args.ftype,
        "data": "synthetic (random-init GGUF weights, random token ids)",
        "config": {"model": f"{args.model} {args.ftype} GGUF", "global_batch": 1, "seq_len": args.steps + args.warmup + 1,
                   "parallelism": (f"tp{emu}-emulated-rank0 (1 GPU, no NVLink hop; NOT a headline number)" if emu > 1 else
                                   f"tp{grid.tp}" + (f"-{args.tp_impl}" if grid.tp > 1 else "")), "instrument": "tg (single-stream decode)",
                   "l2": "inputs larger than L2 (weights >> 126 MB stream from HBM every step)"},
        "clocks": clk,
        "e2e": {"value": round(e2e_tok_s, 2), "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": 4},
        "gpu_launches": int(launches),
        "prefill_tok_s": round(prefill_tok_s, 1), "prefill_tokens": args.prefill, "ttft_ms_p50": round(t_pre.item(), 3),
        "weights_gb": round(weights.nbytes() / 1e9, 3), "load_s": round(load_s, 1),
        "blocks_ms": [round(x, 3) for x in blocks_ms], "timed": f"median of {n_blocks} device-timed blocks of {args.steps} steps",
        "tp_check": tp_check, "symm": getattr(comm, "symm_backend", None), "nvls_multicast": bool(getattr(comm, "mc_base", 0)),
        "hbm_frac_of_measured": None,
    }
    try:
        peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
        res["hbm_frac_of_measured"] = round(weights.nbytes() / (ms_per_step / 1e3) / 1e9 / peaks["hbm_gbs"], 3)
    except Exception:
        pass
    return res if rank == 0 else {}


# ------------------------------------------------------------------------------------------------
def run_reference(args) -> dict:
    """The unmodified reference: llama-bench from baseline/_ref on the same GGUF, its own public CLI."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        if world > 1:  # keep the torchrun group alive until rank 0 is done
            import torch.distributed as dist
            dist.init_process_group("gloo")
            dist.barrier()
        return {}
    out: dict = {"impl": "reference"}
    try:
        sys.path.insert(0, os.path.join(REPO, "baseline"))
        import reference as refarm
        why = refarm.ensure(allow_build=True, build_timeout_s=float(os.environ.get("DLP_REFERENCE_BUILD_TIMEOUT_S", "2400")))
        if why is not None:
            return {"impl": "reference", "unavailable": why + " (pip install impossible: the reference has no setup.py/pyproject)"}
        exe = refarm.exe("llama-bench")
        path = ensure_model(args.model, args.ftype, 0, lambda: None)
        env = refarm.env()
        env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(args.gpus))
        base = [exe, "-m", path, "-ngl", "99", "-fa", "1", "-p", str(args.prefill), "-n", str(args.steps), "-r", "3", "-o", "json"]
        modes = [None] if args.gpus == 1 else ["row", "layer"]   # N > 1: the reference's two multi-GPU modes; the better one is the bar
        clocks = ClockSampler(0)
        clocks.start()
        t0 = time.time()
        runs = {}
        for sm in modes:
            cmd = base + (["-sm", sm] if sm else [])
            p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=3000)
            if p.returncode != 0:
                runs[sm or "none"] = {"error": f"llama-bench exited {p.returncode}: {p.stderr.strip()[-300:]}"}
                continue
            rows = json.loads(p.stdout[p.stdout.index("["):])
            tg = next(r for r in rows if r["n_gen"] > 0 and r["n_prompt"] == 0)
            pp = next((r for r in rows if r["n_prompt"] > 0 and r["n_gen"] == 0), None)
            runs[sm or "none"] = {"decode_tok_s": round(tg["avg_ts"], 2), "prefill_tok_s": round(pp["avg_ts"], 1) if pp else None,
                                  "cmd": " ".join(cmd[1:])}
        clk = clocks.stop()
        good = {k: v for k, v in runs.items() if "decode_tok_s" in v}
        if not good:
            return {"impl": "reference", "unavailable": "; ".join(v["error"] for v in runs.values())[:300]}
        best = max(good, key=lambda k: good[k]["decode_tok_s"])
        value = good[best]["decode_tok_s"]
        out.update({
            "metric": "decode_tok_s", "value": round(value, 2), "unit": "tok/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": 1, "ms_per_step": round(1e3 / value, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "fp16" if args.ftype == "F16" else args.ftype, "data": "synthetic (same random-init GGUF, random token ids)",
            "config": {"model": f"{args.model} {args.ftype} GGUF", "global_batch": 1, "seq_len": args.steps,
                       "parallelism": "1gpu" if args.gpus == 1 else f"-sm {best} x{args.gpus} (better of -sm row / -sm layer)",
                       "instrument": "llama-bench tg (host-timed by the reference)", "cmd": good[best]["cmd"]},
            "clocks": clk, "prefill_tok_s": good[best]["prefill_tok_s"], "prefill_tokens": args.prefill,
            "e2e": {"value": round(value, 2), "unit": "tok/s", "note": "llama-bench times end to end on the host"},
            "split_modes": runs, "wall_s": round(time.time() - t0, 1),
        })
        return out
    except Exception as e:  # never crash the driver's reference arm
        return {"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("gloo")
            dist.barrier()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--ftype", default="F16")
    ap.add_argument("--prefill", type=int, default=512)
    ap.add_argument("--n-ctx", type=int, default=2048)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--blocks", type=int, default=5, help="timed K-step blocks; the median block is reported")
    ap.add_argument("--emulate-tp", type=int, default=0, help="1 GPU only: time rank 0 of an N-way tensor-parallel step (profiling aid)")
    ap.add_argument("--no-tp-check", action="store_true", help="skip the fused-vs-NCCL greedy token comparison")
    ap.add_argument("--tp-impl", default=os.environ.get("DLP_TP_IMPL", "fused"), choices=["fused", "nccl"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if res:
        print(json.dumps(res), flush=True)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        # no NCCL / CUDA teardown: destroy_process_group() and plain exit were seen to block forever on this image
        from distributed_llm_pipeline_b200.parallel.grid import shutdown
        shutdown(0)


if __name__ == "__main__":
    main()
