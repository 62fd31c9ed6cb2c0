#!/usr/bin/env python
"""dlp-bench: prompt-processing / text-generation throughput, llama-bench style.

Parity: `llama-bench` (`llama.cpp/tools/llama-bench/llama-bench.cpp:1962-2010`): `-p` prompt sizes, `-n` generation sizes,
`-r` repetitions, random token ids (valid with random-init weights), outputs md / json / jsonl / csv / sql with avg_ts +- stddev.
Device-timed with CUDA events; under torchrun it runs the pp x tp grid and reports the max over ranks.
"""
import json
import os
import statistics
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from distributed_llm_pipeline_b200.utils.args import COMMON, Opt, apply_loader_flags, build_parser  # noqa: E402

BENCH = [
    Opt(("-p", "--n-prompt"), "comma separated prompt sizes", "512", str),
    Opt(("-n", "--n-gen"), "comma separated generation sizes", "128", str),
    Opt(("-r", "--repetitions"), "repetitions per test", 3, int),
    Opt(("-o", "--output"), "md | json | jsonl | csv | sql", "md", str, choices=("md", "json", "jsonl", "csv", "sql")),
    Opt(("-d", "--n-depth"), "comma separated context depths: tokens already in the KV cache when a test starts (llama-bench -d)", "0", str),
    Opt(("--batch",), "comma separated decode batch sizes (sequences decoded together)", "1", str),
    Opt(("--microbatches",), "pipeline micro-batches per decode round (0 = one per stage when the batch allows)", 0, int),
]


def main(argv=None):
    a = build_parser("dlp-bench", [COMMON, BENCH]).parse_args(argv)
    apply_loader_flags(a)
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.parallel.grid import init_grid, stage_layer_ranges
    from distributed_llm_pipeline_b200.utils.trace import max_over_ranks
    world = int(os.environ.get("WORLD_SIZE", "1"))
    pp = a.pp or (world if a.split_mode == "layer" and not a.tp else 1)
    grid = init_grid(pp=pp, tp=a.tp, device=a.device)
    comm = None
    if grid.tp > 1:
        if a.tp_impl == "fused":
            from distributed_llm_pipeline_b200.parallel.fused import FusedComm
            comm = FusedComm(grid)
        else:
            from distributed_llm_pipeline_b200.parallel.comm import NcclComm
            comm = NcclComm(grid.tp_group, grid.tp, grid.tp_rank)
    from distributed_llm_pipeline_b200.gguf import open_gguf
    from distributed_llm_pipeline_b200.models.config import config_from_gguf
    f = open_gguf(a.model)
    cfg = config_from_gguf(f.metadata, f.tensors)
    split = [float(x) for x in a.tensor_split.split(",")] if a.tensor_split else None
    ranges = stage_layer_ranges(cfg.n_layer, grid.pp, split)
    w = load_weights(f, grid.device, layer_range=ranges[grid.stage], tp_rank=grid.tp_rank, tp_size=grid.tp,
                     first_stage=grid.first_stage, last_stage=grid.last_stage)
    pps = [int(x) for x in a.n_prompt.split(",") if int(x) > 0]
    gens = [int(x) for x in a.n_gen.split(",") if int(x) > 0]
    depths = [max(0, int(x)) for x in str(a.n_depth).split(",")]
    n_ctx = max(depths) + max(pps + [0]) + max(gens + [0]) + 8
    batches = [max(1, int(x)) for x in str(a.batch).split(",")]
    eng = LlamaEngine(w, n_ctx=max(n_ctx, a.ctx_size), max_seqs=max(batches), page_size=64, max_batch_tokens=max(a.ubatch_size, max(batches), 16),
                      comm=comm, first_stage=grid.first_stage, last_stage=grid.last_stage)
    stage = None
    if grid.pp > 1:
        from distributed_llm_pipeline_b200.parallel.pp import PipelineStage
        stage = PipelineStage(eng, grid)
    cuda = grid.device.type == "cuda"
    g = torch.Generator().manual_seed(0)
    rows = []

    def timed(fn):
        if cuda:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            return max_over_ranks(e0.elapsed_time(e1), grid.device)
        import time
        t = time.perf_counter()
        fn()
        return max_over_ranks((time.perf_counter() - t) * 1e3, grid.device)

    def fresh(n):
        for s in list(eng.seqs):
            eng.free_seq(s)
        return [eng.alloc_seq() for _ in range(n)]

    def fill(sid, depth):      # untimed: `depth` tokens of context in front of the measured work (llama-bench.cpp:318,357 `n_depth`)
        if depth > 0:
            ctx = torch.randint(3, cfg.vocab_size, (depth,), generator=g, dtype=torch.int32)
            (stage.prefill(sid, ctx, chunk=a.ubatch_size) if stage else eng.prefill(sid, ctx, chunk=a.ubatch_size))

    for depth in depths:
      tag = f" @ d{depth}" if depth else ""
      for n_p in pps:
        ts = []
        for rep in range(a.repetitions + 1):
            sid = fresh(1)[0]
            fill(sid, depth)
            toks = torch.randint(3, cfg.vocab_size, (n_p,), generator=g, dtype=torch.int32)
            ms = timed(lambda: (stage.prefill(sid, toks, chunk=a.ubatch_size) if stage else eng.prefill(sid, toks, chunk=a.ubatch_size)))
            if rep:
                ts.append(n_p / ms * 1e3)
        rows.append({"test": f"pp{n_p}{tag}", "n_prompt": n_p, "n_gen": 0, "n_depth": depth, "avg_ts": statistics.mean(ts), "stddev_ts": statistics.pstdev(ts),
                     "samples_ts": ts, "n_tok": n_p})
    for depth in depths:
     tag = f" @ d{depth}" if depth else ""
     for n_g in gens:
      for B in batches:
        # pipeline stages overlap only across micro-batches: a batch is cut into one sequence group per stage (north star:
        # "send/recv overlapped with the next micro-batch's compute"); a single sequence runs the stages serially
        n_mb = min(B, a.microbatches or grid.pp) if stage else 1
        ts = []
        for rep in range(a.repetitions + 1):
            sids = fresh(B)
            for s in sids:
                fill(s, depth - 1)
                (stage.prefill(s, [1]) if stage else eng.prefill(s, [1]))
            first = torch.randint(3, cfg.vocab_size, (B,), generator=g, dtype=torch.int32)
            cut = [B * i // n_mb for i in range(n_mb + 1)]
            groups = [sids[cut[i]:cut[i + 1]] for i in range(n_mb)]

            def run():
                if stage:
                    stage.decode_round(groups, [first[cut[i]:cut[i + 1]] for i in range(n_mb)], feedback=n_g > 1)
                    for i in range(n_g - 1):
                        stage.decode_round(groups, None, feedback=i < n_g - 2)
                else:
                    eng.decode(sids, first)
                    for _ in range(n_g - 1):
                        eng.decode(sids)
            ms = timed(run)
            if rep:
                ts.append(n_g * B / ms * 1e3)
        rows.append({"test": f"tg{n_g}" + (f"@b{B}" if B > 1 else "") + tag, "n_prompt": 0, "n_gen": n_g, "n_depth": depth, "batch": B, "microbatches": n_mb,
                     "avg_ts": statistics.mean(ts), "stddev_ts": statistics.pstdev(ts), "samples_ts": ts, "n_tok": n_g * B})
    if grid.rank == 0:
        # every row carries llama-bench's own columns under its names (llama-bench.cpp `test::get_fields`), so that scripts written against
        # `llama-bench -o json|jsonl|csv|sql` read these files too, plus this tool's extras (pp, tp, tp_impl, batch, microbatches, test)
        import datetime
        import platform
        from distributed_llm_pipeline_b200.gguf.constants import FTYPE
        ftype = {v: k for k, v in FTYPE.items()}.get(int(f.metadata.get("general.file_type", 1)), "?")
        size = sum(int(t.nbytes) for t in f.tensors.values()) if all(hasattr(t, "nbytes") for t in f.tensors.values()) else os.path.getsize(a.model)
        gpu = torch.cuda.get_device_name(grid.device) if cuda else ""
        cpu = next((ln.split(":", 1)[1].strip() for ln in open("/proc/cpuinfo") if ln.startswith("model name")), platform.processor()) \
            if os.path.exists("/proc/cpuinfo") else platform.processor()
        ref_meta = {"build_commit": "dlp_b200", "build_number": 0, "cpu_info": cpu, "gpu_info": ", ".join([gpu] * grid.world) if gpu else "",
                    "backends": "CUDA" if cuda else "CPU", "model_filename": a.model, "model_type": f"{cfg.arch} {cfg.n_params() / 1e9:.2f}B {ftype}",
                    "model_size": size, "model_n_params": cfg.n_params(), "n_batch": a.batch_size if hasattr(a, "batch_size") else 2048, "n_ubatch": a.ubatch_size,
                    "n_threads": torch.get_num_threads(), "cpu_mask": "0x0", "cpu_strict": False, "poll": 50, "type_k": "f16", "type_v": "f16",
                    "n_gpu_layers": cfg.n_layer + 1 if cuda else 0, "n_cpu_moe": 0, "split_mode": "row" if grid.tp > 1 else "layer", "main_gpu": 0,
                    "no_kv_offload": False, "flash_attn": True, "devices": "auto", "tensor_split": a.tensor_split or "0.00", "tensor_buft_overrides": "none",
                    "use_mmap": not getattr(a, "no_mmap", False), "use_direct_io": bool(getattr(a, "direct_io", False)), "embeddings": False,
                    "no_op_offload": 0, "no_host": False}
        meta = {**ref_meta, "model": os.path.basename(a.model), "params": cfg.n_params(), "size_bytes": w.nbytes(), "pp": grid.pp, "tp": grid.tp,
                "tp_impl": a.tp_impl if grid.tp > 1 else None, "backend": "dlp_b200"}
        now = datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")
        for r in rows:
            ns = [r["n_tok"] / t * 1e9 for t in r["samples_ts"]]
            r.update(test_time=now, avg_ns=int(statistics.mean(ns)), stddev_ns=int(statistics.pstdev(ns)), samples_ns=[int(x) for x in ns])
            r.pop("n_tok")
        ref_cols = list(ref_meta) + ["n_prompt", "n_gen", "n_depth", "test_time", "avg_ns", "stddev_ns", "avg_ts", "stddev_ts"]
        if a.output == "json":
            print(json.dumps([dict(meta, **r) for r in rows], indent=1))
        elif a.output == "jsonl":
            for r in rows:
                print(json.dumps(dict(meta, **r)))
        elif a.output == "sql":      # llama-bench -o sql: the table definition plus one INSERT per test, its columns first
            cols = ref_cols + ["pp", "tp", "tp_impl", "test"]
            kinds = {"build_number": "INTEGER", "model_size": "INTEGER", "model_n_params": "INTEGER", "n_batch": "INTEGER", "n_ubatch": "INTEGER", "n_threads": "INTEGER",
                     "poll": "INTEGER", "n_gpu_layers": "INTEGER", "n_cpu_moe": "INTEGER", "main_gpu": "INTEGER", "no_op_offload": "INTEGER", "n_prompt": "INTEGER",
                     "n_gen": "INTEGER", "n_depth": "INTEGER", "avg_ns": "INTEGER", "stddev_ns": "INTEGER", "avg_ts": "REAL", "stddev_ts": "REAL", "pp": "INTEGER",
                     "tp": "INTEGER"}
            print("CREATE TABLE IF NOT EXISTS test (\n" + ",\n".join(f"  {c} {kinds.get(c, 'TEXT')}" for c in cols) + "\n);\n")
            for r in rows:
                d = dict(meta, **r)
                vals = ", ".join("NULL" if d[k] is None else f"'{int(d[k]) if isinstance(d[k], bool) else d[k]}'" for k in cols)
                print(f"INSERT INTO test ({', '.join(cols)}) VALUES ({vals});")
        elif a.output == "csv":
            cols = ref_cols + ["pp", "tp", "test"]
            print(",".join(cols))
            for r in rows:
                d = dict(meta, **r)
                print(",".join('"' + (str(int(d[k])) if isinstance(d[k], bool) else f"{d[k]:.6f}" if isinstance(d[k], float) else str(d[k])) + '"' for k in cols))
        else:       # the markdown table of llama-bench (model | size | params | backend | ngl | test | t/s) with the grid in a column of its own
            print(f"| {'model':<30} | {'size':>10} | {'params':>10} | {'backend':<10} | {'ngl':>3} | {'pp x tp':>7} | {'test':>15} | {'t/s':>20} |")
            print(f"| {'-' * 30} | {'-' * 9}: | {'-' * 9}: | {'-' * 10} | {'-' * 2}: | {'-' * 6}: | {'-' * 14}: | {'-' * 19}: |")
            sz = f"{size / 2**30:.2f} GiB" if size >= 2**30 else f"{size / 2**20:.2f} MiB"
            pr = f"{cfg.n_params() / 1e9:.2f} B" if cfg.n_params() >= 1e9 else f"{cfg.n_params() / 1e6:.2f} M"
            for r in rows:
                ts = f"{r['avg_ts']:.2f} ± {r['stddev_ts']:.2f}"
                print(f"| {meta['model_type']:<30} | {sz:>10} | {pr:>10} | {meta['backends']:<10} | {meta['n_gpu_layers']:>3} | {f'{grid.pp} x {grid.tp}':>7} | "
                      f"{r['test']:>15} | {ts:>20} |")
            print("\nbuild: dlp_b200")
    if grid.world > 1:
        from distributed_llm_pipeline_b200.parallel.grid import shutdown
        shutdown(0)


if __name__ == "__main__":
    main()
