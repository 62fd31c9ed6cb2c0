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
        rows.append({"test": f"pp{n_p}{tag}", "n_prompt": n_p, "n_gen": 0, "n_depth": depth, "avg_ts": statistics.mean(ts), "stddev_ts": statistics.pstdev(ts)})
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
                     "avg_ts": statistics.mean(ts), "stddev_ts": statistics.pstdev(ts)})
    if grid.rank == 0:
        meta = {"model": os.path.basename(a.model), "params": cfg.n_params(), "size_bytes": w.nbytes(), "pp": grid.pp, "tp": grid.tp,
                "tp_impl": a.tp_impl if grid.tp > 1 else None, "backend": "dlp_b200"}
        if a.output == "json":
            print(json.dumps([dict(meta, **r) for r in rows], indent=1))
        elif a.output == "jsonl":
            for r in rows:
                print(json.dumps(dict(meta, **r)))
        elif a.output == "sql":      # same idea as llama-bench -o sql: a table definition plus one INSERT per test
            print("CREATE TABLE IF NOT EXISTS test (model TEXT, params INTEGER, size_bytes INTEGER, pp INTEGER, tp INTEGER, tp_impl TEXT, "
                  "backend TEXT, test TEXT, n_prompt INTEGER, n_gen INTEGER, avg_ts REAL, stddev_ts REAL);")
            for r in rows:
                d = dict(meta, **r)
                vals = ", ".join("NULL" if d[k] is None else (f"'{d[k]}'" if isinstance(d[k], str) else f"{d[k]}")
                                 for k in ("model", "params", "size_bytes", "pp", "tp", "tp_impl", "backend", "test", "n_prompt", "n_gen", "avg_ts", "stddev_ts"))
                print(f"INSERT INTO test VALUES ({vals});")
        elif a.output == "csv":
            print("model,pp,tp,test,avg_ts,stddev_ts")
            for r in rows:
                print(f"{meta['model']},{grid.pp},{grid.tp},{r['test']},{r['avg_ts']:.2f},{r['stddev_ts']:.2f}")
        else:
            print("| model | size | pp x tp | test | t/s |\n|---|---:|---|---|---:|")
            for r in rows:
                print(f"| {meta['model']} | {w.nbytes() / 2**30:.2f} GiB/rank | {grid.pp} x {grid.tp} | {r['test']} | {r['avg_ts']:.2f} ± {r['stddev_ts']:.2f} |")
    if grid.world > 1:
        from distributed_llm_pipeline_b200.parallel.grid import shutdown
        shutdown(0)


if __name__ == "__main__":
    main()
