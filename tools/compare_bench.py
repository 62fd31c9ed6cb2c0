#!/usr/bin/env python
"""Compare two sets of benchmark results (dlp_bench / llama-bench `-o json|jsonl` files, or bench.py JSON lines).

Parity: `llama.cpp/scripts/compare-llama-bench.py` (baseline vs. compare commits); here any two result files: rows are matched
on (model, test) and the speed-up is printed as a markdown table.  Works across the two tools, so the same script compares
this engine against the reference's own `llama-bench` output.
"""
import argparse
import json
import os
import sys


def load(path):
    txt = open(path).read().strip()
    rows = []
    try:
        data = json.loads(txt)
        data = data if isinstance(data, list) else [data]
    except json.JSONDecodeError:
        data = [json.loads(l) for l in txt.splitlines() if l.strip().startswith("{")]
    for d in data:
        if "avg_ts" in d:                                      # dlp_bench / llama-bench
            test = d.get("test") or (f"pp{d['n_prompt']}" if d.get("n_prompt") else f"tg{d.get('n_gen')}")
            model = os.path.basename(str(d.get("model") or d.get("model_filename") or "?"))
            rows.append(((model, test), float(d["avg_ts"])))
        elif "metric" in d and "value" in d:                   # bench.py line
            model = str(d.get("config", {}).get("model", "?"))
            rows.append(((model, f"{d['metric']}@{d.get('n_gpus', 1)}gpu"), float(d["value"])))
            if "prefill_tok_s" in d:
                rows.append(((model, f"prefill_tok_s@{d.get('n_gpus', 1)}gpu"), float(d["prefill_tok_s"])))
    return dict(rows)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("baseline")
    ap.add_argument("compare")
    ap.add_argument("--any-model", action="store_true", help="match on the test name only (different file names for the same model)")
    a = ap.parse_args(argv)
    b, c = load(a.baseline), load(a.compare)
    if a.any_model:
        b = {("*", k[1]): v for k, v in b.items()}
        c = {("*", k[1]): v for k, v in c.items()}
    keys = [k for k in b if k in c]
    if not keys:
        sys.exit("no common (model, test) rows")
    print(f"| model | test | t/s {os.path.basename(a.baseline)} | t/s {os.path.basename(a.compare)} | speedup |\n|---|---|---:|---:|---:|")
    for k in keys:
        print(f"| {k[0]} | {k[1]} | {b[k]:.2f} | {c[k]:.2f} | {c[k] / b[k]:.2f} |")


if __name__ == "__main__":
    main()
