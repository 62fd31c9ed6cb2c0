#!/usr/bin/env python
"""dlp-gguf-split: split a GGUF file into `<prefix>-0000i-of-0000N.gguf` shards or merge shards back.

Parity: `llama-gguf-split` (llama.cpp/tools/gguf-split); shard naming as `llama.cpp/src/llama.cpp:1094-1100`.  Shard 1 carries all
metadata (+ `split.no`, `split.count`, `split.tensors.count`), later shards only the split keys.  `open_gguf()` opens a whole set
from its first shard.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGUFWriter, ValueType, open_gguf  # noqa: E402


def _copy_meta(f, w, skip_split=True):
    for key, val in f.metadata.items():
        if skip_split and key.startswith("split."):
            continue
        if isinstance(val, (list, np.ndarray)):
            w.add(key, val)
        else:
            w.add(key, val, f.files[0].metadata_types[key])


def plan(f, max_tensors=0, max_bytes=0, no_tensor_first_split=False):
    """Tensor names per shard: at most `max_tensors` tensors, or at most `max_bytes` of tensor data (a tensor larger than that gets a shard of
    its own); with `no_tensor_first_split` the first shard carries the metadata only (gguf-split.cpp split_strategy)."""
    groups, cur, size = [], [], 0
    for n, t in f.tensors.items():
        full = (max_bytes and cur and size + t.nbytes > max_bytes) or (not max_bytes and len(cur) >= max(1, max_tensors))
        if full:
            groups.append(cur)
            cur, size = [], 0
        cur.append(n)
        size += t.nbytes
    if cur:
        groups.append(cur)
    return ([[]] if no_tensor_first_split else []) + groups


def split(src, prefix, max_tensors=128, max_bytes=0, no_tensor_first_split=False, dry_run=False):
    f = open_gguf(src)
    names = list(f.tensors)
    groups = plan(f, max_tensors, max_bytes, no_tensor_first_split)
    out = []
    for i, grp in enumerate(groups, 1):
        path = f"{prefix}-{i:05d}-of-{len(groups):05d}.gguf"
        out.append(path)
        if dry_run:
            print(f"split {i:05d}: n_tensors = {len(grp)}, total_size = {sum(f.tensors[n].nbytes for n in grp) // 1000000}M")
            continue
        w = GGUFWriter(path)
        if i == 1:
            _copy_meta(f, w)
        else:
            w.add("general.architecture", f.metadata.get("general.architecture", "llama"))
        w.add("split.no", i - 1, ValueType.UINT16)
        w.add("split.count", len(groups), ValueType.UINT16)
        w.add("split.tensors.count", len(names), ValueType.INT32)
        for n in grp:
            t = f.tensors[n]
            w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
        w.write()
    return out


def merge(first_shard, dst):
    f = open_gguf(first_shard)
    w = GGUFWriter(dst)
    _copy_meta(f, w)
    for n, t in f.tensors.items():
        w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
    w.write()
    return dst


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--split", action="store_true", help="split (the default)")
    ap.add_argument("--merge", action="store_true")
    ap.add_argument("--split-max-tensors", type=int, default=128, help="tensors per shard")
    ap.add_argument("--split-max-size", default=None, metavar="N(M|G)", help="tensor bytes per shard, e.g. 500M or 2G (overrides --split-max-tensors)")
    ap.add_argument("--no-tensor-first-split", action="store_true", help="the first shard holds the metadata only")
    ap.add_argument("--dry-run", action="store_true", help="print the plan, write nothing")
    ap.add_argument("src")
    ap.add_argument("dst")
    a = ap.parse_args(argv)
    if a.merge:
        print(merge(a.src, a.dst))
        return
    max_bytes = 0
    if a.split_max_size:
        unit = {"M": 1000 ** 2, "G": 1000 ** 3}.get(a.split_max_size[-1].upper())
        if unit is None or not a.split_max_size[:-1].isdigit():
            ap.error("--split-max-size takes a number followed by M or G")
        max_bytes = int(a.split_max_size[:-1]) * unit
    for p in split(a.src, a.dst, a.split_max_tensors, max_bytes, a.no_tensor_first_split, a.dry_run):
        if not a.dry_run:
            print(p)


if __name__ == "__main__":
    main()
