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


def split(src, prefix, max_tensors):
    f = open_gguf(src)
    names = list(f.tensors)
    groups = [names[i:i + max_tensors] for i in range(0, len(names), max_tensors)]
    out = []
    for i, grp in enumerate(groups, 1):
        path = f"{prefix}-{i:05d}-of-{len(groups):05d}.gguf"
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
        out.append(path)
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
    ap.add_argument("--split", action="store_true")
    ap.add_argument("--merge", action="store_true")
    ap.add_argument("--split-max-tensors", type=int, default=128)
    ap.add_argument("src")
    ap.add_argument("dst")
    a = ap.parse_args(argv)
    if a.merge:
        print(merge(a.src, a.dst))
    else:
        for p in split(a.src, a.dst, a.split_max_tensors):
            print(p)


if __name__ == "__main__":
    main()
