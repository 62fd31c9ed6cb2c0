#!/usr/bin/env python
"""dlp-export-lora: fold LoRA adapters into a base GGUF and write the merged model (parity: `llama-export-lora`).

    dlp_export_lora.py -m base.gguf --lora a.gguf --lora-scaled b.gguf 0.5 -o merged.gguf

Merged tensors are written as F16 (targets of an adapter) - everything else keeps its bytes.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf, quantize  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import LoraAdapter  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("--lora", action="append", default=[])
    ap.add_argument("--lora-scaled", action="append", nargs=2, default=[], metavar=("FILE", "SCALE"))
    a = ap.parse_args(argv)
    adapters = [LoraAdapter(p) for p in a.lora] + [LoraAdapter(p, float(s)) for p, s in a.lora_scaled]
    if not adapters:
        sys.exit("no adapters given")
    f = open_gguf(a.model)
    w = GGUFWriter(a.output)
    for key, val in f.metadata.items():
        if isinstance(val, (list, np.ndarray)):
            w.add(key, val)
        else:
            w.add(key, val, f.files[0].metadata_types[key])
    n_merged = 0
    for name, info in f.tensors.items():
        deltas = [d for d in (ad.delta(name) for ad in adapters) if d is not None]
        if deltas:
            n_merged += 1

            def produce(name=name, info=info, deltas=deltas):
                dense = np.array(f.array(name), dtype=np.float32).reshape(info.shape)
                for d in deltas:
                    dense = dense + d.reshape(dense.shape)
                return quantize(dense.reshape(-1, info.shape[-1]), GGMLType.F16)
            w.add_tensor(name, info.shape, GGMLType.F16, produce)
        else:
            w.add_tensor(name, info.shape, info.ggml_type, lambda name=name: np.ascontiguousarray(f.raw(name)).reshape(-1))
    w.write()
    print(f"merged {n_merged} tensors from {len(adapters)} adapter(s) into {a.output}")


if __name__ == "__main__":
    main()
