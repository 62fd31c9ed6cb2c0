#!/usr/bin/env python
"""dlp-quantize: re-encode a GGUF checkpoint to another file type (F16 / BF16 / Q8_0 / Q6_K / Q4_K_M / Q5_K_M).

Parity: `llama-quantize` (`llama.cpp/tools/quantize`, mix rules `src/llama-quant.cpp:302-405`).  Tensor-type selection is the
same recipe `gguf/synth.py` uses; encoders are the closed-form ones of `gguf/quants.py`; `--imatrix file.gguf` (from `dlp_imatrix.py` or the reference's
`llama-imatrix`) switches the K-quant encoders to error-weighted fits (`llama-quant.cpp:880-960` imatrix lookup per tensor).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf, quantize  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import FTYPE, BLOCK_GEOMETRY  # noqa: E402
from distributed_llm_pipeline_b200.gguf.synth import tensor_type_for  # noqa: E402
from distributed_llm_pipeline_b200.models.config import config_from_gguf  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("src")
    ap.add_argument("dst")
    ap.add_argument("ftype", choices=sorted(FTYPE))
    ap.add_argument("--fast", action="store_true", help="closed-form K-quant scales instead of the error-minimising search (about 10 %% more RMS error)")
    ap.add_argument("--imatrix", default=None, help="importance matrix GGUF: weigh quantisation error by activation statistics")
    a = ap.parse_args(argv)
    imat = {}
    if a.imatrix:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from dlp_imatrix import load as load_imatrix
        imat = load_imatrix(a.imatrix)
    f = open_gguf(a.src)
    cfg = config_from_gguf(f.metadata, f.tensors)
    w = GGUFWriter(a.dst)
    src_file = f.files[0]
    for key, val in f.metadata.items():
        if key == "general.file_type":
            continue
        w.add(key, val, src_file.metadata_types[key]) if not isinstance(val, (list, np.ndarray)) else w.add(key, val)
    w.add("general.file_type", FTYPE[a.ftype])
    n_in = n_out = 0
    for name, info in f.tensors.items():
        il = int(name.split(".")[1]) if name.startswith("blk.") else 0
        want = tensor_type_for(name, a.ftype, il, cfg) if len(info.shape) > 1 else GGMLType.F32
        per, _ = BLOCK_GEOMETRY[want]
        if info.shape[-1] % per:
            want = GGMLType.F16          # rows that do not fill a block stay dense (as llama-quantize does)
        n_in += info.nbytes

        def produce(name=name, want=want, info=info):
            if want == info.ggml_type:
                return np.ascontiguousarray(f.raw(name)).reshape(-1)
            imp = imat.get(name)
            x = f.array(name)
            if imp is None or imp.shape[-1] != info.shape[-1]:
                # no statistics for this tensor: the same search with every column weighted alike (closed form with --fast)
                return quantize(x.reshape(-1, info.shape[-1]), want, None if a.fast else np.ones(info.shape[-1], np.float32))
            if imp.shape[0] == 1:
                return quantize(x.reshape(-1, info.shape[-1]), want, imp[0])
            # stacked experts: each expert's rows with that expert's statistics (unseen experts fall back to the plain encoder)
            parts = [quantize(x[e].reshape(-1, info.shape[-1]), want, imp[e] if float(imp[e].sum()) > 0 else None) for e in range(x.shape[0])]
            return np.concatenate(parts)

        w.add_tensor(name, info.shape, want, produce)
    w.write()
    n_out = os.path.getsize(a.dst)
    if imat:
        print(f"imatrix: {sum(1 for n in f.tensors if n in imat)} of {len(f.tensors)} tensors weighted")
    print(f"{a.src} -> {a.dst}  [{a.ftype}]  {n_in / 2**20:.1f} MiB -> {n_out / 2**20:.1f} MiB")


if __name__ == "__main__":
    main()
