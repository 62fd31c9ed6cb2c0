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
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0],
                                 usage="dlp_quantize.py [options] model-f32.gguf [model-quant.gguf] type [nthreads]")
    ap.add_argument("positional", nargs="+", help="input file, optional output file, file type (name in any case, or llama_ftype number), optional thread count")
    ap.add_argument("--fast", action="store_true", help="closed-form K-quant scales instead of the error-minimising search (about 10 %% more RMS error)")
    ap.add_argument("--imatrix", default=None, help="importance matrix GGUF: weigh quantisation error by activation statistics")
    # llama-quantize's own switches (tools/quantize/quantize.cpp usage)
    ap.add_argument("--allow-requantize", action="store_true", help="accepted: tensors that are already quantised are always decoded and re-encoded here")
    ap.add_argument("--leave-output-tensor", action="store_true", help="keep output.weight as it is")
    ap.add_argument("--pure", action="store_true", help="no mixtures: every matrix gets the file type's own tensor type")
    ap.add_argument("--output-tensor-type", default=None, help="tensor type for output.weight")
    ap.add_argument("--token-embedding-type", default=None, help="tensor type for token_embd.weight")
    ap.add_argument("--tensor-type", action="append", default=[], metavar="PATTERN=TYPE", help="tensor type for tensors whose name matches the regular expression (may be repeated)")
    ap.add_argument("--include-weights", action="append", default=[], help="use the importance matrix only for tensors whose name contains this")
    ap.add_argument("--exclude-weights", action="append", default=[], help="do not use the importance matrix for tensors whose name contains this")
    ap.add_argument("--keep-split", action="store_true", help="accepted: the output is one file (split it with dlp_gguf_split.py)")
    ap.add_argument("--override-kv", action="append", default=[], metavar="KEY=TYPE:VALUE", help="replace a metadata value in the output")
    ap.add_argument("--dry-run", action="store_true", help="print the tensor types that would be written and the resulting size, write nothing")
    a = ap.parse_args(argv)
    alias = {"Q4_K": "Q4_K_M", "Q5_K": "Q5_K_M", "Q3_K_S": "Q3_K", "Q2_K_S": "Q2_K", "COPY": None}
    by_id = {v: k for k, v in FTYPE.items()}

    def as_ftype(tok):
        t = str(tok).upper()
        if t.isdigit() and int(t) in by_id:
            return by_id[int(t)]
        t = alias.get(t, t)
        return t if (t in FTYPE or t is None) else False
    pos = list(a.positional)
    if len(pos) >= 2 and pos[-1].isdigit() and as_ftype(pos[-2]) is not False and not (len(pos) == 2):
        pos.pop()                               # trailing thread count: the encoders are numpy, one process
    if len(pos) < 2 or as_ftype(pos[-1]) is False:
        ap.error(f"expected: in.gguf [out.gguf] type; known types: {', '.join(sorted(FTYPE))} (Q4_K = Q4_K_M, Q5_K = Q5_K_M, Q3_K_S = Q3_K, COPY)")
    a.ftype = as_ftype(pos[-1])
    a.src = pos[0]
    copy_only = a.ftype is None
    if copy_only:
        a.ftype = by_id.get(int(open_gguf(a.src).metadata.get("general.file_type", 1)), "F16")
    a.dst = pos[1] if len(pos) >= 3 else os.path.join(os.path.dirname(os.path.abspath(a.src)), f"ggml-model-{a.ftype}.gguf")

    def as_type(name):
        try:
            return GGMLType[str(name).upper()]
        except KeyError:
            ap.error(f"unknown tensor type {name!r}")
    import re
    rules = []
    for spec in a.tensor_type:
        pat, _, tname = spec.rpartition("=")
        if not pat:
            ap.error(f"--tensor-type {spec!r}: expected PATTERN=TYPE")
        rules.append((re.compile(pat), as_type(tname)))
    base_type = {"Q4_K_M": GGMLType.Q4_K, "Q5_K_M": GGMLType.Q5_K, "MXFP4_MOE": GGMLType.MXFP4}.get(a.ftype) or GGMLType[a.ftype] if a.ftype in GGMLType.__members__ or a.ftype in ("Q4_K_M", "Q5_K_M", "MXFP4_MOE") else None
    imat = {}
    if a.imatrix:
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from dlp_imatrix import load as load_imatrix
        imat = load_imatrix(a.imatrix)
    f = open_gguf(a.src)
    cfg = config_from_gguf(f.metadata, f.tensors)
    w = GGUFWriter(a.dst if not a.dry_run else os.devnull)
    src_file = f.files[0]
    from distributed_llm_pipeline_b200.utils.args import parse_override_kv
    overrides = dict(parse_override_kv(spec, cfg.arch) for spec in a.override_kv)
    for key, val in f.metadata.items():
        if key == "general.file_type" or key in overrides:
            continue
        w.add(key, val, src_file.metadata_types[key]) if not isinstance(val, (list, np.ndarray)) else w.add(key, val)
    for key, val in overrides.items():
        w.add(key, val)
    w.add("general.file_type", FTYPE[a.ftype])
    n_in = n_out = 0
    planned = 0
    for name, info in f.tensors.items():
        il = int(name.split(".")[1]) if name.startswith("blk.") else 0
        want = tensor_type_for(name, a.ftype, il, cfg) if len(info.shape) > 1 else GGMLType.F32
        if len(info.shape) > 1:
            if copy_only or (a.leave_output_tensor and name == "output.weight"):
                want = info.ggml_type
            elif a.pure and base_type is not None and not name.endswith(("ffn_gate_inp.weight", "position_embd.weight")):
                want = base_type
            if a.output_tensor_type and name == "output.weight":
                want = as_type(a.output_tensor_type)
            if a.token_embedding_type and name == "token_embd.weight":
                want = as_type(a.token_embedding_type)
            for rx, t in rules:
                if rx.search(name):
                    want = t
        per, _ = BLOCK_GEOMETRY[want]
        if info.shape[-1] % per:
            want = GGMLType.F16          # rows that do not fill a block stay dense (as llama-quantize does)
        n_in += info.nbytes
        if a.dry_run:
            per, nb = BLOCK_GEOMETRY[want]
            n_el = int(np.prod(info.shape))
            planned += n_el // per * nb
            print(f"{name:48s} {str(tuple(info.shape)):>24s}  {GGMLType(info.ggml_type).name:>8s} -> {want.name}")
            continue

        def produce(name=name, want=want, info=info):
            if want == info.ggml_type:
                return np.ascontiguousarray(f.raw(name)).reshape(-1)
            imp = imat.get(name)
            if (a.include_weights and not any(k in name for k in a.include_weights)) or any(k in name for k in a.exclude_weights):
                imp = None
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
    if a.dry_run:
        print(f"{a.src} -> [{a.ftype}]  {n_in / 2**20:.1f} MiB -> {planned / 2**20:.1f} MiB of tensor data (nothing written)")
        return
    w.write()
    n_out = os.path.getsize(a.dst)
    if imat:
        print(f"imatrix: {sum(1 for n in f.tensors if n in imat)} of {len(f.tensors)} tensors weighted")
    print(f"{a.src} -> {a.dst}  [{a.ftype}]  {n_in / 2**20:.1f} MiB -> {n_out / 2**20:.1f} MiB")


if __name__ == "__main__":
    main()
