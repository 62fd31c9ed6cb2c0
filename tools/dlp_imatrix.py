#!/usr/bin/env python
"""dlp-imatrix: importance matrix of a model over a calibration text - for every weight matrix, the sum over tokens of the
squared activations entering each input column (per expert for routed experts).

Parity: `llama-imatrix` (`llama.cpp/tools/imatrix/imatrix.cpp`): same statistic (`collect_imatrix :225-400`), same GGUF layout
(`save_imatrix :497-600`: `general.type = "imatrix"`, `imatrix.datasets / chunk_count / chunk_size`, one `<weight>.in_sum2`
[n_mat, K] and one `<weight>.counts` [n_mat, 1] tensor per matrix), so the file feeds either quantizer.  The forward pass is the
fp32 graph of `models/reference.py` (every weight mat-mul goes through `ReferenceModel.mm`, which this tool observes) in chunks of
`-c` tokens, the way the reference tool walks its dataset.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import ValueType  # noqa: E402


class Collector:
    def __init__(self, model, process_output: bool = False):
        self.model = model
        self.process_output = process_output     # the output matrix is only measured on request (imatrix.cpp `--process-output`)
        self.sum2: dict[str, np.ndarray] = {}
        self.counts: dict[str, np.ndarray] = {}

    def __call__(self, name: str, x: torch.Tensor, expert):
        if not name.startswith("blk.") and not self.process_output:
            return
        w = self.model.w[name]
        n_mat = w.shape[0] if w.dim() == 3 else 1
        K = w.shape[-1]
        if name not in self.sum2:
            self.sum2[name] = np.zeros((n_mat, K), np.float64)
            self.counts[name] = np.zeros((n_mat, 1), np.float64)
        x2 = x.reshape(-1, K).double()
        e = 0 if expert is None else int(expert)
        self.sum2[name][e] += (x2 * x2).sum(0).numpy()
        self.counts[name][e, 0] += x2.shape[0]


def collect(model_path: str, token_ids: list[int], n_ctx: int, n_chunks: int, log=print, process_output: bool = False):
    from distributed_llm_pipeline_b200.gguf import open_gguf as _open
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel
    m = ReferenceModel.from_gguf(model_path)
    md = _open(model_path).metadata
    bos = int(md.get("tokenizer.ggml.bos_token_id", 1)) if md.get("tokenizer.ggml.add_bos_token", md.get("tokenizer.ggml.model", "llama") == "llama") else None
    col = Collector(m, process_output)
    m.observer = col
    n_chunks = min(n_chunks if n_chunks > 0 else 1 << 30, len(token_ids) // n_ctx)
    if n_chunks < 1:
        raise SystemExit(f"the calibration text has {len(token_ids)} tokens: fewer than one chunk of {n_ctx}")
    for c in range(n_chunks):
        chunk = list(token_ids[c * n_ctx:(c + 1) * n_ctx])
        if bos is not None:
            chunk[0] = bos           # every window starts with BOS, as the reference tool arranges it (imatrix.cpp compute_imatrix)
        m.forward(chunk, m.new_cache(n_ctx))
        log(f"[{c + 1}/{n_chunks}] {len(col.sum2)} matrices")
    m.observer = None
    return col, n_chunks


def save(path: str, col: Collector, n_chunks: int, n_ctx: int, dataset: str):
    w = GGUFWriter(path)
    w.add("general.type", "imatrix")
    w.add("imatrix.datasets", [dataset], etype=ValueType.STRING)
    w.add("imatrix.chunk_count", int(n_chunks), ValueType.UINT32)
    w.add("imatrix.chunk_size", int(n_ctx), ValueType.UINT32)
    for name in sorted(col.sum2):
        s2 = col.sum2[name].astype("<f4")
        cn = col.counts[name].astype("<f4")
        w.add_tensor(name + ".in_sum2", s2.shape, GGMLType.F32, lambda a=s2: a.view(np.uint8).reshape(-1))
        w.add_tensor(name + ".counts", cn.shape, GGMLType.F32, lambda a=cn: a.view(np.uint8).reshape(-1))
    w.write()
    return path


def load(path: str) -> dict[str, np.ndarray]:
    """imatrix GGUF -> {weight name: mean squared activation per input column, [n_mat, K]} (what the quantizer weighs errors by)."""
    f = open_gguf(path)
    if f.metadata.get("general.type") != "imatrix":
        raise ValueError(f"{path} is not an imatrix file")
    out = {}
    for name in f.names():
        if name.endswith(".in_sum2"):
            base = name[: -len(".in_sum2")]
            s2 = np.array(f.array(name), np.float32)
            cn = np.array(f.array(base + ".counts"), np.float32).reshape(-1, 1)
            out[base] = np.where(cn > 0, s2 / np.maximum(cn, 1), 0).astype(np.float32)
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("-f", "--file", default=None, help="calibration text (default: synthetic token ids)")
    ap.add_argument("-o", "--output", default="imatrix.gguf")
    ap.add_argument("-c", "--ctx-size", type=int, default=128)
    ap.add_argument("--chunks", type=int, default=-1)
    ap.add_argument("--process-output", action="store_true", help="also collect statistics for output.weight")
    # llama-imatrix's other switches (tools/imatrix/imatrix.cpp usage)
    ap.add_argument("--chunk", "--from-chunk", type=int, default=0, dest="from_chunk", help="skip this many chunks of the text first")
    ap.add_argument("--in-file", action="append", default=[], help="existing imatrix GGUF to start from / to combine (may be repeated; without -f: just merge them)")
    ap.add_argument("--parse-special", action="store_true", help="special-token text in the calibration file stands for the tokens")
    ap.add_argument("--show-statistics", action="store_true", help="print per-matrix statistics of an imatrix (with --in-file) and exit")
    ap.add_argument("--no-ppl", action="store_true", help="accepted: no perplexity is computed along the way here")
    ap.add_argument("--output-frequency", type=int, default=10, help="accepted: the file is written once at the end")
    ap.add_argument("--save-frequency", type=int, default=0, help="accepted: no intermediate snapshots")
    ap.add_argument("--output-format", default="gguf", choices=("gguf",), help="only the GGUF layout is written (the legacy .dat layout is not)")
    a = ap.parse_args(argv)
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer

    def raw_sums(path):      # {name: (in_sum2, counts)}, chunk count
        fi = open_gguf(path)
        if fi.metadata.get("general.type") != "imatrix":
            raise SystemExit(f"{path} is not an imatrix file")
        return ({n[:-len(".in_sum2")]: (np.array(fi.array(n), np.float64), np.array(fi.array(n[:-len(".in_sum2")] + ".counts"), np.float64))
                 for n in fi.names() if n.endswith(".in_sum2")}, int(fi.metadata.get("imatrix.chunk_count", 0)), int(fi.metadata.get("imatrix.chunk_size", a.ctx_size)))
    if a.show_statistics:
        if not a.in_file:
            ap.error("--show-statistics needs --in-file")
        sums, _, _ = raw_sums(a.in_file[0])
        print(f"{'tensor':48s} {'sum(act^2)':>14s} {'min':>12s} {'max':>12s} {'mean':>12s} {'std':>12s} {'entropy':>9s}")
        for name in sorted(sums):
            s2, cn = sums[name]
            v = (s2 / np.maximum(cn, 1)).reshape(-1)
            pr = v / v.sum() if v.sum() > 0 else np.full_like(v, 1.0 / v.size)
            ent = float(-(pr[pr > 0] * np.log2(pr[pr > 0])).sum())
            print(f"{name:48s} {v.sum():14.4f} {v.min():12.6f} {v.max():12.6f} {v.mean():12.6f} {v.std():12.6f} {ent:9.4f}")
        return
    f = open_gguf(a.model)
    col = None
    n, n_ctx = 0, a.ctx_size
    if a.file or not a.in_file:
        if a.file:
            ids = Tokenizer(f.metadata).encode(open(a.file, encoding="utf-8").read(), parse_special=a.parse_special)
        else:
            g = torch.Generator().manual_seed(0)
            ids = torch.randint(3, len(f.metadata["tokenizer.ggml.tokens"]), (a.ctx_size * (a.from_chunk + max(1, a.chunks if a.chunks > 0 else 4)),), generator=g).tolist()
        ids = ids[a.from_chunk * a.ctx_size:]
        col, n = collect(a.model, ids, a.ctx_size, a.chunks, process_output=a.process_output)
    for path in a.in_file:          # statistics are sums: files combine by adding them (imatrix.cpp load_imatrix)
        sums, chunks, n_ctx = raw_sums(path)
        if col is None:
            col = Collector.__new__(Collector)
            col.sum2, col.counts = {}, {}
        for name, (s2, cn) in sums.items():
            if name in col.sum2 and col.sum2[name].shape == s2.shape:
                col.sum2[name] = col.sum2[name] + s2
                col.counts[name] = col.counts[name] + cn.reshape(col.counts[name].shape)
            else:
                col.sum2[name], col.counts[name] = s2.copy(), cn.copy()
        n += chunks
    save(a.output, col, n, n_ctx, a.file or ("synthetic" if not a.in_file else ",".join(a.in_file)))
    print(f"{a.output}: {len(col.sum2)} matrices, {n} chunks of {n_ctx} tokens")


if __name__ == "__main__":
    main()
