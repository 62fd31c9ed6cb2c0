#!/usr/bin/env python
"""dlp-cvector: control-vector generator - one direction per layer that separates "positive" from "negative" prompts.

Parity: `llama-cvector-generator` (`llama.cpp/tools/cvector-generator/cvector-generator.cpp`).  For each (positive, negative)
prompt pair, both padded to the same token count, the layer outputs (`l_out`, through the engine's eval callback) of every layer
but the last are recorded and subtracted token by token; rows that are exactly zero (shared prefix) are dropped (`:98-160`).
Per layer the direction is the first principal component of those difference rows (`pca.hpp`, here an exact SVD instead of the
power iteration) or their mean (`mean.hpp`), normalised to unit length, and is written as `direction.<layer + 1>` in a GGUF of
architecture `controlvector` (`export_gguf :353-372`) - the file `LlamaEngine.set_control_vector` / `--control-vector` load.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import ValueType  # noqa: E402


def layer_outputs(eng, ids: list[int]) -> list[torch.Tensor]:
    """Hidden state after every layer for one prompt: list of [T, E] fp32 CPU tensors."""
    got: dict[int, torch.Tensor] = {}
    eng.eval_callback = lambda name, il, t: got.__setitem__(il, t.detach().float().cpu().clone()) if name == "l_out" else None
    sid = eng.alloc_seq()
    try:
        eng.prefill(sid, ids)
    finally:
        eng.free_seq(sid)
        eng.eval_callback = None
    return [got[il] for il in sorted(got)]


def directions(diffs: list[torch.Tensor], method: str) -> list[torch.Tensor]:
    out = []
    for d in diffs:                       # [m, E] difference rows of one layer
        if d.shape[0] == 0:
            out.append(torch.zeros(d.shape[1]))
            continue
        if method == "mean":
            v = d.mean(0)
        else:
            _, _, vh = torch.linalg.svd(d.double(), full_matrices=False)
            v = vh[0].float()
            if float((d @ v).mean()) < 0:      # a principal axis has no sign: point it from negative to positive
                v = -v
        out.append(v / v.norm().clamp_min(1e-12))
    return out


def difference_rows(eng, tok, positive: list[str], negative: list[str], log=print) -> list[torch.Tensor]:
    """Per layer (all but the last, which the reference ignores too): the non-zero rows of (positive - negative) layer outputs, [m, E]."""
    if len(positive) != len(negative) or not positive:
        raise SystemExit("need the same (non-zero) number of positive and negative prompts")
    n_layer = eng.cfg.n_layer
    rows: list[list[torch.Tensor]] = [[] for _ in range(n_layer - 1)]
    pad = tok.encode(" ", add_bos=False)[-1:] or [0]
    for i, (p, n) in enumerate(zip(positive, negative)):
        a, b = tok.encode(p), tok.encode(n)
        T = max(len(a), len(b))
        a, b = a + pad * (T - len(a)), b + pad * (T - len(b))
        ha, hb = layer_outputs(eng, a), layer_outputs(eng, b)
        for il in range(n_layer - 1):
            d = ha[il] - hb[il]
            rows[il].append(d[d.abs().sum(-1) > 0])
        log(f"pair {i + 1}/{len(positive)}: {T} tokens")
    return [torch.cat(r) for r in rows]


def generate(eng, tok, positive: list[str], negative: list[str], method: str = "pca", log=print) -> list[torch.Tensor]:
    return directions(difference_rows(eng, tok, positive, negative, log), method)


def save(path: str, dirs: list[torch.Tensor], model_hint: str) -> str:
    w = GGUFWriter(path)
    w.add("general.architecture", "controlvector")
    w.add("controlvector.model_hint", model_hint)
    w.add("controlvector.layer_count", len(dirs), ValueType.INT32)
    for il, v in enumerate(dirs):
        arr = v.numpy().astype("<f4")
        w.add_tensor(f"direction.{il + 1}", arr.shape, GGMLType.F32, lambda a=arr: a.view(np.uint8).reshape(-1))
    w.write()
    return path


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("--positive-file", required=True, help="one prompt per line")
    ap.add_argument("--negative-file", required=True, help="one prompt per line, paired with the positive file")
    ap.add_argument("-o", "--output", default="control_vector.gguf")
    ap.add_argument("--method", default="pca", choices=["pca", "mean"])
    ap.add_argument("-c", "--ctx-size", type=int, default=512)
    ap.add_argument("--device", default="cuda")
    a = ap.parse_args(argv)
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    dev = a.device if (a.device != "cuda" or torch.cuda.is_available()) else "cpu"
    f = open_gguf(a.model)
    eng = LlamaEngine(load_weights(f, dev), n_ctx=a.ctx_size, max_seqs=1, page_size=64 if dev != "cpu" else 16, max_batch_tokens=a.ctx_size)
    tok = Tokenizer(f.metadata)
    lines = lambda p: [l.rstrip("\n").replace("\\n", "\n") for l in open(p, encoding="utf-8") if l.strip()]
    dirs = generate(eng, tok, lines(a.positive_file), lines(a.negative_file), a.method)
    save(a.output, dirs, str(f.metadata.get("general.architecture", "llama")))
    print(f"{a.output}: {len(dirs)} directions of {dirs[0].numel()} ({a.method})")


if __name__ == "__main__":
    main()
