#!/usr/bin/env python
"""dlp-finetune: next-token training of a GGUF model on a text file - all weights, or a LoRA adapter.

Parity: `llama-finetune` (`llama.cpp/examples/training/finetune.cpp`) on top of `ggml-opt` (`ggml/src/ggml-opt.cpp`: dataset of
context-sized windows, forward + backward graph, AdamW, per-epoch train / validation loss) and, for `--lora`, the adapter format
of `llama-adapter.cpp` that `--lora` loads everywhere else in this repository.  The differentiable graph is the fp32 graph of
`models/reference.py` (`forward_grad`) under PyTorch autograd - any device torch has - so this is a small-model / adapter tool,
like the reference example, not a pre-training stack.  Output: a GGUF checkpoint (same metadata, F32 or F16 matrices) or a LoRA GGUF.
"""
import argparse
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf, quantize  # noqa: E402


def windows(ids: list[int], n_ctx: int, stride: int | None = None) -> torch.Tensor:
    stride = stride or n_ctx
    xs = [ids[i:i + n_ctx + 1] for i in range(0, len(ids) - n_ctx, stride)]
    if not xs:
        raise SystemExit(f"the training text has {len(ids)} tokens: fewer than one window of {n_ctx + 1}")
    return torch.tensor(xs, dtype=torch.long)


def loss_on(model, batch: torch.Tensor, n_ctx: int) -> torch.Tensor:
    total = 0.0
    for row in batch:
        logits = model.forward_grad(row[:-1], model.new_cache(n_ctx))
        total = total + torch.nn.functional.cross_entropy(logits.float(), row[1:])
    return total / len(batch)


def train(model, data: torch.Tensor, val: torch.Tensor | None, params: list[torch.Tensor], epochs: int, lr: float, batch: int, n_ctx: int,
          weight_decay: float = 0.0, seed: int = 0, log=print) -> list[dict]:
    opt = torch.optim.AdamW(params, lr=lr, weight_decay=weight_decay)
    g = torch.Generator().manual_seed(seed)
    hist = []
    for ep in range(epochs):
        order = torch.randperm(len(data), generator=g)
        run = 0.0
        for i in range(0, len(order), batch):
            opt.zero_grad(set_to_none=True)
            loss = loss_on(model, data[order[i:i + batch]], n_ctx)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(params, 1.0)
            opt.step()
            run += float(loss.detach()) * len(order[i:i + batch])
        rec = {"epoch": ep + 1, "train_loss": run / len(order)}
        if val is not None and len(val):
            with torch.no_grad():
                rec["val_loss"] = float(loss_on(model, val, n_ctx))
        hist.append(rec)
        log(f"epoch {ep + 1}/{epochs}: train loss {rec['train_loss']:.4f} (ppl {math.exp(min(20, rec['train_loss'])):.1f})"
            + (f", val loss {rec['val_loss']:.4f}" if "val_loss" in rec else ""))
    return hist


def save_model(src, model, dst: str, outtype: str):
    w = GGUFWriter(dst)
    for key, val in src.metadata.items():
        if key == "general.file_type":
            continue
        w.add(key, val, src.files[0].metadata_types[key]) if not isinstance(val, (list, np.ndarray)) else w.add(key, val)
    w.add("general.file_type", {"F32": 0, "F16": 1}[outtype])
    mat = GGMLType.F32 if outtype == "F32" else GGMLType.F16
    for name, info in src.tensors.items():
        arr = model.w[name].detach().cpu().numpy().astype(np.float32)
        t = mat if (arr.ndim > 1 and info.ggml_type != GGMLType.F32) else GGMLType.F32
        w.add_tensor(name, arr.shape, t, lambda arr=arr, t=t: quantize(arr.reshape(-1, arr.shape[-1]) if arr.ndim > 1 else arr, t))
    return w.write()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("-f", "--file", required=True, help="training text")
    ap.add_argument("-o", "--output", required=True, help="output GGUF (model, or adapter with --lora)")
    ap.add_argument("-c", "--ctx-size", type=int, default=64)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("-b", "--batch-size", type=int, default=4)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--weight-decay", type=float, default=0.0)
    ap.add_argument("--val-split", type=float, default=0.05, help="fraction of windows held out for the validation loss (ggml-opt default)")
    ap.add_argument("--outtype", default="F16", choices=["F32", "F16"])
    ap.add_argument("--lora", type=int, default=0, metavar="RANK", help="train a rank-RANK adapter on the attention and FFN matrices instead of the weights")
    ap.add_argument("--lora-alpha", type=float, default=0.0, help="default: 2 x rank")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--device", default="cpu")
    a = ap.parse_args(argv)
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel
    from distributed_llm_pipeline_b200.models.weights import write_lora_gguf
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    torch.manual_seed(a.seed)
    src = open_gguf(a.model)
    model = ReferenceModel.from_gguf(a.model)
    if model.cfg.is_moe and a.lora:
        raise SystemExit("--lora targets dense matrices; routed-expert checkpoints: train the weights instead")
    ids = Tokenizer(src.metadata).encode(open(a.file, encoding="utf-8").read())
    data = windows(ids, a.ctx_size)
    n_val = int(len(data) * a.val_split)
    val, data = (data[-n_val:], data[:-n_val]) if n_val and len(data) > n_val else (None, data)
    if a.lora:
        alpha = a.lora_alpha or 2.0 * a.lora
        model.lora = {}
        for name, w in model.w.items():
            if w.dim() == 2 and name.startswith("blk.") and name.endswith(".weight") and "norm" not in name and "gate_inp" not in name:
                A = (torch.randn(a.lora, w.shape[1]) / math.sqrt(w.shape[1])).requires_grad_()
                B = torch.zeros(w.shape[0], a.lora, requires_grad=True)
                model.lora[name] = (A, B, alpha / a.lora)
        params = [p for A, B, _ in model.lora.values() for p in (A, B)]
    else:
        params = []
        for name, w in model.w.items():
            if name == "rope_freqs.weight":
                continue
            model.w[name] = w.clone().requires_grad_()
            params.append(model.w[name])
    print(f"{len(data)} training windows of {a.ctx_size} tokens, {0 if val is None else len(val)} held out; "
          f"{sum(p.numel() for p in params) / 1e6:.2f} M trainable parameters")
    hist = train(model, data, val, params, a.epochs, a.lr, a.batch_size, a.ctx_size, a.weight_decay, a.seed)
    if a.lora:
        write_lora_gguf(a.output, {n: (A.detach().numpy(), B.detach().numpy()) for n, (A, B, _) in model.lora.items()},
                        a.lora_alpha or 2.0 * a.lora, model.cfg.arch)
    else:
        save_model(src, model, a.output, a.outtype)
    print(f"{a.output}: train loss {hist[0]['train_loss']:.4f} -> {hist[-1]['train_loss']:.4f}")


if __name__ == "__main__":
    main()
