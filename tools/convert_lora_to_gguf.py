#!/usr/bin/env python
"""PEFT LoRA adapter directory (adapter_config.json + adapter_model.safetensors) -> GGUF adapter.

Parity: `llama.cpp/convert_lora_to_gguf.py`: every `...<module>.lora_A.weight` [r, in] / `lora_B.weight` [out, r] pair becomes
`<gguf tensor>.lora_a` / `.lora_b`, `adapter.lora.alpha` carries `lora_alpha`, and the tensor names and the Q/K row permutation of
the base converter (`tools/convert_hf_to_gguf.py`) are applied to the B factors, so the adapter lines up with a base model converted
by that tool (`LoraAdapter` in `models/weights.py`, `--lora` on the CLI and the servers).
"""
import argparse
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np  # noqa: E402

from convert_hf_to_gguf import ARCHS, LAYER_MAP, PERMUTE_QK, TOP_MAP, load_safetensors, permute_qk  # noqa: E402


def convert(adapter_dir: str, base_dir: str, dst: str) -> str:
    from distributed_llm_pipeline_b200.models.weights import write_lora_gguf
    acfg = json.load(open(os.path.join(adapter_dir, "adapter_config.json")))
    if acfg.get("peft_type", "LORA") != "LORA":
        raise SystemExit(f"unsupported peft_type {acfg.get('peft_type')}")
    if acfg.get("use_dora") or acfg.get("bias", "none") != "none":
        raise SystemExit("DoRA and bias-training adapters are not supported (as in the reference converter)")
    bcfg = json.load(open(os.path.join(base_dir, "config.json")))
    hf_arch = (bcfg.get("architectures") or ["?"])[0]
    if hf_arch not in ARCHS:
        raise SystemExit(f"unsupported base architecture {hf_arch}")
    arch = ARCHS[hf_arch]
    H = bcfg["num_attention_heads"]
    Hkv = bcfg.get("num_key_value_heads") or H
    tensors = load_safetensors(adapter_dir)
    pairs: dict[str, dict[str, np.ndarray]] = {}
    for name, t in tensors.items():
        m = re.match(r"(?:base_model\.model\.)?(.+?)\.lora_([AB])(?:\.[^.]+)?\.weight$", name)
        if not m:
            raise SystemExit(f"unexpected tensor {name}")
        mod, which = m.group(1) + ".weight", m.group(2)
        lm = re.match(r"model\.layers\.(\d+)\.(.+)", mod)
        if lm and lm.group(2) in LAYER_MAP:
            out = f"blk.{lm.group(1)}.{LAYER_MAP[lm.group(2)]}"
        elif mod in TOP_MAP:
            out = TOP_MAP[mod]
        else:
            raise SystemExit(f"unmapped LoRA target {mod}")
        arr = t.float().numpy()
        if which == "B" and arch in PERMUTE_QK:         # the permutation acts on output rows: only the B factor has them
            if out.endswith("attn_q.weight"):
                arr = permute_qk(arr, H)
            elif out.endswith("attn_k.weight"):
                arr = permute_qk(arr, Hkv)
        pairs.setdefault(out, {})[which] = arr
    deltas = {}
    for out, ab in sorted(pairs.items()):
        if set(ab) != {"A", "B"}:
            raise SystemExit(f"{out}: adapter has only factor {sorted(ab)}")
        deltas[out] = (ab["A"], ab["B"])
    write_lora_gguf(dst, deltas, float(acfg.get("lora_alpha", acfg.get("r", 1))), arch)
    print(f"{adapter_dir} -> {dst}: {len(deltas)} target matrices, rank {acfg.get('r')}, alpha {acfg.get('lora_alpha')}, arch {arch}")
    return dst


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("adapter_dir")
    ap.add_argument("--base", required=True, help="base model directory (config.json: architecture and head counts)")
    ap.add_argument("--outfile", required=True)
    a = ap.parse_args(argv)
    convert(a.adapter_dir, a.base, a.outfile)


if __name__ == "__main__":
    main()
