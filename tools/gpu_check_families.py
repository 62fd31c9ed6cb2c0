#!/usr/bin/env python
"""One-minute GPU check of the architecture families beyond the flagship: the LayerNorm / bias-activation kernels against torch and,
per family, engine logits (prefill through the GEMM path, decode through the mat-vec path inside a CUDA graph) against the fp32
oracle.  Prints one line per item; exit code 1 on any failure.  (The same comparisons are pytest cases in tests/test_kernels_gpu.py
and tests/test_engine_gpu.py; this script is the short form for a tight GPU budget.)"""
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.llama import LlamaEngine  # noqa: E402
from distributed_llm_pipeline_b200.models.reference import ReferenceModel, nmse  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import load_weights  # noqa: E402
from distributed_llm_pipeline_b200.ops import api as ops  # noqa: E402

FAMILIES = [("tiny-starcoder2", "F16"), ("tiny-starcoder2", "Q8_0"), ("tiny-falcon", "F16"), ("tiny-olmo", "F16"), ("tiny-orion", "F16"),
            ("tiny-command-r", "F16"), ("tiny-gptneox", "F16"), ("tiny-gptneox", "Q8_0"), ("tiny-stablelm", "F16"), ("tiny-nemotron", "F16"),
            ("tiny-arcee", "F16"), ("tiny-smollm3", "F16"), ("tiny-smollm3", "Q8_0"), ("tiny-internlm2", "F16"), ("tiny-exaone", "F16"),
            ("tiny-seed-oss", "F16"), ("tiny-minicpm", "F16"), ("tiny-granitemoe", "F16"), ("tiny-gpt2", "F16"), ("tiny-starcoder", "F16"),
            ("tiny-codeshell", "F16"), ("tiny-qwen", "F16"), ("tiny-olmo2", "F16"), ("tiny-glm4", "F16"), ("tiny-mistral3", "F16"), ("tiny-pangu", "F16"), ("tiny-chatglm", "F16"), ("tiny-plamo", "F16"), ("tiny-deepseek", "F16"), ("tiny-bailingmoe", "F16"), ("tiny-qwen2vl", "F16"), ("tiny-qwen3vl", "F16"), ("tiny-exaone4", "F16"), ("tiny-phi2", "F16"), ("tiny-olmoe", "F16"), ("tiny-qwen3vlmoe", "F16"), ("tiny-glm4moe", "F16"), ("tiny-dots1", "F16"), ("tiny-ernie-moe", "F16"), ("tiny-minimax-m2", "F16"), ("tiny-bailingmoe2", "F16"), ("tiny-bitnet", "F16")]
PROMPT = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 3, 250, 77]


def main():
    t0 = time.time()
    bad = 0
    dev = "cuda"

    def report(name, ok, detail=""):
        nonlocal bad
        bad += 0 if ok else 1
        print(f"{'PASS' if ok else 'FAIL'}  {name}  {detail}", flush=True)
    for affine in ("wb", "w", "none"):
        for T, E in ((1, 256), (7, 1024), (130, 6144)):
            try:
                h, d = torch.randn(T, E, device=dev) * 2 + 0.5, torch.randn(T, E, device=dev)
                w = (torch.rand(E, device=dev) + 0.5) if affine != "none" else None
                b = torch.randn(E, device=dev) * 0.1 if affine == "wb" else None
                ref_h = h + d
                ref = torch.nn.functional.layer_norm(ref_h, (E,), w, b, 1e-5)
                out = ops.layernorm(h, w, b, 1e-5, delta=d)
                e = nmse(ref, out)
                report(f"layernorm {affine} {T}x{E}", torch.allclose(h, ref_h, atol=1e-6) and e < 1e-6, f"nmse {e:.2e}")
            except Exception as ex:  # noqa: BLE001
                report(f"layernorm {affine} {T}x{E}", False, repr(ex)[:200])
    for act in ("gelu", "gelu_erf", "relu2", "silu", "none"):
        try:
            x, bias = torch.randn(9, 1536, device=dev) * 3, torch.randn(1536, device=dev)
            fn = {"none": lambda t: t, "gelu": lambda t: torch.nn.functional.gelu(t, approximate="tanh"), "relu2": lambda t: torch.relu(t) ** 2,
                  "silu": torch.nn.functional.silu, "gelu_erf": torch.nn.functional.gelu}[act]
            e = max(nmse(fn(x + bias), ops.bias_act(x, bias, act)), nmse(fn(x), ops.bias_act(x, None, act)))
            report(f"bias_act {act}", e < 1e-6, f"nmse {e:.2e}")
        except Exception as ex:  # noqa: BLE001
            report(f"bias_act {act}", False, repr(ex)[:200])
    tmp = tempfile.mkdtemp()
    for name, ftype in FAMILIES:
        try:
            path = write_synthetic_gguf(os.path.join(tmp, f"{name}-{ftype}.gguf"), preset(name), ftype, seed=1)
            ref = ReferenceModel.from_gguf(path)
            toks, ref_logits = ref.generate_greedy(PROMPT, 5)
            eng = LlamaEngine(load_weights(path, dev), n_ctx=256, max_seqs=4, page_size=16, max_batch_tokens=64)
            sid = eng.alloc_seq()
            errs = [nmse(ref_logits[0], eng.prefill(sid, PROMPT))]
            for i, t in enumerate(toks[:-1]):
                errs.append(nmse(ref_logits[i + 1], eng.decode([sid], [t], sample=False)[0]))
            graphs = any(k[:2] == (1, False) for k in eng._graphs)
            report(f"{name} {ftype}", max(errs) < 1e-3 and graphs, f"prefill {errs[0]:.2e} decode max {max(errs[1:]):.2e} graph {graphs}")
            del eng
        except Exception as ex:  # noqa: BLE001
            report(f"{name} {ftype}", False, repr(ex)[:300])
    print(f"{'ALL PASS' if not bad else str(bad) + ' FAILED'} in {time.time() - t0:.0f} s", flush=True)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
