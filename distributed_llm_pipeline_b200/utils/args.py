"""Declarative flag table with environment fall-backs, shared by the tools.

Parity: `common_params` + `common_arg{names, help, handler}.set_env("LLAMA_ARG_*")` (`llama.cpp/common/arg.cpp`,
`common/common.h`): every option has CLI aliases and an env var; tools take subsets.  Flag names follow the reference
where the concept exists (`-m -p -n -c -b -ub -np -ngl -sm -ts -mg --rpc -fa --temp --top-k --top-p --min-p --seed
-v --log-file`); `-ngl` / `-fa` are accepted for command-line compatibility (all layers always live on the GPU and
attention is always flash) and reported as such.
"""
from __future__ import annotations

import argparse
import os
from dataclasses import dataclass


@dataclass
class Opt:
    names: tuple
    help: str
    default: object = None
    type: type | None = str
    env: str | None = None
    action: str | None = None
    choices: tuple | None = None


COMMON = [
    Opt(("-m", "--model"), "GGUF checkpoint (first shard of a split set)", env="DLP_ARG_MODEL"),
    Opt(("-c", "--ctx-size"), "context length per sequence", 2048, int, "DLP_ARG_CTX_SIZE"),
    Opt(("-b", "--batch-size"), "max tokens per scheduling step (logical batch)", 2048, int, "DLP_ARG_BATCH"),
    Opt(("-ub", "--ubatch-size"), "prefill chunk size (physical batch)", 512, int, "DLP_ARG_UBATCH"),
    Opt(("-np", "--parallel"), "number of sequence slots", 1, int, "DLP_ARG_N_PARALLEL"),
    Opt(("-ngl", "--gpu-layers", "--n-gpu-layers"), "accepted for compatibility: all layers are always GPU resident", 99, int, "DLP_ARG_N_GPU_LAYERS"),
    Opt(("-sm", "--split-mode"), "none | layer (pipeline stages) | row (tensor parallel)", "layer", str, "DLP_ARG_SPLIT_MODE", choices=("none", "layer", "row")),
    Opt(("-ts", "--tensor-split"), "comma separated stage weights for -sm layer (e.g. 3,1)", None, str, "DLP_ARG_TENSOR_SPLIT"),
    Opt(("--pp",), "pipeline stages (explicit pp x tp grid; default from -sm)", None, int, "DLP_ARG_PP"),
    Opt(("--tp",), "tensor-parallel ranks per stage", None, int, "DLP_ARG_TP"),
    Opt(("--tp-impl",), "fused (in-kernel NVLink all-reduce) | nccl (baseline)", "fused", str, "DLP_TP_IMPL", choices=("fused", "nccl")),
    Opt(("--rpc",), "comma separated host:port list of dlp-worker CPU stages", None, str, "DLP_ARG_RPC"),
    Opt(("-fa", "--flash-attn"), "accepted for compatibility: attention is always flash", "on", str, "DLP_ARG_FLASH_ATTN"),
    Opt(("--lora",), "comma separated LoRA adapter GGUFs, each optionally path:scale (merged at load)", None, str, "DLP_ARG_LORA"),
    Opt(("--device",), "cuda | cpu", "cuda", str, "DLP_ARG_DEVICE"),
    Opt(("--mlock",), "pin the model file's pages in RAM (no swapping)", False, None, "DLP_MLOCK", action="store_true"),
    Opt(("--no-mmap",), "read the model file into memory instead of mapping it", False, None, "DLP_NO_MMAP", action="store_true"),
    Opt(("--direct-io",), "read the model file with O_DIRECT (bypasses the page cache; implies --no-mmap)", False, None, "DLP_DIRECT_IO", action="store_true"),
    Opt(("--list-devices",), "print the available devices and exit", False, None, action="store_true"),
    Opt(("--fit",), "shrink -c / -np until weights + KV cache fit the device memory", False, None, "DLP_ARG_FIT", action="store_true"),
    Opt(("-v", "--verbose"), "debug logging", False, None, "DLP_ARG_VERBOSE", action="store_true"),
    Opt(("--log-file",), "also write logs to this file", None, str, "DLP_ARG_LOG_FILE"),
    Opt(("--trace",), "write a chrome://tracing JSON of the run", None, str, "DLP_ARG_TRACE"),
]
SAMPLING = [
    Opt(("--temp", "--temperature"), "sampling temperature (0 = greedy)", 0.8, float, "DLP_ARG_TEMP"),
    Opt(("--top-k",), "top-k", 40, int, "DLP_ARG_TOP_K"),
    Opt(("--top-p",), "top-p", 0.95, float, "DLP_ARG_TOP_P"),
    Opt(("--min-p",), "min-p", 0.05, float, "DLP_ARG_MIN_P"),
    Opt(("--typical",), "locally typical sampling p", 1.0, float, "DLP_ARG_TYPICAL"),
    Opt(("--repeat-penalty",), "repetition penalty", 1.0, float, "DLP_ARG_REPEAT_PENALTY"),
    Opt(("--presence-penalty",), "presence penalty", 0.0, float, "DLP_ARG_PRESENCE_PENALTY"),
    Opt(("--frequency-penalty",), "frequency penalty", 0.0, float, "DLP_ARG_FREQUENCY_PENALTY"),
    Opt(("-s", "--seed"), "RNG seed", 0, int, "DLP_ARG_SEED"),
    Opt(("--top-nsigma",), "top-n-sigma (<= 0: off)", -1.0, float, "DLP_ARG_TOP_NSIGMA"),
    Opt(("--dynatemp-range",), "dynamic temperature range", 0.0, float, "DLP_ARG_DYNATEMP_RANGE"),
    Opt(("--dynatemp-exp",), "dynamic temperature exponent", 1.0, float, "DLP_ARG_DYNATEMP_EXP"),
    Opt(("--xtc-probability",), "XTC probability", 0.0, float, "DLP_ARG_XTC_PROBABILITY"),
    Opt(("--xtc-threshold",), "XTC threshold", 0.1, float, "DLP_ARG_XTC_THRESHOLD"),
    Opt(("--mirostat",), "0 = off, 2 = mirostat v2", 0, int, "DLP_ARG_MIROSTAT"),
    Opt(("--mirostat-ent",), "mirostat target entropy (tau)", 5.0, float, "DLP_ARG_MIROSTAT_ENT"),
    Opt(("--mirostat-lr",), "mirostat learning rate (eta)", 0.1, float, "DLP_ARG_MIROSTAT_LR"),
    Opt(("--grammar",), "GBNF grammar text constraining the output", "", str, "DLP_ARG_GRAMMAR"),
    Opt(("--grammar-file",), "read the GBNF grammar from a file", None, str),
    Opt(("-j", "--json-schema"), "JSON schema (text) constraining the output", None, str, "DLP_ARG_JSON_SCHEMA"),
]
SPECULATIVE = [
    Opt(("-md", "--model-draft"), "draft GGUF model for speculative decoding (greedy)", None, str, "DLP_ARG_MODEL_DRAFT"),
    Opt(("--lookup",), "speculative decoding with prompt-lookup (n-gram) drafts", False, None, action="store_true"),
    Opt(("--draft-max", "--draft"), "tokens to draft per verification pass", 8, int, "DLP_ARG_DRAFT_MAX"),
]
GENERATION = [
    Opt(("-p", "--prompt"), "prompt text", "", str, "DLP_ARG_PROMPT"),
    Opt(("-f", "--file"), "read the prompt from a file", None, str),
    Opt(("-n", "--predict", "--n-predict"), "tokens to generate", 200, int, "DLP_ARG_N_PREDICT"),
    Opt(("--ignore-eos",), "keep generating after EOS", False, None, action="store_true"),
    Opt(("-st", "--single-turn"), "exit after one reply", False, None, action="store_true"),
    Opt(("--context-shift",), "slide the context window instead of stopping at n_ctx", False, None, "DLP_ARG_CONTEXT_SHIFT", action="store_true"),
    Opt(("--keep",), "prompt tokens kept across a context shift (-1 = all)", 0, int, "DLP_ARG_KEEP"),
]


def build_parser(description: str, groups: list[list[Opt]]) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=description, formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    for g in groups:
        for o in g:
            default = o.default
            if o.env and o.env in os.environ:
                raw = os.environ[o.env]
                default = (raw.lower() not in ("0", "false", "")) if o.action == "store_true" else (o.type(raw) if o.type else raw)
            kw = dict(help=o.help + (f" [env {o.env}]" if o.env else ""), default=default)
            if o.action:
                kw["action"] = o.action
            else:
                kw["type"] = o.type
                if o.choices:
                    kw["choices"] = o.choices
            ap.add_argument(*o.names, **kw)
    return ap


def apply_loader_flags(a):
    """`--mlock / --no-mmap / --direct-io` -> the environment switches `gguf.reader.GGUFFile` reads (so every loader in the
    process, including worker ranks started later, sees them)."""
    for flag, env in (("mlock", "DLP_MLOCK"), ("no_mmap", "DLP_NO_MMAP"), ("direct_io", "DLP_DIRECT_IO")):
        if getattr(a, flag, False):
            os.environ[env] = "1"


def sampling_from_args(a):
    from ..serving.sampler import SamplingParams
    import json
    grammar = open(a.grammar_file).read() if getattr(a, "grammar_file", None) else getattr(a, "grammar", "")
    schema = json.loads(a.json_schema) if getattr(a, "json_schema", None) else None
    return SamplingParams(temperature=a.temp, top_k=a.top_k, top_p=a.top_p, min_p=a.min_p, typical_p=a.typical,
                          repeat_penalty=a.repeat_penalty, presence_penalty=a.presence_penalty,
                          frequency_penalty=a.frequency_penalty, seed=a.seed, top_n_sigma=a.top_nsigma,
                          dynatemp_range=a.dynatemp_range, dynatemp_exponent=a.dynatemp_exp, xtc_probability=a.xtc_probability,
                          xtc_threshold=a.xtc_threshold, mirostat=a.mirostat, mirostat_tau=a.mirostat_ent,
                          mirostat_eta=a.mirostat_lr, grammar=grammar or "", json_schema=schema)


def lora_from_args(a):
    """`--lora a.gguf,b.gguf:0.5` -> [LoraAdapter] (reference flags `--lora` / `--lora-scaled`)."""
    from ..models.weights import LoraAdapter
    out = []
    for item in (getattr(a, "lora", None) or "").split(","):
        if not item:
            continue
        path, _, sc = item.rpartition(":") if item.count(":") and item.rsplit(":", 1)[1].replace(".", "", 1).isdigit() else (item, "", "")
        out.append(LoraAdapter(path or item, float(sc) if sc else 1.0))
    return out
