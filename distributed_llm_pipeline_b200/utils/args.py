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


SAMPLING += [
    Opt(("--repeat-last-n",), "tokens the penalties look back over (0 = off, -1 = all)", 64, int, "DLP_ARG_REPEAT_LAST_N"),
    Opt(("--dry-multiplier",), "DRY repetition penalty multiplier (0 = off)", 0.0, float, "DLP_ARG_DRY_MULTIPLIER"),
    Opt(("--dry-base",), "DRY base", 1.75, float, "DLP_ARG_DRY_BASE"),
    Opt(("--dry-allowed-length",), "DRY: repeats up to this length are free", 2, int, "DLP_ARG_DRY_ALLOWED_LENGTH"),
    Opt(("--dry-penalty-last-n",), "DRY: look-back window (-1 = whole history)", -1, int, "DLP_ARG_DRY_PENALTY_LAST_N"),
    Opt(("--dry-sequence-breaker",), "DRY: a string that ends a repeated span (may be repeated; \"none\" clears the defaults)", None, str, action="append"),
    Opt(("-l", "--logit-bias"), "TOKEN_ID(+/-)BIAS, e.g. 15043+1 or 15043-inf (may be repeated)", None, str, action="append"),
    Opt(("-jf", "--json-schema-file"), "read the JSON schema constraining the output from a file", None, str),
    Opt(("--samplers", "--sampling-seq", "--sampler-seq"), "accepted: the chain order is fixed (penalties, dry, top-n-sigma, top-k, typical, top-p, min-p, xtc, temperature)", None, str),
]
GENERATION += [
    Opt(("-sys", "--system-prompt"), "system message placed in front of the conversation (chat template applied)", None, str, "DLP_ARG_SYSTEM_PROMPT"),
    Opt(("-sysf", "--system-prompt-file"), "read the system message from a file", None, str),
    Opt(("-r", "--reverse-prompt"), "stop generating at this string and hand control back (may be repeated)", None, str, action="append"),
    Opt(("-e", "--escape"), "process \\n \\t \\\" \\\\ escapes in the prompt (the default)", True, None, action="store_true"),
    Opt(("--no-escape",), "leave escapes in the prompt as they are", False, None, action="store_true"),
    Opt(("--in-prefix",), "text put in front of every user input in the conversation loop", "", str),
    Opt(("--in-suffix",), "text put behind every user input in the conversation loop", "", str),
    Opt(("-cnv", "--conversation"), "keep asking for input after each reply (the default on a terminal)", False, None, action="store_true"),
    Opt(("-no-cnv", "--no-conversation"), "one completion, no chat loop (same as -st)", False, None, action="store_true"),
    Opt(("-i", "--interactive"), "same as --conversation", False, None, action="store_true"),
    Opt(("-if", "--interactive-first"), "ask for input before the first generation", False, None, action="store_true"),
    Opt(("--no-display-prompt",), "accepted: the prompt is never echoed here (stdout carries the generated text only)", False, None, action="store_true"),
    Opt(("--display-prompt",), "echo the prompt in front of the generated text", False, None, action="store_true"),
    Opt(("-sp", "--special"), "accepted: control tokens are never written to stdout", False, None, action="store_true"),
    Opt(("--verbose-prompt",), "log the prompt's tokens before generating", False, None, action="store_true"),
    Opt(("--control-vector",), "control vector GGUF added to the residual stream (may be repeated)", None, str, action="append"),
    Opt(("--control-vector-scaled",), "FILE SCALE: a control vector with a strength", None, str, action="append"),
    Opt(("--control-vector-layer-range",), "START END: layers the control vectors apply to", None, str),
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
                if o.action == "append":
                    kw["type"] = o.type
            else:
                kw["type"] = o.type
                if o.choices:
                    kw["choices"] = o.choices
            if "--control-vector-scaled" in o.names or "--control-vector-layer-range" in o.names:
                kw["nargs"] = 2
            ap.add_argument(*o.names, **kw)
    add_compat(ap)
    inner = ap.parse_args

    def parse_args(args=None, namespace=None):       # the compatibility flags act right after parsing, for every tool built on this table
        return finish_compat(ap, inner(args, namespace))
    ap.parse_args = parse_args
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
    if schema is None and getattr(a, "json_schema_file", None):
        schema = json.load(open(a.json_schema_file))
    bias = {}
    for item in getattr(a, "logit_bias", None) or []:       # 15043+1 / 15043-0.5 / 15043-inf
        for sign in ("+", "-"):
            tid, _, val = item.partition(sign)
            if val and tid.strip().isdigit():
                bias[int(tid)] = float(sign + val)
                break
    extra = dict(repeat_last_n=getattr(a, "repeat_last_n", 64), dry_multiplier=getattr(a, "dry_multiplier", 0.0), dry_base=getattr(a, "dry_base", 1.75),
                 dry_allowed_length=getattr(a, "dry_allowed_length", 2), dry_penalty_last_n=getattr(a, "dry_penalty_last_n", -1), logit_bias=bias)
    return SamplingParams(**extra, temperature=a.temp, top_k=a.top_k, top_p=a.top_p, min_p=a.min_p, typical_p=a.typical,
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


# ---- command lines written for the reference keep working -----------------------------------------------------------------------
# (`llama.cpp/common/arg.cpp` knows ~330 flags.)  Those with a meaning here are ordinary options above or in COMPAT_WIRED below; the
# rest fall in two classes: settings of machinery this engine does not have (CPU affinity, ggml back-end placement, KV defragmentation
# ...) are accepted and reported as having no effect, features that do not exist here (multimodal input, model download, TLS, TTS,
# diffusion) are refused with a sentence saying so - never an "unrecognized arguments" for a flag the reference documents.
COMPAT_IGNORED = {   # flag names -> number of values they take
    ("-C", "--cpu-mask"): 1, ("-Cr", "--cpu-range"): 1, ("--cpu-strict",): 1, ("--prio",): 1, ("--poll",): 1, ("-Cb", "--cpu-mask-batch"): 1,
    ("-Crb", "--cpu-range-batch"): 1, ("--cpu-strict-batch",): 1, ("--prio-batch",): 1, ("--poll-batch",): 1, ("--numa",): 1,
    ("-mg", "--main-gpu"): 1, ("-nkvo", "--no-kv-offload"): 0, ("-kvo", "--kv-offload"): 0, ("--no-op-offload",): 0, ("--op-offload",): 0,
    ("-nr", "--no-repack"): 0, ("--repack",): 0, ("--no-host",): 0, ("--mmap",): 0, ("-ndio", "--no-direct-io"): 0, ("-dt", "--defrag-thold"): 1,
    ("-kvu", "--kv-unified"): 0, ("--swa-full",): 0, ("--swa-checkpoints", "--ctx-checkpoints"): 1, ("--check-tensors",): 0,
    ("--log-colors",): "?", ("--log-prefix",): 0, ("--log-timestamps",): 0, ("-lv", "--verbosity", "--log-verbosity"): 1, ("--log-disable",): 0,
    ("--perf",): 0, ("--no-perf",): 0, ("-cmoe", "--cpu-moe"): 0, ("-ncmoe", "--n-cpu-moe"): 1, ("-ot", "--override-tensor"): 1,
    ("-cram", "--cache-ram"): 1, ("--cache-reuse",): 1, ("-sps", "--slot-prompt-similarity"): 1, ("-dev", "--devices"): 1,
    ("--threads-http",): 1, ("-cb", "--cont-batching"): 0, ("-nocb", "--no-cont-batching"): 0, ("--jinja",): 0, ("--warmup",): 0,
    ("--metrics",): 0, ("--slots",): 0, ("--props",): 0, ("--embedding", "--embeddings"): 0, ("--rerank", "--reranking"): 0,
    ("--no-webui",): 0, ("--webui",): 0, ("-fit", "--fit-target"): 1, ("--fit-ctx",): 1, ("--grp-attn-n", "-gan"): 1, ("--grp-attn-w", "-gaw"): 1,
    ("--color", "-co"): "?", ("--simple-io",): 0, ("--backend-sampling",): 0, ("--no-show-timings",): 0, ("--show-timings",): 0,
    ("--offline",): 0, ("--license",): 0,
}
COMPAT_REFUSED = {
    ("-mm", "--mmproj"): "multimodal input is not implemented", ("-mmu", "--mmproj-url"): "multimodal input is not implemented",
    ("--image", "--audio"): "multimodal input is not implemented", ("--image-min-tokens",): "multimodal input is not implemented",
    ("--image-max-tokens",): "multimodal input is not implemented", ("-hf", "-hfr", "--hf-repo"): "there is no network here: pass a local file with -m",
    ("-hff", "--hf-file"): "there is no network here: pass a local file with -m", ("-mu", "--model-url"): "there is no network here: pass a local file with -m",
    ("-dr", "--docker-repo"): "there is no network here: pass a local file with -m", ("--ssl-key-file",): "TLS is not implemented: put a proxy in front",
    ("--ssl-cert-file",): "TLS is not implemented: put a proxy in front", ("-mv", "--model-vocoder"): "text-to-speech is not implemented",
    ("--diffusion-steps",): "diffusion language models are not implemented",
    ("-ctk", "--cache-type-k"): None, ("-ctv", "--cache-type-v"): None,       # f16 only: checked in `finish_compat`
}
COMPAT_WIRED = [
    Opt(("-t", "--threads"), "CPU threads (the CPU path and the host side of sampling)", None, int, "DLP_ARG_THREADS"),
    Opt(("-tb", "--threads-batch"), "accepted: the same pool serves prompt batches", None, int),
    Opt(("--override-kv",), "KEY=TYPE:VALUE (int | float | bool | str) replaces a metadata value of the model file; may be repeated", None, str, action="append"),
    Opt(("--rope-scaling",), "none | linear | yarn: overrides the model's rope scaling type", None, str, choices=("none", "linear", "yarn")),
    Opt(("--rope-scale",), "context scaling factor (1 / rope-freq-scale)", None, float),
    Opt(("--rope-freq-base",), "rotary base frequency", None, float),
    Opt(("--rope-freq-scale",), "rotary frequency scale (linear scaling)", None, float),
    Opt(("--yarn-orig-ctx",), "YaRN: original context length", None, int),
    Opt(("--yarn-ext-factor",), "YaRN: extrapolation mix factor", None, float),
    Opt(("--yarn-attn-factor",), "YaRN: attention magnitude factor", None, float),
    Opt(("--yarn-beta-slow",), "YaRN: high correction dimension", None, float),
    Opt(("--yarn-beta-fast",), "YaRN: low correction dimension", None, float),
]


def add_compat(ap: argparse.ArgumentParser, wired: bool = True):
    """Adds the reference's remaining flags to `ap` (see COMPAT_* above); call `finish_compat(ap, args)` after parsing."""
    have = {s for act in ap._actions for s in act.option_strings}
    if wired:
        for o in COMPAT_WIRED:
            names = [n for n in o.names if n not in have]
            if names:
                kw = dict(help=o.help, default=(o.type(os.environ[o.env]) if o.env and o.env in os.environ and o.type else o.default))
                if o.action:
                    kw["action"] = o.action
                else:
                    kw["type"] = o.type
                    if o.choices:
                        kw["choices"] = o.choices
                ap.add_argument(*names, **kw)
                have.update(names)
    g = ap.add_argument_group("accepted for command lines written for llama.cpp (no effect here, or refused with the reason)")
    for table in (COMPAT_IGNORED, COMPAT_REFUSED):
        for names, spec in table.items():
            names = [n for n in names if n not in have]
            if not names:
                continue
            dest = "compat__" + names[-1].lstrip("-").replace("-", "_")
            nargs = spec if table is COMPAT_IGNORED else 1
            if nargs == 0:
                g.add_argument(*names, dest=dest, action="store_true", default=None, help=argparse.SUPPRESS)
            else:
                kw = dict(nargs="?", const=True) if nargs == "?" else {}
                g.add_argument(*names, dest=dest, default=None, help=argparse.SUPPRESS, **kw)
            have.update(names)
    return ap


def finish_compat(ap: argparse.ArgumentParser, a, log=None):
    """Reports ignored flags, refuses unsupported ones, applies the wired ones that act on the process (threads, metadata overrides)."""
    import sys
    log = log or (lambda s: sys.stderr.write(s + "\n"))
    for table in (COMPAT_IGNORED, COMPAT_REFUSED):
        for names, spec in table.items():
            dest = "compat__" + names[-1].lstrip("-").replace("-", "_")
            v = getattr(a, dest, None)
            if v is None:
                continue
            if table is COMPAT_REFUSED:
                if spec is None:        # KV cache type
                    if str(v).lower() not in ("f16", "fp16"):
                        ap.error(f"{names[-1]} {v}: the paged KV cache holds fp16 only")
                    continue
                ap.error(f"{names[-1]}: {spec}")
            log(f"note: {names[-1]} is accepted for compatibility and has no effect in this engine")
    if getattr(a, "threads", None):
        import torch
        torch.set_num_threads(max(1, int(a.threads)))
        os.environ.setdefault("OMP_NUM_THREADS", str(int(a.threads)))
    ov = list(getattr(a, "override_kv", None) or [])
    rope = {"rope_scaling": "rope.scaling.type=str:{}", "rope_freq_base": "rope.freq_base=float:{}", "yarn_orig_ctx": "rope.scaling.original_context_length=int:{}",
            "yarn_ext_factor": "rope.scaling.yarn_ext_factor=float:{}", "yarn_attn_factor": "rope.scaling.attn_factor=float:{}",
            "yarn_beta_slow": "rope.scaling.yarn_beta_slow=float:{}", "yarn_beta_fast": "rope.scaling.yarn_beta_fast=float:{}"}
    for attr, fmt in rope.items():
        if getattr(a, attr, None) is not None:
            ov.append("{arch}." + fmt.format(getattr(a, attr)))
    scale = getattr(a, "rope_scale", None) or (1.0 / a.rope_freq_scale if getattr(a, "rope_freq_scale", None) else None)
    if scale is not None:
        ov.append("{arch}.rope.scaling.factor=float:" + repr(float(scale)))
        if getattr(a, "rope_scaling", None) is None:
            ov.append("{arch}.rope.scaling.type=str:linear")
    if ov:
        os.environ["DLP_OVERRIDE_KV"] = ";".join(([os.environ["DLP_OVERRIDE_KV"]] if os.environ.get("DLP_OVERRIDE_KV") else []) + ov)
    return a


def parse_override_kv(spec: str, arch: str = "llama"):
    """`KEY=TYPE:VALUE` -> (key, value); `{arch}` in the key stands for the file's architecture (common/arg.cpp `--override-kv`)."""
    key, _, rest = spec.partition("=")
    kind, _, val = rest.partition(":")
    key = key.strip().replace("{arch}", arch)
    if kind == "int":
        return key, int(val)
    if kind == "float":
        return key, float(val)
    if kind == "bool":
        return key, val.strip().lower() in ("1", "true", "yes", "on")
    if kind == "str":
        return key, val
    raise ValueError(f"--override-kv {spec!r}: the type must be int, float, bool or str")
