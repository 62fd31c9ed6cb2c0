"""Files this repository WRITES, loaded by the reference: a LoRA adapter (`write_lora_gguf`, the format `dlp_finetune --lora` and
`convert_lora_to_gguf.py` emit), a control vector (`tools/dlp_cvector.py`) and checkpoints with YaRN / linear rope scaling.  The unmodified
`llama-perplexity` applies them with its own loaders (`--lora`, `--control-vector-scaled`, rope keys from the GGUF) and the whole-vocabulary
log-probabilities must match the fp32 oracle applying the same thing.  (llama-adapter.cpp; llama-context.cpp:82-132 YaRN set-up)"""
import dataclasses
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from baseline import reference as ref_build  # noqa: E402
from test_oracle_vs_ggml import N_CTX, read_logits_file  # noqa: E402

from distributed_llm_pipeline_b200.gguf import open_gguf, write_synthetic_gguf  # noqa: E402
from distributed_llm_pipeline_b200.models.config import preset  # noqa: E402
from distributed_llm_pipeline_b200.models.reference import ReferenceModel  # noqa: E402
from distributed_llm_pipeline_b200.models.weights import write_lora_gguf  # noqa: E402


@pytest.fixture(scope="module")
def ref(tmp_path_factory):
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-perplexity"):
        pytest.skip(f"reference binaries unavailable: {why}")
    scratch = str(tmp_path_factory.mktemp("adapters"))
    env = ref_build.cpu_env(scratch)
    rnd = np.random.default_rng(0)
    words = ["alpha", "beta", "gamma", "delta", "the", "of", "and", "to", "in", "is", "it", "pipe", "line", "gpu", "token", "model"]
    text = os.path.join(scratch, "text.txt")
    open(text, "w").write(" ".join(words[i] for i in rnd.integers(0, len(words), 1500)))

    def run(model, *extra, n_ctx=N_CTX):
        out = os.path.join(scratch, f"o{abs(hash((model,) + extra))}.logits")
        p = subprocess.run([ref_build.exe("llama-perplexity"), "-m", model, "-f", text, "-c", str(n_ctx), "-b", str(n_ctx), "--chunks", "2", "-ngl", "0", "-t", "4",
                            "--save-all-logits", out, *extra], env=env, capture_output=True, text=True, timeout=600)
        assert p.returncode == 0 and os.path.exists(out), (p.stderr or p.stdout)[-800:]
        return read_logits_file(out)
    return scratch, run


def _oracle_logprobs(model: ReferenceModel, path: str, toks, n_ctx, hook=None):
    md = open_gguf(path).metadata
    out = []
    for c in range(toks.shape[0]):
        t = toks[c].tolist()
        if md.get("tokenizer.ggml.add_bos_token"):
            t[0] = int(md["tokenizer.ggml.bos_token_id"])
        logits = hook(t) if hook else model.forward(torch.tensor(t), model.new_cache(n_ctx))
        out.append(torch.log_softmax(logits.float(), -1)[n_ctx // 2:n_ctx - 1].numpy())
    return np.stack(out)


def test_lora_adapter_file_is_applied_the_same_way(ref):
    scratch, run = ref
    path = write_synthetic_gguf(os.path.join(scratch, "base.gguf"), preset("tiny"), "F16", seed=1)
    model = ReferenceModel.from_gguf(path)
    rng = np.random.default_rng(3)
    r, alpha, deltas = 4, 8.0, {}
    for il in range(model.cfg.n_layer):
        for name in ("attn_q", "attn_v", "attn_output", "ffn_gate", "ffn_down"):
            w = model.w[f"blk.{il}.{name}.weight"]
            deltas[f"blk.{il}.{name}.weight"] = ((rng.standard_normal((r, w.shape[1])) * 0.05).astype(np.float32), (rng.standard_normal((w.shape[0], r)) * 0.05).astype(np.float32))
    ad = write_lora_gguf(os.path.join(scratch, "lora.gguf"), deltas, alpha, "llama")
    n_ctx, toks, plain = run(path)
    _, toks2, with_lora = run(path, "--lora", ad)
    assert np.array_equal(toks, toks2) and np.abs(plain - with_lora).max() > 0.05        # the adapter does something in the reference
    for n, (a, b) in deltas.items():
        model.w[n] = model.w[n] + (alpha / r) * torch.from_numpy(b @ a)
    mine = _oracle_logprobs(model, path, toks, n_ctx)
    assert np.abs(mine - with_lora).max() < 6e-3 and np.abs(mine - with_lora).mean() < 1e-3
    # half strength through --lora-scaled
    _, _, half = run(path, "--lora-scaled", f"{ad}:0.5")
    for n, (a, b) in deltas.items():
        model.w[n] = model.w[n] - 0.5 * (alpha / r) * torch.from_numpy(b @ a)
    assert np.abs(_oracle_logprobs(model, path, toks, n_ctx) - half).max() < 6e-3


def test_control_vector_file_is_applied_the_same_way(ref):
    scratch, run = ref
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import dlp_cvector
    cfg = preset("tiny-smollm3")                      # 4 layers -> directions for layers 1..3
    path = write_synthetic_gguf(os.path.join(scratch, "cv-base.gguf"), cfg, "F16", seed=1)
    model = ReferenceModel.from_gguf(path)
    g = torch.Generator().manual_seed(5)
    dirs = [torch.nn.functional.normalize(torch.randn(cfg.n_embd, generator=g), dim=0) for _ in range(cfg.n_layer - 1)]
    cv = dlp_cvector.save(os.path.join(scratch, "cv.gguf"), dirs, "smollm3")
    n_ctx, toks, plain = run(path)
    _, _, steered = run(path, "--control-vector-scaled", f"{cv}:3.0")
    assert np.abs(plain - steered).max() > 0.05

    def hook(t):       # h += 3 * direction[il] after layer il, for il = 1 .. n_layer - 1 (llama-adapter.cpp: tensors[il] on l_out of layer il)
        cache = model.new_cache(n_ctx)
        pos = torch.arange(len(t))
        h = model.embed(torch.tensor(t), pos)
        for il in range(cfg.n_layer):
            h = model.layer(il, h, pos, cache)
            if 1 <= il <= len(dirs):
                h = h + 3.0 * dirs[il - 1]
        return model.head(h)
    mine = _oracle_logprobs(model, path, toks, n_ctx, hook)
    assert np.abs(mine - steered).max() < 6e-3 and np.abs(mine - steered).mean() < 1e-3


@pytest.mark.parametrize("kind", ["linear", "yarn", "rope_freqs"])
def test_rope_scaling_keys_mean_the_same(ref, kind):
    """Context extension written into the file: linear position scaling, YaRN (interpolation ramp + attention magnitude), and the
    per-frequency divisors of Llama-3.1 (`rope_freqs.weight`), evaluated beyond the original context."""
    scratch, run = ref
    if kind == "linear":
        cfg = preset("tiny", n_ctx_train=32, rope_scale_linear=0.5)
    elif kind == "yarn":
        cfg = preset("tiny", n_ctx_train=128, yarn_factor=4.0, yarn_orig_ctx=32)
    else:
        cfg = preset("tiny", n_ctx_train=128)
    path = write_synthetic_gguf(os.path.join(scratch, f"rope-{kind}.gguf"), cfg, "F16", seed=1, rope_freqs=kind == "rope_freqs")
    if kind == "rope_freqs":      # make the divisors non-trivial: rewrite the tensor with a ramp 1 .. 8
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter
        f = open_gguf(path)
        w = GGUFWriter(path + ".2")
        for k, v in f.metadata.items():
            w.add(k, v, f.files[0].metadata_types[k]) if not isinstance(v, (list, np.ndarray)) else w.add(k, v)
        ramp = np.linspace(1.0, 8.0, cfg.rope_dim // 2).astype("<f4")
        for n, t in f.tensors.items():
            w.add_tensor(n, t.shape, t.ggml_type, (lambda: ramp.view(np.uint8).reshape(-1)) if n == "rope_freqs.weight" else (lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1)))
        w.write()
        path = path + ".2"
    model = ReferenceModel.from_gguf(path)
    n_ctx, toks, ref_lp = run(path)
    mine = _oracle_logprobs(model, path, toks, n_ctx)
    assert np.abs(mine - ref_lp).max() < 6e-3 and np.abs(mine - ref_lp).mean() < 1e-3
    plain = ReferenceModel.from_gguf(write_synthetic_gguf(os.path.join(scratch, "rope-none.gguf"), preset("tiny"), "F16", seed=1))
    assert np.abs(_oracle_logprobs(plain, path, toks, n_ctx) - ref_lp).max() > 0.02        # the scaling is not a no-op at these positions
