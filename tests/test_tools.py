"""Aux tools: quantize, gguf-split/merge, tokenize, perplexity, bench, cli (CPU, tiny model)."""
import os
import subprocess
import sys

import numpy as np

from distributed_llm_pipeline_b200.gguf import GGMLType, open_gguf

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(tool, *args):
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", tool), *args], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return r.stdout


def test_quantize_tool(tmp_models, tmp_path):
    dst = str(tmp_path / "q.gguf")
    _run("dlp_quantize.py", tmp_models[("tiny", "F16")], dst, "Q4_K_M")
    f = open_gguf(dst)
    assert f.tensors["blk.0.ffn_gate.weight"].ggml_type == GGMLType.Q4_K
    assert f.tensors["output.weight"].ggml_type == GGMLType.Q6_K
    assert f.tensors["blk.0.attn_norm.weight"].ggml_type == GGMLType.F32
    src = open_gguf(tmp_models[("tiny", "F16")])
    a, b = src.array("blk.0.ffn_up.weight"), f.array("blk.0.ffn_up.weight")
    assert ((a - b) ** 2).sum() / (a ** 2).sum() < 1e-2
    assert f.metadata["tokenizer.ggml.tokens"] == src.metadata["tokenizer.ggml.tokens"]


def test_gguf_split_and_merge(tmp_models, tmp_path):
    src = tmp_models[("tiny", "Q8_0")]
    out = _run("dlp_gguf_split.py", "--split", "--split-max-tensors", "8", src, str(tmp_path / "shard")).split()
    assert len(out) == 3 and out[0].endswith("-00001-of-00003.gguf")
    whole, parts = open_gguf(src), open_gguf(out[0])
    assert set(whole.tensors) == set(parts.tensors)
    for n in whole.tensors:
        assert np.array_equal(whole.raw(n), parts.raw(n))
    merged = str(tmp_path / "merged.gguf")
    _run("dlp_gguf_split.py", "--merge", out[0], merged)
    m = open_gguf(merged)
    assert set(m.tensors) == set(whole.tensors) and "split.no" not in m.metadata


def test_tokenize_perplexity_bench_cli(tmp_models):
    model = tmp_models[("tiny", "F16")]
    out = _run("dlp_tokenize.py", "-m", model, "-p", "hello world", "--ids")
    assert out.strip().startswith("[1,")
    out = _run("dlp_perplexity.py", "-m", model, "--device", "cpu", "-c", "32", "--chunks", "1")
    ppl = float(out.strip().split("PPL = ")[-1].split()[0])
    assert 50 < ppl < 5000                 # random-init model: PPL of the order of the vocabulary size (512)
    out = _run("dlp_bench.py", "-m", model, "--device", "cpu", "-p", "16", "-n", "4", "-r", "1", "-o", "json")
    assert '"test": "pp16"' in out and '"test": "tg4"' in out
    out = _run("dlp_cli.py", "-m", model, "--device", "cpu", "-p", "hi", "-n", "4", "-st", "--temp", "0")
    assert len(out) > 0


def test_fit_params_and_list_devices(tmp_models):
    from distributed_llm_pipeline_b200.models.config import config_from_gguf
    from distributed_llm_pipeline_b200.utils.fit import fit_params, list_devices, model_bytes
    f = open_gguf(tmp_models[("tiny", "F16")])
    cfg = config_from_gguf(f.metadata, f.tensors)
    mb = model_bytes(f)
    kv_tok = 2 * cfg.n_layer * cfg.kv_dim * 2
    big = fit_params(f, cfg, 512, 4, 10 << 30)
    assert big.fits and (big.n_ctx, big.n_slots) == (512, 4)
    tight = fit_params(f, cfg, 4096, 4, int((mb + (64 << 20) + kv_tok * 1024) / 0.92))
    assert tight.fits and tight.n_ctx * tight.n_slots * kv_tok <= kv_tok * 1100 and tight.n_ctx % 64 == 0 and tight.n_slots >= 1
    none = fit_params(f, cfg, 512, 1, mb // 2)
    assert not none.fits and none.min_tp >= 2 and "--tp" in none.note
    assert any(d.startswith("CPU") for d in list_devices())
    out = _run("dlp_cli.py", "--list-devices")
    assert "Available devices" in out


def test_serve_bench_against_orchestrator(tmp_models):
    """The load generator (scripts/server-bench.py counterpart) drives /completion (token-array prompts, streaming) and /chat."""
    import asyncio
    import json
    import socket
    import subprocess
    import sys
    import threading
    import time
    from distributed_llm_pipeline_b200.serving.orchestrator import Orchestrator, build_service
    sched, info, logs = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=128, n_slots=4, n_batch=32)
    sched.start()
    orch = Orchestrator(sched, info, n_predict=6, startup_logs=logs)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    loop = asyncio.new_event_loop()
    threading.Thread(target=lambda: loop.run_until_complete(orch.serve("127.0.0.1", port)), daemon=True).start()
    for _ in range(100):
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.2).close()
            break
        except OSError:
            time.sleep(0.05)
    try:
        for route in ("/completion", "/chat"):
            r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "serve_bench.py"), "--port", str(port), "--route", route, "-c", "1,3",
                                "--prompt-tokens", "12", "-n", "6", "--vocab", "500"], capture_output=True, text=True, timeout=120)
            assert r.returncode == 0, r.stderr
            rows = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
            assert [x["concurrency"] for x in rows] == [1, 3]
            assert all(x["ok"] == x["concurrency"] and x["agg_tok_s"] > 0 and x["ttft_ms_p50"] > 0 for x in rows), rows
    finally:
        sched.stop()


def test_imatrix_tool_and_weighted_quantize(tmp_models, tmp_path):
    """`dlp_imatrix.py` (llama-imatrix): the statistic, the GGUF layout the reference writes, and its use by `dlp_quantize --imatrix`."""
    import torch
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel, rms_norm
    model = tmp_models[("tiny-moe", "F16")]
    imat = str(tmp_path / "imatrix.gguf")
    out = _run("dlp_imatrix.py", "-m", model, "-c", "32", "--chunks", "3", "-o", imat)
    assert "3 chunks of 32 tokens" in out
    f = open_gguf(imat)
    assert f.metadata["general.type"] == "imatrix" and f.metadata["imatrix.chunk_count"] == 3 and f.metadata["imatrix.chunk_size"] == 32
    ref = ReferenceModel.from_gguf(model)
    c = ref.cfg
    # dense matrices: one row, 96 tokens each; routed experts: one row per expert, n_expert_used visits per token in total
    assert f.array("blk.0.attn_q.weight.in_sum2").shape == (1, c.n_embd) and float(f.array("blk.0.attn_q.weight.counts")[0, 0]) == 96
    cn = f.array("blk.1.ffn_down_exps.weight.counts")
    assert cn.shape == (c.n_expert, 1) and float(cn.sum()) == 96 * c.n_expert_used
    assert f.array("blk.1.ffn_down_exps.weight.in_sum2").shape == (c.n_expert, c.n_ff)
    # the statistic itself, recomputed by hand for the first layer's Q projection: sum over tokens of rms_norm(embed)^2
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(3, c.vocab_size, (32 * 3,), generator=g)
    xn = rms_norm(ref.embed(ids), ref.w["blk.0.attn_norm.weight"], c.rms_eps)
    assert np.allclose(f.array("blk.0.attn_q.weight.in_sum2")[0], (xn * xn).sum(0).numpy(), rtol=1e-4)
    # the quantizer weighs its error by it: per-column error weighted by the mean squared activation goes down
    plain, weighted = str(tmp_path / "plain.gguf"), str(tmp_path / "weighted.gguf")
    _run("dlp_quantize.py", model, plain, "Q4_K_M")
    log = _run("dlp_quantize.py", model, weighted, "Q4_K_M", "--imatrix", imat)
    assert "tensors weighted" in log
    src, fp, fw = open_gguf(model), open_gguf(plain), open_gguf(weighted)
    better = 0
    for name in ("blk.0.attn_q.weight", "blk.0.attn_output.weight", "blk.1.attn_k.weight"):
        imp = f.array(name + ".in_sum2")[0] / 96.0
        a = src.array(name)
        e_plain, e_w = (imp * (fp.array(name) - a) ** 2).sum(), (imp * (fw.array(name) - a) ** 2).sum()
        assert e_w < e_plain * 1.02, (name, e_plain, e_w)
        better += e_w < 0.95 * e_plain
    assert better >= 2
    assert fw.tensors["blk.0.ffn_gate_exps.weight"].ggml_type == GGMLType.Q4_K
    # the weighted file is still a model: it loads and tracks the F16 logits at least as well as a 4-bit file should
    lw = ReferenceModel.from_gguf(weighted).forward(ids[:32], ReferenceModel.from_gguf(weighted).new_cache(32))
    l16 = ref.forward(ids[:32], ref.new_cache(32))
    assert float(((lw - l16) ** 2).sum() / (l16 ** 2).sum()) < 0.05
