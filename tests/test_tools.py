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
    md = open_gguf(model).metadata
    if md.get("tokenizer.ggml.add_bos_token", md.get("tokenizer.ggml.model", "llama") == "llama"):
        ids[0::32] = int(md.get("tokenizer.ggml.bos_token_id", 1))          # every window starts with BOS, as in the reference tool
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
        assert e_w < e_plain, (name, e_plain, e_w)      # (the unweighted fit already searches its scales: where the importance is nearly flat,
        better += e_w < 0.95 * e_plain                  # as for Q / K behind the first norm of a random model, the gain is a percent or two)
    assert better >= 1
    assert fw.tensors["blk.0.ffn_gate_exps.weight"].ggml_type == GGMLType.Q4_K
    # the weighted file is still a model: it loads and tracks the F16 logits at least as well as a 4-bit file should
    lw = ReferenceModel.from_gguf(weighted).forward(ids[:32], ReferenceModel.from_gguf(weighted).new_cache(32))
    l16 = ref.forward(ids[:32], ref.new_cache(32))
    assert float(((lw - l16) ** 2).sum() / (l16 ** 2).sum()) < 0.05


def test_cvector_generator(tmp_models, tmp_path):
    """`dlp_cvector.py` (llama-cvector-generator): layer-output differences -> unit directions -> a file the engine applies."""
    import torch
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    model = tmp_models[("tiny-smollm3", "F16")]            # 4 layers -> 3 directions
    pos, neg = tmp_path / "pos.txt", tmp_path / "neg.txt"
    pos.write_text("I am very happy today\nthis is wonderful news\n")
    neg.write_text("I am very sad today\nthis is terrible news\n")
    out = str(tmp_path / "cv.gguf")
    for method in ("mean", "pca"):
        log = _run("dlp_cvector.py", "-m", model, "--positive-file", str(pos), "--negative-file", str(neg), "-o", out, "--method", method, "--device", "cpu")
        assert "3 directions of 256" in log
        f = open_gguf(out)
        assert f.metadata["general.architecture"] == "controlvector" and f.metadata["controlvector.layer_count"] == 3
        assert f.metadata["controlvector.model_hint"] == "smollm3" and sorted(f.tensors) == ["direction.1", "direction.2", "direction.3"]
        for n in f.tensors:
            assert abs(float(np.linalg.norm(f.array(n))) - 1.0) < 1e-4
    # "mean", recomputed with the oracle for the first layer
    _run("dlp_cvector.py", "-m", model, "--positive-file", str(pos), "--negative-file", str(neg), "-o", out, "--method", "mean", "--device", "cpu")
    ref, tok = ReferenceModel.from_gguf(model), Tokenizer.from_gguf(model)
    rows = []
    for p, n in zip(pos.read_text().splitlines(), neg.read_text().splitlines()):
        a, b = tok.encode(p), tok.encode(n)
        T = max(len(a), len(b))
        padt = tok.encode(" ", add_bos=False)[-1:]
        a, b = a + padt * (T - len(a)), b + padt * (T - len(b))
        ha = ref.forward(a, ref.new_cache(64), layers=range(0, 1), return_hidden=True)
        hb = ref.forward(b, ref.new_cache(64), layers=range(0, 1), return_hidden=True)
        d = ha - hb
        rows.append(d[d.abs().sum(-1) > 0])
    want = torch.cat(rows).mean(0)
    want = want / want.norm()
    assert np.allclose(open_gguf(out).array("direction.1"), want.numpy(), atol=2e-4)
    # applying the file moves the residual stream along the direction: the projection of layer 1's output grows with the scale
    eng = LlamaEngine(load_weights(model, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=32)
    ids = tok.encode("this is news")
    d1 = torch.from_numpy(np.array(open_gguf(out).array("direction.1")))

    def proj(scale):
        eng.set_control_vector(out if scale else None, scale or 1.0)
        got = {}
        eng.eval_callback = lambda name, il, t: got.__setitem__(il, t.float().clone())
        sid = eng.alloc_seq()
        eng.prefill(sid, ids)
        eng.free_seq(sid)
        eng.eval_callback = None
        return float((got[1] @ d1).mean())
    base, plus = proj(0), proj(4.0)
    assert abs((plus - base) - 4.0) < 1e-3          # unit direction x scale, added after layer 1


def test_finetune_full_and_lora(tmp_models, tmp_path):
    """`dlp_finetune.py` (examples/training/finetune.cpp over ggml-opt): loss falls, the checkpoint / adapter it writes load in the engine."""
    import random
    import torch
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.reference import ReferenceModel
    from distributed_llm_pipeline_b200.models.weights import LoraAdapter, load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    model = tmp_models[("tiny", "F16")]
    r = random.Random(1)
    phrases = ["the cat sat on the mat", "a dog ran in the park", "the sun is hot today"]
    text = tmp_path / "train.txt"
    text.write_text(". ".join(r.choice(phrases) for _ in range(120)))
    ids = Tokenizer.from_gguf(model).encode(text.read_text())[:33]

    def nll(logits):
        return float(torch.nn.functional.cross_entropy(logits.float(), torch.tensor(ids[1:])))
    base = ReferenceModel.from_gguf(model)
    nll0 = nll(base.forward(ids[:-1], base.new_cache(64)))
    # all weights
    out = str(tmp_path / "ft.gguf")
    log = _run("dlp_finetune.py", "-m", model, "-f", str(text), "-o", out, "-c", "32", "--epochs", "1", "-b", "4", "--lr", "2e-3", "--outtype", "F32")
    assert "train loss" in log and "held out" in log
    ft = ReferenceModel.from_gguf(out)
    assert ft.cfg == base.cfg
    nll1 = nll(ft.forward(ids[:-1], ft.new_cache(64)))
    assert nll1 < 0.5 * nll0, (nll0, nll1)
    eng = LlamaEngine(load_weights(out, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=64)
    lg = eng.prefill(eng.alloc_seq(), ids[:-1], all_logits=True)
    assert abs(nll(lg) - nll1) < 1e-3
    # adapter only
    ad = str(tmp_path / "lora.gguf")
    _run("dlp_finetune.py", "-m", model, "-f", str(text), "-o", ad, "-c", "32", "--epochs", "2", "-b", "4", "--lr", "1e-2", "--lora", "4")
    f = open_gguf(ad)
    assert f.metadata["general.type"] == "adapter" and f.metadata["adapter.lora.alpha"] == 8.0
    assert f.array("blk.0.attn_q.weight.lora_a").shape == (4, base.cfg.n_embd) and "blk.1.ffn_down.weight.lora_b" in f.tensors
    eng2 = LlamaEngine(load_weights(model, "cpu", lora=[LoraAdapter(ad)]), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=64)
    nll2 = nll(eng2.prefill(eng2.alloc_seq(), ids[:-1], all_logits=True))
    assert nll2 < 0.8 * nll0, (nll0, nll2)


def test_gguf_dump_hash_and_metadata_edit(tmp_models, tmp_path):
    """`dlp_gguf.py`: gguf_dump / gguf_hash / gguf_set_metadata / gguf_new_metadata."""
    import json
    import shutil
    src = tmp_models[("tiny", "Q8_0")]
    out = _run("dlp_gguf.py", "dump", src)
    assert "general.architecture = 'llama'" in out and "blk.0.attn_q.weight" in out and "Q8_0" in out
    d = json.loads(_run("dlp_gguf.py", "dump", src, "--json"))
    assert d["metadata"]["llama.block_count"] == 2 and d["tensors"]["output_norm.weight"]["type"] == "F32"
    a = str(tmp_path / "a.gguf")
    shutil.copy(src, a)
    h1 = _run("dlp_gguf.py", "hash", "-q", a).split()[1]
    # in place: same-size values only
    _run("dlp_gguf.py", "set-metadata", a, "llama.context_length", "4096")
    assert open_gguf(a).metadata["llama.context_length"] == 4096
    name = open_gguf(a).metadata["general.name"]
    _run("dlp_gguf.py", "set-metadata", a, "general.name", name[::-1])
    assert open_gguf(a).metadata["general.name"] == name[::-1]
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "dlp_gguf.py"), "set-metadata", a, "general.name", "longer-than-before-xyz"], capture_output=True, text=True)
    assert r.returncode != 0 and "new-metadata" in r.stderr
    # rewrite: add / replace / remove keys, tensor bytes untouched
    b = str(tmp_path / "b.gguf")
    _run("dlp_gguf.py", "new-metadata", a, b, "--set", "general.name=renamed model", "--set", "general.description=test file", "--set", "llama.rope.freq_base=500000.0",
         "--set", "custom.answer:uint32=42", "--remove", "tokenizer.ggml.pre")
    fb = open_gguf(b)
    assert fb.metadata["general.name"] == "renamed model" and fb.metadata["general.description"] == "test file" and fb.metadata["custom.answer"] == 42
    assert abs(fb.metadata["llama.rope.freq_base"] - 500000.0) < 1 and "tokenizer.ggml.pre" not in fb.metadata
    assert _run("dlp_gguf.py", "hash", "-q", b).split()[1] == h1          # metadata edits do not touch the weights
    assert "same tensor data" in _run("dlp_gguf.py", "hash", "-q", b, "--check", src)
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "dlp_gguf.py"), "hash", "-q", b, "--check", tmp_models[("tiny", "F16")]], capture_output=True, text=True)
    assert r.returncode == 1 and "DIFFERENT" in r.stdout


def test_perplexity_tool_matches_llama_perplexity(tmp_models, tmp_path):
    """`dlp_perplexity.py` against the reference tool on the same file and text: same chunking (second half of every n_ctx window), same number."""
    import re
    sys.path.insert(0, REPO)
    from baseline import reference as ref_build
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-perplexity"):
        import pytest
        pytest.skip(f"reference binaries unavailable: {why}")
    rnd = np.random.default_rng(2)
    words = ["alpha", "beta", "gamma", "delta", "the", "of", "and", "to", "in", "is"]
    text = tmp_path / "t.txt"
    text.write_text(" ".join(words[i] for i in rnd.integers(0, len(words), 1200)))
    model = tmp_models[("tiny", "Q8_0")]
    p = subprocess.run([ref_build.exe("llama-perplexity"), "-m", model, "-f", str(text), "-c", "64", "-b", "64", "--chunks", "3", "-ngl", "0", "-t", "4"],
                       env=ref_build.cpu_env(str(tmp_path)), capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-500:]
    ref_ppl = float(re.search(r"Final estimate: PPL = ([0-9.]+)", p.stdout + p.stderr).group(1))
    out = _run("dlp_perplexity.py", "-m", model, "-f", str(text), "-c", "64", "--chunks", "3", "--device", "cpu")
    mine = float(re.search(r"Final estimate: PPL = ([0-9.]+)", out).group(1))
    assert abs(mine - ref_ppl) / ref_ppl < 0.02, (mine, ref_ppl)
    # the same running line ("[1]ppl,[2]ppl,...") and the same error estimate behind "+/-"
    ref_err = float(re.search(r"Final estimate: PPL = [0-9.]+ \+/- ([0-9.]+)", p.stdout + p.stderr).group(1))
    my_err = float(re.search(r"Final estimate: PPL = [0-9.]+ \+/- ([0-9.]+)", out).group(1))
    assert abs(my_err - ref_err) / ref_err < 0.05, (my_err, ref_err)
    run_ref, run_mine = re.findall(r"\[(\d+)\]([0-9.]+),", p.stdout), re.findall(r"\[(\d+)\]([0-9.]+),", out)
    assert [i for i, _ in run_ref] == [i for i, _ in run_mine] == ["1", "2", "3"]
    assert all(abs(float(a) - float(b)) / float(b) < 0.02 for (_, a), (_, b) in zip(run_mine, run_ref))


def test_imatrix_files_combine_by_addition(tmp_models, tmp_path):
    """`dlp_imatrix.py --chunk N` (start further into the text) and `--in-file a --in-file b` (combine): two halves of a text measured apart and
    merged equal one run over the whole text; `--show-statistics` prints one line per matrix."""
    model = tmp_models[("tiny", "F16")]
    rnd = np.random.default_rng(3)
    words = ["alpha", "beta", "gamma", "delta", "the", "of", "and", "to", "in", "is"]
    text = tmp_path / "t.txt"
    text.write_text(" ".join(words[i] for i in rnd.integers(0, len(words), 1500)))
    a, b, ab, whole = (str(tmp_path / n) for n in ("a.gguf", "b.gguf", "ab.gguf", "whole.gguf"))
    _run("dlp_imatrix.py", "-m", model, "-f", str(text), "-c", "64", "--chunks", "2", "-o", a)
    _run("dlp_imatrix.py", "-m", model, "-f", str(text), "-c", "64", "--chunks", "2", "--chunk", "2", "-o", b)
    _run("dlp_imatrix.py", "-m", model, "--in-file", a, "--in-file", b, "-o", ab)
    _run("dlp_imatrix.py", "-m", model, "-f", str(text), "-c", "64", "--chunks", "4", "-o", whole)
    fa, fw = open_gguf(ab), open_gguf(whole)
    assert fa.metadata["imatrix.chunk_count"] == fw.metadata["imatrix.chunk_count"] == 4 and set(fa.names()) == set(fw.names())
    for n in fw.names():
        assert np.allclose(fa.array(n), fw.array(n), rtol=1e-5, atol=1e-6), n
    out = _run("dlp_imatrix.py", "-m", model, "--in-file", whole, "--show-statistics")
    assert "blk.0.attn_q.weight" in out and "entropy" in out
