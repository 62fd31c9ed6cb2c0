"""Tool interoperability with the reference's own command-line tools (CPU builds of `llama-quantize`, `llama-imatrix`, `llama-gguf-split`,
`llama-export-lora`; looked for in `baseline/_ref/bin_cpu`, `$DLP_REF_CPU_TOOLS` or `/tmp/refcpu/bin` - build them with
`tools/build_reference.sh cpu-tools`; the tests skip where they are absent).  Files move in both directions: what one side writes the other
side reads, and the numbers agree."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

from distributed_llm_pipeline_b200.gguf import GGMLType, open_gguf  # noqa: E402


def tool(name: str) -> str:
    for d in (os.environ.get("DLP_REF_CPU_TOOLS"), os.path.join(ROOT, "baseline", "_ref", "bin_cpu"), "/tmp/refcpu/bin"):
        if d and os.path.exists(os.path.join(d, name)):
            return os.path.join(d, name)
    pytest.skip(f"reference tool {name} is not built here")


def run(cmd, **kw):
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, **kw)
    assert p.returncode == 0, (cmd, (p.stderr or p.stdout)[-800:])
    return p.stdout + p.stderr


def ours(script, *args):
    return run([sys.executable, os.path.join(ROOT, "tools", script), *args], env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))


@pytest.mark.parametrize("ftype", ["Q4_K_M", "Q5_K_M", "Q8_0", "Q6_K", "Q3_K", "IQ4_NL"])
def test_quantizer_picks_the_same_types_and_comparable_error(tmp_models, tmp_path, ftype):
    """Per-tensor storage types equal to `llama-quantize`'s for the same file type (the Q4_K_M / Q5_K_M mix rules included), RMS error of the
    own encoders within a few percent of the reference's, Q8_0 bytes equal up to rounding of exact ties."""
    src = tmp_models[("tiny", "F16")]
    ref_out, our_out = str(tmp_path / "ref.gguf"), str(tmp_path / "ours.gguf")
    run([tool("llama-quantize"), src, ref_out, {"Q3_K": "Q3_K_S"}.get(ftype, ftype)])
    ours("dlp_quantize.py", src, our_out, ftype)
    s, r, o = open_gguf(src), open_gguf(ref_out), open_gguf(our_out)
    assert set(r.tensors) == set(o.tensors)
    for n in s.tensors:
        tr, to = GGMLType(r.tensors[n].ggml_type), GGMLType(o.tensors[n].ggml_type)
        assert tr == to, (n, tr.name, to.name)
        if tr in (GGMLType.F32, GGMLType.F16):
            continue
        a = s.array(n)
        er, eo = float(((r.array(n) - a) ** 2).mean() ** 0.5), float(((o.array(n) - a) ** 2).mean() ** 0.5)
        assert eo < er * (1.06 if tr in (GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K, GGMLType.Q8_0) else 1.6), (n, tr.name, er, eo)
        if tr == GGMLType.Q8_0:      # same scales; a few values that sit exactly between two steps may round the other way
            assert float((np.asarray(r.raw(n)) != np.asarray(o.raw(n))).mean()) < 2e-3, n


def test_imatrix_files_travel_both_ways(tmp_models, tmp_path):
    import dlp_imatrix
    model = tmp_models[("tiny-moe", "F16")]
    text = tmp_path / "calib.txt"
    rnd = np.random.default_rng(0)
    words = ["alpha", "beta", "gamma", "delta", "the", "of", "and", "to", "in", "is", "it", "pipe", "line", "gpu", "token", "model"]
    text.write_text(" ".join(words[i] for i in rnd.integers(0, len(words), 1500)))
    ref_im, our_im = str(tmp_path / "ref_im.gguf"), str(tmp_path / "our_im.gguf")
    run([tool("llama-imatrix"), "-m", model, "-f", str(text), "-o", ref_im, "-c", "64", "--chunks", "4", "-t", "4"])
    ours("dlp_imatrix.py", "-m", model, "-f", str(text), "-o", our_im, "-c", "64", "--chunks", "4")
    fr, fo = open_gguf(ref_im), open_gguf(our_im)
    assert {k: fr.metadata[k] for k in ("general.type", "imatrix.chunk_count", "imatrix.chunk_size")} == {k: fo.metadata[k] for k in ("general.type", "imatrix.chunk_count", "imatrix.chunk_size")}
    assert set(fr.tensors) == set(fo.tensors)                       # the same matrices were measured, routers included
    for n in fr.tensors:
        a, b = np.asarray(fr.array(n)).reshape(-1), np.asarray(fo.array(n)).reshape(-1)
        if "_exps." in n or "gate_inp" in n:                       # routed experts: a near-tied routing decision may move a token between experts
            assert np.abs(a - b).max() <= 0.05 * np.abs(a).max() + 4, n
        else:
            assert np.abs(a - b).max() <= 2e-3 * np.abs(a).max(), n
    # each quantizer accepts the other side's file
    run([tool("llama-quantize"), "--imatrix", our_im, model, str(tmp_path / "q_ref.gguf"), "Q4_K_M"])
    log = ours("dlp_quantize.py", model, str(tmp_path / "q_ours.gguf"), "Q4_K_M", "--imatrix", ref_im)
    assert "tensors weighted" in log
    assert set(dlp_imatrix.load(ref_im)) == set(dlp_imatrix.load(our_im))


def test_split_files_travel_both_ways(tmp_models, tmp_path):
    import dlp_gguf
    src = tmp_models[("tiny", "Q8_0")]
    _, whole = dlp_gguf.tensor_hashes(src)
    # ours splits, the reference merges
    ours("dlp_gguf_split.py", "--split", "--split-max-tensors", "7", src, str(tmp_path / "ours"))
    shards = sorted(p for p in os.listdir(tmp_path) if p.startswith("ours-"))
    assert len(shards) >= 3
    run([tool("llama-gguf-split"), "--merge", str(tmp_path / shards[0]), str(tmp_path / "merged_by_ref.gguf")])
    assert dlp_gguf.tensor_hashes(str(tmp_path / "merged_by_ref.gguf"))[1] == whole
    # the reference splits, ours reads the set and merges it
    run([tool("llama-gguf-split"), "--split", "--split-max-tensors", "7", src, str(tmp_path / "ref")])
    first = sorted(p for p in os.listdir(tmp_path) if p.startswith("ref-"))[0]
    f = open_gguf(str(tmp_path / first))
    assert set(f.tensors) == set(open_gguf(src).tensors)
    ours("dlp_gguf_split.py", "--merge", str(tmp_path / first), str(tmp_path / "merged_by_ours.gguf"))
    assert dlp_gguf.tensor_hashes(str(tmp_path / "merged_by_ours.gguf"))[1] == whole


def test_export_lora_merges_like_the_reference(tmp_models, tmp_path):
    from distributed_llm_pipeline_b200.models.weights import write_lora_gguf
    base = tmp_models[("tiny", "F16")]
    s = open_gguf(base)
    rng = np.random.default_rng(1)
    deltas = {}
    for n in ("blk.0.attn_q.weight", "blk.1.ffn_down.weight", "blk.1.attn_v.weight"):
        shape = s.tensors[n].shape
        deltas[n] = ((rng.standard_normal((4, shape[1])) * 0.05).astype(np.float32), (rng.standard_normal((shape[0], 4)) * 0.05).astype(np.float32))
    ad = write_lora_gguf(str(tmp_path / "ad.gguf"), deltas, 8.0, "llama")
    run([tool("llama-export-lora"), "-m", base, "--lora", ad, "-o", str(tmp_path / "ref_merged.gguf")])
    ours("dlp_export_lora.py", "-m", base, "--lora", ad, "-o", str(tmp_path / "our_merged.gguf"))
    r, o = open_gguf(str(tmp_path / "ref_merged.gguf")), open_gguf(str(tmp_path / "our_merged.gguf"))
    for n in s.tensors:
        a, b = r.array(n), o.array(n)
        assert np.abs(a - b).max() < 2e-3, n
        if n in deltas:
            assert np.abs(a - s.array(n)).max() > 1e-2           # the adapter really was merged


@pytest.mark.parametrize("method", ["mean", "pca"])
def test_cvector_generator_finds_the_same_directions(tmp_path, method):
    """Same prompt pairs, same model: `dlp_cvector.py` and `llama-cvector-generator` write the same unit directions for `mean`; for `pca` the
    reference runs a randomly started power iteration, so the comparison is by what counts - the share of the difference energy along the
    direction: the exact SVD here must capture at least as much as the reference's vector."""
    import torch

    import dlp_cvector
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.config import preset
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    model = write_synthetic_gguf(str(tmp_path / "m.gguf"), preset("tiny-smollm3"), "F16", seed=1)
    pos = ["I am very happy today", "this is wonderful news", "what a lovely bright morning"]
    neg = ["I am very sad today", "this is terrible news", "what a gloomy dark evening"]
    (tmp_path / "pos.txt").write_text("\n".join(pos) + "\n")
    (tmp_path / "neg.txt").write_text("\n".join(neg) + "\n")
    run([tool("llama-cvector-generator"), "-m", model, "--positive-file", str(tmp_path / "pos.txt"), "--negative-file", str(tmp_path / "neg.txt"), "--method", method,
         "-o", str(tmp_path / "ref.gguf"), "-t", "2", "--pca-iter", "2000", "--pca-batch", "100"], cwd=str(tmp_path))
    ours("dlp_cvector.py", "-m", model, "--positive-file", str(tmp_path / "pos.txt"), "--negative-file", str(tmp_path / "neg.txt"), "--method", method,
         "-o", str(tmp_path / "ours.gguf"), "--device", "cpu")
    r, o = open_gguf(str(tmp_path / "ref.gguf")), open_gguf(str(tmp_path / "ours.gguf"))
    assert sorted(r.tensors) == sorted(o.tensors) == ["direction.1", "direction.2", "direction.3"]
    assert r.metadata["controlvector.model_hint"] == o.metadata["controlvector.model_hint"] and r.metadata["controlvector.layer_count"] == o.metadata["controlvector.layer_count"]
    eng = LlamaEngine(load_weights(model, "cpu"), n_ctx=64, max_seqs=1, page_size=16, max_batch_tokens=64)
    rows = dlp_cvector.difference_rows(eng, Tokenizer.from_gguf(model), pos, neg, log=lambda s: None)
    for il, n in enumerate(sorted(r.tensors)):
        a, b = torch.from_numpy(np.asarray(r.array(n)).reshape(-1).copy()), torch.from_numpy(np.asarray(o.array(n)).reshape(-1).copy())
        assert abs(float(b.norm()) - 1) < 1e-3
        a = a / a.norm()             # the reference's power iteration does not always return a unit vector
        if method == "mean":
            assert float(a @ b) > 0.99999, (n, float(a @ b))
        else:
            ea, eb = float((rows[il] @ a).pow(2).sum()), float((rows[il] @ b).pow(2).sum())
            assert eb >= 0.999 * ea, (n, ea, eb)       # ours is the optimum; the reference's randomly started iteration lands at or below it


def test_bench_output_has_the_columns_of_llama_bench(tmp_models, tmp_path):
    """`dlp_bench.py -o json|jsonl|csv|md` against the reference's `llama-bench` on the same file: every column of the reference's records is
    present under its name, the identifying values (file, parameter count, size on disk, test shape) are equal, the CSV header starts with
    the reference's header, and the markdown table has its columns."""
    import csv
    import io
    import json
    from baseline import reference as ref_build
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-bench"):
        pytest.skip(f"reference binaries unavailable: {why}")
    model = tmp_models[("tiny", "F16")]
    env = ref_build.cpu_env(str(tmp_path))
    common = ["-m", model, "-p", "16", "-n", "4", "-r", "2"]

    def ref(fmt):
        p = subprocess.run([ref_build.exe("llama-bench"), *common, "-ngl", "0", "-t", "2", "-o", fmt], env=env, capture_output=True, text=True, timeout=300)
        assert p.returncode == 0, p.stderr[-400:]
        return p.stdout
    a, b = json.loads(ours("dlp_bench.py", *common, "--device", "cpu", "-o", "json")), json.loads(ref("json"))
    assert len(a) == len(b) == 2
    for x, y in zip(a, b):
        assert set(y) - set(x) == set(), sorted(set(y) - set(x))
        for k in ("model_filename", "model_size", "model_n_params", "n_prompt", "n_gen", "n_depth", "type_k", "type_v"):
            assert x[k] == y[k], (k, x[k], y[k])
        assert len(x["samples_ns"]) == len(x["samples_ts"]) == 2 and x["avg_ts"] > 0 and abs(x["avg_ns"] * x["avg_ts"] / 1e9 / (x["n_prompt"] + x["n_gen"]) - 1) < 0.5
    ha = next(csv.reader(io.StringIO(ours("dlp_bench.py", *common, "--device", "cpu", "-o", "csv"))))
    hb = next(csv.reader(io.StringIO(ref("csv"))))
    assert ha[:len(hb)] == hb
    ma, mb = ours("dlp_bench.py", *common, "--device", "cpu", "-o", "md").splitlines()[0], ref("md").splitlines()[0]
    cols = lambda line: [c.strip() for c in line.strip("|").split("|")]      # noqa: E731
    assert set(cols(mb)) - {"threads"} <= set(cols(ma))
    jl = [json.loads(ln) for ln in ours("dlp_bench.py", *common, "--device", "cpu", "-o", "jsonl").splitlines() if ln.startswith("{")]
    assert [r["test"] for r in jl] == ["pp16", "tg4"]


@pytest.mark.parametrize("args", [[], ["--ids"], ["--show-count", "--no-bos"], ["--no-parse-special"]])
def test_tokenize_tool_prints_what_llama_tokenize_prints(tmp_models, args):
    """`dlp_tokenize.py` and `llama-tokenize` on the same file: byte-identical stdout (ids, pieces - a lone byte token as that byte, control
    tokens by name -, escapes in the prompt, the count line), with special tokens in the middle of the text."""
    model = tmp_models[("tiny", "F16")]
    prompt = "hello wörld\\nx <s> y</s> z"
    ref = subprocess.run([tool("llama-tokenize"), "-m", model, "-p", prompt, "--log-disable", *args], capture_output=True, timeout=120)
    assert ref.returncode == 0, ref.stderr[-300:]
    mine = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dlp_tokenize.py"), "-m", model, "-p", prompt, *args], capture_output=True, timeout=120,
                          env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert mine.returncode == 0, mine.stderr[-300:]
    assert mine.stdout == ref.stdout


def test_quantizer_switches_select_the_same_tensor_types(tmp_models, tmp_path):
    """`--pure`, `--token-embedding-type`, `--output-tensor-type`, `--leave-output-tensor`, `--tensor-type PATTERN=TYPE`, lower-case / numeric type
    names and the optional thread count: the same command line gives the same per-tensor types as `llama-quantize`."""
    src = tmp_models[("tiny", "F16")]
    for k, flags in enumerate([["--pure"], ["--token-embedding-type", "q8_0", "--output-tensor-type", "q5_k"], ["--leave-output-tensor"],
                               ["--tensor-type", "attn_v=q6_k", "--tensor-type", "ffn_down=q5_k"]]):
        ref_out, our_out = str(tmp_path / f"ref{k}.gguf"), str(tmp_path / f"ours{k}.gguf")
        run([tool("llama-quantize"), *flags, src, ref_out, "q4_k_m", "2"])
        ours("dlp_quantize.py", *flags, src, our_out, "q4_k_m", "2", "--fast")
        r, o = open_gguf(ref_out), open_gguf(our_out)
        assert {n: GGMLType(t.ggml_type).name for n, t in r.tensors.items()} == {n: GGMLType(t.ggml_type).name for n, t in o.tensors.items()}, flags
    ours("dlp_quantize.py", src, str(tmp_path / "n.gguf"), "15", "--fast")          # llama_ftype number of Q4_K_M
    assert open_gguf(str(tmp_path / "n.gguf")).metadata["general.file_type"] == 15
    out = ours("dlp_quantize.py", "--dry-run", src, "q8_0")
    assert "nothing written" in out and "blk.0.attn_q.weight" in out


def test_split_by_size_with_an_empty_first_shard_is_merged_by_the_reference(tmp_models, tmp_path):
    """`--split-max-size` + `--no-tensor-first-split` (shards cut by bytes, the first one metadata only) and `--dry-run`: the reference's
    `llama-gguf-split --merge` reassembles the set bit for bit."""
    src = tmp_models[("tiny", "F16")]
    plan = ours("dlp_gguf_split.py", "--split", "--split-max-size", "1M", "--no-tensor-first-split", "--dry-run", src, str(tmp_path / "dry"))
    assert "split 00001: n_tensors = 0" in plan and not list(tmp_path.glob("dry*"))
    shards = ours("dlp_gguf_split.py", "--split", "--split-max-size", "1M", "--no-tensor-first-split", src, str(tmp_path / "sh")).split()
    assert len(shards) >= 3 and len(open_gguf(shards[0]).files[0].tensors) == 0
    merged = str(tmp_path / "merged.gguf")
    run([tool("llama-gguf-split"), "--merge", shards[0], merged])
    a, b = open_gguf(src), open_gguf(merged)
    assert set(a.tensors) == set(b.tensors) and all(np.array_equal(a.raw(n), b.raw(n)) for n in a.tensors)
