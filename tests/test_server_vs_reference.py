"""The HTTP API side by side with the reference's `llama-server` (unmodified, CPU back-end) on the same tiny checkpoint: tokenisation,
greedy completions with token probabilities, logit bias, grammar-constrained decoding, stop strings, the OpenAI dialect's envelope.
(llama.cpp/tools/server/server.cpp routes :164-197; tests of the same behaviours: tools/server/tests/unit/test_completion.py, test_tokenize.py)"""
import asyncio
import json
import os
import socket
import subprocess
import sys
import threading
import time
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline import reference as ref_build  # noqa: E402

from distributed_llm_pipeline_b200.serving.orchestrator import Orchestrator, build_service  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _post(port, path, body):
    r = urllib.request.urlopen(urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode(), {"Content-Type": "application/json"}), timeout=120)
    return json.loads(r.read())


def _get(port, path):
    return json.loads(urllib.request.urlopen(f"http://127.0.0.1:{port}{path}", timeout=30).read())


@pytest.fixture(scope="module")
def pair(tmp_models, tmp_path_factory):
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-server"):
        pytest.skip(f"reference binaries unavailable: {why}")
    model = tmp_models[("tiny", "F16")]
    env = ref_build.cpu_env(str(tmp_path_factory.mktemp("refsrv")))
    rp = _free_port()
    proc = subprocess.Popen([ref_build.exe("llama-server"), "-m", model, "-ngl", "0", "-c", "512", "-np", "2", "--host", "127.0.0.1", "--port", str(rp), "-t", "4",
                             "--no-warmup", "--metrics", "--slots", "--slot-save-path", str(tmp_path_factory.mktemp("refslots")) + "/"], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    for _ in range(200):
        try:
            _get(rp, "/health")
            break
        except Exception:
            time.sleep(0.1)
    else:
        proc.terminate()
        pytest.skip("the reference server did not come up here")
    sched, info, logs = build_service(model, device="cpu", n_ctx=512, n_slots=2, n_batch=64)
    sched.start()
    orch = Orchestrator(sched, info, n_predict=16, startup_logs=logs, slot_save_path=str(tmp_path_factory.mktemp("ourslots")))
    op = _free_port()
    loop = asyncio.new_event_loop()
    threading.Thread(target=lambda: loop.run_until_complete(orch.serve("127.0.0.1", op)), daemon=True).start()
    for _ in range(100):
        try:
            socket.create_connection(("127.0.0.1", op), timeout=0.2).close()
            break
        except OSError:
            time.sleep(0.05)
    yield op, rp
    sched.stop()
    proc.terminate()
    proc.wait()


TEXTS = ["hello there world", "The quick brown fox jumps over 13 lazy dogs!", "  leading spaces and\nnew lines\n", "ünïcödé 你好 🙂", "",
         "a <s> b</s>c <s>", "<s>x"]          # special tokens inside the text: the SPM space prefix returns behind each of them


def test_tokenize_and_detokenize_agree(pair):
    ours, ref = pair
    for t in TEXTS:
        a, b = _post(ours, "/tokenize", {"content": t})["tokens"], _post(ref, "/tokenize", {"content": t})["tokens"]
        assert a == b, t
        assert _post(ours, "/tokenize", {"content": t, "add_special": True})["tokens"] == _post(ref, "/tokenize", {"content": t, "add_special": True})["tokens"]
        if a:
            assert _post(ours, "/detokenize", {"tokens": a})["content"] == _post(ref, "/detokenize", {"tokens": a})["content"]


def _greedy(port, prompt, n=8, **extra):
    d = _post(port, "/completion", {"prompt": prompt, "n_predict": n, "temperature": 0, "n_probs": 3, "cache_prompt": False, **extra})
    return d, [p["id"] for p in d["completion_probabilities"]], [p["logprob"] for p in d["completion_probabilities"]]


@pytest.mark.parametrize("prompt", ["hello there", "one two three four", "the cat sat on the"])
def test_greedy_completion_and_probabilities_agree(pair, prompt):
    """Same tokens and log-probabilities as the reference (fp32 vs ggml's fp16 KV: 1e-2); where the two pick different tokens the reference's
    own top two candidates must be a near tie - a random-init model has many of those."""
    ours, ref = pair
    do, ido, lpo = _greedy(ours, prompt)
    dr, idr, lpr = _greedy(ref, prompt)
    assert do["tokens_evaluated"] == dr["tokens_evaluated"] and len(ido) == len(idr) == 8
    for i, (a, b) in enumerate(zip(ido, idr)):
        if a != b:
            top = dr["completion_probabilities"][i]["top_logprobs"]
            assert abs(top[0]["logprob"] - top[1]["logprob"]) < 2e-2, (i, ido, idr)
            break
        assert abs(lpo[i] - lpr[i]) < 1e-2
        tr, to = dr["completion_probabilities"][i]["top_logprobs"], do["completion_probabilities"][i]["top_logprobs"]
        assert [t["id"] for t in to][:1] == [t["id"] for t in tr][:1] and to[0]["token"] == tr[0]["token"]
    else:
        assert do["content"] == dr["content"]
    assert i >= 3          # at least the first tokens line up on every prompt


def test_logit_bias_grammar_and_stop_agree(pair):
    ours, ref = pair
    body = {"prompt": "hello", "n_predict": 5, "temperature": 0, "logit_bias": [[300, 100.0]], "cache_prompt": False}
    a, b = _post(ours, "/completion", body), _post(ref, "/completion", body)
    assert a["content"] == b["content"] != ""            # token 300 five times on both
    g = {"prompt": "answer:", "n_predict": 12, "temperature": 0, "cache_prompt": False, "grammar": 'root ::= ("yes" | "no") " " [a-c] [a-c] [a-c]'}
    a, b = _post(ours, "/completion", g), _post(ref, "/completion", g)
    assert a["content"] == b["content"] and a["content"][:3] in ("yes", "no ")
    js = {"prompt": "json:", "n_predict": 160, "temperature": 0, "cache_prompt": False,
          "json_schema": {"type": "object", "properties": {"ok": {"type": "boolean"}, "n": {"enum": [1, 2, 3]}}, "required": ["ok", "n"]}}
    a, b = _post(ours, "/completion", js), _post(ref, "/completion", js)
    for d in (a, b):          # a random-init model may spend its budget on the white space a JSON grammar allows: complete -> valid, else a prefix
        if d["stop_type"] in ("eos", "stop"):
            v = json.loads(d["content"])
            assert isinstance(v["ok"], bool) and v["n"] in (1, 2, 3)
        else:
            assert d["content"].lstrip().startswith("{")
    assert a["stop_type"] == "eos" and json.loads(a["content"])["n"] in (1, 2, 3)       # ours bounds every repetition, so it always closes
    # a stop string cuts both answers at the same place and is reported the same way
    base = _post(ref, "/completion", {"prompt": "hello there", "n_predict": 8, "temperature": 0, "cache_prompt": False})["content"]
    stop = base[3:5]
    a = _post(ours, "/completion", {"prompt": "hello there", "n_predict": 8, "temperature": 0, "cache_prompt": False, "stop": [stop]})
    b = _post(ref, "/completion", {"prompt": "hello there", "n_predict": 8, "temperature": 0, "cache_prompt": False, "stop": [stop]})
    assert a["content"] == b["content"] == base[:base.index(stop)] and a["stop_type"] == b["stop_type"] == "word"
    assert a["stopping_word"] == b["stopping_word"] == stop and a["truncated"] is b["truncated"] is False


def test_openai_envelope_and_meta_routes_agree(pair):
    ours, ref = pair
    body = {"messages": [{"role": "user", "content": "hi"}], "max_tokens": 4, "temperature": 0}
    a, b = _post(ours, "/v1/chat/completions", body), _post(ref, "/v1/chat/completions", body)
    for d in (a, b):
        assert d["object"] == "chat.completion" and d["choices"][0]["message"]["role"] == "assistant" and d["choices"][0]["finish_reason"] == "length"
        assert set(d["usage"]) >= {"prompt_tokens", "completion_tokens", "total_tokens"} and d["usage"]["completion_tokens"] == 4
    assert a["usage"]["prompt_tokens"] == b["usage"]["prompt_tokens"]            # the same template was rendered and tokenised
    assert a["choices"][0]["message"]["content"] == b["choices"][0]["message"]["content"]
    ta = _post(ours, "/apply-template", {"messages": body["messages"]})["prompt"]
    assert ta == _post(ref, "/apply-template", {"messages": body["messages"]})["prompt"]
    ma, mb = _get(ours, "/v1/models"), _get(ref, "/v1/models")
    assert ma["object"] == mb["object"] == "list" and ma["data"][0]["object"] == mb["data"][0]["object"] == "model"
    pa, pb = _get(ours, "/props"), _get(ref, "/props")
    assert pa["total_slots"] == pb["total_slots"] == 2 and "default_generation_settings" in pa and "default_generation_settings" in pb
    assert _get(ours, "/health")["status"] == _get(ref, "/health")["status"] == "ok"


@pytest.mark.parametrize("pooling", ["mean", "last"])
def test_embeddings_agree_with_reference_server(pair, tmp_models, tmp_path, pooling):
    """`/embedding` (final-norm hidden states pooled over the sequence, L2-normalised) against `llama-server --embedding --pooling X`."""
    ours, _ = pair
    env = ref_build.cpu_env(str(tmp_path))
    port = _free_port()
    p = subprocess.Popen([ref_build.exe("llama-server"), "-m", tmp_models[("tiny", "F16")], "-ngl", "0", "-c", "256", "--host", "127.0.0.1", "--port", str(port), "-t", "2",
                          "--no-warmup", "--embedding", "--pooling", pooling], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    try:
        for _ in range(200):
            try:
                _get(port, "/health")
                break
            except Exception:
                time.sleep(0.1)
        else:
            pytest.skip("the reference server did not come up here")
        for text in ["hello there world", "The quick brown fox jumps over the lazy dog"]:
            r = _post(port, "/embedding", {"content": text})
            rv = r[0]["embedding"] if isinstance(r, list) else r["embedding"]
            rv = rv[0] if isinstance(rv[0], list) else rv
            ov = _post(ours, "/embedding", {"content": text, "pooling": pooling})["embedding"]
            assert len(rv) == len(ov)
            dot = sum(a * b for a, b in zip(rv, ov))
            assert dot > 0.9999 and abs(sum(a * a for a in ov) - 1.0) < 1e-3, (pooling, text, dot)
        # the OpenAI route: the same keys (usage included), one entry per input, the same vectors
        body = {"input": ["hello there world", "second text"], "pooling": pooling}
        a, b = _post(ours, "/v1/embeddings", body), _post(port, "/v1/embeddings", {"input": body["input"]})
        assert _keys(b) - _keys(a) == set(), sorted(_keys(b) - _keys(a))
        assert a["usage"] == b["usage"] and [e["index"] for e in a["data"]] == [e["index"] for e in b["data"]] == [0, 1]
        for x, y in zip(a["data"], b["data"]):
            assert sum(p_ * q for p_, q in zip(x["embedding"], y["embedding"])) > 0.9999
        # the native route with several inputs, and a rerank request (this server scores by embedding similarity: the shape is what is compared)
        ra, rb = _post(ours, "/embedding", {"content": ["a b", "c d"], "pooling": pooling}), _post(port, "/embedding", {"content": ["a b", "c d"]})
        assert isinstance(ra, list) and isinstance(rb, list) and len(ra) == len(rb) == 2 and [e["index"] for e in ra] == [e["index"] for e in rb]
    finally:
        p.terminate()
        p.wait()


@pytest.mark.parametrize("extra", [{"repeat_penalty": 1.8, "repeat_last_n": 64}, {"presence_penalty": 1.5, "frequency_penalty": 0.7, "repeat_last_n": -1},
                                   {"dry_multiplier": 2.0, "dry_base": 1.75, "dry_allowed_length": 1, "dry_penalty_last_n": -1, "dry_sequence_breakers": ["\n"]},
                                   {"logit_bias": [[300, 2.5], [301, -100]], "repeat_penalty": 1.3}],
                         ids=["repeat", "presence-frequency", "dry", "bias+repeat"])
def test_deterministic_sampler_stages_agree(pair, extra):
    """Penalties, DRY and logit bias change which token is the arg-max: at temperature 0 the whole chain is deterministic, so the two servers
    must walk the same path (up to the near ties of a random-init model, detected through the reference's own top-2 gap)."""
    ours, ref = pair
    prompt = "one two one two one two one two one"
    do, ido, _ = _greedy(ours, prompt, n=10, **extra)
    dr, idr, _ = _greedy(ref, prompt, n=10, **extra)
    base = _greedy(ref, prompt, n=10)[1]
    agree = 0
    for i, (a, b) in enumerate(zip(ido, idr)):
        if a != b:
            break
        agree += 1
    assert agree >= 6 or agree == len(idr), (ido, idr)
    assert idr != base or "logit_bias" in extra          # the stage under test actually changed the reference's output


@pytest.mark.parametrize("params", [{"top_k": 10, "top_p": 1.0, "min_p": 0.0, "temperature": 1.0}, {"top_k": 0, "top_p": 0.3, "min_p": 0.0, "temperature": 0.7},
                                    {"top_k": 0, "top_p": 1.0, "min_p": 0.4, "temperature": 1.5}, {"top_k": 40, "top_p": 0.9, "min_p": 0.05, "temperature": 0.8},
                                    {"top_k": 0, "top_p": 1.0, "min_p": 0.0, "typical_p": 0.2, "temperature": 1.0}],
                         ids=["top-k", "top-p", "min-p", "default-chain", "typical"])
def test_truncation_samplers_leave_the_same_candidates(pair, params):
    """top-k / typical / top-p / min-p / temperature: the reference reports the distribution AFTER its sampler chain
    (`post_sampling_probs`); the same chain applied here (`filter_candidates` + temperature) to the same logits - taken from the
    reference's own full-vocabulary probabilities - must keep the same tokens with the same probabilities."""
    import math

    import torch

    from distributed_llm_pipeline_b200.serving.sampler import filter_candidates
    _, ref = pair
    prompt = "the quick brown fox"
    body = {"prompt": prompt, "n_predict": 2, "n_probs": 512, "seed": 1, "cache_prompt": False,
            "samplers": ["top_k", "typ_p", "top_p", "min_p", "temperature"], "typical_p": 1.0, **params}
    # the same request twice: once for the distribution in front of the chain, once for the one behind it
    top = _post(ref, "/completion", body)["completion_probabilities"][0]["top_logprobs"]
    assert len(top) == 512
    logits = torch.full((512,), -1e9)
    for t in top:
        logits[t["id"]] = t["logprob"]
    post = _post(ref, "/completion", {**body, "post_sampling_probs": True})["completion_probabilities"][0]["top_probs"]
    theirs = {t["id"]: t["prob"] for t in post if t["prob"] > 0}
    vals, ids = filter_candidates(logits, params["top_k"], params["top_p"], params["min_p"], params.get("typical_p", 1.0))
    probs = torch.softmax(vals / params["temperature"], -1)
    mine = {int(i): float(p) for i, p in zip(ids, probs)}
    assert set(mine) == set(theirs), (sorted(set(mine) ^ set(theirs)), len(mine), len(theirs))
    assert all(math.isclose(mine[i], theirs[i], rel_tol=2e-3, abs_tol=1e-5) for i in mine)


def test_infill_prompt_is_built_the_same_way(tmp_models, tmp_path):
    """`/infill`: <FIM_PRE> prefix <FIM_SUF> suffix <FIM_MID> behind BOS (server-common.cpp `format_infill`); both servers report how many
    tokens they evaluated and echo the prompt, and greedy decoding from it gives the same text."""
    import numpy as np

    from distributed_llm_pipeline_b200.gguf import GGUFWriter, open_gguf
    from distributed_llm_pipeline_b200.gguf.constants import ValueType
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-server"):
        pytest.skip(f"reference binaries unavailable: {why}")
    f = open_gguf(tmp_models[("tiny", "F16")])
    model = str(tmp_path / "fim.gguf")
    w = GGUFWriter(model)
    for k, v in f.metadata.items():
        w.add(k, v, f.files[0].metadata_types[k]) if not isinstance(v, (list, np.ndarray)) else w.add(k, v)
    for key, tid in (("pre", 500), ("suf", 501), ("mid", 502)):
        w.add(f"tokenizer.ggml.fim_{key}_token_id", tid, ValueType.UINT32)
    for n, t in f.tensors.items():
        w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
    w.write()
    env = ref_build.cpu_env(str(tmp_path))
    rp = _free_port()
    proc = subprocess.Popen([ref_build.exe("llama-server"), "-m", model, "-ngl", "0", "-c", "512", "--host", "127.0.0.1", "--port", str(rp), "-t", "2", "--no-warmup"],
                            env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    sched, info, logs = build_service(model, device="cpu", n_ctx=512, n_slots=1, n_batch=64)
    sched.start()
    try:
        for _ in range(200):
            try:
                _get(rp, "/health")
                break
            except Exception:
                time.sleep(0.1)
        else:
            pytest.skip("the reference server did not come up here")
        orch = Orchestrator(sched, info, n_predict=8, startup_logs=logs)
        op = _free_port()
        loop = asyncio.new_event_loop()
        threading.Thread(target=lambda: loop.run_until_complete(orch.serve("127.0.0.1", op)), daemon=True).start()
        time.sleep(0.5)
        body = {"input_prefix": "def add(a, b):\n    return ", "input_suffix": "\n\nprint(add(1, 2))\n", "n_predict": 6, "temperature": 0, "cache_prompt": False,
                "samplers": ["top_k"], "top_k": 1}
        a, b = _post(op, "/infill", body), _post(rp, "/infill", body)
        assert a["tokens_evaluated"] == b["tokens_evaluated"] > 20
        assert a["content"] == b["content"]
    finally:
        sched.stop()
        proc.terminate()
        proc.wait()


def test_context_shift_keeps_generating_like_the_reference(tmp_models, tmp_path):
    """Generation past the end of the context with `--context-shift`: both servers keep `n_keep` tokens, drop half of the rest and re-rotate
    the cached keys (server-context.cpp:1944-2003, llama-kv-cache.cpp K-shift); the greedy continuation after the shift must be the same."""
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-server"):
        pytest.skip(f"reference binaries unavailable: {why}")
    model = tmp_models[("tiny", "F16")]
    env = ref_build.cpu_env(str(tmp_path))
    rp = _free_port()
    proc = subprocess.Popen([ref_build.exe("llama-server"), "-m", model, "-ngl", "0", "-c", "256", "-np", "1", "--host", "127.0.0.1", "--port", str(rp), "-t", "2", "--no-warmup",
                             "--context-shift"], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    sched, info, logs = build_service(model, device="cpu", n_ctx=256, n_slots=1, n_batch=64)        # 256 is the smallest window the reference makes
    sched.context_shift = True
    sched.start()
    try:
        for _ in range(200):
            try:
                _get(rp, "/health")
                break
            except Exception:
                time.sleep(0.1)
        else:
            pytest.skip("the reference server did not come up here")
        orch = Orchestrator(sched, info, n_predict=8, startup_logs=logs)
        op = _free_port()
        loop = asyncio.new_event_loop()
        threading.Thread(target=lambda: loop.run_until_complete(orch.serve("127.0.0.1", op)), daemon=True).start()
        time.sleep(0.5)
        # a +100 bias pins the generated token on both servers, so the two caches hold the same text all the way; what is compared is the
        # model's own (pre-bias) distribution at every step - before, at and after the shift
        body = {"prompt": "one two three four five six seven", "n_predict": 260, "temperature": 0, "cache_prompt": False, "n_keep": 4, "ignore_eos": True,
                "n_probs": 3, "logit_bias": [[300, 100.0]]}
        a, b = _post(op, "/completion", body), _post(rp, "/completion", body)
        pa, pb = a["completion_probabilities"], b["completion_probabilities"]
        assert len(pa) == len(pb) == 260 and a["tokens_evaluated"] == b["tokens_evaluated"]
        assert all(x["id"] == y["id"] == 300 for x, y in zip(pa, pb))
        first_shift = 256 - a["tokens_evaluated"]
        assert 0 < first_shift < 240
        worst = 0.0
        for i, (x, y) in enumerate(zip(pa, pb)):
            tx, ty = x["top_logprobs"][0], y["top_logprobs"][0]
            gap = abs(y["top_logprobs"][0]["logprob"] - y["top_logprobs"][1]["logprob"])
            assert tx["id"] == ty["id"] or gap < 2e-2, (i, first_shift, tx, ty)
            worst = max(worst, abs(x["logprob"] - y["logprob"]))
        assert worst < 2e-2, (worst, first_shift)
        # and the shift is visible at all: the forced token's log-probability moves when half of the context disappears
        jump = abs(pb[first_shift + 1]["logprob"] - pb[first_shift - 2]["logprob"])
        assert jump > 1e-3, jump
    finally:
        sched.stop()
        proc.terminate()
        proc.wait()


def test_zz_metrics_and_slots_agree(pair):
    """Runs last on the shared pair: both servers have then served the same requests, so the Prometheus names must be the same set (ours adds
    the two KV-pool gauges) and the counters that do not depend on wall time must be equal after one more identical request on each side."""
    ours, ref = pair

    def metrics(port):
        text = urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=30).read().decode()
        vals = {ln.split()[0]: float(ln.split()[1]) for ln in text.splitlines() if ln and not ln.startswith("#")}
        kinds = {ln.split()[2]: ln.split()[3] for ln in text.splitlines() if ln.startswith("# TYPE")}
        return vals, kinds

    before = {p: metrics(p)[0] for p in (ours, ref)}
    body = {"prompt": "one two three four five six", "n_predict": 5, "temperature": 0, "cache_prompt": False, "logit_bias": [[7, 100]]}
    for p in (ours, ref):
        _post(p, "/completion", body)
    (va, ka), (vb, kb) = metrics(ours), metrics(ref)
    assert set(vb) <= set(va) and set(va) - set(vb) == {"llamacpp:kv_cache_pages_free", "llamacpp:kv_cache_pages_total"}
    assert all(ka[k] == kb[k] for k in kb)
    for name in ("prompt_tokens_total", "tokens_predicted_total", "n_decode_total"):
        k = "llamacpp:" + name
        assert va[k] - before[ours][k] == vb[k] - before[ref][k] > 0, name
    assert va["llamacpp:requests_processing"] == vb["llamacpp:requests_processing"] == 0
    sa, sb = _get(ours, "/slots"), _get(ref, "/slots")
    assert len(sa) == len(sb) == 2
    served_a, served_b = [x for x in sa if "params" in x], [x for x in sb if "params" in x]
    assert served_a and served_b
    for x in sa + sb:
        assert {"id", "n_ctx", "speculative", "is_processing"} <= set(x) and x["is_processing"] is False
    # the parameter record of a served slot carries the reference's sampling fields under the same names
    shared = set(served_a[0]["params"]) & set(served_b[0]["params"])
    assert {"temperature", "top_k", "top_p", "min_p", "typical_p", "repeat_last_n", "repeat_penalty", "presence_penalty", "frequency_penalty",
            "dry_multiplier", "dry_base", "dry_allowed_length", "mirostat", "mirostat_tau", "mirostat_eta", "xtc_probability", "xtc_threshold",
            "top_n_sigma", "dynatemp_range", "dynatemp_exponent", "n_predict", "max_tokens", "n_keep", "ignore_eos", "n_probs", "stream", "lora",
            "speculative.n_max"} <= shared
    assert set(served_a[0]["next_token"][0]) == set(served_b[0]["next_token"][0])


def _keys(d, pre=""):
    out = set()
    if isinstance(d, dict):
        for k, v in d.items():
            out.add(pre + k)
            out |= _keys(v, pre + k + ".")
    elif isinstance(d, list) and d:
        out |= _keys(d[0], pre + "[].")
    return out


def _status(port, path, body):
    r = urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode(), {"Content-Type": "application/json"})
    try:
        return 200, json.loads(urllib.request.urlopen(r, timeout=60).read())
    except urllib.error.HTTPError as e:
        return e.code, json.loads(e.read())


def test_response_envelopes_carry_every_field_of_the_reference(pair):
    """Field for field: whatever key the reference puts into an answer (closing record of /completion with its generation_settings, stream events,
    token probabilities before and after the sampler chain, both OpenAI dialects, /tokenize with pieces, an array of prompts) is in ours too, the
    deterministic values are equal, and malformed requests are refused with the same status, code and type instead of a dropped connection."""
    ours, ref = pair
    cases = [("/completion", {"prompt": "one two three", "n_predict": 4, "temperature": 0, "return_tokens": True, "cache_prompt": False}),
             ("/completion", {"prompt": [5, 9, 200, 33], "n_predict": 3, "temperature": 0, "cache_prompt": False}),
             ("/completion", {"prompt": ["one two", "three four"], "n_predict": 3, "temperature": 0, "cache_prompt": False}),
             ("/completion", {"prompt": "one", "n_predict": 3, "temperature": 0, "n_probs": 2, "cache_prompt": False}),
             ("/completion", {"prompt": "one", "n_predict": 3, "temperature": 0, "n_probs": 2, "post_sampling_probs": True, "cache_prompt": False}),
             ("/completion", {"prompt": "one two three", "n_predict": 0, "temperature": 0, "cache_prompt": False}),       # still one token, as there
             ("/v1/completions", {"prompt": "one two three", "max_tokens": 4, "temperature": 0}),
             ("/v1/completions", {"prompt": "one two three", "max_tokens": 4, "temperature": 0, "logprobs": 2}),
             ("/v1/chat/completions", {"messages": [{"role": "user", "content": "hi"}], "max_tokens": 4, "temperature": 0, "logprobs": True, "top_logprobs": 2}),
             ("/completion", {"prompt": "one two", "n_predict": 2, "temperature": 0, "cache_prompt": False,
                              "response_fields": ["content", "timings/prompt_n", "generation_settings/n_predict", "nope/x"]}),
             ("/tokenize", {"content": "hello wörld", "with_pieces": True}), ("/tokenize", {"content": "hello", "add_special": True})]
    for path, body in cases:
        (sa, a), (sb, b) = _status(ours, path, body), _status(ref, path, body)
        assert sa == sb == 200, (path, body, a, b)
        assert _keys(b) - _keys(a) == set(), (path, body, sorted(_keys(b) - _keys(a)))
        if "response_fields" in body:
            assert set(a) == set(b) == {"content", "timings/prompt_n", "generation_settings/n_predict"} and a["generation_settings/n_predict"] == 2
        if isinstance(b, list):       # an array of prompts: one record each (the reference lists them as they finish: pair them by prompt)
            assert len(a) == len(b) == len(body["prompt"]) and sorted(x["index"] for x in a) == sorted(y["index"] for y in b)
            a, b = sorted(a, key=lambda r: r["prompt"]), sorted(b, key=lambda r: r["prompt"])
        for x, y in zip(a if isinstance(a, list) else [a], b if isinstance(b, list) else [b]):
            for k in ("content", "tokens", "stop", "stop_type", "stopping_word", "truncated", "tokens_predicted", "tokens_evaluated", "prompt", "has_new_line") + \
                    (() if isinstance(b, list) else ("index",)):
                if k in y:
                    assert x[k] == y[k], (path, body, k, x[k], y[k])
            if "completion_probabilities" in y:
                for p, q in zip(x["completion_probabilities"], y["completion_probabilities"]):
                    assert (p["id"], p["token"], p["bytes"]) == (q["id"], q["token"], q["bytes"])
                    tp, tq = p.get("top_logprobs", p.get("top_probs")), q.get("top_logprobs", q.get("top_probs"))
                    assert [t["id"] for t in tp] == [t["id"] for t in tq]
                    for t, u in zip(tp, tq):
                        assert abs(t.get("logprob", t.get("prob")) - u.get("logprob", u.get("prob"))) < 2e-2
            if path.startswith("/v1/"):
                assert x["choices"][0]["finish_reason"] == y["choices"][0]["finish_reason"] and x["usage"] == y["usage"] and x["object"] == y["object"]
                la, lb = x["choices"][0].get("logprobs"), y["choices"][0].get("logprobs")
                assert (la is None) == (lb is None)
                if lb:
                    assert [e["id"] for e in la["content"]] == [e["id"] for e in lb["content"]]
        if path == "/tokenize":
            assert a == b
    # streams: the same keys in the token events and in the closing event
    for path, body in [("/completion", {"prompt": "one", "n_predict": 2, "temperature": 0, "stream": True, "cache_prompt": False}),
                       ("/v1/chat/completions", {"messages": [{"role": "user", "content": "hi"}], "max_tokens": 2, "temperature": 0, "stream": True})]:
        ev = {}
        for port in (ours, ref):
            r = urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode(), {"Content-Type": "application/json"})
            ev[port] = [json.loads(ln[6:]) for ln in urllib.request.urlopen(r, timeout=60).read().decode().split("\n") if ln.startswith("data: {")]
        assert len(ev[ours]) == len(ev[ref]) >= 3
        for x, y in zip(ev[ours], ev[ref]):
            assert _keys(y) - _keys(x) == set(), (path, sorted(_keys(y) - _keys(x)))
            if "content" in y:
                assert (x["content"], x["tokens"], x["stop"]) == (y["content"], y["tokens"], y["stop"])
            else:
                assert x["choices"][0]["delta"] == y["choices"][0]["delta"] and x["choices"][0]["finish_reason"] == y["choices"][0]["finish_reason"]
    # malformed requests
    for path, body in [("/completion", {"n_predict": 3}), ("/completion", "notadict"), ("/v1/chat/completions", {"messages": "hi"}),
                       ("/v1/chat/completions", {"max_tokens": 3})]:
        (sa, a), (sb, b) = _status(ours, path, body), _status(ref, path, body)
        assert sa == sb == 400 and a["error"]["code"] == b["error"]["code"] == 400 and a["error"]["type"] == b["error"]["type"], (path, body, a, b)
    # values of the wrong type fall back to the defaults on both sides; the request is served
    for body in [{"prompt": "one", "n_predict": 2, "temperature": "hot", "top_k": None, "seed": "x"}, {"prompt": "one", "n_predict": "many", "max_tokens": 2}]:
        (sa, a), (sb, b) = _status(ours, "/completion", body), _status(ref, "/completion", body)
        assert sa == sb == 200 and a["stop"] is b["stop"] is True


def test_the_remaining_routes_answer_like_the_reference(pair):
    """Every route `llama-server` registers (tools/server/server.cpp) exists here: the aliases without `/v1`, `/v1/health`, the combined model
    list, Ollama's `/api/show`, `/lora-adapters`, and the Anthropic token counter - same keys, same counts."""
    ours, ref = pair

    def call(port, method, path, body=None):
        r = urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode() if body is not None else None,
                                   {"Content-Type": "application/json"}, method=method)
        try:
            return 200, json.loads(urllib.request.urlopen(r, timeout=60).read())
        except urllib.error.HTTPError as e:
            return e.code, json.loads(e.read())
    msgs = [{"role": "user", "content": "hello there"}]
    for method, path, body in [("GET", "/v1/health", None), ("GET", "/models", None), ("GET", "/v1/models", None), ("POST", "/api/show", {"name": "x"}),
                               ("GET", "/lora-adapters", None), ("POST", "/lora-adapters", []), ("POST", "/v1/messages/count_tokens", {"messages": msgs}),
                               ("POST", "/chat/completions", {"messages": msgs, "max_tokens": 2, "temperature": 0}),
                               ("POST", "/api/chat", {"messages": msgs, "max_tokens": 2, "temperature": 0}),
                               ("POST", "/v1/messages", {"messages": msgs, "max_tokens": 3, "temperature": 0})]:
        (sa, a), (sb, b) = call(ours, method, path, body), call(ref, method, path, body)
        assert sa == sb == 200, (path, a, b)
        assert _keys(b) - _keys(a) - {"data.[].meta.vocab_type", "data.[].meta.size"} == set(), (path, sorted(_keys(b) - _keys(a)))
    assert call(ours, "POST", "/v1/messages/count_tokens", {"messages": msgs})[1] == call(ref, "POST", "/v1/messages/count_tokens", {"messages": msgs})[1]
    a, b = call(ours, "POST", "/v1/messages", {"messages": msgs, "max_tokens": 3, "temperature": 0})[1], call(ref, "POST", "/v1/messages", {"messages": msgs, "max_tokens": 3, "temperature": 0})[1]
    assert a["content"] == b["content"] and a["usage"] == b["usage"] and a["stop_reason"] == b["stop_reason"]
    ma, mb = call(ours, "GET", "/models")[1]["data"][0]["meta"], call(ref, "GET", "/models")[1]["data"][0]["meta"]
    assert all(ma[k] == mb[k] for k in ("n_vocab", "n_ctx_train", "n_embd", "n_params"))


def test_props_slot_actions_and_errors_agree(pair):
    """`/props` key for key, slot save / restore / erase answers (counts equal: the same prompt is resident on both sides), the refusals of a bad
    slot id / action / unknown route with the same status and error type, and `add_generation_prompt: false` on `/apply-template`."""
    ours, ref = pair

    def call(port, method, path, body=None):
        r = urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode() if body is not None else None,
                                   {"Content-Type": "application/json"}, method=method)
        try:
            return 200, json.loads(urllib.request.urlopen(r, timeout=60).read())
        except urllib.error.HTTPError as e:
            return e.code, json.loads(e.read())
    (sa, a), (sb, b) = call(ours, "GET", "/props"), call(ref, "GET", "/props")
    assert sa == sb == 200 and _keys(b) - _keys(a) == set(), sorted(_keys(b) - _keys(a))
    for port in (ours, ref):
        assert call(port, "POST", "/completion", {"prompt": "one two three four", "n_predict": 3, "temperature": 0, "id_slot": 0})[0] == 200
    counts = {}
    for action, body in [("save", {"filename": "s0.bin"}), ("restore", {"filename": "s0.bin"}), ("erase", None)]:
        (sa, a), (sb, b) = call(ours, "POST", f"/slots/0?action={action}", body), call(ref, "POST", f"/slots/0?action={action}", body)
        assert sa == sb == 200 and _keys(b) - _keys(a) == set(), (action, a, b)
        key = {"save": "n_saved", "restore": "n_restored", "erase": "n_erased"}[action]
        assert a[key] == b[key] > 0, (action, a, b)
        counts[action] = a[key]
    assert counts["save"] == counts["restore"] == counts["erase"]
    for method, path, body, status in [("POST", "/slots/9?action=erase", None, 400), ("POST", "/slots/0?action=bogus", None, 400), ("GET", "/nonexistent", None, 404),
                                       ("POST", "/nonexistent", {}, 404)]:
        (sa, a), (sb, b) = call(ours, method, path, body), call(ref, method, path, body)
        assert sa == sb == status and a["error"]["type"] == b["error"]["type"] and a["error"]["code"] == b["error"]["code"], (path, a, b)
    body = {"messages": [{"role": "user", "content": "u"}], "add_generation_prompt": False}
    assert call(ours, "POST", "/apply-template", body)[1] == call(ref, "POST", "/apply-template", body)[1]


def test_event_streams_agree(pair):
    """The three streaming dialects event by event: `/v1/completions`, `/v1/chat/completions` with `stream_options.include_usage` (one more chunk
    without choices carrying the usage) and Anthropic `/v1/messages` (event names, every key of every event, the deltas themselves)."""
    ours, ref = pair
    msgs = [{"role": "user", "content": "hi"}]

    def events(port, path, body):
        r = urllib.request.Request(f"http://127.0.0.1:{port}{path}", json.dumps(body).encode(), {"Content-Type": "application/json"})
        text = urllib.request.urlopen(r, timeout=60).read().decode()
        names = [ln[7:] for ln in text.split("\n") if ln.startswith("event: ")]
        return names, [json.loads(ln[6:]) for ln in text.split("\n") if ln.startswith("data: {")], text.rstrip().endswith("data: [DONE]")
    for path, body in [("/v1/completions", {"prompt": "one", "max_tokens": 2, "temperature": 0, "stream": True}),
                       ("/v1/chat/completions", {"messages": msgs, "max_tokens": 2, "temperature": 0, "stream": True, "stream_options": {"include_usage": True}}),
                       ("/v1/messages", {"messages": msgs, "max_tokens": 2, "temperature": 0, "stream": True})]:
        (na, ea, da), (nb, eb, db) = events(ours, path, body), events(ref, path, body)
        assert na == nb and len(ea) == len(eb) and da == db, (path, na, nb, len(ea), len(eb))
        for x, y in zip(ea, eb):
            assert _keys(y) - _keys(x) == set(), (path, sorted(_keys(y) - _keys(x)))
            for k in ("choices", "delta", "usage", "type", "index", "content_block"):
                if k in y and k != "usage":
                    assert x[k] == y[k], (path, k, x[k], y[k])
            if "usage" in y and "timings" not in y.get("usage", {}):
                assert x["usage"] == y["usage"], (path, x["usage"], y["usage"])
