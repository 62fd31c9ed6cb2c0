"""Wire-level tests of the /chat SSE API against the reference's format (`orchestrator/src/main.rs:18-27,97,100-111`):
real sockets, real tiny model on CPU, no mocks (the reference's server tests work the same way)."""
import asyncio
import json
import socket
import threading
import time

import pytest

from distributed_llm_pipeline_b200.serving.orchestrator import Orchestrator, build_service, sse_event
from distributed_llm_pipeline_b200.serving.scheduler import Request
from distributed_llm_pipeline_b200.serving.sampler import SamplingParams
from distributed_llm_pipeline_b200.utils.tokenizer import StreamDecoder, Tokenizer


@pytest.fixture(scope="module")
def server(tmp_models):
    sched, info, logs = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=128, n_slots=2, n_batch=32)
    sched.start()
    orch = Orchestrator(sched, info, n_predict=12, startup_logs=logs)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    loop = asyncio.new_event_loop()
    t = threading.Thread(target=lambda: loop.run_until_complete(orch.serve("127.0.0.1", port)), daemon=True)
    t.start()
    for _ in range(100):
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.2).close()
            break
        except OSError:
            time.sleep(0.05)
    yield port, sched
    sched.stop()


def _http(port, method, path, body=b"", headers=None, read_all=True):
    s = socket.create_connection(("127.0.0.1", port), timeout=30)
    h = {"Host": "x", "Content-Length": str(len(body)), "Connection": "close"}
    h.update(headers or {})
    s.sendall((f"{method} {path} HTTP/1.1\r\n" + "".join(f"{k}: {v}\r\n" for k, v in h.items()) + "\r\n").encode() + body)
    data = b""
    while True:
        chunk = s.recv(65536)
        if not chunk:
            break
        data += chunk
    s.close()
    head, _, payload = data.partition(b"\r\n\r\n")
    lines = head.decode().split("\r\n")
    hdrs = {l.split(":", 1)[0].lower(): l.split(":", 1)[1].strip() for l in lines[1:] if ":" in l}
    return lines[0], hdrs, payload


def test_event_wire_format_matches_reference_serialisation():
    # serde_json of StreamMessage{msg_type, content}: field order, compact separators, UTF-8 kept verbatim
    assert sse_event("token", "hé\n") == 'data: {"msg_type":"token","content":"hé\\n"}\n\n'.encode()


def test_chat_streams_log_and_token_events(server):
    port, _ = server
    status, hdrs, payload = _http(port, "POST", "/chat", json.dumps({"prompt": "Once upon a time"}).encode(),
                                  {"Content-Type": "application/json"})
    assert status.startswith("HTTP/1.1 200")
    assert hdrs["content-type"] == "text/event-stream" and hdrs["cache-control"] == "no-cache"
    assert hdrs["access-control-allow-origin"] == "*"
    events = []
    for block in payload.decode().split("\n\n"):
        if block.startswith("data: "):
            events.append(json.loads(block[6:]))
        else:
            assert block == "" or block.startswith(":")     # keep-alive comments only
    kinds = {e["msg_type"] for e in events}
    assert kinds == {"log", "token"}
    assert all(set(e) == {"msg_type", "content"} for e in events)
    assert any("offloaded" in e["content"] for e in events if e["msg_type"] == "log")   # the UI's highlight hook
    assert sum(e["msg_type"] == "token" for e in events) >= 1


def test_chat_rejects_bad_body_and_serves_static_and_cors(server):
    port, _ = server
    status, _, _ = _http(port, "POST", "/chat", b'{"nope": 1}', {"Content-Type": "application/json"})
    assert "422" in status
    status, hdrs, body = _http(port, "GET", "/")
    assert status.startswith("HTTP/1.1 200") and b"/chat" in body and hdrs["content-type"].startswith("text/html")
    status, _, _ = _http(port, "GET", "/../../etc/passwd")
    assert "404" in status
    status, hdrs, _ = _http(port, "OPTIONS", "/chat", headers={"Origin": "http://x", "Access-Control-Request-Method": "POST"})
    assert status.startswith("HTTP/1.1 200") and hdrs["access-control-allow-origin"] == "*"


def test_aux_routes(server):
    port, sched = server
    _, _, body = _http(port, "GET", "/health")
    assert json.loads(body)["status"] == "ok"
    _, _, body = _http(port, "POST", "/completion", json.dumps({"prompt": "a b c", "n_predict": 5, "temperature": 0}).encode())
    d = json.loads(body)
    assert d["tokens_predicted"] <= 5 and "timings" in d and d["stop"] is True
    _, _, body = _http(port, "POST", "/v1/chat/completions",
                       json.dumps({"messages": [{"role": "user", "content": "hi"}], "max_tokens": 4, "temperature": 0}).encode())
    d = json.loads(body)
    assert d["object"] == "chat.completion" and d["usage"]["completion_tokens"] <= 4
    _, _, body = _http(port, "GET", "/metrics")
    text = body.decode()
    assert "llamacpp:prompt_tokens_total" in text and "llamacpp:tokens_predicted_total" in text
    _, _, body = _http(port, "GET", "/slots")
    assert len(json.loads(body)) == sched.n_slots
    _, _, body = _http(port, "POST", "/tokenize", json.dumps({"content": "hello world"}).encode())
    toks = json.loads(body)["tokens"]
    _, _, body = _http(port, "POST", "/detokenize", json.dumps({"tokens": toks}).encode())
    assert json.loads(body)["content"].strip() == "hello world"


def test_concurrent_requests_share_slots_and_queue(server):
    port, sched = server
    results = []

    def one(i):
        _, _, body = _http(port, "POST", "/completion", json.dumps({"prompt": f"request {i}", "n_predict": 6, "temperature": 0}).encode())
        results.append(json.loads(body))

    ts = [threading.Thread(target=one, args=(i,)) for i in range(5)]     # 5 requests, 2 slots -> 3 deferred
    [t.start() for t in ts]
    [t.join(60) for t in ts]
    assert len(results) == 5 and all(r["stop"] for r in results)


def test_greedy_is_deterministic_and_prefix_cache_is_used(server):
    _, sched = server
    p = "the quick brown fox jumps over the lazy dog"
    a = sched.generate(p, n_predict=8, sampling=SamplingParams(temperature=0.0))
    b = sched.generate(p, n_predict=8, sampling=SamplingParams(temperature=0.0))
    assert a.output_ids == b.output_ids
    assert b.n_prompt_cached > 0 and b.n_prompt_processed < len(b.prompt_ids)      # LCP reuse


def test_tokenizer_roundtrip_and_stream_decoder(tmp_models):
    tok = Tokenizer.from_gguf(tmp_models[("tiny", "F16")])
    for text in ["hello world", "a  b", "naïve café ☕", "line1\nline2"]:
        ids = tok.encode(text)
        assert ids[0] == tok.bos_id
        assert tok.decode(ids[1:]) == text
    ids = tok.encode("☕ok", add_bos=False)   # multi-byte char goes through byte tokens in the synthetic vocab
    sd = StreamDecoder(tok)
    pieces = [sd.push(i) for i in ids]
    # (the pieces of a stream keep the space the SPM prefix rule put in front of the text, as the reference's token_to_piece does)
    assert "�" not in "".join(pieces) and "".join(pieces) + sd.flush() == " ☕ok"


def test_stop_string_and_n_predict(server):
    _, sched = server
    r = sched.generate("abc", n_predict=3, sampling=SamplingParams(temperature=0.0))
    assert len(r.output_ids) <= 3 and r.finish_reason in ("length", "stop")
    full = sched.generate("abc", n_predict=10, sampling=SamplingParams(temperature=0.0), cache_prompt=False)
    if len(full.text) > 3:
        needle = full.text[2:4]
        cut = sched.generate("abc", n_predict=10, sampling=SamplingParams(temperature=0.0), stop=[needle])
        assert needle not in cut.text and cut.finish_reason == "stop"


def test_embedding_routes(server):
    port, sched = server
    _, _, body = _http(port, "POST", "/embedding", json.dumps({"content": "hello world"}).encode())
    e = json.loads(body)["embedding"]
    assert len(e) == sched.eng.cfg.n_embd and abs(sum(x * x for x in e) - 1.0) < 1e-3
    _, _, body = _http(port, "POST", "/v1/embeddings", json.dumps({"input": ["hello world", "other text"]}).encode())
    d = json.loads(body)
    assert d["object"] == "list" and len(d["data"]) == 2
    a, b = d["data"][0]["embedding"], d["data"][1]["embedding"]
    assert max(abs(x - y) for x, y in zip(a, e)) < 1e-5 and max(abs(x - y) for x, y in zip(a, b)) > 1e-3
    # generation still works afterwards (the embedding ran in a scratch sequence between batches)
    r = sched.generate("abc", n_predict=3, sampling=SamplingParams(temperature=0.0))
    assert r.finish_reason in ("length", "stop")


def test_grammar_constrained_completion(server):
    port, sched = server
    gbnf = 'root ::= ("ka" | "ri" | "to")+ "."'
    _, _, body = _http(port, "POST", "/completion",
                       json.dumps({"prompt": "say", "n_predict": 12, "temperature": 0.7, "seed": 5, "grammar": gbnf}).encode())
    text = json.loads(body)["content"]
    from distributed_llm_pipeline_b200.serving.grammar import Grammar
    g = Grammar(gbnf)
    body_txt = text.strip()
    assert body_txt and (g.matches(body_txt) or all(g.clone().accepts(body_txt[:i]) for i in range(1, len(body_txt) + 1)))
    assert set(body_txt) <= set("kario.t")
    r = sched.generate("json:", n_predict=24, sampling=SamplingParams(temperature=0.0, json_schema={"type": "boolean"}))
    assert r.text.strip() in ("true", "false") and r.finish_reason == "stop"


def test_context_shift_keeps_generating_past_n_ctx(tmp_models):
    sched, _, _ = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=48, n_slots=1, n_batch=32)
    logs = []
    r = sched.generate("a b c d e", n_predict=70, sampling=SamplingParams(temperature=0.0), ignore_eos=True)
    assert r.finish_reason == "length" and len(r.output_ids) < 48            # default: stop when the context is full
    sched.context_shift = True
    r = sched.generate("a b c d e", n_predict=70, sampling=SamplingParams(temperature=0.0), ignore_eos=True, cache_prompt=False,
                       n_keep=-1, on_event=lambda k, c: logs.append(c) if k == "log" else None)
    assert len(r.output_ids) == 70 and any("context shift" in x for x in logs)
    assert sched.metrics.c["context_shifts_total"] >= 1


def _serve(orch):
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
    return port


def test_api_key_slot_save_restore_and_templates(tmp_models, tmp_path):
    from distributed_llm_pipeline_b200.serving.orchestrator import apply_chat_template
    sched, info, logs = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=128, n_slots=1, n_batch=32)
    sched.start()
    port = _serve(Orchestrator(sched, info, n_predict=8, startup_logs=logs, api_key="s3cret", slot_save_path=str(tmp_path)))
    try:
        body = json.dumps({"prompt": "a b c d e f", "n_predict": 4, "temperature": 0}).encode()
        st, _, _ = _http(port, "POST", "/completion", body)
        assert " 401 " in st
        st, _, _ = _http(port, "GET", "/health")
        assert " 200 " in st
        auth = {"Authorization": "Bearer s3cret"}
        _, _, out = _http(port, "POST", "/completion", body, auth)
        first = json.loads(out)["content"]
        _, _, out = _http(port, "POST", "/slots/0?action=save", json.dumps({"filename": "s.bin"}).encode(), auth)
        saved = json.loads(out)
        assert saved["n_saved"] > 0 and (tmp_path / "s.bin").exists()
        _, _, out = _http(port, "POST", "/slots/0?action=erase", b"{}", auth)
        assert json.loads(out)["n_erased"] == saved["n_saved"]
        _, _, out = _http(port, "POST", "/slots/0?action=restore", json.dumps({"filename": "s.bin"}).encode(), auth)
        assert json.loads(out)["n_restored"] == saved["n_saved"]
        _, _, out = _http(port, "POST", "/completion", body, auth)       # served from the restored cache: same text, prefix reused
        d = json.loads(out)
        assert d["content"] == first and d["timings"]["cache_n"] > 0
        _, _, out = _http(port, "POST", "/apply-template", json.dumps({"messages": [{"role": "user", "content": "hi"}]}).encode(), auth)
        assert json.loads(out)["prompt"] == "<|im_start|>user\nhi<|im_end|>\n<|im_start|>assistant\n"      # no template in the file: ChatML, like llama-server
        st, _, _ = _http(port, "POST", "/infill", json.dumps({"input_prefix": "a", "input_suffix": "b"}).encode(), auth)
        assert " 501 " in st                                             # synthetic vocab has no FIM tokens
    finally:
        sched.stop()
    msgs = [{"role": "system", "content": "S"}, {"role": "user", "content": "U"}]
    assert apply_chat_template({"tokenizer.chat_template": "<|im_start|>"}, msgs).endswith("<|im_start|>assistant\n")
    assert apply_chat_template({"tokenizer.chat_template": "[INST] <<SYS>>"}, msgs) == "[INST] <<SYS>>\nS\n<</SYS>>\n\nU [/INST]"
    assert "<start_of_turn>model" in apply_chat_template({"tokenizer.chat_template": "<start_of_turn>"}, msgs)
    assert apply_chat_template({"tokenizer.chat_template": "<|start_header_id|>"}, msgs).count("<|eot_id|>") == 2


def test_model_router_loads_on_demand_and_unloads_lru(tmp_models):
    from distributed_llm_pipeline_b200.serving.orchestrator import ModelRouter
    models = {"llama-tiny": tmp_models[("tiny", "F16")], "qwen-tiny": tmp_models[("tiny-qwen2", "F16")]}
    router = ModelRouter(models, dict(device="cpu", n_ctx=64, n_slots=1, n_batch=32), max_loaded=1, sleep_idle_s=1.0, n_predict=4)
    port = _serve(router)
    _, _, body = _http(port, "GET", "/v1/models")
    assert {m["id"]: m["status"] for m in json.loads(body)["data"]} == {"llama-tiny": "unloaded", "qwen-tiny": "unloaded"}
    req = lambda m: json.dumps({"model": m, "prompt": "a b c", "n_predict": 3, "temperature": 0}).encode()  # noqa: E731
    _, _, body = _http(port, "POST", "/completion", req("llama-tiny"))
    first = json.loads(body)["content"]
    _, _, body = _http(port, "GET", "/health")
    assert json.loads(body)["models_loaded"] == ["llama-tiny"]
    _, _, body = _http(port, "POST", "/completion", req("qwen-tiny"))          # second model evicts the first (max_loaded = 1)
    assert "content" in json.loads(body)
    _, _, body = _http(port, "GET", "/health")
    assert json.loads(body)["models_loaded"] == ["qwen-tiny"]
    _, _, body = _http(port, "POST", "/completion", req("llama-tiny"))          # reload: same deterministic output
    assert json.loads(body)["content"] == first
    st, _, _ = _http(port, "POST", "/completion", req("nope"))
    assert " 404 " in st
    # explicit unload / load (llama-server router mode: POST /models/unload, /models/load)
    assert json.loads(_http(port, "POST", "/models/unload", json.dumps({"model": "llama-tiny"}).encode())[2]) == {"success": True}
    assert json.loads(_http(port, "GET", "/health")[2])["models_loaded"] == []
    assert json.loads(_http(port, "POST", "/models/load", json.dumps({"model": "qwen-tiny"}).encode())[2]) == {"success": True}
    assert json.loads(_http(port, "GET", "/v1/health")[2])["models_loaded"] == ["qwen-tiny"]
    assert " 404 " in _http(port, "POST", "/models/load", json.dumps({"model": "nope"}).encode())[0]
    _http(port, "POST", "/completion", req("llama-tiny"))
    _, _, body = _http(port, "GET", "/health")
    assert json.loads(body)["loads"] == 5
    deadline = time.time() + 10                                                   # --sleep-idle-seconds: unloaded when idle
    while time.time() < deadline:
        _, _, body = _http(port, "GET", "/health")
        if not json.loads(body)["models_loaded"]:
            break
        time.sleep(0.3)
    assert json.loads(body)["models_loaded"] == []


def test_model_router_spreads_requests_over_replicas(tmp_models):
    """`--replicas 2`: two copies of the model, each with one slot; six concurrent requests are split between them by the number of requests in
    flight, both copies serve some, and every answer is the same deterministic text."""
    import concurrent.futures
    from distributed_llm_pipeline_b200.serving.orchestrator import ModelRouter
    router = ModelRouter({"m": tmp_models[("tiny", "F16")]}, dict(device="cpu", n_ctx=64, n_slots=1, n_batch=32), max_loaded=1, n_predict=4, replicas=2)
    port = _serve(router)
    req = json.dumps({"model": "m", "prompt": "a b c d e", "n_predict": 12, "temperature": 0, "cache_prompt": False}).encode()
    first = json.loads(_http(port, "POST", "/completion", req)[2])["content"]          # loads the group
    with concurrent.futures.ThreadPoolExecutor(6) as ex:
        outs = list(ex.map(lambda _: json.loads(_http(port, "POST", "/completion", req)[2])["content"], range(6)))
    assert outs == [first] * 6
    h = json.loads(_http(port, "GET", "/health")[2])
    assert h["replicas"] == 2 and h["loads"] == 1 and h["in_flight"] == {"m": [0, 0]}
    served = [o.sched.metrics.c["requests_total"] for o in router.loaded["m"]]
    assert sum(served) == 7 and min(served) >= 2, served


def test_ollama_and_anthropic_dialects(server):
    port, _ = server
    _, _, body = _http(port, "POST", "/api/generate", json.dumps({"prompt": "a b c", "stream": False, "options": {"num_predict": 4, "temperature": 0}}).encode())
    d = json.loads(body)
    assert d["done"] is True and isinstance(d["response"], str) and d["eval_count"] <= 4
    _, hdr, body = _http(port, "POST", "/api/chat", json.dumps({"messages": [{"role": "user", "content": "hi"}], "options": {"num_predict": 3, "temperature": 0}}).encode())
    lines = [json.loads(l) for l in body.decode().splitlines() if l.strip()]          # default: NDJSON stream
    assert lines[-1]["done"] is True and all("message" in l for l in lines)
    _, _, body = _http(port, "GET", "/api/tags")
    assert json.loads(body)["models"]
    _, _, body = _http(port, "POST", "/v1/messages", json.dumps({"system": "be brief", "max_tokens": 4, "temperature": 0,
                                                                   "messages": [{"role": "user", "content": [{"type": "text", "text": "hello"}]}]}).encode())
    d = json.loads(body)
    assert d["type"] == "message" and d["content"][0]["type"] == "text" and d["usage"]["output_tokens"] <= 4
    assert d["stop_reason"] in ("max_tokens", "end_turn")
    _, _, body = _http(port, "POST", "/v1/messages", json.dumps({"max_tokens": 3, "temperature": 0, "stream": True,
                                                                   "messages": [{"role": "user", "content": "hello"}]}).encode())
    text = body.decode()
    assert "event: message_start" in text and "event: content_block_delta" in text and text.rstrip().endswith('{"type": "message_stop"}')


def test_scheduler_speculative_lookup_matches_plain_greedy(tmp_models):
    """n_draft > 0: prompt-lookup drafts verified in one forward inside the serving loop - same text, fewer model passes."""
    sched, _, _ = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=256, n_slots=2, n_batch=64)
    prompt = "ka ri to ka ri to ka ri to ka ri"
    plain = sched.generate(prompt, n_predict=40, sampling=SamplingParams(temperature=0.0), cache_prompt=False, ignore_eos=True)
    n_dec_plain = sched.metrics.c["n_decode_total"]
    spec = sched.generate(prompt, n_predict=40, sampling=SamplingParams(temperature=0.0), cache_prompt=False, ignore_eos=True, n_draft=8)
    assert spec.output_ids == plain.output_ids and spec.text == plain.text and len(spec.output_ids) == 40
    n_dec_spec = sched.metrics.c["n_decode_total"] - n_dec_plain
    assert sched.metrics.c.get("draft_tokens_accepted_total", 0) > 0 and n_dec_spec < n_dec_plain
    # the cache stayed consistent: the same prompt again continues identically
    again = sched.generate(prompt, n_predict=12, sampling=SamplingParams(temperature=0.0), ignore_eos=True, n_draft=4)
    assert again.output_ids == plain.output_ids[:12]
    # two concurrent requests, one speculative and one sampled, share the loop
    import threading
    out = {}
    ts = [threading.Thread(target=lambda: out.__setitem__("a", sched.generate(prompt, n_predict=16, sampling=SamplingParams(temperature=0.0), ignore_eos=True, n_draft=6))),
          threading.Thread(target=lambda: out.__setitem__("b", sched.generate("x y z", n_predict=16, sampling=SamplingParams(temperature=0.9, seed=3), ignore_eos=True)))]
    sched.start()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    sched.stop()
    assert out["a"].output_ids == plain.output_ids[:16] and len(out["b"].output_ids) == 16


def test_scheduler_draft_model_speculation(tmp_models):
    """`-md`: a draft engine proposes blocks, the target verifies them in one forward - identical greedy text, fewer target
    passes; a quantised copy of the model drafts (high but not total agreement), slots keep separate draft sequences."""
    target, draft = tmp_models[("tiny", "F16")], tmp_models[("tiny", "Q8_0")]
    plain_s, _, _ = build_service(target, device="cpu", n_ctx=256, n_slots=2, n_batch=64)
    prompt = "one two three four five"
    plain = plain_s.generate(prompt, n_predict=32, sampling=SamplingParams(temperature=0.0), cache_prompt=False, ignore_eos=True)
    n_plain = plain_s.metrics.c["n_decode_total"]
    sched, info, logs = build_service(target, device="cpu", n_ctx=256, n_slots=2, n_batch=64, model_draft=draft, draft_max=6)
    assert info["draft_model"].endswith("Q8_0.gguf") and any("draft model" in l for l in logs)
    spec = sched.generate(prompt, n_predict=32, sampling=SamplingParams(temperature=0.0), cache_prompt=False, ignore_eos=True)
    assert spec.output_ids == plain.output_ids and spec.text == plain.text
    c = sched.metrics.c
    assert c["draft_tokens_total"] > 0 and c["draft_tokens_accepted_total"] > 0.5 * c["draft_tokens_total"]
    assert c["n_decode_total"] < n_plain
    assert not sched._drafters and not sched.draft_eng.seqs             # the draft sequence was released with the request
    # a sampled request is not speculated; n_draft < 0 switches it off for one greedy request
    before = c["draft_tokens_total"]
    sched.generate(prompt, n_predict=8, sampling=SamplingParams(temperature=0.8, seed=1), ignore_eos=True)
    off = sched.generate(prompt, n_predict=8, sampling=SamplingParams(temperature=0.0), ignore_eos=True, cache_prompt=False, n_draft=-1)
    assert sched.metrics.c["draft_tokens_total"] == before and off.output_ids == plain.output_ids[:8]
    # two greedy requests at once: one draft sequence each
    import threading
    out = {}
    ts = [threading.Thread(target=lambda k=k, p=p: out.__setitem__(k, sched.generate(p, n_predict=16, sampling=SamplingParams(temperature=0.0), ignore_eos=True, cache_prompt=False)))
          for k, p in (("a", prompt), ("b", "six seven eight"))]
    sched.start()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    sched.stop()
    assert out["a"].output_ids == plain.output_ids[:16]
    ref_b = plain_s.generate("six seven eight", n_predict=16, sampling=SamplingParams(temperature=0.0), ignore_eos=True, cache_prompt=False)
    assert out["b"].output_ids == ref_b.output_ids
    # a draft model with another vocabulary is refused
    with pytest.raises(ValueError):
        build_service(target, device="cpu", n_ctx=64, n_slots=1, n_batch=32, model_draft=tmp_models[("tiny128", "F16")])
