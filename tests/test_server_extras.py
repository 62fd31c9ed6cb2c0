"""Server features beyond plain completion: tool calling on the OpenAI chat route, reranking.
(reference: llama.cpp/common/chat.cpp + chat-parser.cpp, tools/server/server.cpp rerank handler)"""
import asyncio
import json
import socket
import threading
import time

import pytest

from distributed_llm_pipeline_b200.serving import toolcall
from distributed_llm_pipeline_b200.serving.grammar import Grammar, json_schema_to_gbnf
from distributed_llm_pipeline_b200.serving.orchestrator import Orchestrator, build_service

WEATHER = {"type": "function", "function": {"name": "get_weather", "description": "weather for a city",
                                            "parameters": {"type": "object", "properties": {"city": {"type": "string"}, "unit": {"enum": ["c", "f"]}},
                                                           "required": ["city"]}}}
ADD = {"type": "function", "function": {"name": "add", "parameters": {"type": "object", "properties": {"a": {"type": "integer"}, "b": {"type": "integer"}},
                                                                         "required": ["a", "b"]}}}


def _accepts(gbnf: str, text: str) -> bool:
    return Grammar(gbnf).matches(text)


def test_parse_recognises_every_syntax():
    tools = [WEATHER, ADD]
    m = toolcall.parse('{"tool_call": {"name": "get_weather", "arguments": {"city": "Oslo"}}}', tools)
    assert m.content == "" and m.tool_calls[0]["function"] == {"name": "get_weather", "arguments": '{"city": "Oslo"}'}
    assert m.tool_calls[0]["id"].startswith("call_") and m.to_openai()["content"] is None
    m = toolcall.parse('{"tool_calls": [{"name": "add", "arguments": {"a": 1, "b": 2}}, {"name": "get_weather", "arguments": {"city": "Rome"}}]}', tools)
    assert [c["function"]["name"] for c in m.tool_calls] == ["add", "get_weather"]
    assert toolcall.parse('{"response": "hello"}', tools).content == "hello"
    m = toolcall.parse('Let me check.\n<tool_call>\n{"name": "get_weather", "arguments": {"city": "Paris"}}\n</tool_call>', tools)           # Hermes / Qwen
    assert m.content == "Let me check." and json.loads(m.tool_calls[0]["function"]["arguments"]) == {"city": "Paris"}
    m = toolcall.parse('<|python_tag|>{"name": "add", "parameters": {"a": 3, "b": 4}}', tools)                                               # Llama 3.x
    assert m.tool_calls[0]["function"]["name"] == "add" and json.loads(m.tool_calls[0]["function"]["arguments"]) == {"a": 3, "b": 4}
    m = toolcall.parse('[TOOL_CALLS][{"name": "add", "arguments": {"a": 5, "b": 6}}]', tools)                                                # Mistral
    assert len(m.tool_calls) == 1 and m.content == ""
    m = toolcall.parse('<function=get_weather>{"city": "Lima"}</function>', tools)                                                          # Functionary
    assert m.tool_calls[0]["function"]["name"] == "get_weather"
    # not tool calls: unknown function name, ordinary prose, JSON that is not a call
    assert not toolcall.parse('{"tool_call": {"name": "rm_rf", "arguments": {}}}', tools).tool_calls
    assert toolcall.parse("The weather is fine.", tools).content == "The weather is fine."
    assert not toolcall.parse('{"name": "add"}', tools).tool_calls


def test_prepare_builds_prompt_and_schema():
    msgs = [{"role": "system", "content": "Be brief."}, {"role": "user", "content": "weather in Oslo?"},
            {"role": "assistant", "content": None, "tool_calls": [{"id": "call_1", "type": "function", "function": {"name": "get_weather", "arguments": '{"city": "Oslo"}'}}]},
            {"role": "tool", "tool_call_id": "call_1", "content": "4 C, rain"}]
    st = toolcall.prepare(msgs, [WEATHER, ADD], "auto")
    assert st.active and "get_weather" in st.messages[0]["content"] and st.messages[0]["content"].startswith("Be brief.")
    assert st.messages[2]["role"] == "assistant" and '"tool_calls"' in st.messages[2]["content"]
    assert st.messages[3] == {"role": "user", "content": "Tool result (call_1): 4 C, rain"}
    g = json_schema_to_gbnf(st.json_schema)
    assert _accepts(g, '{"tool_call": {"name": "get_weather", "arguments": {"city": "Oslo"}}}')
    assert _accepts(g, '{"tool_call": {"name": "add", "arguments": {"a": 1, "b": 2}}}')
    assert _accepts(g, '{"response": "hi"}')
    assert not _accepts(g, '{"tool_call": {"name": "delete_all", "arguments": {}}}')
    assert not _accepts(g, '{"tool_call": {"name": "add", "arguments": {"a": "x", "b": 2}}}')
    req = toolcall.prepare(msgs[:2], [WEATHER, ADD], "required")
    assert not _accepts(json_schema_to_gbnf(req.json_schema), '{"response": "hi"}')
    one = toolcall.prepare(msgs[:2], [WEATHER, ADD], {"type": "function", "function": {"name": "add"}})
    g1 = json_schema_to_gbnf(one.json_schema)
    assert _accepts(g1, '{"tool_call": {"name": "add", "arguments": {"a": 1, "b": 2}}}') and not _accepts(g1, '{"tool_call": {"name": "get_weather", "arguments": {"city": "x"}}}')
    par = toolcall.prepare(msgs[:2], [WEATHER, ADD], "required", parallel_tool_calls=True)
    assert _accepts(json_schema_to_gbnf(par.json_schema), '{"tool_calls": [{"name": "add", "arguments": {"a": 1, "b": 2}}, {"name": "add", "arguments": {"a": 3, "b": 4}}]}')
    off = toolcall.prepare(msgs[:2], [WEATHER], "none")
    assert not off.active and off.json_schema is None and len(off.messages) == 2


def _http(port, method, path, body=b""):
    s = socket.create_connection(("127.0.0.1", port), timeout=120)
    s.sendall((f"{method} {path} HTTP/1.1\r\nHost: x\r\nContent-Length: {len(body)}\r\nConnection: close\r\n\r\n").encode() + body)
    data = b""
    while True:
        chunk = s.recv(65536)
        if not chunk:
            break
        data += chunk
    s.close()
    head, _, payload = data.partition(b"\r\n\r\n")
    return head.split(b"\r\n")[0].decode(), payload


@pytest.fixture(scope="module")
def server(tmp_models):
    sched, info, logs = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=1024, n_slots=2, n_batch=64)
    sched.start()
    orch = Orchestrator(sched, info, n_predict=12, startup_logs=logs)
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
    yield port
    sched.stop()


def test_chat_completions_with_required_tool_call(server):
    """A random-init model knows nothing about tools: the grammar alone must force a well-formed call to one of them."""
    body = {"messages": [{"role": "user", "content": "add two and three"}], "tools": [ADD], "tool_choice": "required", "max_tokens": 260,
            "temperature": 0.7, "seed": 5}
    status, payload = _http(server, "POST", "/v1/chat/completions", json.dumps(body).encode())
    assert status.endswith("200 OK"), payload[:300]
    ch = json.loads(payload)["choices"][0]
    # every repetition in the envelope's grammar is bounded (white space <= 8, integers <= 16 digits): it closes well within max_tokens
    assert ch["finish_reason"] == "tool_calls", ch
    call = ch["message"]["tool_calls"][0]
    assert call["type"] == "function" and call["function"]["name"] == "add" and ch["message"]["content"] is None
    args = json.loads(call["function"]["arguments"])
    assert isinstance(args["a"], int) and isinstance(args["b"], int)
    # streaming: the text is held back and the call arrives as one delta
    body["stream"] = True
    status, payload = _http(server, "POST", "/v1/chat/completions", json.dumps(body).encode())
    events = [json.loads(l[6:]) for l in payload.decode().split("\n") if l.startswith("data: {")]
    assert events and events[-1]["choices"][0]["finish_reason"] == "tool_calls"
    calls = [e["choices"][0]["delta"]["tool_calls"] for e in events if "tool_calls" in e["choices"][0]["delta"]]
    assert len(calls) == 1 and calls[0][0]["index"] == 0 and calls[0][0]["function"]["name"] == "add"
    assert not any(e["choices"][0]["delta"].get("content") for e in events)
    assert payload.rstrip().endswith(b"data: [DONE]")


def test_rerank_route(server):
    docs = ["the cat sat on the mat", "quantum chromodynamics", "a cat on a mat"]
    status, payload = _http(server, "POST", "/v1/rerank", json.dumps({"query": "the cat sat on the mat", "documents": docs, "top_n": 2,
                                                                       "return_documents": True}).encode())
    assert status.endswith("200 OK"), payload[:300]
    d = json.loads(payload)
    assert d["object"] == "list" and len(d["results"]) == 2 and d["usage"]["prompt_tokens"] > 0
    assert d["results"][0]["index"] == 0 and d["results"][0]["relevance_score"] > 0.999      # the identical text ranks first
    assert d["results"][0]["relevance_score"] >= d["results"][1]["relevance_score"] and d["results"][0]["document"]["text"] == docs[0]
    status, _ = _http(server, "POST", "/rerank", json.dumps({"query": "x"}).encode())
    assert "400" in status


def test_token_probabilities_on_every_route(server):
    """`n_probs` on /completion, `logprobs` on /v1/completions, `logprobs` + `top_logprobs` on /v1/chat/completions, streamed or not:
    log-probabilities of the full softmax, the sampled token's own entry consistent with the candidates listed."""
    import math
    body = {"prompt": "the quick brown", "n_predict": 5, "temperature": 0, "n_probs": 3, "cache_prompt": False}
    status, payload = _http(server, "POST", "/completion", json.dumps(body).encode())
    d = json.loads(payload)
    cp = d["completion_probabilities"]
    assert len(cp) == d["tokens_predicted"] == 5
    for e in cp:
        assert len(e["top_logprobs"]) == 3 and e["top_logprobs"][0]["logprob"] >= e["top_logprobs"][1]["logprob"] >= e["top_logprobs"][2]["logprob"]
        assert e["logprob"] <= 0 and e["id"] == e["top_logprobs"][0]["id"]          # greedy: the sampled token is the most likely one
        assert sum(math.exp(t["logprob"]) for t in e["top_logprobs"]) <= 1.0 + 1e-5
    # the same request without probabilities gives the same text (the host path samples like the device path at temperature 0)
    plain = json.loads(_http(server, "POST", "/completion", json.dumps({**body, "n_probs": 0}).encode())[1])
    assert plain["content"] == d["content"] and "completion_probabilities" not in plain
    # streamed: the entries arrive with the token events and add up to the same list
    status, payload = _http(server, "POST", "/completion", json.dumps({**body, "stream": True}).encode())
    events = [json.loads(l[6:]) for l in payload.decode().split("\n") if l.startswith("data: {")]
    streamed = [p for e in events for p in e.get("completion_probabilities", [])]
    assert [p["id"] for p in streamed] == [p["id"] for p in cp]
    # OpenAI dialects
    o = json.loads(_http(server, "POST", "/v1/completions", json.dumps({"prompt": "the quick brown", "max_tokens": 4, "temperature": 0, "logprobs": 2}).encode())[1])
    lp = o["choices"][0]["logprobs"]["content"]         # llama-server answers /v1/completions with the chat-style `content` list as well
    assert len(lp) == 4 and all(len(e["top_logprobs"]) == 2 and {"id", "token", "bytes", "logprob"} <= set(e) for e in lp)
    c = json.loads(_http(server, "POST", "/v1/chat/completions", json.dumps({"messages": [{"role": "user", "content": "hi"}], "max_tokens": 3, "temperature": 0,
                                                                              "logprobs": True, "top_logprobs": 2}).encode())[1])
    content = c["choices"][0]["logprobs"]["content"]
    assert len(content) == 3 and all(len(e["top_logprobs"]) == 2 and e["bytes"] == list(e["token"].encode()) for e in content)
    assert "logprobs" not in json.loads(_http(server, "POST", "/v1/chat/completions", json.dumps({"messages": [{"role": "user", "content": "hi"}], "max_tokens": 2}).encode())[1])["choices"][0]
