"""The C++ front-end (runtime/orchestrator.cpp) + the Python engine service it supervises, over real sockets."""
import json
import os
import signal
import socket
import subprocess
import sys
import time

import pytest

from distributed_llm_pipeline_b200.runtime import native

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _dechunk(b: bytes) -> bytes:
    out = b""
    while b:
        line, _, rest = b.partition(b"\r\n")
        n = int(line, 16)
        if n == 0:
            break
        out += rest[:n]
        b = rest[n + 2:]
    return out


@pytest.fixture(scope="module")
def front(tmp_models):
    native.build()
    fport, bport = _free_port(), _free_port()
    spawn = (f"{sys.executable} -m distributed_llm_pipeline_b200.serving.backend -m {tmp_models[('tiny', 'F16')]} "
             f"--device cpu --port {bport} -c 128 -np 2 -b 32")
    env = dict(os.environ, PYTHONPATH=REPO, CUDA_VISIBLE_DEVICES="")
    p = subprocess.Popen([native.path("dlp-orchestrator"), "--host", "127.0.0.1", "--port", str(fport), "--backend", f"127.0.0.1:{bport}",
                          "--static", os.path.join(REPO, "distributed_llm_pipeline_b200", "serving", "static"), "--spawn", spawn, "-n", "10"],
                         env=env, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, preexec_fn=os.setsid)
    p.stdout.readline()
    yield fport, bport
    os.killpg(os.getpgid(p.pid), signal.SIGTERM)


def _post_chat(port, body: bytes):
    s = socket.create_connection(("127.0.0.1", port), timeout=120)
    s.sendall(b"POST /chat HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: " + str(len(body)).encode() + b"\r\n\r\n" + body)
    data = b""
    while True:
        c = s.recv(65536)
        if not c:
            break
        data += c
    s.close()
    head, _, payload = data.partition(b"\r\n\r\n")
    return head.decode(), payload


def test_native_chat_stream(front):
    fport, _ = front
    head, payload = _post_chat(fport, json.dumps({"prompt": "Once upon a time ☕"}).encode())
    assert head.startswith("HTTP/1.1 200")
    low = head.lower()
    assert "content-type: text/event-stream" in low and "cache-control: no-cache" in low
    assert "access-control-allow-origin: *" in low and "transfer-encoding: chunked" in low
    text = _dechunk(payload).decode()
    events = [json.loads(b[6:]) for b in text.split("\n\n") if b.startswith("data: ")]
    assert {e["msg_type"] for e in events} == {"log", "token"}
    assert any("offloaded" in e["content"] for e in events if e["msg_type"] == "log")
    assert all(b == "" or b.startswith("data: ") or b.startswith(":") for b in text.split("\n\n"))


def test_native_rejects_bad_json_and_serves_static(front):
    fport, _ = front
    head, _ = _post_chat(fport, b'{"prompt": 5}')
    assert " 422 " in head
    s = socket.create_connection(("127.0.0.1", fport), timeout=10)
    s.sendall(b"GET / HTTP/1.1\r\nHost: x\r\n\r\n")
    data = b""
    while True:
        c = s.recv(65536)
        if not c:
            break
        data += c
    assert b"200 OK" in data and b"/chat" in data and b"text/html" in data


def _http(port, method, target, body=b""):
    s = socket.create_connection(("127.0.0.1", port), timeout=120)
    s.sendall(f"{method} {target} HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: {len(body)}\r\n\r\n".encode() + body)
    data = b""
    while True:
        c = s.recv(65536)
        if not c:
            break
        data += c
    s.close()
    head, _, payload = data.partition(b"\r\n\r\n")
    return head.decode(), payload


def test_native_front_end_tunnels_the_api_routes(front):
    """Routes the C++ front-end does not implement itself reach the engine service through the raw HTTP tunnel."""
    fport, _ = front
    _post_chat(fport, json.dumps({"prompt": "warm"}).encode())      # make sure the engine service is up
    head, body = _http(fport, "POST", "/completion", json.dumps({"prompt": "a b c", "n_predict": 4, "temperature": 0}).encode())
    assert " 200 " in head and json.loads(body)["tokens_predicted"] <= 4
    head, body = _http(fport, "POST", "/v1/chat/completions",
                       json.dumps({"messages": [{"role": "user", "content": "hi"}], "max_tokens": 3, "temperature": 0}).encode())
    assert json.loads(body)["object"] == "chat.completion"
    head, body = _http(fport, "POST", "/tokenize", json.dumps({"content": "hello world"}).encode())
    toks = json.loads(body)["tokens"]
    head, body = _http(fport, "POST", "/detokenize", json.dumps({"tokens": toks}).encode())
    assert json.loads(body)["content"].strip() == "hello world"
    head, body = _http(fport, "POST", "/embedding", json.dumps({"content": "hello"}).encode())
    assert len(json.loads(body)["embedding"]) == 256
    head, body = _http(fport, "GET", "/props")
    assert json.loads(body)["total_slots"] == 2
    head, body = _http(fport, "POST", "/slots/0?action=erase", b"{}")
    assert " 200 " in head and "n_erased" in json.loads(body)
    # streamed completion: the service's chunked SSE response is relayed verbatim
    head, body = _http(fport, "POST", "/completion", json.dumps({"prompt": "x y", "n_predict": 3, "temperature": 0, "stream": True}).encode())
    assert "text/event-stream" in head.lower() and b"data: " in body


def test_native_supervisor_restarts_engine_service(front):
    fport, bport = front
    # kill the engine service; the front-end must bring it back and serve the next request
    psutil = pytest.importorskip("psutil")
    pids = [c.pid for c in psutil.net_connections("tcp") if c.status == "LISTEN" and c.laddr.port == bport and c.pid]
    if not pids:
        pytest.skip("cannot find the engine-service pid")
    os.kill(pids[0], signal.SIGKILL)
    time.sleep(1.5)
    head, payload = _post_chat(fport, json.dumps({"prompt": "again"}).encode())
    text = _dechunk(payload).decode()
    events = [json.loads(b[6:]) for b in text.split("\n\n") if b.startswith("data: ")]
    assert any(e["msg_type"] == "token" for e in events)
    assert any("restart" in e["content"] for e in events if e["msg_type"] == "log")


def test_native_front_over_tensor_parallel_backend(tmp_models):
    """dlp-orchestrator --spawn "<engine service> --tp 2": the C++ front-end supervises a service that is itself a grid of
    processes (it re-starts under torchrun; rank 0 serves, rank 1 replays).  The reference's equivalent is the hard-coded
    `llama-cli --rpc a,b` child of `orchestrator/src/main.rs:35-57`."""
    native.build()
    fport, bport = _free_port(), _free_port()
    spawn = (f"{sys.executable} -m distributed_llm_pipeline_b200.serving.backend -m {tmp_models[('tiny', 'F16')]} "
             f"--device cpu --port {bport} -c 128 -np 2 -b 32 --tp 2")
    env = dict(os.environ, PYTHONPATH=REPO, CUDA_VISIBLE_DEVICES="", DLP_MASTER_PORT=str(_free_port()))
    p = subprocess.Popen([native.path("dlp-orchestrator"), "--host", "127.0.0.1", "--port", str(fport), "--backend", f"127.0.0.1:{bport}",
                          "--static", os.path.join(REPO, "distributed_llm_pipeline_b200", "serving", "static"), "--spawn", spawn, "-n", "8"],
                         env=env, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, preexec_fn=os.setsid)
    try:
        p.stdout.readline()
        head, payload = _post_chat(fport, json.dumps({"prompt": "a b c"}).encode())
        assert head.startswith("HTTP/1.1 200")
        events = [json.loads(b[6:]) for b in _dechunk(payload).decode().split("\n\n") if b.startswith("data: ")]
        logs = "".join(e["content"] for e in events if e["msg_type"] == "log")
        assert "grid: 1 pipeline stage(s) x 2 tensor-parallel rank(s)" in logs and logs.count("load_tensors: stage") == 2
        assert sum(e["msg_type"] == "token" for e in events) >= 1
    finally:
        os.killpg(os.getpgid(p.pid), signal.SIGTERM)
