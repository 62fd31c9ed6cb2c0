"""The native runtime under sanitizers (reference: LLAMA_SANITIZE_* CI matrix): the worker and the orchestrator
front-end are rebuilt with -fsanitize and driven with real requests; any report fails the test."""
import json
import os
import signal
import socket
import subprocess
import sys
import time

import pytest

from distributed_llm_pipeline_b200.parallel.rpc import RpcEngine, free_port_run, launch_local_workers
from distributed_llm_pipeline_b200.runtime import native

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BAD = ("ERROR: AddressSanitizer", "ERROR: LeakSanitizer", "runtime error:", "WARNING: ThreadSanitizer", "ERROR: ThreadSanitizer")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("kind", ["address"])   # (TSan cannot see libgomp's futex barriers: OpenMP regions are all false positives)
def test_worker_under_sanitizer(tmp_models, tmp_path, kind):
    try:
        native.build(sanitize=kind)
    except RuntimeError as e:
        pytest.skip(f"sanitizer runtime not available: {str(e)[:200]}")
    log = open(tmp_path / f"worker-{kind}.log", "w+")
    os.environ.setdefault("ASAN_OPTIONS", "detect_leaks=1")
    procs, eps = launch_local_workers(tmp_models[("tiny", "Q8_0")], 2, base_port=free_port_run(2), n_ctx=64, threads=2, sanitize=kind, stderr=log)
    try:
        eng = RpcEngine(eps, max_seqs=2)
        a, b = eng.alloc_seq(), eng.alloc_seq()
        eng.prefill(a, list(range(3, 20)))
        eng.prefill(b, [5, 6, 7])
        for _ in range(4):
            eng.decode([a, b], [9, 11], sample=False)
        eng.truncate_seq(a, 4)
        eng.decode([a], [3], sample=False)
        eng.free_seq(b)
        for w in eng.workers:
            w.shutdown()
        for p in procs:
            assert p.wait(timeout=60) == 0
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    log.seek(0)
    text = log.read()
    assert not any(b in text for b in BAD), text[-3000:]


@pytest.mark.parametrize("kind", ["address", "thread"])
def test_orchestrator_front_end_under_sanitizer(tmp_models, tmp_path, kind):
    """The front-end is the multi-threaded part (thread per connection, supervisor, log ring): ASan+UBSan and TSan, with
    concurrent clients."""
    try:
        native.build(sanitize=kind)
    except RuntimeError as e:
        pytest.skip(f"sanitizer runtime not available: {str(e)[:200]}")
    fport, bport = _free_port(), _free_port()
    spawn = (f"{sys.executable} -m distributed_llm_pipeline_b200.serving.backend -m {tmp_models[('tiny', 'F16')]} "
             f"--device cpu --port {bport} -c 64 -np 1 -b 32")
    log = open(tmp_path / "front.log", "w+")
    p = subprocess.Popen([native.path("dlp-orchestrator", kind), "--host", "127.0.0.1", "--port", str(fport), "--backend", f"127.0.0.1:{bport}",
                          "--static", os.path.join(REPO, "distributed_llm_pipeline_b200", "serving", "static"), "--spawn", spawn, "-n", "6"],
                         env=dict(os.environ, PYTHONPATH=REPO, CUDA_VISIBLE_DEVICES=""), stdout=subprocess.PIPE, stderr=log, preexec_fn=os.setsid)
    try:
        p.stdout.readline()

        def http(method, target, body=b""):
            s = socket.create_connection(("127.0.0.1", fport), timeout=120)
            s.sendall(f"{method} {target} HTTP/1.1\r\nHost: x\r\nContent-Length: {len(body)}\r\n\r\n".encode() + body)
            data = b""
            while True:
                c = s.recv(65536)
                if not c:
                    break
                data += c
            s.close()
            return data
        assert b"msg_type" in http("POST", "/chat", json.dumps({"prompt": "hé ☕ \\u00e9"}).encode())
        assert b" 422 " in http("POST", "/chat", b'{"prompt": 5}')
        assert b" 422 " in http("POST", "/chat", b'{"prompt": "unterminated')
        assert b"tokens" in http("POST", "/tokenize", json.dumps({"content": "a b"}).encode())
        assert b"200 OK" in http("GET", "/")
        assert b"404" in http("GET", "/../../etc/passwd") or b"403" in http("GET", "/../../etc/passwd") or b"400" in http("GET", "/../../etc/passwd")
        import threading
        ts = [threading.Thread(target=http, args=("POST", "/chat", json.dumps({"prompt": f"p{i}"}).encode())) for i in range(3)]
        ts += [threading.Thread(target=http, args=("GET", "/health")) for _ in range(3)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
    finally:
        os.killpg(os.getpgid(p.pid), signal.SIGTERM)
        time.sleep(0.5)
    log.seek(0)
    text = log.read()
    assert not any(b in text for b in BAD), text[-3000:]
