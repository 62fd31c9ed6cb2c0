"""The distributed engine behind the front-ends: orchestrator, engine service and CLI driving a pp x tp grid of processes
(gloo on CPU here; the same commands run NCCL + the fused kernels on GPUs) or a chain of dlp-worker CPU stages.

The reference has no such test (nothing starts an rpc-server or a multi-GPU split); its orchestrator simply
assumes two rpc-servers are up (`orchestrator/src/main.rs:47-48`).  Real processes, real sockets, tiny real model."""
import json
import os
import signal
import socket
import subprocess
import sys
import time

import pytest

from distributed_llm_pipeline_b200.serving.orchestrator import build_service
from distributed_llm_pipeline_b200.serving.sampler import SamplingParams

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROMPT = "a b c d e f"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _http(port, method, path, body=b"", timeout=120):
    s = socket.create_connection(("127.0.0.1", port), timeout=timeout)
    s.sendall((f"{method} {path} HTTP/1.1\r\nHost: x\r\nContent-Length: {len(body)}\r\nConnection: close\r\n\r\n").encode() + body)
    data = b""
    while True:
        chunk = s.recv(65536)
        if not chunk:
            break
        data += chunk
    s.close()
    head, _, payload = data.partition(b"\r\n\r\n")
    return head.decode().split("\r\n")[0], payload


class _Proc:
    def __init__(self, args, port, env=None):
        e = dict(os.environ, PYTHONPATH=REPO, DLP_MASTER_PORT=str(_free_port()), **(env or {}))
        self.p = subprocess.Popen([sys.executable] + args, cwd=REPO, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                  text=True, start_new_session=True)
        self.port = port

    def wait_ready(self, timeout=180):
        t0 = time.time()
        while time.time() - t0 < timeout:
            if self.p.poll() is not None:
                raise RuntimeError("server exited: " + (self.p.stdout.read() or "")[-3000:])
            try:
                socket.create_connection(("127.0.0.1", self.port), timeout=0.3).close()
                return
            except OSError:
                time.sleep(0.2)
        raise TimeoutError("server did not come up")

    def stop(self):
        if self.p.poll() is None:
            try:
                os.killpg(self.p.pid, signal.SIGINT)      # exactly the process group this test started
                self.p.wait(timeout=20)
            except (subprocess.TimeoutExpired, ProcessLookupError):
                try:
                    os.killpg(self.p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass


@pytest.fixture(scope="module")
def expected(tmp_models):
    """Greedy continuation from the single-process engine (the oracle for every distributed variant)."""
    sched, _, _ = build_service(tmp_models[("tiny", "F16")], device="cpu", n_ctx=128, n_slots=2, n_batch=32)
    r = sched.generate(PROMPT, n_predict=10, sampling=SamplingParams(temperature=0.0), ignore_eos=True)
    return r.output_ids, r.text


def _completion(port, n=10):
    _, body = _http(port, "POST", "/completion", json.dumps({"prompt": PROMPT, "n_predict": n, "temperature": 0, "ignore_eos": True,
                                                             "cache_prompt": False}).encode())
    return json.loads(body)


def _chat_events(port, prompt):
    status, payload = _http(port, "POST", "/chat", json.dumps({"prompt": prompt}).encode())
    assert status.startswith("HTTP/1.1 200")
    return [json.loads(b[6:]) for b in payload.decode().split("\n\n") if b.startswith("data: ")]


@pytest.mark.parametrize("grid", [("--tp", "2"), ("--pp", "2"), ("--pp", "2", "--tp", "2")], ids=["tp2", "pp2", "pp2xtp2"])
def test_orchestrator_over_process_grid_cpu(tmp_models, expected, grid):
    port = _free_port()
    srv = _Proc(["-m", "distributed_llm_pipeline_b200.serving.orchestrator", "-m", tmp_models[("tiny", "F16")], "--device", "cpu",
                 "--port", str(port), "-c", "128", "-np", "2", "-b", "32", "-n", "8", *grid], port)
    try:
        srv.wait_ready()
        d = _completion(port)
        assert d["content"] == expected[1], (d["content"], expected[1])
        # two concurrent requests share the batched decode step across all ranks
        import threading
        out = {}
        ts = [threading.Thread(target=lambda k=k: out.__setitem__(k, _completion(port))) for k in range(2)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert out[0]["content"] == expected[1] and out[1]["content"] == expected[1]
        ev = _chat_events(port, PROMPT)
        logs = "".join(e["content"] for e in ev if e["msg_type"] == "log")
        assert "offloaded" in logs and "grid:" in logs            # the UI's highlight hook, per stage / rank
        assert sum(e["msg_type"] == "token" for e in ev) >= 1
        world = (2 if "--tp" in grid else 1) * (2 if "--pp" in grid else 1)
        assert logs.count("load_tensors: stage") == world
    finally:
        srv.stop()


def test_orchestrator_over_cpu_stage_workers(tmp_models, expected):
    """BASELINE config 1 as one command: two dlp-worker CPU stages launched and supervised by the orchestrator."""
    from distributed_llm_pipeline_b200.runtime import native
    native.build()
    port, base = _free_port(), _free_port()
    srv = _Proc(["-m", "distributed_llm_pipeline_b200.serving.orchestrator", "-m", tmp_models[("tiny", "F16")], "--device", "cpu",
                 "--port", str(port), "-c", "128", "-np", "2", "-n", "8", "--workers", "2", "--worker-base-port", str(base)], port)
    try:
        srv.wait_ready()
        d = _completion(port)
        assert d["content"] == expected[1], (d["content"], expected[1])
        ev = _chat_events(port, PROMPT)
        logs = "".join(e["content"] for e in ev if e["msg_type"] == "log")
        assert logs.count("RPC worker") == 2
    finally:
        srv.stop()


def test_cli_over_process_grid_cpu(tmp_models, expected):
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "dlp_cli.py"), "-m", tmp_models[("tiny", "F16")], "--device", "cpu",
                        "-c", "128", "-p", PROMPT, "-n", "10", "--temp", "0", "--ignore-eos", "-st", "--tp", "2"],
                       cwd=REPO, env=dict(os.environ, PYTHONPATH=REPO, DLP_MASTER_PORT=str(_free_port())), capture_output=True, text=True,
                       timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.strip() == expected[1].strip(), (r.stdout, expected[1])
    assert "grid: 1 pipeline stage(s) x 2 tensor-parallel rank(s)" in r.stderr


def test_control_channel_roundtrip():
    """The shared-memory mailbox: ordered delivery to two followers, large and small messages, back-pressure."""
    import threading
    from distributed_llm_pipeline_b200.parallel.spmd import ShmChannel
    lead = ShmChannel(None, 2, None)
    got = {0: [], 1: []}

    def follower(i):
        ch = ShmChannel(lead.name, 2, i)
        while True:
            m = ch.recv(timeout_s=20)
            if m == "stop":
                break
            got[i].append(m)
        ch.close()
    ts = [threading.Thread(target=follower, args=(i,)) for i in range(2)]
    [t.start() for t in ts]
    msgs = [("decode", ([0, 1], [5, 6], True), {}), ("prefill", (0, list(range(3000)), False), {})] + [("n", (i,), {}) for i in range(200)]
    for m in msgs:
        lead.send(m)
    lead.send("stop")
    [t.join(timeout=30) for t in ts]
    lead.close()
    assert got[0] == msgs and got[1] == msgs
