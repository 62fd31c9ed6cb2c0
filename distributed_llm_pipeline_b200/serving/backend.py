"""Engine service: exposes the scheduler over a local TCP socket with newline-delimited JSON.

This is the process boundary behind the native front-end (`runtime/orchestrator.cpp`).  The reference's boundary is
`tokio::process::Command` -> one `llama-cli` per request with stdout/stderr pipes (`orchestrator/src/main.rs:35-57`);
here the engine is a long-lived service:

    -> {"op":"generate","prompt":"...","n_predict":200[,"temperature":..]}\\n
    <- {"t":"log","c":"..."}\\n  {"t":"token","c":"..."}\\n ...  {"t":"done","c":"stop|length|error|cancelled"}\\n
    -> {"op":"get","path":"/health|/metrics|/slots"}\\n      <- {"c":"<body>"}\\n

Closing the connection cancels the request.  Everything the service prints to stderr is forwarded to browsers as `log`
events by the front-end (the reference's "system log" pane).
"""
from __future__ import annotations

import argparse
import json
import socketserver
import sys
import threading

from .orchestrator import add_parallel_args, build_service, maybe_reexec, service_kwargs, stop_service
from ..utils.args import lora_from_args
from .sampler import SamplingParams
from .scheduler import Request


class _Handler(socketserver.StreamRequestHandler):
    def handle(self):
        srv = self.server
        line = self.rfile.readline()
        if not line:
            return
        try:
            msg = json.loads(line)
        except json.JSONDecodeError:
            self._send({"t": "error", "c": "bad request"})
            return
        if msg.get("op") == "get":
            path = msg.get("path", "/health")
            if path == "/metrics":
                body = srv.sched.metrics.prometheus()
            elif path == "/slots":
                body = json.dumps(srv.sched.slots_info())
            else:
                body = json.dumps({"status": "ok"})
            self._send({"c": body})
            return
        if msg.get("op") == "http":
            # raw tunnel for every route the native front-end does not implement itself: the request line / headers arrive as
            # JSON, the body as `body_len` raw bytes; the full HTTP response (SSE streams included) is written back verbatim
            body = self.rfile.read(int(msg.get("body_len", 0))) if int(msg.get("body_len", 0)) > 0 else b""
            self._tunnel(msg, body)
            return
        if msg.get("op") != "generate":
            self._send({"t": "error", "c": "unknown op"})
            return
        done = threading.Event()
        lock = threading.Lock()
        alive = [True]

        def on_event(kind, content):
            with lock:
                if not alive[0]:
                    return
                try:
                    self._send({"t": kind, "c": content})
                except OSError:
                    alive[0] = False
                    req.cancelled = True
            if kind == "done":
                done.set()

        sp = SamplingParams(temperature=float(msg.get("temperature", 0.8)), top_k=int(msg.get("top_k", 40)),
                            top_p=float(msg.get("top_p", 0.95)), min_p=float(msg.get("min_p", 0.05)), seed=int(msg.get("seed", 0)))
        req = Request(prompt=msg.get("prompt", ""), n_predict=int(msg.get("n_predict", 200)), sampling=sp, on_event=on_event)
        for l in srv.startup_logs:
            self._send({"t": "log", "c": l})
        srv.sched.submit(req)
        # watch for the peer closing the socket while we stream
        self.connection.settimeout(0.25)
        while not done.is_set():
            try:
                if self.connection.recv(1) == b"":
                    req.cancelled = True
                    alive[0] = False
                    break
            except (TimeoutError, OSError) as e:
                if not isinstance(e, TimeoutError) and "timed out" not in str(e):
                    req.cancelled = True
                    alive[0] = False
                    break
        done.wait(timeout=5)

    def _send(self, obj):
        self.wfile.write(json.dumps(obj, ensure_ascii=False).encode() + b"\n")
        self.wfile.flush()

    def _tunnel(self, msg, body: bytes):
        import asyncio
        wfile = self.wfile

        class _Writer:                       # the two StreamWriter methods Orchestrator uses
            def write(self, data: bytes):
                wfile.write(data)

            async def drain(self):
                wfile.flush()

            def close(self):
                pass

        orch = self.server.orch
        target = str(msg.get("target", "/"))
        path, _, query = target.partition("?")
        headers = {str(k).lower(): str(v) for k, v in (msg.get("headers") or {}).items()}

        async def run():
            orch._query = dict(kv.partition("=")[::2] for kv in query.split("&") if kv)
            await orch.route(str(msg.get("method", "GET")).upper(), path, headers, body, None, _Writer())

        try:
            asyncio.run(run())
            wfile.flush()
        except (OSError, ValueError):
            pass


class EngineService(socketserver.ThreadingTCPServer):
    allow_reuse_address = True
    daemon_threads = True

    def __init__(self, addr, sched, startup_logs, orch=None):
        super().__init__(addr, _Handler)
        self.sched, self.startup_logs = sched, startup_logs
        self.orch = orch                     # serving.orchestrator.Orchestrator: route table for the raw HTTP tunnel


def main(argv=None):
    ap = argparse.ArgumentParser(description="dlp_b200 engine service (backend of dlp-orchestrator)")
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=3006)
    ap.add_argument("-c", "--ctx-size", type=int, default=2048)
    ap.add_argument("-np", "--parallel", type=int, default=4)
    ap.add_argument("-b", "--batch-size", type=int, default=2048)
    ap.add_argument("--device", default="cuda")
    ap.add_argument("-n", "--predict", type=int, default=200)
    ap.add_argument("--lora", default=None)
    ap.add_argument("--api-key", default=None)
    ap.add_argument("--max-queue", type=int, default=0)
    ap.add_argument("--slot-save-path", default=None)
    ap.add_argument("--context-shift", action="store_true")
    add_parallel_args(ap)
    a = ap.parse_args(argv)
    maybe_reexec(a, "distributed_llm_pipeline_b200.serving.backend", argv)
    sched, info, logs = build_service(a.model, a.device, a.ctx_size, a.parallel, a.batch_size, lora=lora_from_args(a), **service_kwargs(a))
    if not a.no_warmup:
        logs.append(f"warm-up: decode graphs for 1..{sched.n_slots} sequences captured in {sched.warmup():.2f} s\n")
    for l in logs:
        sys.stderr.write(l)
    sys.stderr.flush()
    sched.context_shift = a.context_shift
    sched.start()
    from .orchestrator import Orchestrator
    orch = Orchestrator(sched, info, n_predict=a.predict, startup_logs=logs, api_key=a.api_key, max_queue=a.max_queue,
                        slot_save_path=a.slot_save_path)
    srv = EngineService((a.host, a.port), sched, logs, orch)
    sys.stderr.write(f"engine service listening on {a.host}:{srv.server_address[1]}\n")
    sys.stderr.flush()
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        stop_service(sched)


if __name__ == "__main__":
    main()
