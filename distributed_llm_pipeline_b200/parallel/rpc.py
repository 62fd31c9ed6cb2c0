"""Client side of the native CPU workers (`runtime/worker.cpp`): chain N workers into a layer-split pipeline.

Parity: this is the role `llama-cli --rpc host:port,host:port` plays in the reference (`orchestrator/src/main.rs:47-48`,
`ggml-rpc.cpp`), with a compact protocol that ships activations only - no tensors, no graphs.  `RpcEngine` offers the
same interface as `LlamaEngine` (alloc/free/prefill/decode/truncate), so the scheduler and both orchestrators can
serve through workers unchanged.
"""
from __future__ import annotations

import os
import socket
import struct
import subprocess
import time

import numpy as np
import torch

PROTO_VERSION = 1
HELLO, RUN, RESET, TRUNCATE, SHUTDOWN = 1, 2, 3, 4, 5
_HDR = struct.Struct("<BIIIIQ")
_INFO = struct.Struct("<9I")


class WorkerError(RuntimeError):
    pass


class WorkerClient:
    def __init__(self, host: str, port: int, timeout: float = 600.0, connect_retries: int = 100):
        last = None
        for _ in range(connect_retries):
            try:
                self.sock = socket.create_connection((host, port), timeout=timeout)
                break
            except OSError as e:
                last = e
                time.sleep(0.1)
        else:
            raise WorkerError(f"cannot reach worker {host}:{port}: {last}")
        self.sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        self.addr = f"{host}:{port}"
        raw = self._call(HELLO, 0, 0, 0, 0, struct.pack("<I", PROTO_VERSION))
        (self.version, self.n_embd, self.n_vocab, self.n_layer, self.l0, self.l1, self.first, self.last, self.n_ctx) = _INFO.unpack(raw)

    def _recv(self, n: int) -> bytes:
        buf = bytearray()
        while len(buf) < n:
            chunk = self.sock.recv(min(n - len(buf), 1 << 20))
            if not chunk:
                raise WorkerError(f"worker {self.addr} closed the connection")   # failure detection: clean error, no abort
            buf += chunk
        return bytes(buf)

    def _call(self, cmd, seq, n_tokens, pos0, flags, payload: bytes = b"") -> bytes:
        try:
            self.sock.sendall(_HDR.pack(cmd, seq, n_tokens, pos0, flags, len(payload)) + payload)
            status, n = struct.unpack("<IQ", self._recv(12))
            data = self._recv(n) if n else b""
        except OSError as e:
            raise WorkerError(f"worker {self.addr}: {e}") from e
        if status != 0:
            raise WorkerError(f"worker {self.addr}: {data.decode(errors='replace')}")
        return data

    def run(self, seq: int, n_tokens: int, pos0: int, payload: bytes, all_logits: bool = False) -> np.ndarray:
        return np.frombuffer(self._call(RUN, seq, n_tokens, pos0, 1 if all_logits else 0, payload), dtype=np.float32)

    def reset(self, seq: int):
        self._call(RESET, seq, 0, 0, 0)

    def truncate(self, seq: int, length: int):
        self._call(TRUNCATE, seq, 0, length, 0)

    def shutdown(self):
        try:
            self._call(SHUTDOWN, 0, 0, 0, 0)
        except WorkerError:
            pass
        self.sock.close()

    def close(self):
        self.sock.close()


class _KvStub:
    n_pages = 0
    free: list = []

    def bytes(self):
        return 0


class RpcEngine:
    """LlamaEngine-shaped facade over a chain of workers (stage i feeds stage i+1)."""

    def __init__(self, endpoints: list[str], n_ctx: int | None = None, max_seqs: int = 4, recover: bool = False,
                 recover_wait_s: float = 30.0):
        """`recover=True`: when a worker dies mid-request (connection reset / closed) the engine waits up to `recover_wait_s`
        for the endpoint to come back (a supervisor such as `LocalWorkerPool` restarts it), reconnects the whole chain, replays
        the token history of every live sequence to rebuild the workers' KV caches and repeats the failed call - the caller
        only sees a slower step.  The reference aborts the client process in this situation (`ggml-rpc.cpp:66`)."""
        self.endpoints = list(endpoints)
        self.recover, self.recover_wait_s = recover, recover_wait_s
        self.history: dict[int, list[int]] = {}
        self.n_recoveries = 0
        self.workers = []
        for ep in endpoints:
            host, _, port = ep.rpartition(":")
            self.workers.append(WorkerClient(host or "127.0.0.1", int(port)))
        w0, wl = self.workers[0], self.workers[-1]
        if not w0.first or not wl.last:
            raise WorkerError("worker chain must start at layer 0 and end at the last layer")
        for a, b in zip(self.workers, self.workers[1:]):
            if a.l1 != b.l0:
                raise WorkerError(f"layer ranges are not contiguous: [{a.l0},{a.l1}) then [{b.l0},{b.l1})")
        self.n_ctx = min([w.n_ctx for w in self.workers] + ([n_ctx] if n_ctx else []))
        self.max_seqs, self.max_tokens = max_seqs, self.n_ctx
        self.vocab = wl.n_vocab
        self.device = torch.device("cpu")
        self.kv = _KvStub()
        self.lens: dict[int, int] = {}
        self._free = list(range(max_seqs - 1, -1, -1))
        self.sampling: dict = {}

    def describe(self) -> list[str]:
        return [f"RPC worker {w.addr}: layers [{w.l0}, {w.l1}) of {w.n_layer}{' +embedding' if w.first else ''}{' +lm_head' if w.last else ''}\n"
                for w in self.workers]

    def alloc_seq(self) -> int:
        sid = self._free.pop()
        self.lens[sid] = 0
        self.history[sid] = []
        return sid

    def free_seq(self, sid: int):
        self._guard(lambda: [w.reset(sid) for w in self.workers], skip_sid=sid)
        del self.lens[sid]
        self.history.pop(sid, None)
        self._free.append(sid)

    # ---- failure handling ---------------------------------------------------------------------------------------
    def _reconnect_and_replay(self, skip_sid: int | None = None):
        deadline = time.time() + self.recover_wait_s
        for w in self.workers:
            try:
                w.close()
            except OSError:
                pass
        workers = []
        for ep in self.endpoints:
            host, _, port = ep.rpartition(":")
            while True:
                try:
                    workers.append(WorkerClient(host or "127.0.0.1", int(port), connect_retries=5))
                    break
                except WorkerError:
                    if time.time() > deadline:
                        raise
                    time.sleep(0.2)
        self.workers = workers
        self.n_recoveries += 1
        for sid, hist in self.history.items():          # surviving workers still hold stale caches: reset everywhere, then replay
            for w in self.workers:
                w.reset(sid)
            n = self.lens.get(sid, 0) if sid != skip_sid else 0
            self.lens[sid] = 0
            if n:
                self._forward_once(sid, hist[:n])

    def _guard(self, fn, skip_sid: int | None = None):
        try:
            return fn()
        except WorkerError:
            if not self.recover:
                raise
            self._reconnect_and_replay(skip_sid)
            return fn()

    def seq_len(self, sid: int) -> int:
        return self.lens[sid]

    def truncate_seq(self, sid: int, new_len: int):
        new_len = max(0, min(new_len, self.lens[sid]))
        self.lens[sid] = new_len                          # (set first: a recovery replays exactly the kept prefix)
        del self.history[sid][new_len:]
        self._guard(lambda: [w.truncate(sid, new_len) for w in self.workers])

    def _forward(self, sid: int, tokens, all_logits=False) -> torch.Tensor:
        toks = [int(t) for t in np.asarray(tokens, dtype=np.int32).reshape(-1)]
        out = self._guard(lambda: self._forward_once(sid, toks, all_logits))
        self.history[sid][self.lens[sid] - len(toks):] = toks
        return out

    def _forward_once(self, sid: int, tokens, all_logits=False) -> torch.Tensor:
        toks = np.asarray(tokens, dtype=np.int32).reshape(-1)
        T, pos0 = toks.size, self.lens[sid]
        data = toks.tobytes()
        for w in self.workers:
            out = w.run(sid, T, pos0, data, all_logits)
            data = out.tobytes()
        self.lens[sid] = pos0 + T
        return torch.from_numpy(out.copy()).view(-1, self.vocab)

    @torch.no_grad()
    def prefill(self, sid: int, tokens, hidden_in=None, chunk=None) -> torch.Tensor:
        return self._forward(sid, tokens)[-1]

    @torch.no_grad()
    def decode(self, sids: list[int], tokens=None, hidden_in=None, sample: bool = True, use_graph: bool = True):
        toks = torch.as_tensor(tokens, dtype=torch.int32).tolist()
        logits = torch.stack([self._forward(s, [t])[-1] for s, t in zip(sids, toks)])
        if sample:
            return logits.argmax(-1).to(torch.int32)
        return logits

    def close(self):
        for w in self.workers:
            w.close()


def free_port_run(n: int = 2, host: str = "127.0.0.1") -> int:
    """First port of a run of `n` consecutive ports that are free right now (the local launchers number their workers base, base+1, ...)."""
    for _ in range(200):
        probe = socket.socket()
        probe.bind((host, 0))
        base = probe.getsockname()[1]
        probe.close()
        socks = []
        try:
            for i in range(n):
                s = socket.socket()
                socks.append(s)
                s.bind((host, base + i))
            return base
        except OSError:
            continue
        finally:
            for s in socks:
                s.close()
    raise WorkerError(f"no run of {n} free ports on {host}")


def launch_local_workers(model: str, n_workers: int, base_port: int = 50052, n_ctx: int = 2048, n_layer: int | None = None,
                         threads: int | None = None, verbose: bool = False, sanitize: str | None = None,
                         stderr=None) -> tuple[list[subprocess.Popen], list[str]]:
    """Start `n_workers` dlp-worker processes on localhost with an even layer split (the reference has no such script:
    `rpc-server` instances must be started by hand).  Returns (processes, endpoints)."""
    from ..gguf import open_gguf
    from ..models.config import config_from_gguf
    from ..runtime import native
    from .grid import stage_layer_ranges
    native.build(sanitize=sanitize)
    if n_layer is None:
        f = open_gguf(model)
        n_layer = config_from_gguf(f.metadata, f.tensors).n_layer
        f.close()
    procs, eps = [], []
    env = dict(os.environ)
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    for i, r in enumerate(stage_layer_ranges(n_layer, n_workers)):
        port = base_port + i
        cmd = [native.path("dlp-worker", sanitize), "-m", model, "--layers", f"{r.start}:{r.stop}", "-p", str(port), "-c", str(n_ctx)]
        if verbose:
            cmd.append("-v")
        procs.append(subprocess.Popen(cmd, env=env, stderr=stderr))
        eps.append(f"127.0.0.1:{port}")
    return procs, eps


class LocalWorkerPool:
    """Supervised local workers ("auto-heal", proposed but not implemented in the reference's design report, PDF p.7): a monitor
    thread restarts any dlp-worker that exits, with the same layer range and port, so that `RpcEngine(recover=True)` can
    reconnect and replay."""

    def __init__(self, model: str, n_workers: int, base_port: int = 50052, n_ctx: int = 2048, threads: int | None = None):
        import threading
        from ..gguf import open_gguf
        from ..models.config import config_from_gguf
        from ..runtime import native
        from .grid import stage_layer_ranges
        native.build()
        f = open_gguf(model)
        n_layer = config_from_gguf(f.metadata, f.tensors).n_layer
        f.close()
        self.env = dict(os.environ)
        if threads:
            self.env["OMP_NUM_THREADS"] = str(threads)
        self.cmds, self.endpoints = [], []
        for i, r in enumerate(stage_layer_ranges(n_layer, n_workers)):
            self.cmds.append([native.path("dlp-worker"), "-m", model, "--layers", f"{r.start}:{r.stop}", "-p", str(base_port + i), "-c", str(n_ctx)])
            self.endpoints.append(f"127.0.0.1:{base_port + i}")
        self.procs = [subprocess.Popen(c, env=self.env) for c in self.cmds]
        self.restarts = 0
        self._stop = False
        self._thread = threading.Thread(target=self._monitor, daemon=True)
        self._thread.start()

    def _monitor(self):
        while not self._stop:
            for i, p in enumerate(self.procs):
                if p.poll() is not None and not self._stop:
                    self.procs[i] = subprocess.Popen(self.cmds[i], env=self.env)
                    self.restarts += 1
            time.sleep(0.1)

    def stop(self):
        self._stop = True
        self._thread.join(timeout=2)
        for p in self.procs:
            if p.poll() is None:
                p.terminate()
        for p in self.procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
