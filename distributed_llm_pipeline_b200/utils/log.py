"""Logging: levels, timestamps, optional file sink, in-memory ring (feeds the orchestrator's `log` SSE events).

Parity: `common_log` (`llama.cpp/common/log.cpp:120-160,396-423`): asynchronous worker + ring of entries, levels
OUTPUT/ERROR/WARN/INFO/DEBUG, `--log-file`, `-v/--verbose`, `-lv N`, timestamps / prefixes.  The ring doubles as the source
of the orchestrator's system-log pane (the reference forwards llama-cli's stderr, `orchestrator/src/main.rs:62-80`).
"""
from __future__ import annotations

import queue
import sys
import threading
import time
from collections import deque

LEVELS = {"none": 0, "error": 1, "warn": 2, "info": 3, "debug": 4}


class Logger:
    def __init__(self, verbosity: int = 3, timestamps: bool = True, prefix: bool = True, ring: int = 512):
        self.verbosity, self.timestamps, self.prefix = verbosity, timestamps, prefix
        self.ring: deque[str] = deque(maxlen=ring)
        self.sinks: list = [sys.stderr]
        self.listeners: list = []
        self._q: queue.Queue = queue.Queue()
        self._t0 = time.time()
        self._thread = threading.Thread(target=self._worker, name="dlp-log", daemon=True)
        self._thread.start()

    def add_file(self, path: str):
        self.sinks.append(open(path, "a", buffering=1))

    def subscribe(self, fn):
        """fn(line) is called from the log thread for every entry (used to mirror logs into SSE streams)."""
        self.listeners.append(fn)

    def _worker(self):
        while True:
            line = self._q.get()
            if line is None:
                break
            self.ring.append(line)
            for s in self.sinks:
                try:
                    s.write(line)
                    s.flush()
                except (OSError, ValueError):
                    pass
            for fn in list(self.listeners):
                try:
                    fn(line)
                except Exception:
                    self.listeners.remove(fn)

    def log(self, level: str, msg: str):
        if LEVELS[level] > self.verbosity:
            return
        head = ""
        if self.timestamps:
            head += f"{time.time() - self._t0:9.3f} "
        if self.prefix:
            head += {"error": "E ", "warn": "W ", "info": "I ", "debug": "D "}.get(level, "")
        self._q.put(head + msg + ("" if msg.endswith("\n") else "\n"))

    def error(self, m): self.log("error", m)
    def warn(self, m): self.log("warn", m)
    def info(self, m): self.log("info", m)
    def debug(self, m): self.log("debug", m)

    def flush(self, timeout: float = 1.0):
        t = time.time()
        while not self._q.empty() and time.time() - t < timeout:
            time.sleep(0.005)

    def close(self):
        self._q.put(None)


_default: Logger | None = None


def get_logger() -> Logger:
    global _default
    if _default is None:
        _default = Logger()
    return _default
