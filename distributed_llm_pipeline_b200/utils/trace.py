"""Tracing / profiling hooks: CUDA-event phase timers, NVTX ranges, chrome-trace export, perf counters.

Parity: the reference exposes wall-clock perf counters (`llama_perf_context`: load / prompt-eval / eval ms, token counts,
`llama.h:1473-1498`), a scheduler debug dump (`GGML_SCHED_DEBUG`) and an eval callback (`ggml-backend.cpp:1584-1616`);
it has no NVTX / chrome-trace output.  Here every phase is timed on the device with CUDA events (max over
ranks when asked), optionally wrapped in NVTX ranges for Nsight, and can be exported as a chrome://tracing JSON.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from collections import defaultdict

import torch


class PerfCounters:
    """Equivalent of llama_perf_context_data."""

    def __init__(self):
        self.t_load_ms = 0.0
        self.t_p_eval_ms = 0.0
        self.t_eval_ms = 0.0
        self.n_p_eval = 0
        self.n_eval = 0
        self.n_graph_replays = 0

    def report(self) -> str:
        def rate(n, ms):
            return f"{n / ms * 1e3:10.2f} tokens per second" if ms > 0 else "       n/a"
        return (f"perf:        load time = {self.t_load_ms:10.2f} ms\n"
                f"perf: prompt eval time = {self.t_p_eval_ms:10.2f} ms / {self.n_p_eval:5d} tokens ({rate(self.n_p_eval, self.t_p_eval_ms)})\n"
                f"perf:        eval time = {self.t_eval_ms:10.2f} ms / {self.n_eval:5d} runs   ({rate(self.n_eval, self.t_eval_ms)})\n"
                f"perf:    graph replays = {self.n_graph_replays}\n")


class Tracer:
    def __init__(self, enabled: bool | None = None, nvtx: bool | None = None):
        self.enabled = bool(int(os.environ.get("DLP_TRACE", "0"))) if enabled is None else enabled
        self.nvtx = bool(int(os.environ.get("DLP_NVTX", "0"))) if nvtx is None else nvtx
        self.events: list[dict] = []
        self._pending: list[tuple] = []
        self.totals: dict[str, float] = defaultdict(float)
        self.counts: dict[str, int] = defaultdict(int)
        self._t0 = time.perf_counter()

    @contextlib.contextmanager
    def span(self, name: str, device_timed: bool = True, **args):
        if not self.enabled and not self.nvtx:
            yield
            return
        cuda = device_timed and torch.cuda.is_available()
        if self.nvtx and cuda:
            torch.cuda.nvtx.range_push(name)
        t_host = time.perf_counter()
        e0 = e1 = None
        if self.enabled and cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        try:
            yield
        finally:
            if e1 is not None:
                e1.record()
                self._pending.append((name, t_host, e0, e1, args))
            elif self.enabled:
                dur = time.perf_counter() - t_host
                self._add(name, t_host, dur * 1e3, args)
            if self.nvtx and cuda:
                torch.cuda.nvtx.range_pop()

    def _add(self, name, t_host, dur_ms, args):
        self.events.append({"name": name, "ph": "X", "pid": os.getpid(), "tid": 0, "ts": (t_host - self._t0) * 1e6, "dur": dur_ms * 1e3,
                            "args": args})
        self.totals[name] += dur_ms
        self.counts[name] += 1

    def resolve(self):
        """Turn pending CUDA-event pairs into durations (synchronises)."""
        if self._pending:
            torch.cuda.synchronize()
            for name, t_host, e0, e1, args in self._pending:
                self._add(name, t_host, e0.elapsed_time(e1), args)
            self._pending.clear()

    def summary(self) -> dict:
        self.resolve()
        return {k: {"total_ms": round(v, 3), "count": self.counts[k], "mean_ms": round(v / self.counts[k], 4)} for k, v in self.totals.items()}

    def export_chrome(self, path: str):
        self.resolve()
        with open(path, "w") as f:
            json.dump({"traceEvents": self.events, "displayTimeUnit": "ms"}, f)
        return path


def max_over_ranks(ms: float, device=None) -> float:
    """Device-timed multi-GPU numbers are reported as the max over ranks."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device or ("cuda" if torch.cuda.is_available() else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
