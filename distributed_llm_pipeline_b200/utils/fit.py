"""Fit the run-time sizes to the device memory (`--fit`) and list devices (`--list-devices`).

Parity: `llama_params_fit` (`llama.cpp/src/llama.cpp:161-745`, flag `arg.cpp:2369-2389`) shrinks `-c` / `-ngl` / `-ts` until the model
and its KV cache fit; here every layer is always device resident, so the free variables are the context length, the number of
sequence slots and - when even that is not enough - the tensor-parallel degree that would be needed.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass
class FitResult:
    model_bytes: int
    kv_bytes_per_token: int
    n_ctx: int
    n_slots: int
    fits: bool
    min_tp: int          # smallest tensor-parallel degree (power of two <= n_head_kv) whose per-rank share fits
    note: str


def model_bytes(files) -> int:
    return sum(t.nbytes for t in files.tensors.values())


def fit_params(files, cfg, n_ctx: int, n_slots: int, free_bytes: int, reserve: float = 0.08, page: int = 64) -> FitResult:
    """Largest (n_ctx, n_slots) <= the requested ones such that weights + paged KV + activations fit into `free_bytes`."""
    mb = model_bytes(files)
    kv_tok = 2 * cfg.n_layer * cfg.kv_dim * 2                      # K and V, fp16, all layers
    act = 64 << 20                                                  # activation / logits / workspace head-room
    budget = int(free_bytes * (1.0 - reserve)) - mb - act
    want_ctx, want_slots = n_ctx, n_slots
    note = "requested sizes fit"
    if budget < kv_tok * page:
        tp = 1
        while tp < max(1, cfg.n_head_kv) and int(free_bytes * (1.0 - reserve)) - mb // (tp * 2) - act < kv_tok * page // (tp * 2):
            tp *= 2
        return FitResult(mb, kv_tok, 0, 0, False, tp * 2, f"weights ({mb / 2**30:.1f} GiB) do not fit next to a KV page: use --tp {tp * 2} or a smaller file type")
    while n_slots > 1 and kv_tok * n_ctx * n_slots > budget:
        n_slots -= 1
    if kv_tok * n_ctx * n_slots > budget:
        n_ctx = max(page, budget // (kv_tok * n_slots) // page * page)
    if (n_ctx, n_slots) != (want_ctx, want_slots):
        note = f"reduced to n_ctx {n_ctx} x {n_slots} slots (KV {kv_tok * n_ctx * n_slots / 2**30:.2f} GiB, weights {mb / 2**30:.2f} GiB)"
    return FitResult(mb, kv_tok, n_ctx, n_slots, True, 1, note)


def list_devices() -> list[str]:
    import torch
    out = []
    if torch.cuda.is_available():
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            free, total = torch.cuda.mem_get_info(i)
            out.append(f"CUDA{i}: {p.name} (sm_{p.major}{p.minor}, {p.multi_processor_count} SMs, {total / 2**30:.1f} GiB, {free / 2**30:.1f} GiB free)")
    import os
    out.append(f"CPU: {os.cpu_count()} cores (dlp-worker stages, --rpc)")
    return out
