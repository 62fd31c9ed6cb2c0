"""Tensor-parallel communicators plugged into LlamaEngine.

`NcclComm` is the *baseline* path (separate GEMM kernel, then an NCCL collective) that the north star tells us
to keep for A/B runs; `FusedComm` (parallel/fused.py) replaces the row-parallel GEMM + all-reduce pair with
kernels that push partial tiles straight into peer HBM over NVLink.  Reference counterpart: the hub-and-spoke
`cudaMemcpyPeerAsync` broadcast/gather around every split matmul (`ggml-cuda.cu:1680-1744`).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class NcclComm:
    fused = False

    def __init__(self, group, tp_size: int, tp_rank: int):
        self.group, self.tp_size, self.tp_rank = group, tp_size, tp_rank
        self._gather_buf: torch.Tensor | None = None

    def allreduce_delta(self, delta: torch.Tensor, il: int, which: int) -> torch.Tensor:
        dist.all_reduce(delta, group=self.group)
        return delta

    def reserve_logits(self, rows: int, v_local: int, dtype, device):
        """Persistent gather buffer for up to `rows` decode rows.  Captured CUDA graphs keep its address, so it is allocated once
        (by the engine, for its largest decode batch) and never replaced; larger one-off gathers use a temporary."""
        need = self.tp_size * max(rows, 8) * v_local
        if self._gather_buf is None or self._gather_buf.numel() < need or self._gather_buf.dtype != dtype:
            self._gather_buf = torch.empty(need, dtype=dtype, device=device)

    def allgather_logits(self, logits: torch.Tensor, vocab_range, vocab_size: int) -> torch.Tensor:
        v0, v1 = vocab_range
        if v1 - v0 == vocab_size:
            return logits
        B, V = logits.shape[0], v1 - v0
        if self._gather_buf is None:
            self.reserve_logits(B, V, logits.dtype, logits.device)
        n = self.tp_size * B * V
        if n <= self._gather_buf.numel() and self._gather_buf.dtype == logits.dtype:
            flat = self._gather_buf[:n]              # a view: the same address for a given batch size, step after step
        else:       # more rows than any decode batch (per-position logits of a prompt): never touch the persistent buffer
            flat = torch.empty(n, dtype=logits.dtype, device=logits.device)
        dist.all_gather_into_tensor(flat, logits.contiguous().view(-1), group=self.group)
        return flat.view(self.tp_size, B, V).permute(1, 0, 2).reshape(B, vocab_size)
