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

    def allgather_logits(self, logits: torch.Tensor, vocab_range, vocab_size: int) -> torch.Tensor:
        v0, v1 = vocab_range
        if v1 - v0 == vocab_size:
            return logits
        B = logits.shape[0]
        if self._gather_buf is None or self._gather_buf.shape[1] < B:
            self._gather_buf = torch.empty(self.tp_size, max(B, 8), v1 - v0, dtype=logits.dtype, device=logits.device)
        buf = self._gather_buf[:, :B].contiguous() if self._gather_buf.shape[1] != B else self._gather_buf
        dist.all_gather_into_tensor(buf.view(-1), logits.contiguous().view(-1), group=self.group)
        return buf.permute(1, 0, 2).reshape(B, vocab_size)
