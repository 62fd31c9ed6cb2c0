"""Pipeline-parallel stage runner: activation hand-off over NCCL p2p, overlapped with the next micro-batch.

Parity: the reference's layer split (`-sm layer`) turns device changes into scheduler splits whose inputs are
copied with `cudaMemcpyPeerAsync` + events (`ggml-cuda.cu:2792-2845`) and, over RPC, through two synchronous TCP
hops via the client (`ggml-backend.cpp:396-409`); micro-batch overlap exists only between ubatches of one
prefill (`llama-context.cpp:317-341`).  Here every stage is a process (or a TP group of processes); hidden
states [T, n_embd] travel rank-to-rank with `isend` / `irecv` (NCCL on GPUs, gloo on CPU), receives for all
micro-batches are posted before compute starts and sends are not waited on until the round ends, so stage s
computes micro-batch i+1 while micro-batch i is in flight to stage s+1 - for prefill chunks *and* for decode
rounds with >= 2 sequence groups.  Sampled tokens return from the last stage to the first over the same fabric.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class PipelineStage:
    WIRE = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}

    def __init__(self, engine, grid, wire_dtype: torch.dtype | None = None, max_microbatches: int = 4):
        """`wire_dtype`: what crosses the stage boundary.  Default fp16 (half the bytes of the reference's F32 hop,
        `ggml-cuda.cu:2826`; values beyond the fp16 range are clamped, not turned into inf); env DLP_PP_WIRE=f32|f16|bf16."""
        import os
        if wire_dtype is None:
            wire_dtype = self.WIRE[os.environ.get("DLP_PP_WIRE", "f16")]
        self.eng, self.grid = engine, grid
        self.prev, self.next = grid.pp_prev, grid.pp_next
        self.first, self.last = grid.first_stage, grid.last_stage
        self.wire_dtype = wire_dtype
        dev, E = engine.device, engine.cfg.n_embd
        M = engine.max_tokens
        self.rx = [torch.zeros(M, E, dtype=wire_dtype, device=dev) for _ in range(max_microbatches)]
        self.tx = [torch.zeros(M, E, dtype=wire_dtype, device=dev) for _ in range(max_microbatches)]
        self.tok_buf = torch.zeros(engine.max_seqs, dtype=torch.int32, device=dev)
        # rank of the same tp-rank in the last / first stage (token feedback ring)
        self.last_rank = (grid.pp - 1) * grid.tp + grid.tp_rank
        self.first_rank = grid.tp_rank

    # ---- helpers ----------------------------------------------------------------------------------
    def _send_hidden(self, h: torch.Tensor, slot: int):
        buf = self.tx[slot][: h.shape[0]]
        if self.wire_dtype == torch.float16 and h.dtype == torch.float32:
            h = h.clamp(-65504.0, 65504.0)
        buf.copy_(h)
        return dist.isend(buf, dst=self.next)

    def _post_recv(self, T: int, slot: int):
        return dist.irecv(self.rx[slot][:T], src=self.prev)

    # ---- prefill: chunks of one prompt are the micro-batches -----------------------------------------
    @torch.no_grad()
    def prefill(self, sid: int, tokens, chunk: int | None = None):
        """Run a prompt through all stages; returns last-token logits on the last stage (None elsewhere)."""
        eng = self.eng
        n = len(tokens)
        chunk = min(chunk or eng.max_tokens, eng.max_tokens)
        bounds = [(c0, min(n, c0 + chunk)) for c0 in range(0, n, chunk)]
        sends, logits = [], None
        nslots = len(self.rx)
        recvs = {}
        if not self.first:  # keep up to nslots receives in flight ahead of compute
            for i in range(min(nslots, len(bounds))):
                recvs[i] = self._post_recv(bounds[i][1] - bounds[i][0], i % nslots)
        for i, (c0, c1) in enumerate(bounds):
            T = c1 - c0
            hidden = None
            if not self.first:
                recvs.pop(i).wait()
                hidden = self.rx[i % nslots][:T].float()
            out = eng.prefill(sid, tokens[c0:c1], hidden_in=hidden, chunk=chunk)
            if not self.first and i + nslots < len(bounds):
                j = i + nslots
                recvs[j] = self._post_recv(bounds[j][1] - bounds[j][0], j % nslots)
            if not self.last:
                if len(sends) >= nslots:
                    sends.pop(0).wait()
                sends.append(self._send_hidden(out, i % nslots))
            else:
                logits = out
        for w in sends:
            w.wait()
        return logits

    # ---- decode: one token for each micro-batch (group of sequences) -------------------------------------
    @torch.no_grad()
    def decode_round(self, groups: list[list[int]], tokens: list | None = None, sample: bool = True, feedback: bool | None = None):
        """One decode step for every group.  `tokens[i]` (host ids) feeds group i on the first stage; when None the
        first stage receives the ids the last stage sampled in the previous round - which must have been run with
        `feedback=True` (default: feedback exactly when no tokens are given; pass feedback=True on a round that is
        seeded with host tokens but followed by self-feeding rounds).  Returns, on the last stage, a list of per-group
        results (token ids if sample else logits); elsewhere a list of None."""
        if feedback is None:
            feedback = tokens is None
        eng = self.eng
        nslots = len(self.rx)
        assert len(groups) <= nslots, "more micro-batches than p2p slots"
        recvs, sends, outs = [], [], []
        if not self.first:
            recvs = [self._post_recv(len(g), i) for i, g in enumerate(groups)]
        for i, g in enumerate(groups):
            B = len(g)
            if self.first:
                if tokens is not None:
                    t_in = tokens[i]
                else:  # ids sampled on the last stage for this group
                    if self.grid.pp > 1:
                        dist.recv(self.tok_buf[:B], src=self.last_rank)
                    t_in = self.tok_buf[:B] if self.grid.pp > 1 else None
                out = eng.decode(g, t_in, sample=sample)
            else:
                recvs[i].wait()
                out = eng.decode(g, hidden_in=self.rx[i][:B].float(), sample=sample)
            if not self.last:
                sends.append(self._send_hidden(out, i))
                outs.append(None)
            else:
                res = out.clone()
                outs.append(res)
                if sample and feedback and self.grid.pp > 1:
                    sends.append(dist.isend(res, dst=self.first_rank))
        for w in sends:
            w.wait()
        return outs

    def decode(self, sids: list[int], tokens=None, sample: bool = True):
        """Single micro-batch convenience wrapper (returns logits / ids on the last stage, None elsewhere)."""
        res = self.decode_round([sids], None if tokens is None else [tokens], sample=sample)[0]
        return res
