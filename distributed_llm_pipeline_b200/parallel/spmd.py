"""One serving engine out of a pp x tp grid of processes: a leader that owns the scheduler, followers that replay its calls.

The reference puts its orchestrator in front of a *distributed* model: `llama-cli --rpc a,b -ngl 99` per request
(`orchestrator/src/main.rs:35-57`), or `llama-server` driving whatever `-sm layer|row` split is loaded
(`llama.cpp/tools/server/server-context.cpp:1917-2860`, `common/arg.cpp:2310-2331`).  There one process owns the whole graph and
the workers are remote-execution servers.  Here every (stage, tensor-parallel rank) is a process with its own engine shard, so
the serving core needs all of them to make the same engine call at the same time:

* the **leader** is the first rank of the LAST pipeline stage - the rank that ends up with logits / sampled ids - and runs the
  scheduler (and the HTTP front-end) against a :class:`DistributedEngine`, which has the method surface of `LlamaEngine`;
* every call that touches engine state (`prefill`, `decode`, `alloc_seq`, `truncate_seq`, ...) is first published as a small step
  descriptor on a control channel and then executed locally; **followers** sit in :func:`follower_loop`, read the descriptors
  and make the same call on their shard.  Token ids travel inside the descriptor (they pass through the host anyway for
  streaming and stop detection), activations travel stage to stage over NCCL / gloo p2p (`parallel/pp.py`), and tensor-parallel
  partial sums through the fused kernels or NCCL (`parallel/fused.py`, `parallel/comm.py`).

Control channel: a shared-memory mailbox for single-node grids (a few microseconds per step - a decode step under TP = 8 is about
a millisecond, so a 100 us gloo broadcast per step would cost 10 %), `torch.distributed` object broadcast otherwise.
"""
from __future__ import annotations

import os
import pickle
import struct
import sys
import time

import torch
import torch.distributed as dist


# ---- control channels --------------------------------------------------------------------------------------------------
class ShmChannel:
    """Single-producer mailbox in POSIX shared memory.  Layout: u64 seq | u64 nbytes | per-follower u64 ack (one cache line
    each) | payload.  The leader waits until every follower has copied message n-1 out before it writes message n."""
    HDR, LINE, CAP = 64, 64, 4 << 20

    def __init__(self, name: str | None, n_followers: int, index: int | None):
        from multiprocessing import shared_memory
        self.n, self.index = n_followers, index
        size = self.HDR + self.LINE * max(1, n_followers) + self.CAP
        if name is None:
            self.shm = shared_memory.SharedMemory(create=True, size=size)
            self.shm.buf[:self.HDR + self.LINE * max(1, n_followers)] = bytes(self.HDR + self.LINE * max(1, n_followers))
            self.owner = True
        else:
            self.shm = shared_memory.SharedMemory(name=name)
            self.owner = False
            try:   # the creator unlinks it; keep Python's resource tracker of this process from unlinking it as well
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:  # noqa: BLE001
                pass
        self.name = self.shm.name
        self.buf = self.shm.buf
        self.seq = 0
        self.pay0 = self.HDR + self.LINE * max(1, n_followers)

    def _u64(self, off: int) -> int:
        return struct.unpack_from("<Q", self.buf, off)[0]

    def send(self, obj):
        data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        if len(data) > self.CAP:
            raise ValueError(f"control message of {len(data)} bytes exceeds the mailbox")
        t0 = time.monotonic()
        for f in range(self.n):                      # everyone has taken the previous message
            while self._u64(self.HDR + self.LINE * f) < self.seq:
                if time.monotonic() - t0 > 120:
                    raise TimeoutError(f"follower {f} stopped reading the control channel")
        self.buf[self.pay0:self.pay0 + len(data)] = data
        struct.pack_into("<Q", self.buf, 8, len(data))
        self.seq += 1
        struct.pack_into("<Q", self.buf, 0, self.seq)     # publish (x86 keeps the store order)

    def recv(self, timeout_s: float | None = None):
        want = self.seq + 1
        t0, spins = time.monotonic(), 0
        while self._u64(0) < want:
            spins += 1
            if spins > 2000:                          # idle server: stop burning the core
                time.sleep(0.0002)
                if timeout_s is not None and time.monotonic() - t0 > timeout_s:
                    raise TimeoutError("no control message")
        n = self._u64(8)
        obj = pickle.loads(bytes(self.buf[self.pay0:self.pay0 + n]))
        self.seq = want
        struct.pack_into("<Q", self.buf, self.HDR + self.LINE * self.index, want)
        return obj

    def close(self):
        try:
            self.buf = None
            self.shm.close()
            if self.owner:
                self.shm.unlink()
        except Exception:  # noqa: BLE001
            pass


class DistChannel:
    """Object broadcast over a torch.distributed (gloo) group: works across nodes, ~100 us per message."""

    def __init__(self, group, src: int):
        self.group, self.src = group, src

    def send(self, obj):
        dist.broadcast_object_list([obj], src=self.src, group=self.group)

    def recv(self, timeout_s=None):
        box = [None]
        dist.broadcast_object_list(box, src=self.src, group=self.group)
        return box[0]

    def close(self):
        pass


def make_channel(grid, leader: int, kind: str | None = None):
    """Every rank calls this once (collective).  kind: "shm" (default on one node) | "dist"."""
    kind = kind or os.environ.get("DLP_CTRL", "shm")
    if grid.world == 1:
        return None
    if kind == "shm":
        followers = [r for r in range(grid.world) if r != leader]
        box = [None]
        if grid.rank == leader:
            ch = ShmChannel(None, len(followers), None)
            box[0] = ch.name
        dist.broadcast_object_list(box, src=leader)
        if grid.rank != leader:
            ch = ShmChannel(box[0], len(followers), followers.index(grid.rank))
        dist.barrier()
        return ch
    return DistChannel(dist.new_group(list(range(grid.world)), backend="gloo"), leader)


# ---- the same engine call on every rank -----------------------------------------------------------------------------------
class LocalExec:
    """What one rank does for each engine-level operation (through the pipeline stage runner when pp > 1)."""

    def __init__(self, engine, grid, stage=None):
        self.eng, self.grid, self.stage = engine, grid, stage

    def prefill(self, sid, tokens, all_logits=False):
        if self.stage is not None:
            return self.stage.prefill(sid, tokens)
        return self.eng.prefill(sid, tokens, all_logits=all_logits)

    def decode(self, sids, tokens, sample):
        if self.stage is not None:
            return self.stage.decode_round([sids], [tokens], sample=sample, feedback=False)[0]
        return self.eng.decode(sids, tokens, sample=sample)

    def call(self, name, *args, **kw):
        return getattr(self.eng, name)(*args, **kw)


class DistributedEngine:
    """Leader-side engine with the `LlamaEngine` surface the serving core uses; every state-changing call is replayed by the
    followers (see module docstring)."""
    REPLAYED = ("alloc_seq", "free_seq", "truncate_seq", "copy_seq", "shift_seq", "set_seq_sampling", "set_control_vector")

    def __init__(self, engine, grid, stage, channel):
        self.local = LocalExec(engine, grid, stage)
        self.eng, self.grid, self.ch = engine, grid, channel
        self.cfg, self.device = engine.cfg, engine.device
        self.n_ctx, self.max_seqs, self.max_tokens, self.page_size = engine.n_ctx, engine.max_seqs, engine.max_tokens, engine.page_size
        self.kv = engine.kv
        self.steps = 0

    # bookkeeping that is identical on every rank: answered locally
    @property
    def sampling(self):
        return self.eng.sampling

    @sampling.setter
    def sampling(self, v):
        self._replay("_set_sampling", dict(v))

    @property
    def _free_sids(self):
        return self.eng._free_sids

    def seq_len(self, sid):
        return self.eng.seq_len(sid)

    def memory_breakdown(self):
        return self.eng.memory_breakdown()

    def _replay(self, name, *args, **kw):
        if self.ch is not None:
            self.ch.send((name, args, kw))
        if name == "_set_sampling":
            self.eng.sampling = args[0]
            return None
        return self.local.call(name, *args, **kw)

    def __getattr__(self, name):
        if name in self.REPLAYED:
            return lambda *a, **k: self._replay(name, *a, **k)
        raise AttributeError(name)

    def prefill(self, sid, tokens, hidden_in=None, chunk=None, all_logits=False):
        if all_logits and (self.local.stage is not None or getattr(self.eng, "sp_ok", False)):
            raise NotImplementedError("per-position logits are not available on the pipelined / sequence-parallel prefill path")
        toks = [int(t) for t in (tokens.tolist() if hasattr(tokens, "tolist") else tokens)]
        if self.ch is not None:
            self.ch.send(("prefill", (sid, toks, all_logits), {}))
        return self.local.prefill(sid, toks, all_logits)

    def decode(self, sids, tokens=None, hidden_in=None, sample=True, use_graph=True):
        toks = None if tokens is None else [int(t) for t in (tokens.tolist() if hasattr(tokens, "tolist") else tokens)]
        if toks is None and self.local.stage is not None:
            raise ValueError("a pipelined engine is fed its tokens by the host")
        if self.ch is not None:
            self.ch.send(("decode", (list(sids), toks, bool(sample)), {}))
        self.steps += 1
        return self.local.decode(list(sids), toks, bool(sample))

    def embed(self, *a, **k):
        raise NotImplementedError("embeddings run on a single-process engine")

    def shutdown(self):
        if self.ch is not None:
            self.ch.send(("stop", (), {}))
            self.ch.close()


def follower_loop(engine, grid, stage, channel, log=None) -> int:
    """Replay the leader's calls until it says stop.  Returns the number of steps executed."""
    local = LocalExec(engine, grid, stage)
    n = 0
    while True:
        name, args, kw = channel.recv()
        if name == "stop":
            break
        try:
            if name == "prefill":
                local.prefill(*args)
            elif name == "decode":
                local.decode(*args)
            elif name == "_set_sampling":
                engine.sampling = args[0]
            else:
                local.call(name, *args, **kw)
        except (RuntimeError, ValueError, NotImplementedError) as e:   # the leader hit the same error on identical state
            if log:
                log(f"rank {grid.rank}: {name} failed like on the leader: {e}\n")
        n += 1
    channel.close()
    return n


# ---- construction ------------------------------------------------------------------------------------------------------
def leader_rank(pp: int, tp: int) -> int:
    return (pp - 1) * tp


def build_engine(model: str, grid, n_ctx: int, max_seqs: int, n_batch: int, page_size: int, tp_impl: str = "fused", lora=None,
                 tensor_split=None):
    """This rank's shard + engine (+ pipeline stage runner).  Returns (engine, stage, comm, load_seconds)."""
    from ..gguf import open_gguf
    from ..models.config import config_from_gguf
    from ..models.llama import LlamaEngine
    from ..models.weights import load_weights
    from .grid import stage_layer_ranges
    f = open_gguf(model) if isinstance(model, str) else model
    cfg = config_from_gguf(f.metadata, f.tensors)
    comm = None
    if grid.tp > 1:
        if tp_impl == "fused" and grid.device.type == "cuda" and not cfg.is_moe:
            from .fused import FusedComm
            comm = FusedComm(grid)
        else:
            from .comm import NcclComm
            comm = NcclComm(grid.tp_group, grid.tp, grid.tp_rank)
    ranges = stage_layer_ranges(cfg.n_layer, grid.pp, tensor_split)
    t0 = time.time()
    w = load_weights(f, grid.device, layer_range=ranges[grid.stage], tp_rank=grid.tp_rank, tp_size=grid.tp,
                     first_stage=grid.first_stage, last_stage=grid.last_stage, lora=lora)
    eng = LlamaEngine(w, n_ctx=n_ctx, max_seqs=max_seqs, page_size=page_size, max_batch_tokens=n_batch, comm=comm,
                      first_stage=grid.first_stage, last_stage=grid.last_stage)
    stage = None
    if grid.pp > 1:
        from .pp import PipelineStage
        stage = PipelineStage(eng, grid)
    return eng, stage, comm, time.time() - t0, ranges


def reexec_under_torchrun(world: int, argv: list[str], module: str | None = None, script: str | None = None):
    """A plain `python -m ... --tp 2` start becomes `torchrun --nproc-per-node world ...` of the same command."""
    port = int(os.environ.get("DLP_MASTER_PORT", "0")) or (29500 + os.getpid() % 2000)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port)]
    cmd += (["-m", module] if module else [script]) + list(argv)
    os.execv(sys.executable, cmd)
