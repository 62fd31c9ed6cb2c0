"""Process grid (pp x tp) and layer->stage assignment.

Parity: the reference assigns layer `il` to the device whose cumulative split fraction first exceeds
(il / n_layer) - `llama.cpp/src/llama-model.cpp:2492-2537` (fractions from `--tensor-split` or free memory) - and
can only express "all GPUs row-split" for TP (`-sm row`).  Here the grid is explicit: world = pp x tp, ranks of
one stage are contiguous (so a TP group sits on one NVSwitch domain), and stage boundaries follow the same
upper-bound rule over user weights (default: even).
"""
from __future__ import annotations

import bisect
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


def stage_layer_ranges(n_layer: int, pp: int, split: list[float] | None = None) -> list[range]:
    """Contiguous layer ranges per stage; `split` are relative weights like the reference's --tensor-split."""
    w = [1.0] * pp if not split else [float(x) for x in split]
    if len(w) != pp or any(x < 0 for x in w) or sum(w) <= 0:
        raise ValueError(f"bad stage split {split} for pp={pp}")
    tot, acc, cum = sum(w), 0.0, []
    for x in w:
        acc += x
        cum.append(acc / tot)
    cum[-1] = 1.0
    owner = [min(pp - 1, bisect.bisect_right(cum, il / n_layer)) for il in range(n_layer)]
    ranges = []
    for s in range(pp):
        ls = [il for il in range(n_layer) if owner[il] == s]
        ranges.append(range(ls[0], ls[-1] + 1) if ls else range(0, 0))
    return ranges


@dataclass
class Grid:
    rank: int
    world: int
    pp: int
    tp: int
    device: torch.device
    tp_group: object = None
    pp_prev: int | None = None   # global rank of the same tp-rank in the previous stage
    pp_next: int | None = None

    @property
    def stage(self) -> int:
        return self.rank // self.tp

    @property
    def tp_rank(self) -> int:
        return self.rank % self.tp

    @property
    def first_stage(self) -> bool:
        return self.stage == 0

    @property
    def last_stage(self) -> bool:
        return self.stage == self.pp - 1


def init_grid(pp: int = 1, tp: int | None = None, backend: str | None = None, device: str | None = None,
              timeout_s: float | None = None) -> Grid:
    """Initialise torch.distributed from the torchrun environment (or a single-process grid)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if tp is None:
        tp = world // pp
    if pp * tp != world:
        raise ValueError(f"pp({pp}) x tp({tp}) != world size {world}")
    use_cuda = torch.cuda.is_available() and (device is None or device.startswith("cuda"))
    dev = torch.device(f"cuda:{local}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kw = {"device_id": dev} if use_cuda else {}
        timeout_s = timeout_s or float(os.environ.get("DLP_DIST_TIMEOUT_S", "600"))
        import datetime
        kw["timeout"] = datetime.timedelta(seconds=timeout_s)   # a dead / diverged peer surfaces as an error, not a hang
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world, **kw)
    g = Grid(rank, world, pp, tp, dev)
    if world > 1:
        for s in range(pp):  # every rank must create every group
            ranks = list(range(s * tp, (s + 1) * tp))
            grp = dist.new_group(ranks) if tp > 1 else None
            if s == g.stage:
                g.tp_group = grp
        if g.stage > 0:
            g.pp_prev = g.rank - tp
        if g.stage < pp - 1:
            g.pp_next = g.rank + tp
    return g


def shutdown(code: int = 0):
    """Leave a multi-rank job without running NCCL / CUDA teardown.

    Measured on this image (torch 2.11, NCCL 2.28.9, driver 580): `dist.destroy_process_group()` - and plain interpreter
    exit with live NCCL communicators - can block forever after all useful work is done, which turns a 3-second job into a
    hung box.  All ranks synchronise once more, flush their output and then leave with os._exit()."""
    import sys
    try:
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.barrier()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
    except Exception:
        pass
    sys.stdout.flush()
    sys.stderr.flush()
    if dist.is_initialized() and dist.get_world_size() > 1:
        os._exit(code)
    sys.exit(code)
