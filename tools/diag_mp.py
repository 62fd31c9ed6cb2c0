#!/usr/bin/env python
"""Step-by-step multi-GPU diagnostic (run under torchrun): every step prints before/after and a watchdog dumps all stacks
and exits if a step takes longer than its budget - a hang costs seconds of GPU time, not the call's whole limit."""
import faulthandler
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

RANK = int(os.environ.get("RANK", "0"))
T0 = time.time()


def step(name, budget=45):
    class _S:
        def __enter__(self):
            print(f"[{RANK}] {time.time() - T0:7.2f}s  >> {name}", flush=True)
            faulthandler.dump_traceback_later(budget, exit=True)

        def __exit__(self, *exc):
            faulthandler.cancel_dump_traceback_later()
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            print(f"[{RANK}] {time.time() - T0:7.2f}s  << {name} {'FAILED ' + repr(exc[1]) if exc[0] else 'ok'}", flush=True)
            return False
    return _S()


def main():
    from distributed_llm_pipeline_b200.parallel.grid import init_grid
    with step("init_grid (nccl, device_id)"):
        grid = init_grid(pp=1, timeout_s=60)
    dev = grid.device
    with step("all_reduce on the default group"):
        x = torch.ones(4, device=dev)
        dist.all_reduce(x)
        assert float(x[0]) == grid.world
    with step("all_reduce on the tp group (new_group)"):
        x = torch.ones(4, device=dev)
        dist.all_reduce(x, group=grid.tp_group)
    with step("all_gather_object"):
        out = [None] * grid.world
        dist.all_gather_object(out, b"x" * 64, group=grid.tp_group)
    with step("nvidia-smi topo / p2p"):
        if RANK == 0:
            os.system("nvidia-smi topo -m | head -12; df -h /dev/shm | tail -1")
        print(f"[{RANK}] can_access_peer: {[torch.cuda.can_device_access_peer(dev.index, j) for j in range(torch.cuda.device_count()) if j != dev.index]}", flush=True)
    with step("FusedComm.setup (symm alloc + IPC exchange)"):
        from distributed_llm_pipeline_b200.parallel.fused import FusedComm
        comm = FusedComm(grid)
        comm.setup(512, 4)
    with step("fused push + consume (one site, T=1)"):
        from distributed_llm_pipeline_b200.gguf import GGMLType
        from distributed_llm_pipeline_b200.ops.weights import DeviceWeight
        tp, r = grid.tp, grid.tp_rank
        g = torch.Generator().manual_seed(5)
        w_down = torch.randn(512, 1024, generator=g) * 0.05
        w_next = torch.randn(256, 512, generator=g) * 0.05
        act = torch.randn(1, 1024, generator=g).half().to(dev)
        h = torch.randn(1, 512, generator=g).to(dev)
        nw = torch.ones(512, device=dev)
        Fs = 1024 // tp
        ws = DeviceWeight.from_dense(w_down[:, r * Fs:(r + 1) * Fs].contiguous(), GGMLType.F16, dev)
        wn = DeviceWeight.from_dense(w_next, GGMLType.F16, dev)
        for it in range(3):
            comm.begin_step()
            comm.gemv(ws, act[:, r * Fs:(r + 1) * Fs].contiguous(), 4, 1, ar_out=comm.ar_out(0))
            h_out = torch.zeros_like(h)
            y = torch.empty(1, 256, device=dev)
            comm.gemv(wn, h, 1, 1, out=y, norm_w=nw, eps=1e-5, ar_in=comm.ar_in(0, h_out))
            torch.cuda.synchronize()
            ref = h + act.float() @ w_down.half().float().T.to(dev)
            err = float((h_out - ref).abs().max())
            print(f"[{RANK}] iter {it}: max|h_out - ref| = {err:.3e}, error flag {int(comm.error.item())}", flush=True)
    with step("NCCL all_reduce inside a CUDA graph"):
        buf = torch.ones(8, device=dev)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            dist.all_reduce(buf, group=grid.tp_group)
        torch.cuda.current_stream().wait_stream(s)
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph):
            dist.all_reduce(buf, group=grid.tp_group)
        gph.replay()
    with step("tiny model: TP engine nccl, prefill + decode", 90):
        from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
        from distributed_llm_pipeline_b200.models.config import preset
        from distributed_llm_pipeline_b200.models.llama import LlamaEngine
        from distributed_llm_pipeline_b200.models.weights import load_weights
        from distributed_llm_pipeline_b200.parallel.comm import NcclComm
        path = "/tmp/diag_tiny128.gguf"
        if RANK == 0:
            write_synthetic_gguf(path, preset("tiny128"), "F16", seed=1)
        dist.barrier()
        w = load_weights(path, dev, tp_rank=grid.tp_rank, tp_size=grid.tp)
        eng = LlamaEngine(w, n_ctx=128, max_seqs=4, page_size=16, max_batch_tokens=32, comm=NcclComm(grid.tp_group, grid.tp, grid.tp_rank))
        sid = eng.alloc_seq()
        lg = eng.prefill(sid, [1, 5, 9, 200, 33, 7, 8])
        print(f"[{RANK}] prefill ok {float(lg[0]):.4f}", flush=True)
        lg = eng.decode([sid], [11], sample=False, use_graph=False)
        print(f"[{RANK}] eager decode ok {float(lg[0, 0]):.4f}", flush=True)
        lg = eng.decode([sid], [12], sample=False, use_graph=True)
        print(f"[{RANK}] graph decode ok {float(lg[0, 0]):.4f}", flush=True)
    with step("tiny model: TP engine fused, prefill (SP) + decode", 90):
        comm2 = FusedComm(grid)
        eng2 = LlamaEngine(w, n_ctx=128, max_seqs=4, page_size=16, max_batch_tokens=32, comm=comm2)
        sid = eng2.alloc_seq()
        lg2 = eng2.prefill(sid, [1, 5, 9, 200, 33, 7, 8])
        print(f"[{RANK}] fused SP prefill ok {float(lg2[0]):.4f}", flush=True)
        a = eng2.decode([sid], [11], sample=False, use_graph=False).clone()
        print(f"[{RANK}] fused eager decode ok {float(a[0, 0]):.4f}", flush=True)
        b = eng2.decode([sid], [12], sample=False, use_graph=True).clone()
        print(f"[{RANK}] fused graph decode ok {float(b[0, 0]):.4f} err flag {int(comm2.error.item())}", flush=True)
    dist.barrier()
    print(f"[{RANK}] DIAG DONE", flush=True)
    from distributed_llm_pipeline_b200.parallel.grid import shutdown
    shutdown(0)


if __name__ == "__main__":
    main()
