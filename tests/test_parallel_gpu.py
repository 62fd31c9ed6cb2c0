"""Multi-GPU paths on real devices (skipped below 2 GPUs): TP over NCCL, TP with the all-reduce fused into the
mat-vec kernels (NVLink peer stores), and a 2-stage pipeline - each against the fp32 oracle."""
import os
import sys as _sys
_sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _torchrun(n, args, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(HERE, "mp_worker.py")] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=240)


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.parametrize("impl", ["nccl", "fused"])
@pytest.mark.parametrize("batch", [1, 2, 3])
def test_tp2_gpu(tmp_models, impl, batch):
    _need(2)
    r = _torchrun(2, ["--model", tmp_models[("tiny128", "F16")], "--device", "cuda", "--impl", impl, "--batch", str(batch), "--greedy", "12"],
                  29711 + batch)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MP_RESULT" in r.stdout


def test_tp2_fused_quantised(tmp_models):
    _need(2)
    r = _torchrun(2, ["--model", tmp_models[("tiny128", "Q4_K_M")], "--device", "cuda", "--impl", "fused"], 29721)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("symm", ["ipc", "nomc"])
def test_tp2_fused_other_transports(tmp_models, symm, monkeypatch):
    """The same fused step over cudaMalloc + CUDA IPC peer pointers, and over symmetric memory with the multicast alias unused."""
    _need(2)
    if symm == "ipc":
        monkeypatch.setenv("DLP_SYMM", "ipc")
    else:
        monkeypatch.setenv("DLP_SYMM_MC", "0")
    r = _torchrun(2, ["--model", tmp_models[("tiny128", "F16")], "--device", "cuda", "--impl", "fused", "--batch", "2", "--greedy", "12"], 29725)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("n", [4, 8])
def test_tp_wide_fused_gpu(tmp_models, n):
    _need(n)
    r = _torchrun(n, ["--model", tmp_models[("tiny-tp8", "F16")], "--device", "cuda", "--impl", "fused", "--batch", "2", "--greedy", "12"], 29727 + n)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("ftype", ["F16", "Q8_0"])
@pytest.mark.parametrize("batch", [1, 3])
def test_tp2_moe_gpu(tmp_models, ftype, batch):
    """Mixture of experts under tensor parallelism: every expert's n_ff is sharded, routing + grouped GEMM stay on the device,
    the combine's partial sums are all-reduced (the reference cannot split expert weights at all: MUL_MAT_ID asserts,
    ggml-cuda.cu:2274)."""
    _need(2)
    # Q8_0: the decode-sized steps go through the int8-activation mat-vec kernels (the reference's q8_1 numerics class).  On this
    # random tiny model that noise moves a near-tied router decision of one sequence: measured worst NMSE 7.6e-3 with the mat-vec
    # path vs 1.3e-4 with the fp16-activation grouped GEMM (DLP_MOE_GEMV=0) - a property of top-k routing, not of the reduction
    # (F16 weights and batch 1 stay below 1e-3), hence the wider bound for this one case.
    tol = ["--tol", "2e-2"] if ftype == "Q8_0" and batch > 1 else []
    r = _torchrun(2, ["--model", tmp_models[("tiny-moe-tp", ftype)], "--device", "cuda", "--impl", "fused", "--batch", str(batch), "--greedy", "8"] + tol,
                  29751 + batch)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_pp2_gpu(tmp_models):
    _need(2)
    r = _torchrun(2, ["--model", tmp_models[("tiny128", "F16")], "--device", "cuda", "--pp", "2", "--batch", "2"], 29731)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_pp2_tp2_fused_gpu(tmp_models):
    _need(4)
    r = _torchrun(4, ["--model", tmp_models[("tiny128", "F16")], "--device", "cuda", "--pp", "2", "--impl", "fused"], 29741)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("flags", [("--tp", "2"), ("--tp", "2", "--tp-impl", "nccl"), ("--pp", "2")], ids=["tp2-fused", "tp2-nccl", "pp2"])
def test_orchestrator_over_gpu_grid(tmp_models, flags):
    """POST /completion to an orchestrator whose engine is a 2-GPU grid (fused TP, NCCL TP, 2-stage pipeline): same greedy text
    as the single-GPU engine, per-rank "offloaded" lines in the /chat log stream."""
    _need(2)
    import json
    from test_distributed_serving import PROMPT, _Proc, _chat_events, _completion, _free_port
    from distributed_llm_pipeline_b200.serving.orchestrator import build_service
    from distributed_llm_pipeline_b200.serving.sampler import SamplingParams
    model = tmp_models[("tiny128", "F16")]
    sched, _, _ = build_service(model, device="cuda", n_ctx=128, n_slots=2, n_batch=32)
    want = sched.generate(PROMPT, n_predict=8, sampling=SamplingParams(temperature=0.0), ignore_eos=True).text
    port = _free_port()
    srv = _Proc(["-m", "distributed_llm_pipeline_b200.serving.orchestrator", "-m", model, "--device", "cuda", "--port", str(port),
                 "-c", "128", "-np", "2", "-b", "32", "-n", "8", *flags], port)
    try:
        srv.wait_ready(240)
        d = _completion(port, 8)
        assert d["content"] == want, (d["content"], want)
        logs = "".join(e["content"] for e in _chat_events(port, PROMPT) if e["msg_type"] == "log")
        assert logs.count("load_tensors: stage") == 2 and "grid:" in logs
        json.dumps(d)
    finally:
        srv.stop()
