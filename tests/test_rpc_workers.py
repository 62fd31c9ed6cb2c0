"""Config 1 path: two native CPU workers on localhost, each owning half of the layers, chained by the client.
Oracle: the fp32 reference model on the same GGUF (what the reference's rpc-server pair + llama-cli would compute)."""
import socket

import pytest
import torch

from distributed_llm_pipeline_b200.models.reference import ReferenceModel, nmse
from distributed_llm_pipeline_b200.parallel.rpc import RpcEngine, WorkerClient, WorkerError, free_port_run, launch_local_workers
from distributed_llm_pipeline_b200.serving.sampler import SamplingParams
from distributed_llm_pipeline_b200.serving.scheduler import Scheduler
from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _free_base(n=2):
    return free_port_run(n)


@pytest.fixture(scope="module", params=["Q8_0", "Q4_K_M"])
def chain(request, tmp_models):
    path = tmp_models[("tiny", request.param)]
    base = _free_base(2)
    procs, eps = launch_local_workers(path, 2, base_port=base, n_ctx=128, threads=2)
    eng = RpcEngine(eps, max_seqs=2)
    yield path, eng, eps
    for w in eng.workers:
        w.shutdown()
    for p in procs:
        p.wait(timeout=10)


def test_two_worker_pipeline_matches_oracle(chain):
    path, eng, _ = chain
    ref = ReferenceModel.from_gguf(path)
    prompt = [1, 5, 9, 200, 33, 7, 8, 100, 42]
    toks, ref_logits = ref.generate_greedy(prompt, 4)
    sid = eng.alloc_seq()
    # Q8_0 runs on the int8 x int8 AVX2 path (activations quantised per 32 values, like the reference's CPU vec-dot):
    # its own MUL_MAT tolerance is NMSE 5e-4; the fp32-decode path of the other types is exact to 1e-6
    tol = 3e-4 if "Q8_0" in path else 1e-6
    assert nmse(ref_logits[0], eng.prefill(sid, prompt)) < tol
    for i, t in enumerate(toks[:-1]):
        assert nmse(ref_logits[i + 1], eng.decode([sid], [t], sample=False)[0]) < tol
    # truncate = prefix reuse: rewind to the prompt and replay one step
    eng.truncate_seq(sid, len(prompt))
    assert nmse(ref_logits[1], eng.decode([sid], [toks[0]], sample=False)[0]) < tol
    eng.free_seq(sid)
    assert [w.l0 for w in eng.workers] == [0, 1] and eng.workers[0].first and eng.workers[1].last


def test_scheduler_serves_through_workers(chain):
    path, eng, _ = chain
    sched = Scheduler(eng, Tokenizer.from_gguf(path), n_batch=32)
    r = sched.generate("hello there", n_predict=5, sampling=SamplingParams(temperature=0.0))
    assert 1 <= len(r.output_ids) <= 5 and r.finish_reason in ("length", "stop")
    assert any("RPC worker" in l for l in eng.describe())


def test_worker_validates_requests(chain):
    _, eng, eps = chain
    host, port = eps[0].split(":")
    # the worker serves one client at a time (like the reference's rpc-server); reuse the engine's connection
    w = eng.workers[0]
    with pytest.raises(WorkerError, match="does not continue"):
        w.run(7, 1, 5, b"\x01\x00\x00\x00")                 # position gap
    with pytest.raises(WorkerError, match="payload size"):
        w.run(7, 2, 0, b"\x01\x00\x00\x00")                 # 2 tokens announced, 1 sent
    with pytest.raises(WorkerError, match="out of range"):
        w.run(7, 1, 0, (10 ** 6).to_bytes(4, "little"))     # token id beyond the vocabulary
    w.reset(7)


def test_worker_crash_is_healed_by_reconnect_and_replay(tmp_models):
    """Fault injection: SIGKILL one stage worker in the middle of a generation.  The pool restarts it, the engine reconnects,
    replays the sequence history into the fresh KV caches and the token stream is identical to an undisturbed run."""
    import os
    import signal
    from distributed_llm_pipeline_b200.parallel.rpc import LocalWorkerPool
    path = tmp_models[("tiny", "Q8_0")]
    pool = LocalWorkerPool(path, 2, base_port=_free_base(2), n_ctx=128, threads=2)
    try:
        eng = RpcEngine(pool.endpoints, max_seqs=2, recover=True, recover_wait_s=30)
        prompt = [1, 5, 9, 200, 33, 7, 8]

        def run(kill_at):
            sid = eng.alloc_seq()
            logits = eng.prefill(sid, prompt)
            out = [int(logits.argmax())]
            for i in range(10):
                if i == kill_at:
                    os.kill(pool.procs[1].pid, signal.SIGKILL)
                out.append(int(eng.decode([sid], [out[-1]], sample=False)[0].argmax()))
            eng.free_seq(sid)
            return out
        clean = run(-1)
        healed = run(4)
        assert healed == clean
        assert eng.n_recoveries >= 1 and pool.restarts >= 1
        eng.close()                                                         # (a worker serves one client at a time)
        plain = RpcEngine(pool.endpoints, max_seqs=1)                       # without recover=True the failure is a clean error
        sid = plain.alloc_seq()
        plain.prefill(sid, prompt)
        os.kill(pool.procs[0].pid, signal.SIGKILL)
        with pytest.raises(WorkerError):
            for _ in range(3):
                plain.decode([sid], [5], sample=False)
    finally:
        pool.stop()


@pytest.mark.parametrize("key", [("tiny-qwen2", "Q8_0"), ("tiny-qwen3", "F16"), ("tiny-phi3", "Q8_0"), ("tiny-granite", "Q8_0"), ("tiny-minicpm", "F16"),
                                 ("tiny-seed-oss", "F16"), ("tiny-qwen", "F16"), ("tiny-exaone", "F16"), ("tiny-internlm2", "F16")])
def test_other_architectures_on_cpu_workers(tmp_models, key):
    """qwen2 (bias), qwen3 (per-head Q/K norm), phi3 (fused tensors), granite / minicpm (four scalars), seed_oss (its own norm tensor name),
    qwen (fused QKV with bias, separate gate / up), exaone, internlm2 - through two native stage workers."""
    path = tmp_models[key]
    procs, eps = launch_local_workers(path, 2, base_port=_free_base(2), n_ctx=64, threads=2)
    try:
        eng = RpcEngine(eps, max_seqs=1)
        ref = ReferenceModel.from_gguf(path)
        prompt = [1, 5, 9, 200, 33, 7, 8, 100, 42]
        toks, ref_logits = ref.generate_greedy(prompt, 3)
        sid = eng.alloc_seq()
        tol = 3e-4 if key[1] == "Q8_0" else 1e-6
        assert nmse(ref_logits[0], eng.prefill(sid, prompt)) < tol
        for i, t in enumerate(toks[:-1]):
            assert nmse(ref_logits[i + 1], eng.decode([sid], [t], sample=False)[0]) < tol
        for w in eng.workers:
            w.shutdown()
    finally:
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()


def test_q8_0_avx2_and_vnni_kernels(tmp_models, monkeypatch):
    """The worker picks the AVX-512 VNNI mat-mul at run time where the CPU has it; DLP_WORKER_NO_VNNI forces the AVX2 one.
    Both (panel-blocked, 4 tokens per pass, ragged token tail by the row kernel) must match the oracle on a 13-token prompt."""
    path = tmp_models[("tiny", "Q8_0")]
    ref = ReferenceModel.from_gguf(path)
    prompt = [1, 5, 9, 200, 33, 7, 8, 100, 42, 17, 300, 2, 77]
    _, ref_logits = ref.generate_greedy(prompt, 1)
    outs = []
    for no_vnni in (False, True):
        if no_vnni:
            monkeypatch.setenv("DLP_WORKER_NO_VNNI", "1")
        procs, eps = launch_local_workers(path, 2, base_port=_free_base(2), n_ctx=64, threads=2)
        try:
            eng = RpcEngine(eps, max_seqs=1)
            sid = eng.alloc_seq()
            outs.append(eng.prefill(sid, prompt))
            assert nmse(ref_logits[0], outs[-1]) < 3e-4
            for w in eng.workers:
                w.shutdown()
        finally:
            for p in procs:
                try:
                    p.wait(timeout=10)
                except Exception:
                    p.kill()
    assert nmse(outs[0], outs[1]) < 1e-8      # same int8 products, fp32 summation order differs


def test_worker_bench_mode(tmp_models):
    """`dlp-worker --bench T`: the stage timed locally (no sockets) with the per-phase split on stderr."""
    import subprocess
    from distributed_llm_pipeline_b200.runtime import native
    native.build()
    p = subprocess.run([native.path("dlp-worker"), "-m", tmp_models[("tiny", "Q8_0")], "-c", "64", "--bench", "16", "--bench-reps", "1"],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert "bench: prefill 16 tokens" in p.stderr and "matmul" in p.stderr


@pytest.mark.parametrize("ftype", ["Q4_0", "Q4_1", "Q5_0", "Q5_1", "Q2_K", "Q3_K", "IQ4_NL", "IQ4_XS"])
def test_cpu_workers_read_the_other_block_types(tmp_path, ftype):
    """The legacy 4- / 5-bit blocks, the 2- and 3-bit K-quants and the non-linear 4-bit kinds through two native stage workers: the row decoders
    of `runtime/gguf_lite.h` against the numpy codecs behind the oracle (prefill and one decode step)."""
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.config import preset
    path = write_synthetic_gguf(str(tmp_path / f"tiny-{ftype}.gguf"), preset("tiny"), ftype, seed=1)
    procs, eps = launch_local_workers(path, 2, base_port=_free_base(2), n_ctx=64, threads=2)
    eng = None
    try:
        eng = RpcEngine(eps, max_seqs=1)
        ref = ReferenceModel.from_gguf(path)
        prompt = [1, 5, 9, 200, 33, 7, 8, 100, 42]
        sid = eng.alloc_seq()
        got = torch.as_tensor(eng.prefill(sid, prompt)).float()
        assert nmse(got, ref.forward(torch.tensor(prompt), ref.new_cache(64))[-1]) < 1e-9
        step = torch.as_tensor(eng.decode([sid], [7], sample=False)[0]).float()
        assert nmse(step, ref.forward(torch.tensor(prompt + [7]), ref.new_cache(64))[-1]) < 1e-9
    finally:
        for w in (eng.workers if eng is not None else []):
            w.shutdown()
        for p in procs:
            p.wait(timeout=10)
