"""KV/session save-restore (incl. restore into a fragmented page pool, cf. test-state-restore-fragmented.cpp),
tracer and logger."""
import json
import os

import pytest
import torch

from distributed_llm_pipeline_b200.models.llama import LlamaEngine
from distributed_llm_pipeline_b200.models.reference import nmse
from distributed_llm_pipeline_b200.models.weights import load_weights
from distributed_llm_pipeline_b200.utils.log import Logger
from distributed_llm_pipeline_b200.utils.session import load_sequence, save_sequence
from distributed_llm_pipeline_b200.utils.trace import PerfCounters, Tracer


def _run(device, tmp_models, tmp_path):
    path = tmp_models[("tiny", "F16")]
    eng = LlamaEngine(load_weights(path, device), n_ctx=128, max_seqs=3, page_size=16, max_batch_tokens=32)
    a = eng.alloc_seq()
    prompt = list(range(3, 40))
    eng.prefill(a, prompt)
    want = eng.decode([a], [7], sample=False, use_graph=False)[0].clone()
    eng.truncate_seq(a, len(prompt))
    f = str(tmp_path / "seq.bin")
    nbytes = save_sequence(eng, a, f, tokens=prompt)
    assert nbytes > len(prompt) * 2 * eng.Hkv * eng.D * 2
    # fragment the pool: occupy and free pages so the restored sequence lands on different, non-contiguous pages
    b = eng.alloc_seq()
    eng.prefill(b, list(range(50, 90)))
    eng.free_seq(a)
    c = eng.alloc_seq()
    eng.prefill(c, [1, 2, 3])
    d = eng.alloc_seq() if eng._free_sids else a
    hdr = load_sequence(eng, d, f)
    assert hdr["tokens"] == prompt and eng.seq_len(d) == len(prompt)
    got = eng.decode([d], [7], sample=False, use_graph=False)[0]
    assert nmse(want, got) < 1e-6
    # a different model must be refused
    other = LlamaEngine(load_weights(tmp_models[("tiny128", "F16")], device), n_ctx=128, max_seqs=1, page_size=16, max_batch_tokens=32)
    with pytest.raises(ValueError, match="different model"):
        load_sequence(other, other.alloc_seq(), f)


def test_session_roundtrip_cpu(tmp_models, tmp_path):
    _run("cpu", tmp_models, tmp_path)


@pytest.mark.gpu
def test_session_roundtrip_gpu(tmp_models, tmp_path):
    _run("cuda", tmp_models, tmp_path)


def test_tracer_and_perf_counters(tmp_path):
    tr = Tracer(enabled=True, nvtx=False)
    with tr.span("phase_a", device_timed=False, tokens=3):
        sum(range(1000))
    with tr.span("phase_a", device_timed=False):
        pass
    s = tr.summary()
    assert s["phase_a"]["count"] == 2 and s["phase_a"]["total_ms"] >= 0
    out = tr.export_chrome(str(tmp_path / "trace.json"))
    ev = json.load(open(out))["traceEvents"]
    assert len(ev) == 2 and ev[0]["ph"] == "X" and ev[0]["args"] == {"tokens": 3}
    pc = PerfCounters()
    pc.t_eval_ms, pc.n_eval = 100.0, 50
    assert "500.00 tokens per second" in pc.report()


def test_logger_levels_ring_and_file(tmp_path):
    lg = Logger(verbosity=3)
    lg.sinks.clear()
    lg.add_file(str(tmp_path / "log.txt"))
    seen = []
    lg.subscribe(seen.append)
    lg.info("offloaded 3/3 layers")
    lg.debug("hidden")
    lg.error("boom")
    lg.flush()
    text = open(tmp_path / "log.txt").read()
    assert "offloaded 3/3 layers" in text and "boom" in text and "hidden" not in text
    assert len(lg.ring) == 2 and len(seen) == 2
    lg.close()
