"""Tensor- and pipeline-parallel layouts on CPU (gloo, world_size 2): sharded loading, NCCL-baseline reduction path,
stage hand-off - everything except the CUDA kernels themselves."""
import os
import subprocess
import sys

import pytest

from distributed_llm_pipeline_b200.parallel.grid import stage_layer_ranges

HERE = os.path.dirname(os.path.abspath(__file__))


def _torchrun(n, args, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(HERE, "mp_worker.py")] + args
    env = dict(os.environ, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)


@pytest.mark.parametrize("batch", [1, 3])
def test_tp2_gloo_matches_oracle(tmp_models, batch):
    r = _torchrun(2, ["--model", tmp_models[("tiny", "F16")], "--device", "cpu", "--batch", str(batch), "--tol", "1e-5"], 29611 + batch)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MP_RESULT" in r.stdout


@pytest.mark.parametrize("name", ["tiny128", "tiny"])   # block-aligned K shards / unaligned (dense fallback)
def test_tp2_quantised_k_sharding(tmp_models, name):
    r = _torchrun(2, ["--model", tmp_models[(name, "Q4_K_M")], "--device", "cpu", "--tol", "1e-5"], 29621 + len(name))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_pp2_gloo_matches_oracle(tmp_models):
    r = _torchrun(2, ["--model", tmp_models[("tiny", "F16")], "--device", "cpu", "--pp", "2", "--batch", "2", "--tol", "1e-5"], 29631)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_stage_layer_ranges_follow_split_rule():
    assert [list(r) for r in stage_layer_ranges(8, 2)] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert [len(r) for r in stage_layer_ranges(32, 4)] == [8, 8, 8, 8]
    assert [len(r) for r in stage_layer_ranges(22, 2)] == [11, 11]
    assert [len(r) for r in stage_layer_ranges(10, 2, [3, 1])] == [8, 2]     # weighted like --tensor-split 3,1
    rs = stage_layer_ranges(7, 3)
    assert sum(len(r) for r in rs) == 7 and rs[0].start == 0 and rs[-1].stop == 7
    with pytest.raises(ValueError):
        stage_layer_ranges(8, 2, [1, 2, 3])


def test_pp2_self_feeding_decode_rounds(tmp_models):
    """Pipeline decode where the last stage's sampled ids travel back to the first stage (no host tokens after round 0)."""
    script = os.path.join(HERE, "mp_pp_feedback.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29641", script, tmp_models[("tiny", "F16")]]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0 and "FEEDBACK_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_tp4_gloo_with_replicated_kv_heads(tmp_models):
    """tp = 4 > n_head_kv = 2: KV heads are replicated across rank pairs, Q heads / FFN / vocab split four ways."""
    r = _torchrun(4, ["--model", tmp_models[("tiny128", "F16")], "--device", "cpu", "--batch", "2", "--tol", "1e-5"], 29651)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MP_RESULT" in r.stdout


def test_pp2_x_tp2_grid_gloo(tmp_models):
    """The explicit pp x tp grid (2 stages x 2 tensor-parallel ranks, world size 4) the reference cannot express."""
    r = _torchrun(4, ["--model", tmp_models[("tiny128", "F16")], "--device", "cpu", "--pp", "2", "--batch", "2", "--tol", "1e-5"], 29661)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
