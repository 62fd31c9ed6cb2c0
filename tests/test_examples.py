"""Every program under examples/ runs on a tiny random-init checkpoint (CPU) and does what its docstring says."""
import os
import subprocess
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(script, *args, stdin=None):
    r = subprocess.run([sys.executable, os.path.join(REPO, "examples", script), *args], capture_output=True, text=True, timeout=300, input=stdin,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return r.stdout


@pytest.fixture(scope="module")
def model(tmp_models):
    return tmp_models[("tiny", "F16")]


def test_simple_and_eval_callback(model):
    out = run("simple.py", "-m", model, "--device", "cpu", "-n", "8", "-p", "hello there")
    assert out.startswith("hello there") and "8 generated" in out and "tok/s" in out
    out = run("eval_callback.py", "-m", model, "--device", "cpu", "-p", "hello there")
    assert out.count("l_out-") == 2 and "next token:" in out


def test_simple_chat_reuses_the_conversation(model):
    out = run("simple_chat.py", "-m", model, "--device", "cpu", "-n", "6", "--turn", "hi there", "--turn", "and again")
    lines = [l for l in out.splitlines() if "new prompt tokens" in l]
    assert len(lines) == 2 and ", 0 reused]" in lines[0]
    assert int(lines[1].split(",")[1].split()[0]) > 0          # the second turn finds the first one in the cache


def test_batched_and_parallel(model):
    out = run("batched.py", "-m", model, "--device", "cpu", "-np", "3", "-n", "6")
    assert out.count("--- sequence") == 3 and "3 x 6 tokens" in out
    out = run("parallel.py", "-m", model, "--device", "cpu", "-np", "2", "-ns", "5", "-n", "4")
    assert out.count("client ") == 5 and "5 requests over 2 slots: 20 tokens" in out
    assert any("from the cache" in l and "(0 from the cache)" not in l for l in out.splitlines())      # the shared system prompt was reused


def test_embedding_and_retrieval(model, tmp_path):
    out = run("embedding.py", "-m", model, "--device", "cpu", "-p", "alpha beta", "-p", "alpha beta", "-p", "gamma delta epsilon")
    rows = [l.split() for l in out.split("cosine similarity:")[1].strip().splitlines()]
    assert abs(float(rows[0][1]) - 1.0) < 1e-3 and float(rows[0][2]) < 0.999
    doc = tmp_path / "doc.txt"
    doc.write_text("the cat sat on the mat\nquantum chromodynamics of quarks\nbread is baked in an oven\n")
    out = run("retrieval.py", "-m", model, "--device", "cpu", "--context-file", str(doc), "--chunk-size", "10", "--top-k", "1", "-q", "bread is baked in an oven")
    assert "3 chunks embedded" in out and "'bread is baked in an oven'" in out.split("query:")[1]


def test_passkey_slides_the_window(model):
    out = run("passkey.py", "-m", model, "--device", "cpu", "-c", "256", "--junk", "12", "-b", "32", "-n", "4")
    assert "through a 256-token window" in out and "answer:" in out
    n_shift = int(out.split("(")[1].split()[0])
    assert n_shift >= 1                                         # the prompt (several hundred tokens) did not fit: it was shifted through


def test_save_load_state_and_speculative(model, tmp_models):
    assert "identical continuation" in run("save_load_state.py", "-m", model, "--device", "cpu", "-n", "10")
    out = run("speculative.py", "-m", model, "--device", "cpu", "-n", "16")
    assert "same tokens as plain greedy: True" in out
    out = run("speculative.py", "-m", model, "--device", "cpu", "-n", "16", "-md", tmp_models[("tiny", "Q8_0")], "--draft-max", "4")
    assert "same tokens as plain greedy: True" in out and "accepted" in out


def test_gguf_example(tmp_path):
    out = run("gguf_example.py", str(tmp_path / "e.gguf"))
    assert "7 key-value pairs, 4 tensors" in out and "data check passed" in out and "some.parameter.string = 'hello world'" in out
