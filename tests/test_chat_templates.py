"""Built-in chat formats (`serving/chat_templates.py`) against the reference's own formatter: `llama-server --chat-template NAME` asked on
`/apply-template` (llama.cpp/src/llama-chat.cpp).  The answers for all 46 names are kept in `tests/data/chat_templates_reference.json`; when
the reference binaries are present a sample of names is asked again live, so the fixture cannot drift from the binary."""
import json
import os
import subprocess
import sys
import time
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baseline import reference as ref_build  # noqa: E402

from distributed_llm_pipeline_b200.serving import chat_templates as ct  # noqa: E402

FIXTURE = os.path.join(ROOT, "tests", "data", "chat_templates_reference.json")
CONVS = {"full": [{"role": "system", "content": "You are a helpful assistant"}, {"role": "user", "content": "Hello"}, {"role": "assistant", "content": "Hi there"},
                  {"role": "user", "content": "Who are you"}, {"role": "assistant", "content": "   I am an assistant   "}, {"role": "user", "content": "Another question"}],
         "nosys": [{"role": "user", "content": "Hello"}, {"role": "assistant", "content": "Hi there"}, {"role": "user", "content": "Bye"}]}
REF = json.load(open(FIXTURE, encoding="utf-8"))


@pytest.mark.parametrize("name", sorted(REF))
def test_builtin_template_matches_reference_output(name):
    for conv, msgs in CONVS.items():
        assert ct.apply(name, msgs) == REF[name][conv], (name, conv)


def test_every_reference_name_is_covered_and_unknown_names_are_refused():
    assert set(REF) <= set(ct.names()) and len(REF) == 46
    with pytest.raises(KeyError):
        ct.apply("no-such-template", CONVS["nosys"])
    # OpenAI content parts are flattened
    assert ct.apply("chatml", [{"role": "user", "content": [{"type": "text", "text": "a"}, {"type": "text", "text": "b"}]}]) == "<|im_start|>user\nab<|im_end|>\n<|im_start|>assistant\n"
    assert ct.apply("chatml", CONVS["nosys"], add_generation_prompt=False).endswith("<|im_end|>\n")


def test_detection_from_template_text():
    assert ct.detect("{% for m in messages %}<|im_start|>{{ m.role }}\n...<|im_end|>") == "chatml"
    assert ct.detect("...<|start_header_id|>...<|eot_id|>") == "llama3"
    assert ct.detect("{{ '[INST] ' + m.content + ' [/INST]' }} <<SYS>> ") == "llama2-sys"
    assert ct.detect("[SYSTEM_PROMPT] x [/SYSTEM_PROMPT][INST]") == "mistral-v7"
    assert ct.detect("<start_of_turn>user") == "gemma" and ct.detect("<|user|>\n..<|end|>\n<|assistant|>") == "phi3"
    assert ct.detect("<｜User｜>..<｜Assistant｜>") == "deepseek3" and ct.detect("plain text") is None


def test_sample_of_names_against_the_live_reference_server(tmp_models, tmp_path):
    why = ref_build.ensure(allow_build=False)
    if why is not None or not ref_build.have("llama-server"):
        pytest.skip(f"reference binaries unavailable: {why}")
    env = ref_build.cpu_env(str(tmp_path))
    for i, name in enumerate(["chatml", "mistral-v7", "gemma", "deepseek3"]):
        port = 38300 + i
        p = subprocess.Popen([ref_build.exe("llama-server"), "-m", tmp_models[("tiny", "F16")], "-ngl", "0", "-c", "256", "--host", "127.0.0.1", "--port", str(port), "-t", "2",
                              "--chat-template", name, "--no-warmup", "--no-jinja"], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        try:
            for _ in range(150):
                try:
                    urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=1).read()
                    break
                except Exception:
                    time.sleep(0.1)
            else:
                pytest.skip("the reference server did not come up here")
            for conv, msgs in CONVS.items():
                r = urllib.request.urlopen(urllib.request.Request(f"http://127.0.0.1:{port}/apply-template", json.dumps({"messages": msgs}).encode(),
                                                                  {"Content-Type": "application/json"}), timeout=10)
                live = json.loads(r.read())["prompt"]
                assert live == REF[name][conv] == ct.apply(name, msgs)
        finally:
            p.terminate()
            p.wait()


def test_orchestrator_formatter_prefers_name_then_jinja_then_detection():
    from distributed_llm_pipeline_b200.serving.orchestrator import apply_chat_template
    msgs = CONVS["nosys"]
    assert apply_chat_template({"dlp.chat_template_name": "gemma", "tokenizer.chat_template": "<|im_start|>"}, msgs) == REF["gemma"]["nosys"]
    jinja = "{% for m in messages %}<{{ m.role }}>{{ m.content }}{% endfor %}<assistant>"
    assert apply_chat_template({"tokenizer.chat_template": jinja}, msgs) == "<user>Hello<assistant>Hi there<user>Bye<assistant>"
    assert apply_chat_template({"tokenizer.chat_template": "uses <|start_header_id|> and <|eot_id|>"}, msgs) == REF["llama3"]["nosys"]
    assert apply_chat_template({}, msgs) == REF["chatml"]["nosys"]         # no template at all: ChatML, like the reference server
