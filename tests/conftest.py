import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def tmp_models(tmp_path_factory):
    """Tiny random-init GGUF checkpoints shared by the tests."""
    from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
    from distributed_llm_pipeline_b200.models.config import preset
    root = tmp_path_factory.mktemp("models")
    out = {}
    for name, ftype in [("tiny", "F16"), ("tiny", "Q8_0"), ("tiny", "Q4_K_M"), ("tiny128", "F16"), ("tiny128", "Q4_K_M"), ("tiny-moe", "F16"),
                        ("tiny-qwen2", "F16"), ("tiny-qwen2", "Q8_0"), ("tiny-qwen3", "F16"), ("tiny-phi3", "F16"),
                        ("tiny-phi3", "Q8_0"), ("tiny-qwen3moe", "F16"), ("tiny-qwen2moe", "F16"), ("tiny-granite", "F16"),
                        ("tiny-granite", "Q8_0"), ("tiny-granite", "Q4_K_M"), ("tiny-tp8", "F16"), ("tiny-moe-tp", "F16"), ("tiny-moe-tp", "Q8_0")] + \
            [(n, "F16") for n in ("tiny-internlm2", "tiny-xverse", "tiny-ernie", "tiny-exaone", "tiny-seed-oss", "tiny-minicpm", "tiny-granitemoe", "tiny-smollm3",
                                       "tiny-starcoder2", "tiny-falcon", "tiny-olmo", "tiny-orion", "tiny-command-r", "tiny-gptneox", "tiny-stablelm", "tiny-nemotron", "tiny-arcee",
                                       "tiny-gpt2", "tiny-starcoder", "tiny-codeshell", "tiny-qwen", "tiny-olmo2", "tiny-glm4", "tiny-mistral3", "tiny-pangu", "tiny-chatglm", "tiny-plamo", "tiny-deepseek", "tiny-bailingmoe", "tiny-qwen2vl", "tiny-qwen3vl", "tiny-exaone4", "tiny-phi2", "tiny-olmoe", "tiny-qwen3vlmoe", "tiny-glm4moe", "tiny-dots1", "tiny-ernie-moe", "tiny-minimax-m2", "tiny-bailingmoe2", "tiny-bitnet")] + \
            [("tiny-smollm3", "Q8_0"), ("tiny-starcoder2", "Q8_0"), ("tiny-falcon", "Q8_0"), ("tiny-gptneox", "Q8_0")]:
        out[(name, ftype)] = write_synthetic_gguf(str(root / f"{name}-{ftype}.gguf"), preset(name), ftype, seed=1)
    return out
