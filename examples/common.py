"""Shared set-up for the example programs: repository on sys.path, model + engine + tokenizer from one line."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def parser(description: str) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("-m", "--model", required=True, help="GGUF checkpoint")
    ap.add_argument("--device", default="cuda")
    ap.add_argument("-c", "--ctx-size", type=int, default=512)
    return ap


def load(a, max_seqs: int = 1, n_batch: int | None = None):
    """-> (engine, tokenizer).  Falls back to the CPU op implementations when there is no GPU."""
    from distributed_llm_pipeline_b200.gguf import open_gguf
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    dev = a.device if (a.device != "cuda" or torch.cuda.is_available()) else "cpu"
    f = open_gguf(a.model)
    eng = LlamaEngine(load_weights(f, dev), n_ctx=a.ctx_size, max_seqs=max_seqs, page_size=64 if dev != "cpu" else 16,
                      max_batch_tokens=n_batch or min(a.ctx_size, 512))
    return eng, Tokenizer(f.metadata)
