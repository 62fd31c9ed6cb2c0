#!/usr/bin/env python
"""dlp-perplexity: perplexity of a token stream under a GGUF model (chunks of n_ctx tokens, second half of each chunk scored).

Parity: `llama-perplexity` (llama.cpp/tools/perplexity/perplexity.cpp) default strided mode.  Works with random-init models
too (then PPL ~ vocab size) - used as a whole-model numerics check between back-ends / quantisations.
"""
import argparse
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("-f", "--file", default=None, help="text file (tokenised with the model's tokenizer); default: synthetic ids")
    ap.add_argument("-c", "--ctx-size", type=int, default=128)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--device", default="cuda")
    a = ap.parse_args(argv)
    from distributed_llm_pipeline_b200.gguf import open_gguf
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    dev = a.device if (a.device != "cuda" or torch.cuda.is_available()) else "cpu"
    f = open_gguf(a.model)
    w = load_weights(f, dev)
    eng = LlamaEngine(w, n_ctx=a.ctx_size, max_seqs=1, page_size=64 if dev != "cpu" else 16, max_batch_tokens=a.ctx_size)
    if a.file:
        ids = Tokenizer(f.metadata).encode(open(a.file).read())
    else:
        g = torch.Generator().manual_seed(0)
        ids = torch.randint(3, w.cfg.vocab_size, (a.ctx_size * a.chunks,), generator=g).tolist()
    nll, nll2, count = 0.0, 0.0, 0
    n_chunks = min(a.chunks, len(ids) // a.ctx_size)
    tok = Tokenizer(f.metadata)
    half = a.ctx_size // 2
    for c in range(n_chunks):
        chunk = list(ids[c * a.ctx_size:(c + 1) * a.ctx_size])
        if tok.add_bos:                 # every window starts with BOS, as the reference tool arranges it (perplexity.cpp:573-578)
            chunk[0] = tok.bos_id
        sid = eng.alloc_seq()
        logits = eng.prefill(sid, chunk, all_logits=True)          # one forward; logits of every position
        lp = torch.log_softmax(logits.float(), -1)
        # positions n_ctx/2 .. n_ctx-2 predict tokens n_ctx/2+1 .. n_ctx-1: the second half of the window, scored with at least half a
        # window of context (perplexity.cpp:595-640 `first = n_ctx/2`)
        tgt = torch.tensor(chunk[half + 1:], device=lp.device)
        tok_nll = -lp[half:a.ctx_size - 1].gather(1, tgt[:, None]).double()
        nll += float(tok_nll.sum())
        nll2 += float((tok_nll * tok_nll).sum())
        count += a.ctx_size - 1 - half
        eng.free_seq(sid)
        print(f"[{c + 1}]{math.exp(nll / count):.4f},", end="", flush=True)          # the reference's running line: [1]647.7593,[2]561.2521,
    print()
    # the reference's closing line, with its error estimate: standard error of the mean token NLL, carried over to the perplexity
    mean = nll / max(1, count)
    var = nll2 / max(1, count) - mean * mean
    err = math.sqrt(var / (count - 1)) * math.exp(mean) if var > 0 and count > 1 else 0.0
    print(f"Final estimate: PPL = {math.exp(mean):.4f} +/- {err:.5f}")


if __name__ == "__main__":
    main()
