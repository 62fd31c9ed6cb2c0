#!/usr/bin/env python
"""dlp-tokenize: print the token ids (and pieces) of a prompt.  Parity: `llama-tokenize` (llama.cpp/tools/tokenize)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("-m", "--model", required=True)
    ap.add_argument("-p", "--prompt", default=None)
    ap.add_argument("-f", "--file", default=None)
    ap.add_argument("--ids", action="store_true", help="print only the ids as a Python list")
    ap.add_argument("--no-bos", action="store_true")
    a = ap.parse_args(argv)
    text = open(a.file).read() if a.file else (a.prompt or sys.stdin.read())
    tok = Tokenizer.from_gguf(a.model)
    ids = tok.encode(text, add_bos=not a.no_bos)
    if a.ids:
        print(ids)
    else:
        for i in ids:
            print(f"{i:6d} -> {tok.token_bytes(i).decode('utf-8', errors='replace')!r}")
        print(f"Total number of tokens: {len(ids)}")


if __name__ == "__main__":
    main()
