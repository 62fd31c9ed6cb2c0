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
    ap.add_argument("--no-bos", action="store_true", help="do not add BOS even where the model uses one")
    ap.add_argument("--stdin", action="store_true", help="read the prompt from standard input")
    ap.add_argument("--no-parse-special", action="store_true", help="treat special-token text in the prompt as plain text")
    ap.add_argument("--no-escape", action="store_true", help="leave \\n, \\t ... in the prompt as they are")
    ap.add_argument("--show-count", action="store_true", help="print the number of tokens at the end")
    ap.add_argument("--log-disable", action="store_true", help="accepted: nothing but the tokens is written to stdout anyway")
    a = ap.parse_args(argv)
    text = open(a.file).read() if a.file else (sys.stdin.read() if a.stdin or a.prompt is None else a.prompt)
    if not a.no_escape and not a.file and not a.stdin:
        for esc, ch in (("\\n", "\n"), ("\\t", "\t"), ("\\'", "'"), ('\\"', '"'), ("\\\\", "\\")):
            text = text.replace(esc, ch)
    tok = Tokenizer.from_gguf(a.model)
    if a.no_parse_special:
        saved, tok.special = tok.special, {}
    ids = tok.encode(text, add_bos=(tok.add_bos and not a.no_bos))
    if a.no_parse_special:
        tok.special = saved
    if a.ids:
        print("[" + ", ".join(str(i) for i in ids) + "]")
    else:
        for i in ids:       # control tokens by name, other tokens as the text they stand for (llama-tokenize: token_to_piece with special = true)
            piece = tok.tokens[i].encode("utf-8") if tok.types[i] in (3, 4) else tok.token_bytes(i)      # a lone byte token is written as that byte
            sys.stdout.flush()
            sys.stdout.buffer.write(f"{i:6d} -> '".encode() + piece + b"'\n")
    if a.show_count:
        print(f"Total number of tokens: {len(ids)}")


if __name__ == "__main__":
    main()
