#!/usr/bin/env python
"""retrieval: split files into chunks, embed them, answer queries with the top-k most similar chunks.   (llama.cpp/examples/retrieval)"""
import torch
from common import load, parser

ap = parser(__doc__)
ap.add_argument("--context-file", action="append", required=True)
ap.add_argument("--chunk-size", type=int, default=120, help="minimum characters per chunk")
ap.add_argument("--chunk-separator", default="\n")
ap.add_argument("--top-k", type=int, default=3)
ap.add_argument("-q", "--query", action="append", help="queries (default: read lines from stdin)")
a = ap.parse_args()
eng, tok = load(a)
chunks = []
for path in a.context_file:
    buf, pos = "", 0
    for part in open(path, encoding="utf-8").read().split(a.chunk_separator):
        buf += part + a.chunk_separator
        if len(buf) >= a.chunk_size:
            chunks.append((path, pos, buf.strip()))
            pos += len(buf)
            buf = ""
    if buf.strip():
        chunks.append((path, pos, buf.strip()))
mat = torch.stack([eng.embed(tok.encode(c)[: eng.max_tokens]) for _, _, c in chunks])
print(f"{len(chunks)} chunks embedded ({mat.shape[1]} dims)")
import sys
for q in (a.query or (l.strip() for l in sys.stdin)):
    if not q:
        break
    sims = mat @ eng.embed(tok.encode(q))
    print(f"\nquery: {q}")
    for score, idx in zip(*sims.topk(min(a.top_k, len(chunks)))):
        path, pos, text = chunks[int(idx)]
        print(f"  {float(score):.3f}  {path}:{pos}  {text[:70]!r}")
