#!/usr/bin/env python
"""embedding: pooled, normalised sentence embeddings and their cosine-similarity matrix.   (llama.cpp/examples/embedding)"""
from common import load, parser

ap = parser(__doc__)
ap.add_argument("-p", "--prompt", action="append", help="one text per flag")
ap.add_argument("--pooling", default="mean", choices=["mean", "last", "cls"])
a = ap.parse_args()
texts = a.prompt or ["Hello world", "Hello there world", "Quantum chromodynamics"]
eng, tok = load(a)
vecs = [eng.embed(tok.encode(t), a.pooling) for t in texts]
for t, v in zip(texts, vecs):
    print(f"{t[:40]!r:44s} [{', '.join(f'{x:+.4f}' for x in v[:6].tolist())}, ...] ({v.numel()} dims)")
print("\ncosine similarity:")
for i, vi in enumerate(vecs):
    print("  " + " ".join(f"{float(vi @ vj):6.3f}" for vj in vecs))
