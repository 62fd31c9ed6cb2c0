#!/usr/bin/env python
"""simple: load a model, tokenize a prompt, generate greedily, print text and speed.   (llama.cpp/examples/simple)"""
import time

from common import load, parser

ap = parser(__doc__)
ap.add_argument("-p", "--prompt", default="Hello my name is")
ap.add_argument("-n", "--predict", type=int, default=32)
a = ap.parse_args()
eng, tok = load(a)
ids = tok.encode(a.prompt)
t0 = time.time()
out = eng.generate(ids, a.predict, stop_ids=tuple(tok.stop_ids))
dt = time.time() - t0
print(a.prompt + tok.decode(out, strip_leading_space=False))
print(f"\n[{len(ids)} prompt tokens, {len(out)} generated in {dt:.2f} s = {len(out) / dt:.1f} tok/s]")
