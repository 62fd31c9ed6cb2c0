#!/usr/bin/env python
"""save-load-state: generate, save the sequence (KV pages + tokens) to a file, restore it into a different slot of a fresh engine
and check that the continuation is identical.   (llama.cpp/examples/save-load-state)"""
import os
import tempfile

from common import load, parser

from distributed_llm_pipeline_b200.utils.session import load_sequence, save_sequence

ap = parser(__doc__)
ap.add_argument("-p", "--prompt", default="The quick brown fox")
ap.add_argument("-n", "--predict", type=int, default=16)
a = ap.parse_args()
eng, tok = load(a, max_seqs=2)
ids = tok.encode(a.prompt)
sid = eng.alloc_seq()
first = [int(eng.prefill(sid, ids).argmax())]
for _ in range(a.predict // 2 - 1):
    first.append(int(eng.decode([sid], [first[-1]], sample=False)[0].argmax()))
path = os.path.join(tempfile.mkdtemp(), "state.bin")
nbytes = save_sequence(eng, sid, path, tokens=ids + first[:-1])
rest = [first[-1]]
for _ in range(a.predict - len(first)):
    rest.append(int(eng.decode([sid], [rest[-1]], sample=False)[0].argmax()))
eng2, _ = load(a, max_seqs=2)
eng2.alloc_seq()                                            # restore into slot 1, not the slot it was saved from
sid2 = eng2.alloc_seq()
hdr = load_sequence(eng2, sid2, path)
rest2 = [first[-1]]
for _ in range(a.predict - len(first)):
    rest2.append(int(eng2.decode([sid2], [rest2[-1]], sample=False)[0].argmax()))
print(f"saved {nbytes} bytes ({hdr['n_tokens'] if 'n_tokens' in hdr else len(ids) + len(first) - 1} tokens) -> {path}")
print(f"original : {tok.decode(first + rest[1:], strip_leading_space=False)!r}")
print(f"restored : {tok.decode(first + rest2[1:], strip_leading_space=False)!r}")
print("identical continuation" if rest == rest2 else "MISMATCH")
raise SystemExit(0 if rest == rest2 else 1)
