#!/usr/bin/env python
"""passkey: hide a number in a long stretch of filler text and ask for it; the prompt is fed in chunks and, once the window is
full, the engine slides it (`shift_seq`: the first --keep tokens stay, the oldest of the rest are dropped, stored keys are re-rotated)
so a prompt longer than the context can still be processed.   (llama.cpp/examples/passkey)"""
import random

from common import load, parser

ap = parser(__doc__)
ap.add_argument("--junk", type=int, default=40, help="repetitions of the filler sentence")
ap.add_argument("--pos", type=int, default=-1, help="repetition after which the passkey is inserted (default: random)")
ap.add_argument("--keep", type=int, default=32)
ap.add_argument("-b", "--batch", type=int, default=64)
ap.add_argument("-n", "--predict", type=int, default=12)
ap.add_argument("--seed", type=int, default=1)
a = ap.parse_args()
eng, tok = load(a, n_batch=a.batch)
rnd = random.Random(a.seed)
key = rnd.randrange(10000, 99999)
pos = a.pos if a.pos >= 0 else rnd.randrange(a.junk)
filler = "The grass is green. The sky is blue. The sun is yellow. Here we go. There and back again. "
text = "There is an important info hidden inside a lot of irrelevant text. Find it and memorize it. "
for i in range(a.junk):
    if i == pos:
        text += f"The pass key is {key}. Remember it. {key} is the pass key. "
    text += filler
text += "What is the pass key? The pass key is"
ids = tok.encode(text)
sid = eng.alloc_seq()
n_shift = 0
for c0 in range(0, len(ids), a.batch):
    chunk = ids[c0:c0 + a.batch]
    room = eng.n_ctx - a.predict - eng.seq_len(sid)
    if len(chunk) > room:                                   # slide the window: drop half of what is not pinned
        n_discard = max(len(chunk) - room, (eng.seq_len(sid) - a.keep) // 2)
        eng.shift_seq(sid, a.keep, n_discard)
        n_shift += 1
    logits = eng.prefill(sid, chunk)
out = [int(logits.argmax())]
for _ in range(a.predict - 1):
    out.append(int(eng.decode([sid], [out[-1]], sample=False)[0].argmax()))
print(f"prompt: {len(ids)} tokens through a {eng.n_ctx}-token window ({n_shift} shifts, {a.keep} tokens pinned), passkey {key} after filler {pos}")
print(f"answer: {tok.decode(out, strip_leading_space=False)!r}")
