#!/usr/bin/env python
"""simple-chat: a chat loop over the model's template that keeps the conversation in the KV cache and only evaluates the new
turn each time.   (llama.cpp/examples/simple-chat)"""
import sys

from common import load, parser

from distributed_llm_pipeline_b200.serving.orchestrator import apply_chat_template

ap = parser(__doc__)
ap.add_argument("-n", "--predict", type=int, default=64)
ap.add_argument("--turn", action="append", help="scripted user turns instead of stdin")
a = ap.parse_args()
eng, tok = load(a)
meta = {"tokenizer.chat_template": tok.chat_template if hasattr(tok, "chat_template") else ""}
sid, messages, cached = eng.alloc_seq(), [], []
turns = iter(a.turn) if a.turn else (line.rstrip("\n") for line in sys.stdin)
for user in turns:
    if not user:
        break
    messages.append({"role": "user", "content": user})
    ids = tok.encode(apply_chat_template(meta, messages))
    keep = 0
    while keep < min(len(cached), len(ids) - 1) and cached[keep] == ids[keep]:      # the common prefix is already evaluated
        keep += 1
    eng.truncate_seq(sid, keep)
    reply = eng.generate(ids[keep:], a.predict, sid=sid, stop_ids=tuple(tok.stop_ids))
    cached = ids + reply[:-1]
    eng.truncate_seq(sid, len(cached))
    text = tok.decode([t for t in reply if t not in tok.stop_ids], strip_leading_space=False)
    messages.append({"role": "assistant", "content": text})
    print(f"> {user}\n{text}\n[{len(ids) - keep} new prompt tokens, {keep} reused]", flush=True)
