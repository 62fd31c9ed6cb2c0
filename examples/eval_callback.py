#!/usr/bin/env python
"""eval-callback: print a summary of the residual stream after every layer while a prompt is evaluated.
(llama.cpp/examples/eval-callback; the hook is `LlamaEngine.eval_callback`, the counterpart of `ggml_backend_sched_eval_callback`)"""
from common import load, parser

ap = parser(__doc__)
ap.add_argument("-p", "--prompt", default="Hello my name is")
a = ap.parse_args()
eng, tok = load(a)


def show(name, il, t):
    f = t.float()
    print(f"{name}-{il:<3d} shape {tuple(t.shape)}  mean {float(f.mean()):+.5f}  std {float(f.std()):.5f}  max |x| {float(f.abs().max()):.4f}  "
          f"first row [{', '.join(f'{x:+.4f}' for x in f[0, :4].tolist())}, ...]")


eng.eval_callback = show
sid = eng.alloc_seq()
logits = eng.prefill(sid, tok.encode(a.prompt))
eng.eval_callback = None
print(f"next token: {int(logits.argmax())} {tok.decode([int(logits.argmax())], strip_leading_space=False)!r}")
