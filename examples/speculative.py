#!/usr/bin/env python
"""speculative / lookup: greedy generation with a draft model (`-md`) or prompt-lookup n-grams proposing blocks that the target
verifies in one forward; prints the text and the acceptance statistics.   (llama.cpp/examples/speculative, examples/lookup)"""
import time

from common import load, parser

from distributed_llm_pipeline_b200.serving.speculative import ModelDrafter, NgramDrafter, SpecStats, generate_speculative

ap = parser(__doc__)
ap.add_argument("-md", "--model-draft", default=None, help="draft GGUF (same vocabulary); default: prompt-lookup drafts")
ap.add_argument("-p", "--prompt", default="one two three one two three one two")
ap.add_argument("-n", "--predict", type=int, default=32)
ap.add_argument("--draft-max", type=int, default=8)
a = ap.parse_args()
eng, tok = load(a, max_seqs=2)
drafter = NgramDrafter()
if a.model_draft:
    import argparse
    d_eng, _ = load(argparse.Namespace(model=a.model_draft, device=a.device, ctx_size=a.ctx_size), max_seqs=2)
    drafter = ModelDrafter(d_eng)
ids = tok.encode(a.prompt)
t0 = time.time()
plain = eng.generate(ids, a.predict)
t_plain = time.time() - t0
st = SpecStats()
t0 = time.time()
out = generate_speculative(eng, ids, a.predict, drafter, a.draft_max, stats=st)
t_spec = time.time() - t0
print(a.prompt + tok.decode(out, strip_leading_space=False))
print(f"\n[drafted {st.n_drafted}, accepted {st.n_accepted} ({100 * st.acceptance:.0f} %), {st.tokens_per_pass:.2f} tokens per target pass; "
      f"{t_plain:.2f} s plain vs {t_spec:.2f} s speculative; same tokens as plain greedy: {out == plain}]")
