#!/usr/bin/env python
"""parallel: simulated clients submit requests to the serving core at once; slots, continuous batching and prompt-prefix reuse do
the rest.   (llama.cpp/examples/parallel)"""
import threading
import time

from common import load, parser

from distributed_llm_pipeline_b200.serving.sampler import SamplingParams
from distributed_llm_pipeline_b200.serving.scheduler import Scheduler

ap = parser(__doc__)
ap.add_argument("-np", "--parallel", type=int, default=4, help="slots")
ap.add_argument("-ns", "--sequences", type=int, default=8, help="requests in total")
ap.add_argument("-n", "--predict", type=int, default=16)
a = ap.parse_args()
eng, tok = load(a, max_seqs=a.parallel + 1)
sched = Scheduler(eng, tok, n_slots=a.parallel)
sched.start()
system = "You are a helpful assistant. "
questions = ["What is the capital of France?", "Name a prime number.", "How many legs has a spider?", "What colour is the sky?"]
done, t0 = [], time.time()


def client(i):
    r = sched.generate(system + questions[i % len(questions)], n_predict=a.predict, sampling=SamplingParams(temperature=0.0), ignore_eos=True)
    done.append((i, r))


threads = [threading.Thread(target=client, args=(i,)) for i in range(a.sequences)]
[t.start() for t in threads]
[t.join() for t in threads]
sched.stop()
dt = time.time() - t0
for i, r in sorted(done):
    tm = r.timings()
    print(f"client {i}: {tm['prompt_n']} prompt tokens ({tm['cache_n']} from the cache), {tm['predicted_n']} generated | {r.text[:48]!r}")
n = sum(len(r.output_ids) for _, r in done)
print(f"\n[{a.sequences} requests over {a.parallel} slots: {n} tokens in {dt:.2f} s = {n / dt:.1f} tok/s]")
