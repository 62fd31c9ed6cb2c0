import os, sys, time, random, threading
sys.path.insert(0, os.getcwd())
import torch
from distributed_llm_pipeline_b200.gguf import write_synthetic_gguf
from distributed_llm_pipeline_b200.models.config import preset
from distributed_llm_pipeline_b200.serving.orchestrator import build_service
from distributed_llm_pipeline_b200.serving.scheduler import Request
from distributed_llm_pipeline_b200.serving.sampler import SamplingParams
name, nl, ft = sys.argv[1], int(sys.argv[2]), sys.argv[3]
path = f"/tmp/dlp_models/{name}-{nl}L-{ft}.gguf"
os.makedirs("/tmp/dlp_models", exist_ok=True)
if not os.path.exists(path):
    write_synthetic_gguf(path, preset(name, n_layer=nl), ft, seed=0)
sched, info, logs = build_service(path, "cuda", 1024, 16, 2048)
print("warmup", sched.warmup(), flush=True)
sched.start()
rnd = random.Random(0)
for rnd_i in range(3):
    reqs = []
    for i in range(16):
        n = rnd.randrange(150, 420)
        toks = [1] + [rnd.randrange(3, 31000) for _ in range(n)]
        reqs.append(sched.submit(Request(prompt=toks, n_predict=48, sampling=SamplingParams(temperature=0.8, seed=i), ignore_eos=True)))
    for r in reqs:
        r.done.wait(120)
    torch.cuda.synchronize()
    print("round", rnd_i, [r.finish_reason for r in reqs][:4], sum(len(r.output_ids) for r in reqs), flush=True)
sched.stop()
print("ok")
