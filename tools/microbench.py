#!/usr/bin/env python
"""Device-timed micro-benchmarks of the hot ops at Llama-3-8B shapes (CUDA events, warm-up, L2 flushed between reps)."""
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType  # noqa: E402
from distributed_llm_pipeline_b200.ops import api as ops  # noqa: E402
from distributed_llm_pipeline_b200.ops.weights import DeviceWeight  # noqa: E402

DEV = "cuda"
PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def timeit(fn, reps=20, warm=3, flush_l2=True):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        if flush_l2:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def dense(N, K, dtype=torch.float16):
    w = (torch.randn(N, K, device=DEV) * 0.02).to(dtype)
    return DeviceWeight(GGMLType.F16, N, K, [w])


rows = []


def report(name, us, flops=0, bytes_=0):
    tf = flops / us / 1e6 if flops else 0
    gb = bytes_ / us / 1e3 if bytes_ else 0
    rows.append({"op": name, "us": round(us, 2), "tflops": round(tf, 1), "gbs": round(gb, 1),
                 "frac_of_measured": round(max(tf / PEAKS["bf16_tflops"], gb / PEAKS["hbm_gbs"]), 3)})
    print(f"{name:44s} {us:9.2f} us  {tf:8.1f} TFLOP/s  {gb:8.1f} GB/s  ({rows[-1]['frac_of_measured']:.2f} of measured peak)", flush=True)


E, F, QKV, V = 4096, 14336, 6144, 128256
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "gemm"):
    for T in (512, 2048):
        x = torch.randn(T, E, device=DEV).half()
        xf = torch.randn(T, F, device=DEV).half()
        h = torch.zeros(T, E, device=DEV)
        for name, w, xin, mode, up in (("qkv", dense(QKV, E), x, "f16", None), ("o_proj(add)", dense(E, E), x, "add", None),
                                       ("gate_up(swiglu)", dense(F, E), x, "swiglu", dense(F, E)), ("down(add)", dense(E, F), xf, "add", None)):
            out = h if mode == "add" else None
            us = timeit(lambda: ops.linear(w, xin, mode, w_up=up, out=out))
            fl = 2 * T * w.N * w.K * (2 if up is not None else 1)
            report(f"gemm T={T} {name} [{w.N}x{w.K}]", us, flops=fl)
        if T == 512:
            a, b = torch.randn(T, E, device=DEV).half(), torch.randn(QKV, E, device=DEV).half()
            us = timeit(lambda: torch.matmul(a, b.T))
            report(f"cuBLAS fp16 T={T} qkv (reference point)", us, flops=2 * T * QKV * E)
if which in ("all", "gemv"):
    hh = torch.randn(1, E, device=DEV)
    nw = torch.ones(E, device=DEV)
    act = torch.randn(1, F, device=DEV).half()
    att = torch.randn(1, E, device=DEV).half()
    for name, w, xin, mode, up, norm in (("qkv+norm", dense(QKV, E), hh, "f32", None, nw), ("o_proj(add)", dense(E, E), att, "add", None, None),
                                         ("gate_up+norm(swiglu)", dense(F, E), hh, "swiglu", dense(F, E), nw), ("down(add)", dense(E, F), act, "add", None, None),
                                         ("lm_head+norm", dense(V, E), hh, "f32", None, nw)):
        out = torch.zeros(1, w.N, device=DEV) if mode == "add" else None
        us = timeit(lambda: ops.linear(w, xin, mode, w_up=up, norm_w=norm, eps=1e-5, out=out))
        report(f"gemv T=1 {name} [{w.N}x{w.K}]", us, bytes_=w.N * w.K * 2 * (2 if up is not None else 1))
if which in ("all", "small"):
    Hq, Hkv, D, page = 32, 8, 128, 64
    for T in (1, 512):
        qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=DEV)
        pos = torch.arange(T, dtype=torch.int32, device=DEV)
        slot = pos.clone()
        inv = (500000.0 ** (-torch.arange(0, D // 2).double() * 2 / D)).float().to(DEV)
        kc = torch.zeros(64, Hkv, page, D, dtype=torch.float16, device=DEV)
        vc = torch.zeros_like(kc)
        report(f"rope_kv T={T}", timeit(lambda: ops.rope_kv(qkv, pos, slot, inv, kc, vc, Hq, Hkv, D), flush_l2=False))
        h = torch.randn(T, E, device=DEV)
        report(f"rmsnorm T={T}", timeit(lambda: ops.rmsnorm(h, torch.ones(E, device=DEV), 1e-5), flush_l2=False))
    bt = torch.arange(64, dtype=torch.int32, device=DEV).view(1, 64)
    q = torch.randn(512, Hq * D, device=DEV).half()
    report("attn_prefill T=512", timeit(lambda: ops.attn_prefill(q, kc, vc, bt[0], 0, Hq, Hkv, D), flush_l2=False), flops=4 * 512 * 512 * Hq * D / 2)
    for ctx in (128, 4096):
        n_pages = (ctx + page - 1) // page
        kc2 = torch.randn(n_pages, Hkv, page, D, device=DEV).half()
        vc2 = torch.randn_like(kc2)
        bt2 = torch.arange(n_pages, dtype=torch.int32, device=DEV).view(1, -1)
        kvl = torch.tensor([ctx], dtype=torch.int32, device=DEV)
        q1 = torch.randn(1, Hq * D, device=DEV).half()
        report(f"attn_decode ctx={ctx} splits=16", timeit(lambda: ops.attn_paged(q1, kc2, vc2, bt2, kvl, Hq, Hkv, D, n_splits=16), flush_l2=False),
               bytes_=2 * ctx * Hkv * D * 2)
    lg = torch.randn(1, V, device=DEV)
    report("argmax V=128256", timeit(lambda: ops.argmax(lg), flush_l2=False))
    report("sample top-k/top-p V=128256", timeit(lambda: ops.sample(lg, temp=0.8), flush_l2=False))
    tok = torch.tensor([5], dtype=torch.int32, device=DEV)
    emb = dense(V, E)
    report("embed 1 row", timeit(lambda: ops.dequant_rows(emb, tok), flush_l2=False))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open(f"gpurun_out/microbench_{which}.json", "w"), indent=1)
