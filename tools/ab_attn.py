#!/usr/bin/env python
"""A/B of the decode-attention step in isolation: N back-to-back layers inside one CUDA graph (PDL on), per-layer device time.

  old      : rope_kv kernel + split-KV attention with the ticket/atomic combine (global workspace)
  onepass  : attn_decode_rope (RoPE + KV write + attention + cluster/DSMEM combine), cluster size 8 / 16, sincos in-kernel or
             from the per-step table
Each variant is preceded by a mat-vec so the PDL chain looks like the real layer (QKV -> attention -> Wo)."""
import argparse
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from distributed_llm_pipeline_b200.ops import api as ops  # noqa: E402
from distributed_llm_pipeline_b200.ops.weights import DeviceWeight  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--len", type=int, default=600)
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
dev = "cuda"
Hq, Hkv, D, E, page = 32, 8, 128, 4096, 64
H = Hq + 2 * Hkv
n_pages = 64
kc = torch.randn(n_pages, Hkv, page, D, device=dev).half()
vc = torch.randn(n_pages, Hkv, page, D, device=dev).half()
bt = torch.randperm(n_pages).view(1, -1).to(torch.int32).to(dev)
kv_len = torch.tensor([a.len], dtype=torch.int32, device=dev)
pos = kv_len - 1
slot = (bt[0, (pos // page).long()] * page + pos % page).to(torch.int32)
inv = (500000.0 ** (-torch.arange(0, D // 2).double() * 2 / D)).float().to(dev)
ang = pos.float()[:, None] * inv[None]
table = torch.stack([ang.cos(), ang.sin()], -1).contiguous()
h = torch.randn(1, E, device=dev)
norm_w = torch.ones(E, device=dev)
wqkv = DeviceWeight(0, H * D, E, [torch.randn(H * D, E, device=dev).half() * 0.02])
wqkv.wtype = __import__("distributed_llm_pipeline_b200.gguf", fromlist=["GGMLType"]).GGMLType.F16
wo = DeviceWeight(wqkv.wtype, E, Hq * D, [torch.randn(E, Hq * D, device=dev).half() * 0.02])
qkv = torch.zeros(1, H * D, device=dev)
q = torch.zeros(1, Hq * D, dtype=torch.float16, device=dev)
att = torch.zeros(1, Hq * D, dtype=torch.float16, device=dev)
ns = 16
ws = (torch.zeros(Hq * ns * D, device=dev), torch.zeros(Hq * ns * 2, device=dev), torch.zeros(Hq, dtype=torch.int32, device=dev))


def layer(kind):
    ops.linear(wqkv, h, "f32", norm_w=norm_w, eps=1e-5, out=qkv)
    if kind == "old":
        ops.rope_kv(qkv, pos, slot, inv, kc, vc, Hq, Hkv, D, False, q_out=q)
        ops.attn_paged(q, kc, vc, bt, kv_len, Hq, Hkv, D, n_splits=ns, workspace=ws, out=att)
    elif kind == "none":
        pass
    else:
        _, cl, tab = kind.split("-")
        ops.attn_decode_rope(qkv, inv, kc, vc, bt, kv_len, Hq, Hkv, D, n_splits=int(cl), out=att, rope_cs=table if tab == "tab" else None)
    ops.linear(wo, att, "add", out=h)


def bench(kind):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            layer(kind)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(a.layers):
                layer(kind)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / a.reps / a.layers


base = bench("none")
print(f"kv_len {a.len}: mat-vecs only {base:.2f} us/layer")
for kind in ("old", "onepass-8-sincos", "onepass-8-tab", "onepass-16-tab", "onepass-4-tab", "onepass-1-tab"):
    t = bench(kind)
    print(f"  {kind:18s} {t:7.2f} us/layer   (+{t - base:.2f} us over the mat-vecs)")
