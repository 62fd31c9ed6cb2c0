"""Numerics of every CUDA kernel against a plain PyTorch fp32 restatement of the same op.

Metric and thresholds follow the reference's op tests (NMSE; MUL_MAT / FLASH_ATTN_EXT 5e-4:
llama.cpp/tests/test-backend-ops.cpp:273-287,3716-3718,6049-6051)."""
import math

import pytest
import torch

from distributed_llm_pipeline_b200.gguf import GGMLType
from distributed_llm_pipeline_b200.models.reference import apply_rope, nmse
from distributed_llm_pipeline_b200.ops import api as ops
from distributed_llm_pipeline_b200.ops.weights import DeviceWeight

pytestmark = pytest.mark.gpu
DEV = "cuda"
WTYPES = [GGMLType.F16, GGMLType.BF16, GGMLType.Q8_0, GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K]


def _w(N, K, t, seed=0):
    g = torch.Generator().manual_seed(seed)
    dense = torch.randn(N, K, generator=g) * 0.05
    w = DeviceWeight.from_dense(dense, t, DEV)
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    exact = torch.from_numpy(dequantize(quantize(dense.numpy(), t), t, (N, K)).astype(np.float32)).to(DEV)
    return w, exact


@pytest.mark.parametrize("t", WTYPES)
def test_dequant_rows_exact(t):
    w, exact = _w(64, 512, t)
    out = ops.dequant_rows(w, None)
    assert torch.equal(out, exact) or nmse(exact, out) < 1e-9
    rows = torch.tensor([5, 0, 63, 5], dtype=torch.int32, device=DEV)
    assert nmse(exact[rows.long()], ops.dequant_rows(w, rows)) < 1e-9


@pytest.mark.parametrize("t", WTYPES)
@pytest.mark.parametrize("T", [1, 2, 4])
def test_gemv_store_and_add(t, T):
    N, K = 384, 1024
    w, exact = _w(N, K, t, 1)
    x = (torch.randn(T, K, device=DEV) * 2).half()
    ref = x.float() @ exact.T
    y16 = ops.linear(w, x, "f16")
    y32 = ops.linear(w, x, "f32")
    # block-quantised weights use int8 activations + dp4a on the decode path (the reference's q8_1 numerics class;
    # its own MUL_MAT tolerance is 5e-4): allow 5e-5 there, 5e-6 for the dense types
    tol = 5e-5 if t.name.startswith("Q") else 5e-6
    assert nmse(ref, y32) < tol and nmse(ref, y16) < tol
    acc = torch.ones(T, N, device=DEV)
    ops.linear(w, x, "add", out=acc)
    assert nmse(ref + 1, acc) < tol


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_K, GGMLType.Q6_K])
def test_gemv_fused_norm_and_swiglu(t):
    N, K, T = 512, 768, 2
    wg, eg = _w(N, K, t, 2)
    wu, eu = _w(N, K, t, 3)
    h = torch.randn(T, K, device=DEV) * 3
    nw = torch.rand(K, device=DEV) + 0.5
    eps = 1e-5
    xn = (h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps) * nw).half().float()
    ref = torch.nn.functional.silu(xn @ eg.T) * (xn @ eu.T)
    y = ops.linear(wg, h, "swiglu", w_up=wu, norm_w=nw, eps=eps)
    assert nmse(ref, y) < (2e-4 if t.name.startswith("Q") else 2e-5)


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.BF16])
@pytest.mark.parametrize("T", [1, 2, 4])
@pytest.mark.parametrize("N,K", [(6144, 1024), (4736, 2048), (3072, 3584)])
def test_gemv_large_shapes(t, T, N, K):
    """Decode-sized matrices (thousands of rows, the shapes the balanced row ranges / dynamic unit queue are tuned for): every
    epilogue, repeated launches, bit-identical results run to run."""
    w, exact = _w(N, K, t, 5)
    wu, eu = _w(N, K, t, 6)
    x = (torch.randn(T, K, device=DEV) * 2).half()
    ref = x.float() @ exact.T
    for _ in range(3):
        assert nmse(ref, ops.linear(w, x, "f32")) < 5e-6
        assert nmse(ref, ops.linear(w, x, "f16")) < 5e-6
    acc = torch.ones(T, N, device=DEV)
    ops.linear(w, x, "add", out=acc)
    ops.linear(w, x, "add", out=acc)
    assert nmse(2 * ref + 1, acc) < 5e-6
    h = torch.randn(T, K, device=DEV) * 3
    nw = torch.rand(K, device=DEV) + 0.5
    xn = (h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + 1e-5) * nw).half().float()
    refs = torch.nn.functional.silu(xn @ exact.T) * (xn @ eu.T)
    for _ in range(2):
        assert nmse(refs, ops.linear(w, h, "swiglu", w_up=wu, norm_w=nw, eps=1e-5)) < 2e-5
    a, b = ops.linear(w, x, "f32"), ops.linear(w, x, "f32")
    assert torch.equal(a, b)            # chunk partials are summed in a fixed order: bit-identical run to run


@pytest.mark.parametrize("T,N,K", [(5, 128, 64), (16, 256, 256), (37, 384, 1024), (128, 512, 512), (300, 128, 2048), (512, 1024, 4096)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_gemm_tcgen05_store(T, N, K, dtype):
    w = (torch.randn(N, K, device=DEV) * 0.05).to(dtype)
    x = torch.randn(T, K, device=DEV).to(dtype)
    ref = x.float() @ w.float().T
    y = torch.empty(T, N, device=DEV, dtype=torch.float32)
    ops.gemm_f16(w, x, y, epilogue=1)
    assert nmse(ref, y) < 1e-6
    y16 = torch.empty(T, N, device=DEV, dtype=torch.float16)
    ops.gemm_f16(w, x, y16, epilogue=0)
    assert nmse(ref, y16) < 5e-6


def test_gemm_tcgen05_ragged_n_splitk_swiglu():
    T, N, K = 70, 200, 1536   # N not a multiple of 128: OOB rows are zero-filled and masked
    w = (torch.randn(N, K, device=DEV) * 0.05).half()
    u = (torch.randn(N, K, device=DEV) * 0.05).half()
    x = torch.randn(T, K, device=DEV).half()
    ref = x.float() @ w.float().T
    acc = torch.full((T, N), 2.0, device=DEV)
    ops.gemm_f16(w, x, acc, epilogue=2, split_k=4)
    assert nmse(ref + 2, acc) < 1e-6
    sw = torch.empty(T, N, device=DEV, dtype=torch.float16)
    ops.gemm_f16(w, x, sw, epilogue=3, w_up=u)
    assert nmse(torch.nn.functional.silu(ref) * (x.float() @ u.float().T), sw) < 5e-6


@pytest.mark.parametrize("t", [GGMLType.Q8_0, GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K])
@pytest.mark.parametrize("T,N,K", [(48, 256, 512), (300, 200, 1024), (7, 384, 2048)])
def test_linear_quant_dequant_fused_gemm(t, T, N, K):
    """Block-quant decode fused into the tcgen05 GEMM's A-operand producer (no scratch), all epilogues."""
    w, exact = _w(N, K, t, 4)
    u, exact_u = _w(N, K, t, 5)
    x = torch.randn(T, K, device=DEV).half()
    ref = x.float() @ exact.T
    assert nmse(ref, ops.linear(w, x, "f32")) < 5e-6            # weights are rounded to fp16 after decode
    assert nmse(ref, ops.linear(w, x, "f16")) < 5e-6
    acc = torch.ones(T, N, device=DEV)
    ops.linear(w, x, "add", out=acc)
    assert nmse(ref + 1, acc) < 5e-6
    sw = ops.linear(w, x, "swiglu", w_up=u)
    assert nmse(torch.nn.functional.silu(ref) * (x.float() @ exact_u.T), sw) < 2e-5


def test_linear_quant_two_pass_scratch_entry_point():
    """dlp_linear's caller-provided-scratch contract still holds (used when DLP_GEMM_DEQUANT_SCRATCH=1)."""
    import ctypes as C
    from distributed_llm_pipeline_b200.ops import lib as L
    T, N, K = 48, 256, 512
    w, exact = _w(N, K, GGMLType.Q8_0, 4)
    x = torch.randn(T, K, device=DEV).half()
    y = torch.empty(T, N, device=DEV)
    scratch = torch.empty(N * K * 2, dtype=torch.uint8, device=DEV)
    rc = L.load().dlp_linear(w.ref(), None, x.data_ptr(), None, 0.0, y.data_ptr(), N, 1, T, scratch.data_ptr(), scratch.numel(),
                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0 and nmse(x.float() @ exact.T, y) < 5e-6


def test_rmsnorm_with_residual():
    T, E = 7, 1024
    h = torch.randn(T, E, device=DEV)
    d = torch.randn(T, E, device=DEV)
    w = torch.rand(E, device=DEV) + 0.5
    ref_h = h + d
    ref = ref_h * torch.rsqrt(ref_h.pow(2).mean(-1, keepdim=True) + 1e-5) * w
    out = ops.rmsnorm(h, w, 1e-5, delta=d)
    assert torch.allclose(h, ref_h, atol=1e-6) and nmse(ref, out) < 1e-6


@pytest.mark.parametrize("affine", ["wb", "w", "none"])
@pytest.mark.parametrize("T,E", [(1, 256), (7, 1024), (130, 6144)])
def test_layernorm_with_residual(affine, T, E):
    """LayerNorm of the starcoder2 / falcon / olmo / orion / command-r families against torch in fp32: with weight and bias, weight
    only, and no parameters; the residual add into h included."""
    h = torch.randn(T, E, device=DEV) * 2 + 0.5          # non-zero mean: the centring matters
    d = torch.randn(T, E, device=DEV)
    w = (torch.rand(E, device=DEV) + 0.5) if affine != "none" else None
    b = torch.randn(E, device=DEV) * 0.1 if affine == "wb" else None
    ref_h = h + d
    ref = torch.nn.functional.layer_norm(ref_h, (E,), w, b, 1e-5)
    out = ops.layernorm(h, w, b, 1e-5, delta=d)
    assert torch.allclose(h, ref_h, atol=1e-6) and nmse(ref, out) < 1e-6
    h2 = ref_h.clone()
    assert nmse(ref, ops.layernorm(h2, w, b, 1e-5)) < 1e-6 and torch.equal(h2, ref_h)      # without delta h is left alone


@pytest.mark.parametrize("act", ["gelu", "gelu_erf", "relu2", "silu", "none"])
def test_bias_act(act):
    """Bias + activation between the two mat-muls of a non-gated MLP, against torch in fp32."""
    T, F = 9, 1536
    x = torch.randn(T, F, device=DEV) * 3
    bias = torch.randn(F, device=DEV)
    fn = {"none": lambda t: t, "gelu": lambda t: torch.nn.functional.gelu(t, approximate="tanh"), "relu2": lambda t: torch.relu(t) ** 2,
          "silu": torch.nn.functional.silu, "gelu_erf": torch.nn.functional.gelu}[act]
    assert nmse(fn(x + bias), ops.bias_act(x, bias, act)) < 1e-6
    assert nmse(fn(x), ops.bias_act(x, None, act)) < 1e-6


@pytest.mark.parametrize("neox", [False, True])
@pytest.mark.parametrize("D", [64, 128])
def test_rope_kv_write(neox, D):
    T, Hq, Hkv, page = 9, 4, 2, 16
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=DEV)
    pos = torch.arange(30, 30 + T, dtype=torch.int32, device=DEV)
    pages = torch.tensor([3, 1, 2], dtype=torch.int32)
    slot = (pages[(pos.cpu() // page).long()] * page + pos.cpu() % page).to(torch.int32).to(DEV)
    inv = (10000.0 ** (-torch.arange(0, D // 2).double() * 2 / D)).float().to(DEV)
    kc = torch.zeros(4, Hkv, page, D, dtype=torch.float16, device=DEV)
    vc = torch.zeros_like(kc)
    q = ops.rope_kv(qkv, pos, slot, inv, kc, vc, Hq, Hkv, D, neox)
    x = qkv.view(T, Hq + 2 * Hkv, D)
    q_ref = apply_rope(x[:, :Hq].cpu(), pos.cpu(), inv.cpu(), neox) / math.sqrt(D)
    k_ref = apply_rope(x[:, Hq:Hq + Hkv].cpu(), pos.cpu(), inv.cpu(), neox)
    assert nmse(q_ref.reshape(T, -1), q) < 1e-6
    pg, off = (slot.cpu() // page).long(), (slot.cpu() % page).long()
    assert nmse(k_ref, kc.cpu()[pg, :, off]) < 1e-6
    assert nmse(x[:, Hq + Hkv:].cpu(), vc.cpu()[pg, :, off]) < 1e-6


@pytest.mark.parametrize("in_dtype", [torch.float32, torch.float16])
@pytest.mark.parametrize("D,with_bias,with_norm", [(64, True, False), (128, False, True), (64, True, True), (128, True, True)])
def test_rope_kv_bias_and_head_norm(D, with_bias, with_norm, in_dtype):
    """qwen2 bias / qwen3 per-head Q,K RMSNorm fused into the RoPE + KV-write launch, against the fp32 formula."""
    from distributed_llm_pipeline_b200.models.reference import rms_norm
    T, Hq, Hkv, page, eps = 7, 6, 2, 16, 1e-6
    H = Hq + 2 * Hkv
    qkv = torch.randn(T, H * D, device=DEV).to(in_dtype)
    pos = torch.arange(3, 3 + T, dtype=torch.int32, device=DEV)
    slot = (pos + 2 * page).to(torch.int32)
    inv = (10000.0 ** (-torch.arange(0, D // 2).double() * 2 / D)).float().to(DEV)
    kc = torch.zeros(4, Hkv, page, D, dtype=torch.float16, device=DEV)
    vc = torch.zeros_like(kc)
    bias = torch.randn(H * D, device=DEV) * 0.3 if with_bias else None
    qn = 1 + 0.2 * torch.randn(D, device=DEV) if with_norm else None
    kn = 1 + 0.2 * torch.randn(D, device=DEV) if with_norm else None
    q = ops.rope_kv(qkv, pos, slot, inv, kc, vc, Hq, Hkv, D, True, bias=bias, q_norm=qn, k_norm=kn, eps=eps)
    x = qkv.float().cpu()
    if with_bias:
        x = x + bias.cpu()
    x = x.view(T, H, D)
    xq, xk = x[:, :Hq], x[:, Hq:Hq + Hkv]
    if with_norm:
        xq, xk = rms_norm(xq, qn.cpu(), eps), rms_norm(xk, kn.cpu(), eps)
    q_ref = apply_rope(xq, pos.cpu(), inv.cpu(), True) / math.sqrt(D)
    k_ref = apply_rope(xk, pos.cpu(), inv.cpu(), True)
    tol = 1e-6 if in_dtype == torch.float32 else 2e-6
    assert nmse(q_ref.reshape(T, -1), q) < tol
    pg, off = (slot.cpu() // page).long(), (slot.cpu() % page).long()
    assert nmse(k_ref, kc.cpu()[pg, :, off]) < tol
    assert nmse(x[:, Hq + Hkv:], vc.cpu()[pg, :, off]) < tol


def _fill_cache(n_pages, Hkv, page, D, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    kc = torch.randn(n_pages, Hkv, page, D, device=DEV, generator=g).half()
    vc = torch.randn(n_pages, Hkv, page, D, device=DEV, generator=g).half()
    return kc, vc


@pytest.mark.parametrize("D,Hq,Hkv", [(128, 8, 2), (64, 8, 1), (128, 4, 4), (64, 6, 3)])
@pytest.mark.parametrize("n_splits", [1, 4])
def test_attn_paged_decode(D, Hq, Hkv, n_splits):
    page, n_pages, max_pages = 16, 200, 48
    kc, vc = _fill_cache(n_pages, Hkv, page, D)
    lens = torch.tensor([1, 17, 300, 700], dtype=torch.int32, device=DEV)   # short: single CTA; long: split-KV + combine
    nq = lens.numel()
    bt = torch.randperm(n_pages)[: nq * max_pages].view(nq, max_pages).to(torch.int32).to(DEV)
    q = (torch.randn(nq, Hq * D, device=DEV) / math.sqrt(D)).half()
    out = ops.attn_paged(q, kc, vc, bt, lens, Hq, Hkv, D, n_splits=n_splits)
    ref = ops._attn_cpu(q.cpu(), kc.cpu(), vc.cpu(), bt.cpu(), lens.cpu(), Hq, Hkv, D)
    assert nmse(ref, out) < 5e-5
    if n_splits > 1:   # combine fused into the attention kernel (last split CTA merges); twice: counters must self-reset
        ws = (torch.empty(nq * Hq * n_splits * D, device=DEV), torch.empty(nq * Hq * n_splits * 2, device=DEV),
              torch.zeros(nq * Hq, dtype=torch.int32, device=DEV))
        for _ in range(2):
            out2 = ops.attn_paged(q, kc, vc, bt, lens, Hq, Hkv, D, n_splits=n_splits, workspace=ws)
            assert nmse(ref, out2) < 5e-5
        assert int(ws[2].abs().sum()) == 0


@pytest.mark.parametrize("D,Hq,Hkv", [(128, 8, 2), (64, 8, 1), (128, 4, 4), (64, 6, 3)])
@pytest.mark.parametrize("neox", [False, True])
@pytest.mark.parametrize("n_splits", [1, 8, 16])
def test_attn_decode_rope_onepass(D, Hq, Hkv, neox, n_splits):
    """RoPE + K/V page write + split-KV attention + cluster (DSMEM) combine in one launch, against the separate fp32 steps."""
    page, n_pages, max_pages = 16, 400, 80
    kc, vc = _fill_cache(n_pages, Hkv, page, D, seed=3)
    lens = torch.tensor([1, 2, 17, 64, 65, 300, 700, 1200], dtype=torch.int32, device=DEV)   # kv_len includes the new token
    nq = lens.numel()
    bt = torch.randperm(n_pages)[: nq * 40].view(nq, 40).to(torch.int32)
    bt = torch.cat([bt, torch.zeros(nq, max_pages - 40, dtype=torch.int32)], 1).to(DEV)          # 8 x 40 pages of 16 = 640 < 1200?
    lens = torch.minimum(lens, torch.tensor(40 * page, device=DEV)).to(torch.int32)
    H = Hq + 2 * Hkv
    qkv = torch.randn(nq, H * D, device=DEV)
    inv = (10000.0 ** (-torch.arange(0, D // 2).double() * 2 / D)).float().to(DEV)
    # oracle: CPU rope + cache write + attention on copies of the cache
    kc_ref, vc_ref = kc.cpu().clone(), vc.cpu().clone()
    pos = (lens - 1).cpu()
    slot = (bt.cpu()[torch.arange(nq), (pos // page).long()] * page + pos % page).to(torch.int32)
    q_ref = ops.rope_kv(qkv.cpu(), pos.to(torch.int32), slot, inv.cpu(), kc_ref, vc_ref, Hq, Hkv, D, neox)
    ref = ops._attn_cpu(q_ref, kc_ref, vc_ref, bt.cpu(), lens.cpu(), Hq, Hkv, D)
    ang = pos.float()[:, None] * inv.cpu()[None]
    table = torch.stack([ang.cos(), ang.sin()], -1).contiguous().to(DEV)      # the per-step table advance_rope writes
    for tab in (None, table):      # in-kernel sincos, then the table; the second call rewrites the same slots with the same values
        out = ops.attn_decode_rope(qkv, inv, kc, vc, bt, lens, Hq, Hkv, D, n_splits=n_splits, neox=neox, rope_cs=tab)
        assert nmse(ref, out) < 5e-5
    pg, off = (slot // page).long(), (slot % page).long()
    assert nmse(kc_ref[pg, :, off], kc.cpu()[pg, :, off]) < 1e-6
    assert nmse(vc_ref[pg, :, off], vc.cpu()[pg, :, off]) < 1e-6
    untouched = torch.ones(n_pages, dtype=torch.bool)
    untouched[pg] = False
    assert torch.equal(kc.cpu()[untouched], kc_ref[untouched])          # nothing else in the cache was written


@pytest.mark.parametrize("impl", ["tc", "mma"])
@pytest.mark.parametrize("D,Hq,Hkv", [(128, 8, 2), (64, 4, 2)])
@pytest.mark.parametrize("T,pos0,page", [(5, 0, 16), (64, 0, 16), (100, 37, 16), (200, 64, 16), (128, 0, 64), (300, 250, 64), (513, 1000, 64), (1, 400, 8)])
def test_attn_prefill_causal(D, Hq, Hkv, T, pos0, page, impl):
    """Both prefill-attention kernels (tcgen05/TMEM/TMA and the mma.sync FA-2 one) against fp32 attention over the same pages:
    tiles that end inside a page, chunks that start mid-sequence, more than one 128-query tile, several 128-key tiles."""
    total = pos0 + T
    n_need = (total + page - 1) // page
    kc, vc = _fill_cache(n_need + 3, Hkv, page, D, seed=1)
    row = torch.randperm(n_need + 3)[:n_need].to(torch.int32).to(DEV)
    q = (torch.randn(T, Hq * D, device=DEV) / math.sqrt(D)).half()
    out = ops.attn_prefill(q, kc, vc, row, pos0, Hq, Hkv, D, impl=impl)
    kv_len = torch.arange(pos0 + 1, pos0 + T + 1, dtype=torch.int32)
    ref = ops._attn_cpu(q.cpu(), kc.cpu(), vc.cpu(), row.cpu().view(1, -1).expand(T, -1), kv_len, Hq, Hkv, D)
    assert nmse(ref, out) < 5e-4


def test_argmax_and_sampler():
    R, V = 3, 32000
    lg = torch.randn(R, V, device=DEV)
    assert torch.equal(ops.argmax(lg).long(), lg.argmax(-1))
    # greedy through the sampler entry point
    assert torch.equal(ops.sample(lg, temp=0.0).long(), lg.argmax(-1))
    # top_k=1 must return the argmax whatever the temperature
    assert torch.equal(ops.sample(lg, top_k=1, temp=1.5).long(), lg.argmax(-1))
    # draws stay inside the top-k set and follow the distribution roughly
    lg2 = torch.full((1, V), -10.0, device=DEV)
    lg2[0, 7], lg2[0, 99], lg2[0, 1234] = 2.0, 1.0, 0.0
    ctr = torch.zeros(1, dtype=torch.int32, device=DEV)
    counts = {7: 0, 99: 0, 1234: 0}
    for i in range(300):
        ctr.fill_(i)
        t = int(ops.sample(lg2, top_k=3, top_p=1.0, min_p=0.0, temp=1.0, seed=5, step_ctr=ctr)[0])
        assert t in counts
        counts[t] += 1
    assert counts[7] > counts[99] > counts[1234] > 0
    # top-p keeps only the head of the distribution
    for i in range(50):
        ctr.fill_(i)
        assert int(ops.sample(lg2, top_k=40, top_p=0.5, min_p=0.0, temp=1.0, seed=1, step_ctr=ctr)[0]) == 7


def test_advance_bookkeeping():
    page, B, max_pages = 16, 3, 4
    bt = torch.arange(B * max_pages, dtype=torch.int32, device=DEV).view(B, max_pages)
    sl = torch.tensor([0, 15, 16], dtype=torch.int32, device=DEV)
    pos = torch.zeros(B, dtype=torch.int32, device=DEV)
    slot = torch.zeros(B, dtype=torch.int32, device=DEV)
    kvl = torch.zeros(B, dtype=torch.int32, device=DEV)
    ops.advance(sl, None, pos, slot, kvl, bt, page, B)
    assert kvl.tolist() == [1, 16, 17]
    assert pos.tolist() == [0, 15, 16] and sl.tolist() == [1, 16, 17]
    assert slot.tolist() == [0 * 16 + 0, 4 * 16 + 15, 9 * 16 + 0]


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_K])
@pytest.mark.parametrize("T", [1, 2, 4])
def test_fused_allreduce_protocol_virtual_ranks(t, T):
    """Row-parallel GEMV -> self-signalling stores into every rank's buffer; next GEMV's prologue spins on the values, folds
    h += sum(partials) into its RMSNorm and re-arms the other parity.  Four virtual ranks on one GPU, phases launched rank by
    rank (no co-scheduled waiting kernels); five steps so that both parities are reused after being re-armed."""
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    tp, E, F = 4, 512, 1024
    comms = FusedComm.make_virtual(tp, E, n_sites=2)
    g = torch.Generator().manual_seed(11)
    w_down = torch.randn(E, F, generator=g) * 0.05          # row-parallel: K = F sharded
    w_next = torch.randn(256, E, generator=g) * 0.05         # next column-parallel matrix (replicated here)
    nw = (torch.rand(E, generator=g) + 0.5).to(DEV)
    act = (torch.randn(T, F, generator=g)).half().to(DEV)
    h = torch.randn(T, E, generator=g).to(DEV)
    Fs = F // tp
    shards = [DeviceWeight.from_dense(w_down[:, r * Fs:(r + 1) * Fs].contiguous(), t, DEV) for r in range(tp)]
    wn = DeviceWeight.from_dense(w_next, t, DEV)
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    deq = lambda m: torch.from_numpy(dequantize(quantize(m.numpy(), t), t, tuple(m.shape)).astype(np.float32)).to(DEV)
    for step in range(5):                                       # several epochs through the same site: no stale data
        for c in comms:
            c.begin_step()
        for r, c in enumerate(comms):                           # phase 1: every rank pushes its partial rows
            c.gemv(shards[r], act[:, r * Fs:(r + 1) * Fs].contiguous(), 4, T, ar_out=c.ar_out(0))
        partial_sum = sum(act[:, r * Fs:(r + 1) * Fs].float() @ deq(w_down[:, r * Fs:(r + 1) * Fs].contiguous()).T for r in range(tp))
        h_ref = h + partial_sum
        xn = (h_ref * torch.rsqrt(h_ref.pow(2).mean(-1, keepdim=True) + 1e-5) * nw).half().float()
        y_ref = xn @ deq(w_next).T
        for r, c in enumerate(comms):                           # phase 2: every rank consumes
            h_out = torch.zeros_like(h)
            y = torch.empty(T, 256, device=DEV)
            c.gemv(wn, h, 1, T, out=y, norm_w=nw, eps=1e-5, ar_in=c.ar_in(0, h_out))
            assert nmse(h_ref, h_out) < (1e-4 if t.name.startswith("Q") else 1e-6)
            assert nmse(y_ref, y) < (3e-4 if t.name.startswith("Q") else 2e-5)
            c.check_error()
        act = (act.float() * 0.5 + 0.1).half()


def test_fused_allreduce_passes_negative_zero_and_changing_batch():
    """-0.0 is the "empty" pattern on the wire: a partial that is exactly -0.0 must arrive as +0.0, not hang the consumer; and a
    step with fewer token rows after one with more must not see the stale rows (the whole site is re-armed)."""
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    tp, E, F = 2, 256, 512
    comms = FusedComm.make_virtual(tp, E, n_sites=1)
    g = torch.Generator().manual_seed(5)
    w_next = DeviceWeight.from_dense(torch.randn(64, E, generator=g) * 0.05, GGMLType.F16, DEV)
    nw = torch.ones(E, device=DEV)
    zero_w = [DeviceWeight.from_dense(torch.zeros(E, F // tp), GGMLType.F16, DEV) for _ in range(tp)]
    for step, T in enumerate([4, 1, 4, 2, 1]):
        act = -torch.ones(T, F // tp, dtype=torch.float16, device=DEV)       # (-1) * (+0) = -0.0 in every partial
        h = torch.randn(T, E, generator=g).to(DEV)
        for c in comms:
            c.begin_step()
        for r, c in enumerate(comms):
            c.gemv(zero_w[r], act, 4, T, ar_out=c.ar_out(0))
        for c in comms:
            h_out = torch.zeros_like(h)
            y = torch.empty(T, 64, device=DEV)
            c.gemv(w_next, h, 1, T, out=y, norm_w=nw, eps=1e-5, ar_in=c.ar_in(0, h_out))
            torch.cuda.synchronize()
            c.check_error()
            assert torch.equal(h_out, h), step


def test_sharded_argmax_emulated_rank():
    """dlp_argmax_tp on the single-device emulation (rank 0's stores land in every slot of its own buffer): picks the shard's
    best entry, lowest id on ties, offset by the shard's first global id; slots are re-armed across steps."""
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    c = FusedComm.make_emulated(4, 256, 1)
    g = torch.Generator().manual_seed(9)
    for step in range(5):
        B, V, v0 = [1, 2, 4, 1, 37][step], 4099, 1000
        lg = torch.randn(B, V, generator=g).to(DEV)
        lg[0, 77] = lg[0, 300] = 50.0                      # tie: the lower id wins
        out = torch.zeros(B, dtype=torch.int32, device=DEV)
        c.argmax_sharded(lg, v0, out)
        torch.cuda.synchronize()
        c.check_error()
        want = lg.argmax(-1) + v0
        want[0] = 77 + v0
        assert out.tolist() == want.tolist(), step


def test_emulated_tp_engine_fused_equals_unfused(tmp_models):
    """The single-device emulation of a tp-way rank (bench.py --emulate-tp): its fused decode step (stores + spin + sharded
    arg-max inside a CUDA graph) must reproduce the unfused step of the same shard (separate all-reduce stand-in)."""
    from distributed_llm_pipeline_b200.models.llama import LlamaEngine
    from distributed_llm_pipeline_b200.models.weights import load_weights
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    path = tmp_models[("tiny128", "F16")]
    tp = 2
    w = load_weights(path, "cuda", tp_rank=0, tp_size=tp)
    fused = LlamaEngine(w, n_ctx=128, max_seqs=2, page_size=16, max_batch_tokens=32, comm=FusedComm.make_emulated(tp))
    plain_comm = FusedComm.make_emulated(tp)
    plain_comm.fused = False
    plain = LlamaEngine(w, n_ctx=128, max_seqs=2, page_size=16, max_batch_tokens=32, comm=plain_comm)
    assert fused.fused_tp and not plain.fused_tp
    prompt = [1, 5, 9, 200, 33, 7, 8]
    sa, sb = fused.alloc_seq(), plain.alloc_seq()
    la, lb = fused.prefill(sa, prompt), plain.prefill(sb, prompt)
    assert nmse(lb, la) < 1e-5
    tok = int(lb.argmax())
    want = [tok]
    for _ in range(9):
        lb = plain.decode([sb], [want[-1]], sample=False, use_graph=False)[0]
        la = fused.decode([sa], [want[-1]], sample=False)[0]
        assert nmse(lb, la) < 1e-4
        want.append(int(lb.argmax()))
    fused.free_seq(sa)
    assert fused.generate(prompt, 10) == want            # device-fed tokens through the sharded arg-max
    fused.comm.check_error()


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0])
@pytest.mark.parametrize("T", [64, 200])
def test_fused_gemm_reduce_scatter_allgather_virtual_ranks(t, T):
    """Sequence-parallel prefill pair on 4 virtual ranks (one GPU, phases launched rank by rank):
    row-parallel GEMM --RS push--> owners' reduce buffers --rs_norm_ag--> everyone's gather buffer --AG wait--> GEMM."""
    from distributed_llm_pipeline_b200.parallel.fused import FusedComm
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    tp, E, F, N2 = 4, 512, 1024, 384
    Tp = (T + tp - 1) // tp * tp
    rows = Tp // tp
    comms = FusedComm.make_virtual_sp(tp, E, Tp)
    g = torch.Generator().manual_seed(21)
    w_down = torch.randn(E, F, generator=g) * 0.05
    w_next = torch.randn(N2, E, generator=g) * 0.05
    nw = (torch.rand(E, generator=g) + 0.5).to(DEV)
    act = torch.zeros(Tp, F)
    act[:T] = torch.randn(T, F, generator=g)
    act = act.half().to(DEV)
    h_full = torch.zeros(Tp, E)
    h_full[:T] = torch.randn(T, E, generator=g)
    h_full = h_full.to(DEV)
    Fs = F // tp
    deq = lambda m: torch.from_numpy(dequantize(quantize(m.numpy(), t), t, tuple(m.shape)).astype(np.float32)).to(DEV)
    shards = [DeviceWeight.from_dense(w_down[:, r * Fs:(r + 1) * Fs].contiguous(), t, DEV) for r in range(tp)]
    wn = DeviceWeight.from_dense(w_next, t, DEV)
    h_sh = [h_full[r * rows:(r + 1) * rows].clone() for r in range(tp)]
    for rep in range(2):        # twice through the same buffers: epochs advance, nothing stale is read
        for r, c in enumerate(comms):      # phase 1: row-parallel GEMM, partial tiles pushed to the owner ranks
            c.sp_gemm(shards[r], act[:, r * Fs:(r + 1) * Fs].contiguous(), Fs, Tp, rows, 5, rs_to=0)
        for r, c in enumerate(comms):      # phase 2: owners reduce + norm + push their rows into every gather buffer
            c.sp_norm_ag(h_sh[r], rows, nw, 1e-5, g=0, red=0)
        partial = sum(act[:, r * Fs:(r + 1) * Fs].float() @ deq(w_down[:, r * Fs:(r + 1) * Fs].contiguous()).T for r in range(tp))
        h_full = h_full + partial
        xn = (h_full * torch.rsqrt(h_full.pow(2).mean(-1, keepdim=True) + 1e-5) * nw).half().float()
        y_ref = xn @ deq(w_next).T
        for r, c in enumerate(comms):      # phase 3: column-parallel GEMM reads the gathered X (waits per owner flag)
            assert nmse(h_full[r * rows:(r + 1) * rows], h_sh[r]) < 1e-6
            y = torch.empty(Tp, N2, device=DEV)
            c.sp_gemm(wn, c.sp_gather_view(0, Tp), E, Tp, rows, 1, y=y, x_from_g=0)
            assert nmse(y_ref[:T], y[:T]) < 2e-5
            c.check_error()
        act = (act.float() * 0.7).half()


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_K])
def test_moe_route_and_expert_indexed_gemv(t):
    """Device-side MoE decode pieces: router (softmax -> top-2 -> renormalise) and mat-vecs that pick the expert's rows
    from a stacked weight through a device-resident id, with the expert weight applied in the epilogue."""
    nE, k, E, F = 8, 2, 512, 768
    g = torch.Generator().manual_seed(31)
    logits = torch.randn(3, nE, generator=g).to(DEV)
    ids, w = ops.moe_route(logits, k)
    pr = torch.softmax(logits, -1)
    tw, ti = pr.topk(k, -1)
    assert torch.equal(ids.view(3, k).long(), ti)
    assert torch.allclose(w.view(3, k), tw / tw.sum(-1, keepdim=True), atol=1e-6)
    # stacked experts
    gate = torch.randn(nE * F, E, generator=g) * 0.05
    up = torch.randn(nE * F, E, generator=g) * 0.05
    down = torch.randn(nE * E, F, generator=g) * 0.05
    Wg, Wu, Wd = (DeviceWeight.from_dense(m, t, DEV) for m in (gate, up, down))
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    deq = lambda m: torch.from_numpy(dequantize(quantize(m.numpy(), t), t, tuple(m.shape)).astype(np.float32)).to(DEV)
    dg, du, dd = deq(gate), deq(up), deq(down)
    h = torch.randn(1, E, generator=g).to(DEV)
    nw = (torch.rand(E, generator=g) + 0.5).to(DEV)
    xn = (h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + 1e-5) * nw).half().float()
    href = h.clone()
    hout = h.clone()
    act = torch.empty(k, F, dtype=torch.float16, device=DEV)
    for j in range(k):
        e = int(ids[j])
        ops.gemv_moe(Wg, h, act[j:j + 1], "swiglu", ids[j:j + 1], F, w_up=Wu, norm_w=nw, eps=1e-5)
        a_ref = torch.nn.functional.silu(xn @ dg[e * F:(e + 1) * F].T) * (xn @ du[e * F:(e + 1) * F].T)
        assert nmse(a_ref, act[j:j + 1]) < (3e-4 if t.name.startswith("Q") else 2e-5)
        href += float(w[j]) * (act[j:j + 1].float() @ dd[e * E:(e + 1) * E].T)
    for j in range(k):
        ops.gemv_moe(Wd, act[j:j + 1], hout, "add", ids[j:j + 1], E, out_scale=w[j:j + 1])
    assert nmse(href, hout) < (1e-4 if t.name.startswith("Q") else 1e-6)


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_K])
@pytest.mark.parametrize("T,nE,k", [(1, 4, 2), (7, 8, 2), (64, 8, 2), (300, 8, 2), (33, 60, 4), (129, 128, 8)])
def test_moe_grouped_gemm_pipeline(t, T, nE, k):
    """route -> plan -> gather -> grouped SwiGLU GEMM -> grouped down GEMM -> combine, against per-token fp32 math with the
    same (dequantised) expert weights.  Covers experts that receive no token, partial token tiles and > 32 experts."""
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    E, F = 256, 256
    g = torch.Generator().manual_seed(100 + T + nE)
    logits = torch.randn(T, nE, generator=g).to(DEV)
    if T >= 64:
        logits[:, 1] = -30.0                                   # expert 1 never selected: an empty group in the plan
    A = T * k
    ids = torch.zeros(A, dtype=torch.int32, device=DEV)
    w = torch.zeros(A, dtype=torch.float32, device=DEV)
    for renorm in (True, False):
        ops.moe_route_ex(logits, nE, k, renorm, ids, w)
        pr = torch.softmax(logits, -1)
        tw, ti = pr.topk(k, -1)
        assert torch.equal(ids.view(T, k).long(), ti)
        assert torch.allclose(w.view(T, k), tw / tw.sum(-1, keepdim=True) if renorm else tw, atol=1e-6)
    bt = 16 if A // nE <= 16 else 32 if A // nE <= 32 else 64
    plan = ops.MoePlan(A, nE, DEV)
    ops.moe_plan(ids, nE, bt, plan)
    pos, src, bounds = plan.pos.long(), plan.src.long(), plan.bounds.tolist()
    assert bounds[0] == 0 and bounds[-1] == A
    assert torch.equal(src[pos], torch.arange(A, device=DEV))                     # a permutation ...
    for e in range(nE):
        rows = src[bounds[e]:bounds[e + 1]]
        assert bool((ids[rows] == e).all()) and bool((rows[1:] > rows[:-1]).all())   # ... grouped by expert, stable in token order
    nt = int(plan.n_tiles)
    assert nt == sum((bounds[e + 1] - bounds[e] + bt - 1) // bt for e in range(nE))
    assert int(plan.tile_rows[:nt].sum()) == A
    gate = torch.randn(nE * F, E, generator=g) * 0.05
    up = torch.randn(nE * F, E, generator=g) * 0.05
    down = torch.randn(nE * E, F, generator=g) * 0.05
    Wg, Wu, Wd = (DeviceWeight.from_dense(m, t, DEV) for m in (gate, up, down))
    deq = lambda m: torch.from_numpy(dequantize(quantize(m.numpy(), t), t, tuple(m.shape)).astype(np.float32)).to(DEV)
    dg, du, dd = deq(gate), deq(up), deq(down)
    x = torch.randn(T, E, generator=g).half().to(DEV)
    xs = torch.zeros(A, E, dtype=torch.float16, device=DEV)
    ops.moe_gather(x, plan, k, A, xs)
    assert torch.equal(xs, x[src // k])
    act = torch.zeros(A, F, dtype=torch.float16, device=DEV)
    ops.linear_grouped(Wg, xs, act, "swiglu", nE, F, bt, plan, w_up=Wu)
    y = torch.zeros(A, E, dtype=torch.float32, device=DEV)
    ops.linear_grouped(Wd, act, y, "f32", nE, E, bt, plan)
    h = torch.randn(T, E, generator=g).to(DEV)
    h_ref = h.clone()
    ops.moe_combine(y, plan, w, T, k, h)
    tol = 3e-4 if t.name.startswith("Q") else 2e-5
    for e in range(nE):
        sel = (ids == e).nonzero().flatten()
        if sel.numel() == 0:
            continue
        xe = x[sel // k].float()
        a_ref = torch.nn.functional.silu(xe @ dg[e * F:(e + 1) * F].T) * (xe @ du[e * F:(e + 1) * F].T)
        assert nmse(a_ref, act[pos[sel]]) < tol, e
        y_ref = act[pos[sel]].float() @ dd[e * E:(e + 1) * E].T
        assert nmse(y_ref, y[pos[sel]]) < tol, e
        h_ref.index_add_(0, sel // k, w[sel][:, None] * y[pos[sel]])
    assert nmse(h_ref, h) < 1e-6


@pytest.mark.parametrize("t", [GGMLType.F16, GGMLType.Q8_0, GGMLType.Q4_K])
@pytest.mark.parametrize("T", [5, 8, 32, 64, 200])
@pytest.mark.parametrize("mode", ["f16", "f32", "add", "swiglu"])
def test_gemm_split_k_workspace_fixup(t, T, mode):
    """Skinny problems (few 128-row tiles) split K across CTAs through the workspace; the last split adds the partials in split
    order, so the result matches the fp32 reference and is bit-identical from run to run and to the unsplit kernel's rounding
    class.  N = 384 gives 3 tiles -> 8 splits of K = 4096."""
    from distributed_llm_pipeline_b200.gguf.quants import dequantize, quantize
    import numpy as np
    N, K = 384, 4096
    g = torch.Generator().manual_seed(7 + T)
    w = torch.randn(N, K, generator=g) * 0.03
    w2 = torch.randn(N, K, generator=g) * 0.03
    x = torch.randn(T, K, generator=g).half().to(DEV)
    W, W2 = DeviceWeight.from_dense(w, t, DEV), DeviceWeight.from_dense(w2, t, DEV)
    deq = lambda m: torch.from_numpy(dequantize(quantize(m.numpy(), t), t, tuple(m.shape)).astype(np.float32)).to(DEV)
    y_ref = x.float() @ deq(w).T
    if mode == "swiglu":
        y_ref = torch.nn.functional.silu(y_ref) * (x.float() @ deq(w2).T)
    outs = []
    for rep in range(2):
        if mode == "add":
            out = torch.ones(T, N, device=DEV)
            ops.linear(W, x, "add", out=out)
            got = out - 1.0
        else:
            got = ops.linear(W, x, mode, w_up=W2 if mode == "swiglu" else None).float()
        outs.append(got.clone())
    tol = 4e-4 if t.name.startswith("Q") else (1e-5 if mode in ("f32", "add") else 2e-5)
    assert nmse(y_ref, outs[0]) < tol
    if mode != "add":          # (the residual-add epilogue reduces its splits with atomics: order, hence rounding, may differ)
        assert torch.equal(outs[0], outs[1])
    ws = ops.gemm_workspace(torch.device(DEV))
    assert int(ws[:16384].view(torch.int32).abs().sum()) == 0          # every tile counter went back to zero
