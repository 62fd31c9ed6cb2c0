"""Python-facing ops: thin wrappers over the C ABI for CUDA tensors, PyTorch restatements for CPU tensors.

The CUDA path is the product; the CPU path exists so that engine / scheduling / parallel-layout logic can be
unit-tested without a GPU (gloo, world_size > 1).  A CUDA tensor never takes the fallback: if the native
library cannot be loaded the call raises.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import lib as _lib
from .weights import DeviceWeight

OUT_F16, OUT_F32, OUT_ADD_F32, OUT_SWIGLU_F16 = 0, 1, 2, 3
_OUT = {"f16": OUT_F16, "f32": OUT_F32, "add": OUT_ADD_F32, "swiglu": OUT_SWIGLU_F16}

import os as _os
TWO_PASS_DEQUANT = _os.environ.get("DLP_GEMM_DEQUANT_SCRATCH", "0") == "1"
# RoPE kernel warms L2 with the O-projection weights (decode).  Measured slower on B200 (F16 312.8 vs 346.1 tok/s, Q8_0 371 vs
# 390, profiles/r1_l2_prefetch_ab.txt): kept as an experiment switch, off by default.
PREFETCH_NEXT = _os.environ.get("DLP_L2_PREFETCH", "0") == "1"

# launch counter: every native kernel launch issued through this module (bench.py reports it)
launch_count = 0


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: torch.Tensor | None) -> int:
    return 0 if t is None else t.data_ptr()


def _count(n: int = 1):
    global launch_count
    launch_count += n


def native(t: torch.Tensor) -> bool:
    return t.is_cuda


# ------------------------------------------------------------------------------------------------
def dequant_rows(w: DeviceWeight, rows: torch.Tensor | None, out_dtype=torch.float32, out: torch.Tensor | None = None):
    """Decode rows of `w` (all rows when `rows` is None) -> [n, K]; the embedding gather."""
    n = w.N if rows is None else rows.numel()
    if w.dense is not None:
        res = (w.dense if rows is None else w.dense[rows.long()]).to(out_dtype)
        if out is not None:
            out.copy_(res)
            return out
        return res
    if out is None:
        out = torch.empty(n, w.K, dtype=out_dtype, device=w.device)
    L = _lib.load()
    fn = L.dlp_dequant_rows_f32 if out.dtype == torch.float32 else L.dlp_dequant_rows_f16
    if out.dtype not in (torch.float32, torch.float16):
        raise TypeError("dequant_rows: out dtype must be fp32 or fp16")
    _lib.check(fn(w.ref(), _p(rows), n, out.data_ptr(), out.stride(0), _stream()), "dequant_rows")
    _count()
    return out


def rmsnorm(h: torch.Tensor, w: torch.Tensor, eps: float, delta: torch.Tensor | None = None,
            out: torch.Tensor | None = None) -> torch.Tensor:
    """h (fp32 [T,E], updated in place with += delta) -> fp16 RMSNorm(h) * w."""
    T, E = h.shape
    if out is None:
        out = torch.empty(T, E, dtype=torch.float16, device=h.device)
    if not native(h):
        if delta is not None:
            h.add_(delta)
        out.copy_((h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps) * w).to(torch.float16))
        return out
    _lib.check(_lib.load().dlp_rmsnorm(h.data_ptr(), _p(delta), w.data_ptr(), out.data_ptr(), T, E, eps, _stream()), "rmsnorm")
    _count()
    return out


def layernorm(h: torch.Tensor, w: torch.Tensor | None, b: torch.Tensor | None, eps: float, delta: torch.Tensor | None = None,
              out: torch.Tensor | None = None) -> torch.Tensor:
    """h (fp32 [T,E], updated in place with += delta) -> fp16 LayerNorm(h) * w + b (either may be None)."""
    T, E = h.shape
    if out is None:
        out = torch.empty(T, E, dtype=torch.float16, device=h.device)
    if not native(h):
        if delta is not None:
            h.add_(delta)
        out.copy_(torch.nn.functional.layer_norm(h, (E,), w, b, eps).to(torch.float16))
        return out
    _lib.check(_lib.load().dlp_layernorm(h.data_ptr(), _p(delta), _p(w), _p(b), out.data_ptr(), T, E, eps, _stream()), "layernorm")
    _count()
    return out


ACTS = {"none": 0, "gelu": 1, "relu2": 2, "silu": 3, "gelu_erf": 4}


def bias_act(x: torch.Tensor, bias: torch.Tensor | None, act: str, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp16 act(x + bias) for fp32 x [T, F]: the step between the two mat-muls of a non-gated MLP."""
    T, F = x.shape
    if out is None:
        out = torch.empty(T, F, dtype=torch.float16, device=x.device)
    if not native(x) or x.dtype != torch.float32:
        v = x.float() + (bias if bias is not None else 0.0)
        fn = {"none": lambda t: t, "gelu": lambda t: torch.nn.functional.gelu(t, approximate="tanh"), "relu2": lambda t: torch.relu(t) ** 2,
              "silu": torch.nn.functional.silu, "gelu_erf": torch.nn.functional.gelu}[act]
        out.copy_(fn(v).to(torch.float16))
        return out
    _lib.check(_lib.load().dlp_bias_act(x.data_ptr(), _p(bias), out.data_ptr(), T, F, ACTS[act], _stream()), "bias_act")
    _count()
    return out


def _linear_cpu(w, x, w_up, norm_w, eps, out, mode):
    xf = x.float()
    if norm_w is not None:
        xf = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps) * norm_w).to(torch.float16).float()
    y = xf @ w.dense.T
    if mode == OUT_SWIGLU_F16:
        y = torch.nn.functional.silu(y) * (xf @ w_up.dense.T)
    if mode == OUT_ADD_F32:
        out.add_(y)
    else:
        out.copy_(y.to(out.dtype))
    return out


def linear(w: DeviceWeight, x: torch.Tensor, out_mode: str = "f16", w_up: DeviceWeight | None = None,
           norm_w: torch.Tensor | None = None, eps: float = 0.0, out: torch.Tensor | None = None,
           scratch: torch.Tensor | None = None) -> torch.Tensor:
    """y[T,N] (op)= x[T,K] @ W^T with the fused prologue/epilogue options of dlp_gemv / dlp_linear.

    x is fp16, or the fp32 residual stream when `norm_w` (RMSNorm fusion) is given.
    out_mode: "f16" | "f32" store, "add" (out fp32 += y; residual fusion), "swiglu" (needs w_up)."""
    mode = _OUT[out_mode]
    T = x.shape[0]
    if out is None:
        if mode == OUT_ADD_F32:
            raise ValueError("linear: 'add' needs the fp32 destination in `out`")
        out = torch.empty(T, w.N, dtype=torch.float32 if mode == OUT_F32 else torch.float16, device=x.device)
    if not native(x):
        return _linear_cpu(w, x, w_up, norm_w, eps, out, mode)
    L = _lib.load()
    up = w_up.ref() if w_up is not None else None
    if T in (1, 2, 4):
        _lib.check(L.dlp_gemv(w.ref(), up, x.data_ptr(), _p(norm_w), eps, out.data_ptr(), out.stride(0), mode, T,
                              _stream()), "gemv")
        _count()
        return out
    if T == 3:  # 2 + 1
        linear(w, x[:2], out_mode, w_up, norm_w, eps, out[:2])
        linear(w, x[2:], out_mode, w_up, norm_w, eps, out[2:])
        return out
    if norm_w is not None:
        x = rmsnorm(x, norm_w, eps)
    elif x.dtype != torch.float16:
        x = x.to(torch.float16)
    quant = w.wtype.name.startswith("Q")
    if quant and TWO_PASS_DEQUANT:   # A/B path: decode to an fp16 scratch, then the dense GEMM (default: fused in-kernel)
        need = w.N * w.K * 2 * (2 if w_up is not None else 1)
        if scratch is None or scratch.numel() * scratch.element_size() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
        _count(2 if w_up is not None else 1)
    sb = 0 if scratch is None else scratch.numel() * scratch.element_size()
    ws = gemm_workspace(x.device)                 # split-K workspace (skinny problems use every SM)
    L.dlp_linear_ws.restype = C.c_int
    L.dlp_linear_ws.argtypes = [C.POINTER(_lib.dlp_weight), C.POINTER(_lib.dlp_weight), C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                                C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
    _lib.check(L.dlp_linear_ws(w.ref(), up, x.data_ptr(), None, 0.0, out.data_ptr(), out.stride(0), mode, T, _p(scratch), sb,
                               _p(ws), 0 if ws is None else ws.numel(), _stream()), "linear")
    _count()
    return out


_GEMM_WS: dict = {}
GEMM_WS_BYTES = int(_os.environ.get("DLP_GEMM_WS_MB", "64")) << 20


def gemm_workspace(device) -> torch.Tensor | None:
    """Per-device split-K workspace of the tcgen05 GEMM: 16 KB of tile counters (zero once; they reset themselves) + partial tiles.
    One buffer serves every GEMM of a device because GEMM launches on a stream never overlap.  DLP_GEMM_WS_MB=0 disables split-K."""
    if GEMM_WS_BYTES <= 0:
        return None
    key = str(device)
    ws = _GEMM_WS.get(key)
    if ws is None:
        ws = torch.zeros(GEMM_WS_BYTES + 16384, dtype=torch.uint8, device=device)
        _GEMM_WS[key] = ws
    return ws


def gemm_f16(w: torch.Tensor, x: torch.Tensor, out: torch.Tensor, epilogue: int = 0, w_up: torch.Tensor | None = None,
             split_k: int = 0) -> torch.Tensor:
    """Direct access to the tcgen05 GEMM on dense fp16/bf16 tensors (x [T,K], w [N,K])."""
    T, K = x.shape
    N = w.shape[0]
    _lib.check(_lib.load().dlp_gemm_f16(w.data_ptr(), _p(w_up), w.stride(0), x.data_ptr(), x.stride(0), out.data_ptr(),
                                        out.stride(0), T, N, K, epilogue, int(w.dtype == torch.bfloat16), split_k,
                                        _stream()), "gemm_f16")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
def rope_kv(qkv: torch.Tensor, pos: torch.Tensor, slot: torch.Tensor, inv_freq: torch.Tensor, k_cache: torch.Tensor,
            v_cache: torch.Tensor, n_head: int, n_head_kv: int, head_dim: int, neox: bool = False,
            q_out: torch.Tensor | None = None, bias: torch.Tensor | None = None, q_norm: torch.Tensor | None = None,
            k_norm: torch.Tensor | None = None, eps: float = 1e-6, prefetch=None, attn_scale: float = 1.0) -> torch.Tensor:
    """Rotate Q,K; write K,V into the paged cache [pages, Hkv, page, D]; return q (fp16 [T, Hq*D], pre-scaled).
    Optional `bias` (qwen2) is added to q|k|v first; `q_norm`/`k_norm` (qwen3) are per-head RMSNorm weights.
    `prefetch`: a DeviceWeight whose planes (the next mat-vec's weights) the kernel pulls into L2 while it runs."""
    T = qkv.shape[0]
    page_size = k_cache.shape[2]
    n_rot = inv_freq.numel() * 2
    q_scale = attn_scale / math.sqrt(head_dim)     # attn_scale: YaRN magnitude correction folded into the pre-scaled Q
    if q_out is None:
        q_out = torch.empty(T, n_head * head_dim, dtype=torch.float16, device=qkv.device)
    if not native(qkv):
        from ..models.reference import apply_rope
        x = qkv.float()
        if bias is not None:
            x = x + bias
        x = x.view(T, n_head + 2 * n_head_kv, head_dim)
        xq, xk = x[:, :n_head], x[:, n_head:n_head + n_head_kv]
        if q_norm is not None:
            from ..models.reference import rms_norm
            xq, xk = rms_norm(xq, q_norm, eps), rms_norm(xk, k_norm, eps)
        q = apply_rope(xq, pos, inv_freq, neox)
        k = apply_rope(xk, pos, inv_freq, neox)
        v = x[:, n_head + n_head_kv:]
        q_out.copy_((q * q_scale).reshape(T, -1).to(torch.float16))
        ok = slot >= 0
        pg = (slot[ok] // page_size).long()
        off = (slot[ok] % page_size).long()
        k_cache[pg, :, off] = k[ok].to(k_cache.dtype)
        v_cache[pg, :, off] = v[ok].to(v_cache.dtype)
        return q_out
    pf = []
    if prefetch is not None and PREFETCH_NEXT and prefetch.dense is None:
        pf = [(pl.data_ptr(), pl.numel() * pl.element_size()) for pl in prefetch.planes[:2]]
        pf = [(p, n) for p, n in pf if p % 16 == 0 and n % 16 == 0 and n < 2**32]
    if bias is not None or q_norm is not None or pf:
        pf += [(None, 0)] * (2 - len(pf))
        _lib.check(_lib.load().dlp_rope_kv_ex(
            qkv.data_ptr(), int(qkv.dtype == torch.float16), pos.data_ptr(), slot.data_ptr(), inv_freq.data_ptr(),
            q_out.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), T, n_head, n_head_kv, head_dim, n_rot, page_size,
            int(neox), q_scale, bias.data_ptr() if bias is not None else None, q_norm.data_ptr() if q_norm is not None else None,
            k_norm.data_ptr() if k_norm is not None else None, float(eps), pf[0][0], pf[0][1], pf[1][0], pf[1][1], _stream()),
            "rope_kv_ex")
        _count()
        return q_out
    _lib.check(_lib.load().dlp_rope_kv(qkv.data_ptr(), int(qkv.dtype == torch.float16), pos.data_ptr(), slot.data_ptr(),
                                       inv_freq.data_ptr(), q_out.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), T,
                                       n_head, n_head_kv, head_dim, n_rot, page_size, int(neox), q_scale, _stream()),
               "rope_kv")
    _count()
    return q_out


def _attn_cpu(q, k_cache, v_cache, block_table_rows, kv_len, n_head, n_head_kv, head_dim):
    nq = q.shape[0]
    page = k_cache.shape[2]
    out = torch.empty(nq, n_head * head_dim, dtype=torch.float16, device=q.device)
    g = n_head // n_head_kv
    for i in range(nq):
        L = int(kv_len[i])
        if L == 0:
            out[i] = 0
            continue
        pages = block_table_rows[i][: (L + page - 1) // page].long()
        K = k_cache[pages].permute(1, 0, 2, 3).reshape(n_head_kv, -1, head_dim)[:, :L].float()
        V = v_cache[pages].permute(1, 0, 2, 3).reshape(n_head_kv, -1, head_dim)[:, :L].float()
        qi = q[i].float().view(n_head_kv, g, head_dim)
        s = torch.einsum("hgd,hld->hgl", qi, K)
        p = torch.softmax(s, -1)
        out[i] = torch.einsum("hgl,hld->hgd", p, V).reshape(-1).to(torch.float16)
    return out


def attn_paged(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
               kv_len: torch.Tensor, n_head: int, n_head_kv: int, head_dim: int, bt_row: torch.Tensor | None = None,
               n_splits: int = 1, workspace: tuple[torch.Tensor, torch.Tensor] | None = None,
               out: torch.Tensor | None = None) -> torch.Tensor:
    """Attention of nq independent queries (q fp16 [nq, Hq*D], already scaled) over the paged cache."""
    nq = q.shape[0]
    if not native(q):
        rows = block_table if bt_row is None else block_table[bt_row.long()]
        res = _attn_cpu(q, k_cache, v_cache, rows, kv_len, n_head, n_head_kv, head_dim)
        if out is not None:
            out.copy_(res)
            return out
        return res
    if out is None:
        out = torch.empty(nq, n_head * head_dim, dtype=torch.float16, device=q.device)
    po = pm = ctr = None
    if n_splits > 1:
        if workspace is None:
            po = torch.empty(nq * n_head * n_splits * head_dim, dtype=torch.float32, device=q.device)
            pm = torch.empty(nq * n_head * n_splits * 2, dtype=torch.float32, device=q.device)
        elif len(workspace) == 3:   # (partials, lse, counters): combine fused into the attention kernel
            po, pm, ctr = workspace
        else:
            po, pm = workspace
    if ctr is not None:
        _lib.check(_lib.load().dlp_attn_paged_fused(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table.data_ptr(),
                                                    _p(bt_row), kv_len.data_ptr(), out.data_ptr(), _p(po), _p(pm), ctr.data_ptr(), nq,
                                                    n_head, n_head_kv, head_dim, block_table.shape[1], k_cache.shape[2], n_splits,
                                                    _stream()), "attn_paged_fused")
        _count()
        return out
    _lib.check(_lib.load().dlp_attn_paged(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table.data_ptr(),
                                          _p(bt_row), kv_len.data_ptr(), out.data_ptr(), _p(po), _p(pm), nq, n_head,
                                          n_head_kv, head_dim, block_table.shape[1], k_cache.shape[2], n_splits,
                                          _stream()), "attn_paged")
    _count(2 if n_splits > 1 else 1)
    return out


def attn_decode_rope(qkv: torch.Tensor, inv_freq: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                     block_table: torch.Tensor, kv_len: torch.Tensor, n_head: int, n_head_kv: int, head_dim: int,
                     bt_row: torch.Tensor | None = None, n_splits: int = 8, neox: bool = False, attn_scale: float = 1.0,
                     out: torch.Tensor | None = None, rope_cs: torch.Tensor | None = None) -> torch.Tensor:
    """One-launch decode attention: RoPE on Q and the new K, K/V page write at position kv_len-1, split-KV attention with the
    splits combined inside a thread-block cluster.  qkv: fp32 [nq, (Hq+2Hkv)*D] straight from the QKV mat-vec."""
    nq = qkv.shape[0]
    if out is None:
        out = torch.empty(nq, n_head * head_dim, dtype=torch.float16, device=qkv.device)
    if not native(qkv) or qkv.dtype != torch.float32 or inv_freq.numel() * 2 != head_dim:
        ps = k_cache.shape[2]
        pos = (kv_len - 1).to(torch.int32)
        rows = block_table if bt_row is None else block_table[bt_row.long()]
        page = rows[torch.arange(nq, device=rows.device), (pos // ps).long()]
        slot = (page * ps + pos % ps).to(torch.int32)
        q = rope_kv(qkv, pos, slot, inv_freq, k_cache, v_cache, n_head, n_head_kv, head_dim, neox, attn_scale=attn_scale)
        return attn_paged(q, k_cache, v_cache, block_table, kv_len, n_head, n_head_kv, head_dim, bt_row=bt_row, out=out)
    _lib.check(_lib.load().dlp_attn_decode_rope(
        qkv.data_ptr(), inv_freq.data_ptr(), _p(rope_cs), k_cache.data_ptr(), v_cache.data_ptr(), block_table.data_ptr(), _p(bt_row),
        kv_len.data_ptr(), out.data_ptr(), nq, n_head, n_head_kv, head_dim, block_table.shape[1], k_cache.shape[2],
        max(1, min(16, n_splits)), int(neox), attn_scale / math.sqrt(head_dim), _stream()), "attn_decode_rope")
    _count()
    return out


ATTN_PREFILL_IMPL = _os.environ.get("DLP_ATTN_PREFILL", "tc")     # "tc" = tcgen05 kernel, "mma" = the mma.sync FA-2 kernel (A/B, fallback)


def attn_prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table_row: torch.Tensor, pos0: int,
                 n_head: int, n_head_kv: int, head_dim: int, out: torch.Tensor | None = None, impl: str | None = None) -> torch.Tensor:
    """Causal attention of a T-token chunk at positions [pos0, pos0+T) of one sequence."""
    T = q.shape[0]
    if not native(q):
        kv_len = torch.arange(pos0 + 1, pos0 + T + 1, dtype=torch.int32)
        rows = block_table_row.view(1, -1).expand(T, -1)
        res = _attn_cpu(q, k_cache, v_cache, rows, kv_len, n_head, n_head_kv, head_dim)
        if out is not None:
            out.copy_(res)
            return out
        return res
    if out is None:
        out = torch.empty(T, n_head * head_dim, dtype=torch.float16, device=q.device)
    page = k_cache.shape[2]
    impl = impl or ATTN_PREFILL_IMPL
    if (impl == "tc" and page % 8 == 0 and 128 % page == 0 and q.is_contiguous() and q.shape[1] == n_head * head_dim
            and k_cache.is_contiguous() and v_cache.is_contiguous()):
        # tcgen05 / TMEM / TMA flash-prefill (csrc/attention_tc.cu)
        L = _lib.load()
        L.dlp_attn_prefill_tc.restype = C.c_int
        L.dlp_attn_prefill_tc.argtypes = [C.c_void_p] * 5 + [C.c_int] * 8 + [C.c_void_p]
        _lib.check(L.dlp_attn_prefill_tc(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table_row.data_ptr(), out.data_ptr(),
                                         T, pos0, n_head, n_head_kv, head_dim, page, k_cache.shape[0], block_table_row.numel(), _stream()),
                   "attn_prefill_tc")
        _count()
        return out
    _lib.check(_lib.load().dlp_attn_prefill(q.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(), block_table_row.data_ptr(),
                                            out.data_ptr(), T, pos0, n_head, n_head_kv, head_dim, page,
                                            _stream()), "attn_prefill")
    _count()
    return out


def advance(seq_len: torch.Tensor, active: torch.Tensor | None, pos: torch.Tensor, slot: torch.Tensor,
            kv_len: torch.Tensor | None, block_table: torch.Tensor, page_size: int, n: int,
            inv_freq: torch.Tensor | None = None, rope_cs: torch.Tensor | None = None):
    """Device-side bookkeeping for the n batch entries whose sequence slots are active[:n] (see dlp_advance).  With
    `rope_cs` (fp32 [n, D/2, 2]) it also writes the step's (cos, sin) RoPE table for the one-pass attention kernel."""
    if native(seq_len) and rope_cs is not None:
        _lib.check(_lib.load().dlp_advance_rope(seq_len.data_ptr(), _p(active), pos.data_ptr(), slot.data_ptr(), _p(kv_len),
                                                block_table.data_ptr(), block_table.shape[1], page_size, n, inv_freq.data_ptr(),
                                                inv_freq.numel(), rope_cs.data_ptr(), _stream()), "advance_rope")
        _count()
        return
    if not native(seq_len):
        sid = (active[:n] if active is not None else torch.arange(n)).long()
        p = seq_len[sid]
        pos[:n].copy_(p)
        pg = (p // page_size).long()
        ok = pg < block_table.shape[1]
        pgc = pg.clamp_max(block_table.shape[1] - 1)
        s = block_table[sid, pgc] * page_size + p % page_size
        slot[:n].copy_(torch.where(ok, s, torch.full_like(s, -1)))
        if kv_len is not None:
            kv_len[:n].copy_(p + 1)
        seq_len[sid] = p + 1
        return
    _lib.check(_lib.load().dlp_advance(seq_len.data_ptr(), _p(active), pos.data_ptr(), slot.data_ptr(), _p(kv_len),
                                       block_table.data_ptr(), block_table.shape[1], page_size, n, _stream()), "advance")
    _count()


def argmax(logits: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    R, V = logits.shape
    if out is None:
        out = torch.empty(R, dtype=torch.int32, device=logits.device)
    if not native(logits):
        out.copy_(logits.argmax(-1).to(torch.int32))
        return out
    _lib.check(_lib.load().dlp_argmax(logits.data_ptr(), R, V, logits.stride(0), out.data_ptr(), _stream()), "argmax")
    _count()
    return out


def sample(logits: torch.Tensor, top_k: int = 40, top_p: float = 0.95, min_p: float = 0.05, temp: float = 0.8,
           seed: int = 0, step_ctr: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
    """top-k -> top-p -> min-p -> temperature -> draw (device-side; temp <= 0 is greedy)."""
    R, V = logits.shape
    if out is None:
        out = torch.empty(R, dtype=torch.int32, device=logits.device)
    if not native(logits):
        from ..serving.sampler import sample_reference
        ctr = int(step_ctr[0]) if step_ctr is not None else 0
        out.copy_(sample_reference(logits, top_k, top_p, min_p, temp, seed + ctr))
        return out
    _lib.check(_lib.load().dlp_sample(logits.data_ptr(), R, V, logits.stride(0), top_k, top_p, min_p, temp, seed & 0xFFFFFFFF,
                                      _p(step_ctr), out.data_ptr(), _stream()), "sample")
    _count()
    return out


SAMPLE_REC_WORDS = 8   # int32 words per dlp_sample_params record: top_k, top_p, min_p, temp, seed, ctr, 2 reserved


def pack_sample_record(top_k: int, top_p: float, min_p: float, temp: float, seed: int) -> torch.Tensor:
    """Host image (int32 [8]) of one dlp_sample_params record with its draw counter at zero."""
    import struct
    return torch.tensor(struct.unpack("8i", struct.pack("ifffII2I", int(top_k), float(top_p), float(min_p), float(temp),
                                                        seed & 0xFFFFFFFF, 0, 0, 0)), dtype=torch.int32)


def sample_dev(logits: torch.Tensor, params: torch.Tensor, row_slot: torch.Tensor | None, out: torch.Tensor) -> torch.Tensor:
    """Sampler chain with per-sequence records in device memory (`params` int32 [n_slots, 8]; row i uses record row_slot[i]).
    Nothing request-specific is a launch argument, so one captured graph serves every request."""
    R, V = logits.shape
    if not native(logits):
        import struct
        from ..serving.sampler import sample_reference
        for i in range(R):
            slot = int(row_slot[i]) if row_slot is not None else i
            k, tp, mp, t, seed, ctr, _, _ = struct.unpack("ifffII2I", params[slot].cpu().numpy().tobytes())
            if t <= 0:
                out[i] = int(logits[i].argmax())
            else:
                out[i:i + 1].copy_(sample_reference(logits[i:i + 1], k, tp, mp, t, seed + ctr))
                params[slot, 5] += 1
        return out
    lib = _lib.load()
    lib.dlp_sample_dev.restype = C.c_int
    lib.dlp_sample_dev.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    _lib.check(lib.dlp_sample_dev(logits.data_ptr(), R, V, logits.stride(0), params.data_ptr(), _p(row_slot), out.data_ptr(), _stream()),
               "sample_dev")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
def moe_route(logits: torch.Tensor, top_k: int, ids: torch.Tensor | None = None, weights: torch.Tensor | None = None):
    """softmax over all experts -> top-k -> renormalise; returns (ids int32 [T*k], weights fp32 [T*k]) on the device."""
    T, nE = logits.shape
    if ids is None:
        ids = torch.empty(T * top_k, dtype=torch.int32, device=logits.device)
        weights = torch.empty(T * top_k, dtype=torch.float32, device=logits.device)
    if not native(logits):
        pr = torch.softmax(logits.float(), -1)
        w, i = pr.topk(top_k, -1)
        ids.copy_(i.reshape(-1).to(torch.int32))
        weights.copy_((w / w.sum(-1, keepdim=True)).reshape(-1))
        return ids, weights
    _lib.check(_lib.load().dlp_moe_route(logits.data_ptr(), T, nE, top_k, ids.data_ptr(), weights.data_ptr(), _stream()), "moe_route")
    _count()
    return ids, weights


class MoePlan:
    """Device buffers of one routing plan (dlp_moe_plan_bufs) for up to `max_assign` (token, expert) assignments."""

    def __init__(self, max_assign: int, n_expert: int, device, token_tile_min: int = 16):
        i32 = torch.int32
        cap = max_assign // token_tile_min + n_expert + 1
        self.pos = torch.zeros(max_assign, dtype=i32, device=device)
        self.src = torch.zeros(max_assign, dtype=i32, device=device)
        self.bounds = torch.zeros(n_expert + 1, dtype=i32, device=device)
        self.tile_expert = torch.zeros(cap, dtype=i32, device=device)
        self.tile_row0 = torch.zeros(cap, dtype=i32, device=device)
        self.tile_rows = torch.zeros(cap, dtype=i32, device=device)
        self.n_tiles = torch.zeros(1, dtype=i32, device=device)

        class _B(C.Structure):
            _fields_ = [(n, C.c_void_p) for n in ("pos", "src", "bounds", "tile_expert", "tile_row0", "tile_rows", "n_tiles")]
        self.c = _B(self.pos.data_ptr(), self.src.data_ptr(), self.bounds.data_ptr(), self.tile_expert.data_ptr(),
                    self.tile_row0.data_ptr(), self.tile_rows.data_ptr(), self.n_tiles.data_ptr())


def moe_route_ex(logits: torch.Tensor, n_expert: int, top_k: int, renorm: bool, ids: torch.Tensor, weights: torch.Tensor):
    """softmax over the first n_expert columns -> top-k -> optional renormalisation (<= 256 experts), device only."""
    lib = _lib.load()
    lib.dlp_moe_route_ex.restype = C.c_int
    lib.dlp_moe_route_ex.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    _lib.check(lib.dlp_moe_route_ex(logits.data_ptr(), logits.stride(0), logits.shape[0], n_expert, top_k, int(renorm), ids.data_ptr(),
                                    weights.data_ptr(), _stream()), "moe_route_ex")
    _count()


def moe_plan(ids: torch.Tensor, n_expert: int, token_tile: int, plan: MoePlan):
    lib = _lib.load()
    lib.dlp_moe_plan.restype = C.c_int
    lib.dlp_moe_plan.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    _lib.check(lib.dlp_moe_plan(ids.data_ptr(), ids.numel(), n_expert, token_tile, C.byref(plan.c), _stream()), "moe_plan")
    _count()


def moe_gather(x: torch.Tensor, plan: MoePlan, top_k: int, n_assign: int, out: torch.Tensor):
    lib = _lib.load()
    lib.dlp_moe_gather.restype = C.c_int
    lib.dlp_moe_gather.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    _lib.check(lib.dlp_moe_gather(x.data_ptr(), x.stride(0), plan.src.data_ptr(), top_k, n_assign, x.shape[1], out.data_ptr(), _stream()), "moe_gather")
    _count()


def linear_grouped(w: DeviceWeight, xs: torch.Tensor, out: torch.Tensor, out_mode: str, n_expert: int, rows_per_expert: int, token_tile: int,
                   plan: MoePlan, w_up: DeviceWeight | None = None):
    """Grouped tcgen05 GEMM over expert-sorted rows: out[row] (op)= xs[row] . W_expert(row)^T, one launch for all experts."""
    lib = _lib.load()
    lib.dlp_linear_grouped.restype = C.c_int
    lib.dlp_linear_grouped.argtypes = [C.POINTER(_lib.dlp_weight), C.POINTER(_lib.dlp_weight), C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    ws = gemm_workspace(xs.device)
    _lib.check(lib.dlp_linear_grouped(w.ref(), w_up.ref() if w_up is not None else None, xs.data_ptr(), out.data_ptr(), out.stride(0),
                                      _OUT[out_mode], xs.shape[0], n_expert, rows_per_expert, token_tile, C.byref(plan.c), _p(ws),
                                      0 if ws is None else ws.numel(), _stream()), "linear_grouped")
    _count()
    return out


def moe_combine(y: torch.Tensor, plan: MoePlan, weights: torch.Tensor, T: int, top_k: int, h: torch.Tensor):
    lib = _lib.load()
    lib.dlp_moe_combine.restype = C.c_int
    lib.dlp_moe_combine.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    _lib.check(lib.dlp_moe_combine(y.data_ptr(), plan.pos.data_ptr(), weights.data_ptr(), T, top_k, y.shape[1], h.data_ptr(), _stream()), "moe_combine")
    _count()


def gemv_moe(w: DeviceWeight, x: torch.Tensor, out: torch.Tensor, out_mode: str, expert_id: torch.Tensor, rows_per_expert: int,
             w_up: DeviceWeight | None = None, norm_w: torch.Tensor | None = None, eps: float = 0.0,
             out_scale: torch.Tensor | None = None) -> torch.Tensor:
    """One token's mat-vec against expert `expert_id[0]` (device scalar) of a stacked [n_expert*rows, K] weight."""
    _lib.check(_lib.load().dlp_gemv_moe(w.ref(), w_up.ref() if w_up is not None else None, x.data_ptr(), _p(norm_w), eps,
                                        out.data_ptr(), out.stride(0), _OUT[out_mode], expert_id.data_ptr(), rows_per_expert,
                                        _p(out_scale), _stream()), "gemv_moe")
    _count()
    return out
