"""Device-resident weight matrices in kernel layout (+ load-time repack of GGUF blocks).

Parity: replaces the reference's buffer-type machinery for weights (`select_weight_buft`,
`llama.cpp/src/llama-model.cpp:312-323`) and its CPU-side weight repacking (`ggml-cpu/repack.cpp`) with one
explicit rule per type, documented in `csrc/quant.cuh`: Q8_0 and Q6_K are split into 16-byte-aligned
structure-of-arrays planes at load; Q4_K / Q5_K super-blocks are already 16-byte multiples and stay native.
Sharding happens on the raw bytes *before* upload, so each rank only ever touches its own slice of the file.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np
import torch

from ..gguf.constants import BLOCK_GEOMETRY, GGMLType
from ..gguf.quants import dequantize
from .lib import dlp_weight

KERNEL_TYPES = {GGMLType.F16, GGMLType.BF16, GGMLType.Q8_0, GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K}


def slice_raw(raw: np.ndarray, ggml_type: GGMLType, shape: tuple[int, int], rows: tuple[int, int] | None = None,
              cols: tuple[int, int] | None = None) -> tuple[np.ndarray, tuple[int, int]]:
    """Slice a 2-D weight [N, K] held as raw bytes [N, row_bytes] by output rows and/or (block-aligned) K range."""
    N, K = shape
    per, nbytes = BLOCK_GEOMETRY[GGMLType(ggml_type)]
    raw = raw.reshape(N, K // per * nbytes)
    if rows is not None:
        raw = raw[rows[0]:rows[1]]
        N = rows[1] - rows[0]
    if cols is not None:
        k0, k1 = cols
        if k0 % per or k1 % per:
            raise ValueError(f"K range {cols} is not aligned to the {per}-element block of {GGMLType(ggml_type).name}")
        raw = raw[:, k0 // per * nbytes:k1 // per * nbytes]
        K = k1 - k0
    return raw, (N, K)


def scale_raw(raw: np.ndarray, ggml_type: GGMLType, shape: tuple[int, int], s: float) -> tuple[np.ndarray, GGMLType]:
    """Multiply a weight matrix held as GGUF bytes by a scalar without decoding it: block formats only have their fp16 block scales
    rewritten (granite's residual multiplier is folded into Wo / W_down this way at load time).  Returns (bytes, type); formats
    the kernels do not read natively come back as fp16."""
    t = GGMLType(ggml_type)
    N, K = shape
    per, nbytes = BLOCK_GEOMETRY[t]
    raw = np.ascontiguousarray(raw).reshape(-1)

    def rescale(col0, blk_bytes):
        b = raw.reshape(-1, blk_bytes).copy()
        for c in col0:
            d = b[:, c:c + 2].copy().view("<f2").astype(np.float32) * s
            b[:, c:c + 2] = d.astype("<f2").view(np.uint8)
        return b.reshape(-1)
    if t == GGMLType.Q8_0:
        return rescale([0], 34), t
    if t == GGMLType.Q4_K:
        return rescale([0, 2], 144), t
    if t == GGMLType.Q5_K:
        return rescale([0, 2], 176), t
    if t == GGMLType.Q6_K:
        return rescale([208], 210), t
    from ..gguf.quants import quantize
    dense = dequantize(raw, t, (N, K)) * np.float32(s)
    out_t = t if t in (GGMLType.F16, GGMLType.BF16, GGMLType.F32) else GGMLType.F16
    return quantize(dense, out_t), out_t


@dataclass
class DeviceWeight:
    wtype: GGMLType
    N: int
    K: int
    planes: list[torch.Tensor] = field(default_factory=list)  # base, aux0, aux1, aux2 (device memory owners)
    dense: torch.Tensor | None = None                          # fp32 [N, K] (CPU fallback / oracle only)
    _c: dlp_weight | None = None

    @property
    def device(self) -> torch.device:
        return self.planes[0].device if self.planes else self.dense.device

    @property
    def nbytes(self) -> int:
        return sum(p.numel() * p.element_size() for p in self.planes) + (self.dense.numel() * self.dense.element_size() if self.dense is not None else 0)

    @property
    def c(self) -> dlp_weight:
        if self._c is None:
            ptrs = [p.data_ptr() for p in self.planes] + [0] * (4 - len(self.planes))
            self._c = dlp_weight(ptrs[0], ptrs[1], ptrs[2], ptrs[3], int(self.wtype), self.N, self.K, 0)
        return self._c

    def ref(self):
        return C.byref(self.c)

    def to_dense(self, dtype=torch.float32) -> torch.Tensor:
        """Decode back to a dense [N, K] tensor (tests / CPU fallback)."""
        if self.dense is not None:
            return self.dense.to(dtype)
        from . import api
        return api.dequant_rows(self, None, out_dtype=dtype)

    @classmethod
    def from_raw(cls, raw: np.ndarray, ggml_type: GGMLType, shape: tuple[int, int], device) -> "DeviceWeight":
        """raw: uint8 [N, row_bytes] in GGUF layout for a logical [N, K] matrix."""
        t = GGMLType(ggml_type)
        N, K = shape
        device = torch.device(device)
        raw = np.ascontiguousarray(raw).reshape(N, -1)
        if device.type == "cpu":
            dense = torch.from_numpy(dequantize(raw, t, (N, K)).astype(np.float32, copy=True))
            return cls(t if t in KERNEL_TYPES else GGMLType.F16, N, K, [], dense)
        if t == GGMLType.F32:  # no fp32 matmul path on the device: store as fp16
            w = torch.from_numpy(raw.view("<f4").reshape(N, K).astype(np.float16)).to(device)
            return cls(GGMLType.F16, N, K, [w])
        if t not in KERNEL_TYPES:  # legacy quants (Q4_0, ...) are decoded once at load
            w = torch.from_numpy(dequantize(raw, t, (N, K)).astype(np.float16)).to(device)
            return cls(GGMLType.F16, N, K, [w])
        if K % 256:
            raise ValueError(f"K={K} must be a multiple of 256 for the device kernels")
        buf = _upload(raw, device)
        if t in (GGMLType.F16, GGMLType.BF16):
            return cls(t, N, K, [buf.view(torch.float16 if t == GGMLType.F16 else torch.bfloat16).view(N, K)])
        if t in (GGMLType.Q4_K, GGMLType.Q5_K):
            return cls(t, N, K, [buf])
        if t == GGMLType.Q8_0:
            b = buf.view(N, K // 32, 34)
            d = b[:, :, :2].contiguous().view(torch.float16).view(N, K // 32)
            qs = b[:, :, 2:].contiguous().view(torch.int8).view(N, K)
            return cls(t, N, K, [qs, d])
        if t == GGMLType.Q6_K:
            b = buf.view(N, K // 256, 210)
            ql = b[:, :, 0:128].contiguous().view(N, K // 2)
            qh = b[:, :, 128:192].contiguous().view(N, K // 4)
            sc = b[:, :, 192:208].contiguous().view(torch.int8).view(N, K // 16)
            d = b[:, :, 208:210].contiguous().view(torch.float16).view(N, K // 256)
            return cls(t, N, K, [ql, qh, sc, d])
        raise NotImplementedError(t)

    @classmethod
    def from_dense(cls, w: torch.Tensor, ggml_type: GGMLType = GGMLType.F16, device=None) -> "DeviceWeight":
        """Quantise a dense fp32 [N, K] tensor to `ggml_type` and upload (tests, synthetic models)."""
        from ..gguf.quants import quantize
        N, K = w.shape
        raw = quantize(w.detach().float().cpu().numpy(), ggml_type)
        return cls.from_raw(raw, ggml_type, (N, K), device or w.device)


class _Stager:
    """Pinned staging ring for weight uploads: 4 x 64 MiB page-locked buffers with one CUDA event each.  The (read-only, mmap-backed)
    GGUF bytes are copied into a free buffer on the host while the previous buffers are still in flight to the device, so file
    reads and H2D DMA overlap.  Parity: `llama_model_loader::load_all_data` (llama.cpp/src/llama-model-loader.cpp:970-1172: four
    64 MiB pinned staging buffers + events around `ggml_backend_tensor_set_async`)."""
    N_BUF, BYTES = 4, 64 << 20

    def __init__(self):
        self.bufs = [torch.empty(self.BYTES, dtype=torch.uint8).pin_memory() for _ in range(self.N_BUF)]
        self.views = [b.numpy() for b in self.bufs]
        self.events = [torch.cuda.Event() for _ in range(self.N_BUF)]
        self.used = [False] * self.N_BUF
        self.i = 0

    def upload(self, raw: np.ndarray, device) -> torch.Tensor:
        flat = raw.reshape(-1).view(np.uint8)
        n = flat.shape[0]
        out = torch.empty(n, dtype=torch.uint8, device=device)
        for off in range(0, n, self.BYTES):
            m = min(self.BYTES, n - off)
            k = self.i
            self.i = (self.i + 1) % self.N_BUF
            if self.used[k]:
                self.events[k].synchronize()          # the DMA that last read this buffer has finished
            np.copyto(self.views[k][:m], flat[off:off + m])
            out[off:off + m].copy_(self.bufs[k][:m], non_blocking=True)
            self.events[k].record()
            self.used[k] = True
        return out.view(raw.shape)


_STAGER: _Stager | None = None


def _upload(raw: np.ndarray, device) -> torch.Tensor:
    """Host bytes -> device tensor of the same shape (uint8), through the pinned staging ring on CUDA devices."""
    global _STAGER
    if torch.device(device).type != "cuda":
        return torch.from_numpy(np.array(raw, copy=True)).to(device)
    with torch.cuda.device(device):
        if _STAGER is None:
            try:
                _STAGER = _Stager()
            except RuntimeError:                      # no pinned memory to be had: plain synchronous copy
                return torch.from_numpy(np.array(raw, copy=True)).to(device)
        return _STAGER.upload(raw, device)


def release_staging():
    """Free the pinned staging buffers (call after loading when host memory matters)."""
    global _STAGER
    if _STAGER is not None:
        torch.cuda.synchronize()
        _STAGER = None


def stack_rows(weights: list[DeviceWeight]) -> DeviceWeight:
    """Concatenate matrices with equal K and type along N (e.g. Wq|Wk|Wv -> one fused QKV GEMV)."""
    w0 = weights[0]
    if any(w.wtype != w0.wtype or w.K != w0.K for w in weights):
        raise ValueError("stack_rows: mismatched type or K")
    N = sum(w.N for w in weights)
    if w0.dense is not None:
        return DeviceWeight(w0.wtype, N, w0.K, [], torch.cat([w.dense for w in weights], 0))
    planes = [torch.cat([w.planes[i] for w in weights], 0) for i in range(len(w0.planes))]
    return DeviceWeight(w0.wtype, N, w0.K, planes)
