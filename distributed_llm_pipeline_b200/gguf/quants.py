"""Block-quant codecs (numpy, vectorised) for the GGUF tensor types this engine serves.

These are the *host-side oracles* for every dequant-fused CUDA kernel in `csrc/` and the codecs used
by the synthetic-model writer.  The bit layouts are the on-disk contract described by the reference's
block structs (`llama.cpp/ggml/src/ggml-common.h:170-336`); the decode formulas correspond to
`dequantize_row_q8_0` / `get_scale_min_k4` / `dequantize_row_q4_K` / `dequantize_row_q6_K`
(`llama.cpp/ggml/src/ggml-quants.c:401,703,1352,1762`).  The encoders here are our own (closed-form
min/max fits, not the reference's iterative search): any valid block decodes identically everywhere,
which is all the file format requires.

All functions operate on flat `uint8` byte buffers <-> `float32` arrays whose last dimension is a
multiple of the block size.
"""
from __future__ import annotations

import numpy as np

from .constants import BLOCK_GEOMETRY, GGMLType, QK_K

# --------------------------------------------------------------------------------------------
# helpers


def _blocks(raw: np.ndarray, ggml_type: GGMLType) -> np.ndarray:
    per, nbytes = BLOCK_GEOMETRY[ggml_type]
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    if raw.size % nbytes:
        raise ValueError(f"byte length {raw.size} is not a multiple of {nbytes} for {ggml_type!r}")
    return raw.reshape(-1, nbytes)


def _f16(col: np.ndarray) -> np.ndarray:
    """Interpret a [..., 2] uint8 column pair as little-endian fp16 -> float32 [...]."""
    return np.ascontiguousarray(col).view("<f2").astype(np.float32)[..., 0]


def bf16_to_f32(u16: np.ndarray) -> np.ndarray:
    return (u16.astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16 bit pattern (uint16)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    rounded = u + (0x7FFF + ((u >> 16) & 1))
    out = (rounded >> 16).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        out = np.where(nan, np.uint16(0x7FC0), out)
    return out


# --------------------------------------------------------------------------------------------
# K-quant 6-bit scale/min packing (12 bytes hold 8 scales + 8 mins)


def unpack_scale_min_k4(scales: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """scales: [nb, 12] uint8 -> (sc [nb, 8], mn [nb, 8]) uint8 in 0..63."""
    s = scales.astype(np.uint8)
    sc = np.empty(s.shape[:-1] + (8,), np.uint8)
    mn = np.empty_like(sc)
    sc[..., :4] = s[..., 0:4] & 63
    mn[..., :4] = s[..., 4:8] & 63
    sc[..., 4:] = (s[..., 8:12] & 0x0F) | ((s[..., 0:4] >> 6) << 4)
    mn[..., 4:] = (s[..., 8:12] >> 4) | ((s[..., 4:8] >> 6) << 4)
    return sc, mn


def pack_scale_min_k4(sc: np.ndarray, mn: np.ndarray) -> np.ndarray:
    """Inverse of :func:`unpack_scale_min_k4` (inputs 0..63)."""
    sc = sc.astype(np.uint8)
    mn = mn.astype(np.uint8)
    out = np.zeros(sc.shape[:-1] + (12,), np.uint8)
    out[..., 0:4] = (sc[..., :4] & 63) | ((sc[..., 4:] >> 4) << 6)
    out[..., 4:8] = (mn[..., :4] & 63) | ((mn[..., 4:] >> 4) << 6)
    out[..., 8:12] = (sc[..., 4:] & 0x0F) | ((mn[..., 4:] & 0x0F) << 4)
    return out


# --------------------------------------------------------------------------------------------
# dequantise


def dequant_q8_0(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q8_0)
    d = _f16(b[:, 0:2])
    q = b[:, 2:].view(np.int8).astype(np.float32)
    return (d[:, None] * q).reshape(-1)


def dequant_q4_0(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q4_0)
    d = _f16(b[:, 0:2])
    qs = b[:, 2:]
    lo = (qs & 0x0F).astype(np.int8) - 8
    hi = (qs >> 4).astype(np.int8) - 8
    q = np.concatenate([lo, hi], axis=1).astype(np.float32)
    return (d[:, None] * q).reshape(-1)


def dequant_q4_1(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q4_1)
    d, m = _f16(b[:, 0:2]), _f16(b[:, 2:4])
    qs = b[:, 4:]
    q = np.concatenate([qs & 0x0F, qs >> 4], axis=1).astype(np.float32)
    return (d[:, None] * q + m[:, None]).reshape(-1)


def _q5_high_bits(qh_bytes: np.ndarray) -> np.ndarray:
    """[nb, 4] bytes -> [nb, 32] fifth bits (bit j of the little-endian u32 belongs to element j)."""
    qh = qh_bytes.copy().view("<u4").reshape(-1)
    return ((qh[:, None] >> np.arange(32, dtype=np.uint32)[None, :]) & 1).astype(np.uint8)


def dequant_q5_0(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q5_0)
    d = _f16(b[:, 0:2])
    hb = _q5_high_bits(b[:, 2:6])
    qs = b[:, 6:]
    q = np.concatenate([qs & 0x0F, qs >> 4], axis=1) | (hb << 4)
    return (d[:, None] * (q.astype(np.float32) - 16.0)).reshape(-1)


def dequant_q5_1(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q5_1)
    d, m = _f16(b[:, 0:2]), _f16(b[:, 2:4])
    hb = _q5_high_bits(b[:, 4:8])
    qs = b[:, 8:]
    q = np.concatenate([qs & 0x0F, qs >> 4], axis=1) | (hb << 4)
    return (d[:, None] * q.astype(np.float32) + m[:, None]).reshape(-1)


def dequant_q2_k(raw: np.ndarray) -> np.ndarray:
    """block_q2_K: scales[16] (4-bit scale | 4-bit min << 4), qs[64] (2 bits), d, dmin  (ggml-common.h, dequantize_row_q2_K)."""
    b = _blocks(raw, GGMLType.Q2_K)
    nb = b.shape[0]
    sc = b[:, 0:16]
    qs = b[:, 16:80].reshape(nb, 2, 32)                                  # two 128-element halves, 32 bytes each
    d, dmin = _f16(b[:, 80:82]), _f16(b[:, 82:84])
    shifts = np.array([0, 2, 4, 6], np.uint8)
    q = ((qs[:, :, None, :] >> shifts[None, None, :, None]) & 3).astype(np.float32)   # [nb, half, shift, 32]
    q = q.reshape(nb, 16, 16)                                            # 16 sub-blocks of 16 in element order
    dl = d[:, None] * (sc & 0x0F).astype(np.float32)
    ml = dmin[:, None] * (sc >> 4).astype(np.float32)
    return (dl[:, :, None] * q - ml[:, :, None]).reshape(-1)


def dequant_q3_k(raw: np.ndarray) -> np.ndarray:
    """block_q3_K: hmask[32], qs[64] (low 2 bits), scales[12] (16 x 6-bit), d  (dequantize_row_q3_K)."""
    b = _blocks(raw, GGMLType.Q3_K)
    nb = b.shape[0]
    hmask = b[:, 0:32]
    qs = b[:, 32:96].reshape(nb, 2, 32)
    s12 = b[:, 96:108]
    d = _f16(b[:, 108:110])
    lo4 = np.concatenate([s12[:, 0:8] & 0x0F, s12[:, 0:8] >> 4], axis=1)                      # scales 0..15, low 4 bits
    hi2 = np.concatenate([(s12[:, 8:12] >> s) & 3 for s in (0, 2, 4, 6)], axis=1)             # scales 0..15, high 2 bits
    scales = (lo4 | (hi2 << 4)).astype(np.int8).astype(np.float32) - 32.0
    shifts = np.array([0, 2, 4, 6], np.uint8)
    low = ((qs[:, :, None, :] >> shifts[None, None, :, None]) & 3).astype(np.int8)            # [nb, half, shift, 32]
    bit = np.arange(8, dtype=np.uint8).reshape(1, 2, 4, 1)                                    # mask bit = half*4 + shift
    high = ((hmask[:, None, None, :] >> bit) & 1).astype(np.int8)
    q = (low - 4 * (1 - high)).astype(np.float32).reshape(nb, 16, 16)
    return (d[:, None, None] * scales[:, :, None] * q).reshape(-1)


def dequant_q4_k(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q4_K)
    d = _f16(b[:, 0:2])
    dmin = _f16(b[:, 2:4])
    sc, mn = unpack_scale_min_k4(b[:, 4:16])
    qs = b[:, 16:144].reshape(-1, 4, 32)  # 4 chunks of 64 values, 32 bytes each
    lo = (qs & 0x0F).astype(np.float32)  # sub-block 2j
    hi = (qs >> 4).astype(np.float32)  # sub-block 2j+1
    q = np.stack([lo, hi], axis=2).reshape(-1, 8, 32)
    scale = d[:, None] * sc.astype(np.float32)
    minv = dmin[:, None] * mn.astype(np.float32)
    return (scale[:, :, None] * q - minv[:, :, None]).reshape(-1)


def dequant_q5_k(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q5_K)
    d = _f16(b[:, 0:2])
    dmin = _f16(b[:, 2:4])
    sc, mn = unpack_scale_min_k4(b[:, 4:16])
    qh = b[:, 16:48]  # [nb, 32]
    qs = b[:, 48:176].reshape(-1, 4, 32)
    lo = (qs & 0x0F).astype(np.uint8)
    hi = (qs >> 4).astype(np.uint8)
    q4 = np.stack([lo, hi], axis=2).reshape(-1, 8, 32)
    bits = ((qh[:, None, :] >> np.arange(8, dtype=np.uint8)[None, :, None]) & 1).astype(np.uint8)
    q = (q4 + (bits << 4)).astype(np.float32)
    scale = d[:, None] * sc.astype(np.float32)
    minv = dmin[:, None] * mn.astype(np.float32)
    return (scale[:, :, None] * q - minv[:, :, None]).reshape(-1)


def dequant_q6_k(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.Q6_K)
    ql = b[:, 0:128].reshape(-1, 2, 64)
    qh = b[:, 128:192].reshape(-1, 2, 32)
    sc = b[:, 192:208].view(np.int8).astype(np.float32).reshape(-1, 2, 8)
    d = _f16(b[:, 208:210])
    l0 = ql[:, :, :32]
    l1 = ql[:, :, 32:]
    q1 = (l0 & 0x0F) | (((qh >> 0) & 3) << 4)
    q2 = (l1 & 0x0F) | (((qh >> 2) & 3) << 4)
    q3 = (l0 >> 4) | (((qh >> 4) & 3) << 4)
    q4 = (l1 >> 4) | (((qh >> 6) & 3) << 4)
    q = np.stack([q1, q2, q3, q4], axis=2).astype(np.int16) - 32  # [nb, 2, 4, 32]
    # scale index within a 128-half: is = l//16 + 2*k for quarter k
    idx = (np.arange(32) // 16)[None, :] + 2 * np.arange(4)[:, None]  # [4, 32]
    s = sc[:, :, idx]  # [nb, 2, 4, 32]
    return (d[:, None, None, None] * s * q.astype(np.float32)).reshape(-1)


# ---- non-linear 4-bit, ternary and microscaling types (block layouts: ggml-common.h:236-262,416-440; value tables are part of
# the file format: `kvalues_iq4nl`, `kvalues_mxfp4`).  Decoded to fp16 at load by the engine; no kernel reads them directly.

IQ4_VALUES = np.array([-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113], np.float32)
MXFP4_VALUES = np.array([0, 1, 2, 3, 4, 6, 8, 12, 0, -1, -2, -3, -4, -6, -8, -12], np.float32)     # E2M1 values, doubled


def _nibble_halves(qs: np.ndarray) -> np.ndarray:
    """[nb, n/2] bytes -> [nb, n] nibble codes: element j is the low nibble of byte j, element j + n/2 the high nibble."""
    return np.concatenate([qs & 0x0F, qs >> 4], axis=-1)


def dequant_iq4_nl(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ4_NL)
    d = _f16(b[:, 0:2])[:, None]
    return (d * IQ4_VALUES[_nibble_halves(b[:, 2:18])]).reshape(-1)


def dequant_iq4_xs(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ4_XS)
    d = _f16(b[:, 0:2])[:, None]
    sh = b[:, 2:4].copy().view("<u2").astype(np.uint32)                      # [nb, 1]: two high bits per sub-block
    sl = b[:, 4:8]
    lo = np.stack([sl & 0x0F, sl >> 4], axis=-1).reshape(-1, 8).astype(np.uint32)
    hi = (sh >> (2 * np.arange(8, dtype=np.uint32))[None, :]) & 3
    ls = (lo | (hi << 4)).astype(np.float32) - 32.0                          # 6-bit sub-block scale, offset 32
    codes = _nibble_halves(b[:, 8:136].reshape(-1, 8, 16))                   # [nb, 8, 32]
    return ((d * ls)[:, :, None] * IQ4_VALUES[codes]).reshape(-1)


def dequant_tq2_0(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.TQ2_0)
    d = _f16(b[:, 64:66])[:, None]
    qs = b[:, 0:64].reshape(-1, 2, 1, 32)                                    # two groups of 32 bytes, four 2-bit planes each
    q = (qs >> (2 * np.arange(4, dtype=np.uint8)).reshape(1, 1, 4, 1)) & 3
    return (d * (q.reshape(-1, 256).astype(np.float32) - 1.0)).reshape(-1)


_POW3 = np.array([1, 3, 9, 27, 81, 243], np.uint8)


def dequant_tq1_0(raw: np.ndarray) -> np.ndarray:
    """Five trits per byte in base 3 (the byte is the fixed-point fraction trits / 3^5): multiplying by 3^n modulo 256 moves
    trit n to the top, `* 3 >> 8` reads it."""
    b = _blocks(raw, GGMLType.TQ1_0)
    d = _f16(b[:, 52:54])[:, None]

    def trits(bytes_: np.ndarray, n: int) -> np.ndarray:                    # [nb, m] -> [nb, n, m]
        q = (bytes_[:, None, :] * _POW3[:n].reshape(1, n, 1)).astype(np.uint8)
        return ((q.astype(np.uint16) * 3) >> 8).astype(np.float32)

    parts = [trits(b[:, 0:32], 5).reshape(-1, 160), trits(b[:, 32:48], 5).reshape(-1, 80), trits(b[:, 48:52], 4).reshape(-1, 16)]
    return (d * (np.concatenate(parts, axis=1) - 1.0)).reshape(-1)


def e8m0_to_f32_half(e: np.ndarray) -> np.ndarray:
    """Shared MX exponent -> 2^(e - 128) (half the E8M0 value, because the value table holds doubled E2M1 numbers)."""
    e = e.astype(np.uint32)
    bits = np.where(e < 2, np.uint32(0x00200000) << e, (e - 1) << 23).astype(np.uint32)
    return bits.view(np.float32)


def dequant_mxfp4(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.MXFP4)
    d = e8m0_to_f32_half(b[:, 0:1])
    return (d * MXFP4_VALUES[_nibble_halves(b[:, 1:17])]).reshape(-1)


# --------------------------------------------------------------------------------------------
# code-book types (IQ1 / IQ2 / IQ3): groups of 8 (IQ1, IQ2) or 4 (IQ3) weights are one entry of a fixed lattice code book, plus signs and
# a sub-block scale.  Layouts: `block_iq2_xxs` ... `block_iq1_m`, ggml-common.h:338-430; decode formulas: `dequantize_row_iq*`,
# ggml-quants.c:2420-2700.  Decode only: files of these types are produced by `llama-quantize` with an importance matrix.

_IQ_GRIDS: dict[str, np.ndarray] = {}


def iq_grid(name: str) -> np.ndarray:
    """Code book `name` as float32 [n_codes, group] (loaded once from `iq_grids.npz`, written by tools/make_iq_grids.py)."""
    if not _IQ_GRIDS:
        import os
        with np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "iq_grids.npz")) as z:
            for k in z.files:
                _IQ_GRIDS[k] = z[k].astype(np.float32)
    return _IQ_GRIDS[name]


def _sign_rows() -> np.ndarray:
    """[128, 8] of +-1: seven stored sign bits, the eighth chosen so that the number of minus signs is even."""
    i = np.arange(128, dtype=np.uint16)
    parity = np.zeros(128, np.uint16)
    for b in range(7):
        parity ^= (i >> b) & 1
    full = i | (parity << 7)
    return 1.0 - 2.0 * ((full[:, None] >> np.arange(8, dtype=np.uint16)[None, :]) & 1).astype(np.float32)


_SIGNS7 = _sign_rows()
_BIT8 = np.arange(8, dtype=np.uint8)


def _signs8(bytes_: np.ndarray) -> np.ndarray:
    """Explicit sign bytes [...] -> [..., 8] of +-1 (bit j set = element j negative)."""
    return 1.0 - 2.0 * ((bytes_[..., None] >> _BIT8) & 1).astype(np.float32)


def dequant_iq2_xxs(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ2_XXS)
    d = _f16(b[:, 0:2])
    w = b[:, 2:66].copy().view("<u4").reshape(-1, 8, 2)                       # per 32 weights: four code bytes | 4 x 7 sign bits + 4-bit scale
    codes = w[:, :, 0:1].copy().view(np.uint8).reshape(-1, 8, 4)
    aux = w[:, :, 1]
    db = d[:, None] * (0.5 + (aux >> 28).astype(np.float32)) * 0.25
    sgn = _SIGNS7[(aux[:, :, None] >> (7 * np.arange(4, dtype=np.uint32))) & 127]      # [nb, 8, 4, 8]
    return (db[:, :, None, None] * iq_grid("iq2_xxs")[codes] * sgn).reshape(-1)


def dequant_iq2_xs(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ2_XS)
    d = _f16(b[:, 0:2])
    q = b[:, 2:66].copy().view("<u2").reshape(-1, 8, 4)                       # 9-bit code | 7-bit sign row
    sc = b[:, 66:74]
    nib = np.stack([sc & 0x0F, sc >> 4], axis=-1).astype(np.float32)          # [nb, 8, 2]: one scale per 16 weights
    db = d[:, None, None] * (0.5 + nib) * 0.25
    vals = iq_grid("iq2_xs")[q & 511] * _SIGNS7[q >> 9]                      # [nb, 8, 4, 8]
    return (np.repeat(db, 2, axis=2)[..., None] * vals).reshape(-1)


def dequant_iq2_s(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ2_S)
    d = _f16(b[:, 0:2])
    lo = b[:, 2:34].reshape(-1, 8, 4).astype(np.uint16)
    sg = b[:, 34:66].reshape(-1, 8, 4)
    qh = b[:, 66:74].astype(np.uint16)
    sc = b[:, 74:82]
    hi = (qh[:, :, None] >> (2 * np.arange(4, dtype=np.uint16))) & 3          # two more code bits per group of 8
    nib = np.stack([sc & 0x0F, sc >> 4], axis=-1).astype(np.float32)
    db = d[:, None, None] * (0.5 + nib) * 0.25
    vals = iq_grid("iq2_s")[lo | (hi << 8)] * _signs8(sg)
    return (np.repeat(db, 2, axis=2)[..., None] * vals).reshape(-1)


def dequant_iq3_xxs(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ3_XXS)
    d = _f16(b[:, 0:2])
    codes = b[:, 2:66].reshape(-1, 8, 4, 2)                                   # two 4-weight codes per group of 8
    aux = b[:, 66:98].copy().view("<u4").reshape(-1, 8)
    db = d[:, None] * (0.5 + (aux >> 28).astype(np.float32)) * 0.5
    sgn = _SIGNS7[(aux[:, :, None] >> (7 * np.arange(4, dtype=np.uint32))) & 127]      # [nb, 8, 4, 8]
    vals = iq_grid("iq3_xxs")[codes].reshape(-1, 8, 4, 8) * sgn
    return (db[:, :, None, None] * vals).reshape(-1)


def dequant_iq3_s(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ3_S)
    d = _f16(b[:, 0:2])
    lo = b[:, 2:66].reshape(-1, 8, 8).astype(np.uint16)                       # eight 4-weight codes per 32 weights
    qh = b[:, 66:74].astype(np.uint16)
    sg = b[:, 74:106].reshape(-1, 8, 4)
    sc = b[:, 106:110]
    hi = (qh[:, :, None] >> np.arange(8, dtype=np.uint16)) & 1                # ninth code bit
    nib = np.stack([sc & 0x0F, sc >> 4], axis=-1).reshape(-1, 8).astype(np.float32)
    db = d[:, None] * (1.0 + 2.0 * nib)
    vals = iq_grid("iq3_s")[lo | (hi << 8)].reshape(-1, 8, 4, 8) * _signs8(sg)
    return (db[:, :, None, None] * vals).reshape(-1)


IQ1_DELTA = 0.125


def dequant_iq1_s(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ1_S)
    d = _f16(b[:, 0:2])
    lo = b[:, 2:34].reshape(-1, 8, 4).astype(np.uint16)
    qh = b[:, 34:50].copy().view("<u2").reshape(-1, 8)                        # 4 x 3 code bits | 3-bit scale | shift sign
    hi = (qh[:, :, None] >> (3 * np.arange(4, dtype=np.uint16))) & 7
    dl = d[:, None] * (2.0 * ((qh >> 12) & 7).astype(np.float32) + 1.0)
    delta = np.where(qh & 0x8000, -IQ1_DELTA, IQ1_DELTA).astype(np.float32)
    vals = iq_grid("iq1_s")[lo | (hi << 8)] + delta[:, :, None, None]
    return (dl[:, :, None, None] * vals).reshape(-1)


def dequant_iq1_m(raw: np.ndarray) -> np.ndarray:
    b = _blocks(raw, GGMLType.IQ1_M)
    lo = b[:, 0:32].reshape(-1, 8, 4).astype(np.uint16)
    qh = b[:, 32:48].reshape(-1, 8, 2)
    sc = b[:, 48:56].copy().view("<u2").reshape(-1, 4)                        # 4 x (4 x 3-bit scales | one nibble of the fp16 super-scale)
    dbits = ((sc[:, 0] >> 12) | ((sc[:, 1] >> 8) & 0x00F0) | ((sc[:, 2] >> 4) & 0x0F00) | (sc[:, 3] & 0xF000)).astype(np.uint16)
    d = dbits.view("<f2").astype(np.float32)
    s3 = (sc[:, :, None] >> (3 * np.arange(4, dtype=np.uint16))) & 7          # [nb, 4, 4] -> one scale per 16 weights
    dl = d[:, None] * (2.0 * s3.reshape(-1, 16).astype(np.float32) + 1.0)     # [nb, 16]
    nibs = np.stack([qh & 0x0F, qh >> 4], axis=-1).reshape(-1, 8, 4).astype(np.uint16)   # per group of 8: 3 code bits | shift sign
    delta = np.where(nibs & 8, -IQ1_DELTA, IQ1_DELTA).astype(np.float32)
    vals = iq_grid("iq1_s")[lo | ((nibs & 7) << 8)] + delta[..., None]        # [nb, 8, 4, 8]
    return (np.repeat(dl, 2, axis=1).reshape(-1, 8, 4)[..., None] * vals).reshape(-1)



_IQ_DECODERS = {GGMLType.IQ2_XXS: dequant_iq2_xxs, GGMLType.IQ2_XS: dequant_iq2_xs, GGMLType.IQ2_S: dequant_iq2_s, GGMLType.IQ3_XXS: dequant_iq3_xxs,
                GGMLType.IQ3_S: dequant_iq3_s, GGMLType.IQ1_S: dequant_iq1_s, GGMLType.IQ1_M: dequant_iq1_m}


def dequantize(raw: np.ndarray, ggml_type: GGMLType, shape: tuple[int, ...] | None = None) -> np.ndarray:
    """Decode a raw byte buffer of `ggml_type` to float32 (optionally reshaped, row-major)."""
    t = GGMLType(ggml_type)
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    if t == GGMLType.F32:
        out = raw.view("<f4").astype(np.float32)
    elif t == GGMLType.F16:
        out = raw.view("<f2").astype(np.float32)
    elif t == GGMLType.BF16:
        out = bf16_to_f32(raw.view("<u2"))
    elif t == GGMLType.Q8_0:
        out = dequant_q8_0(raw)
    elif t == GGMLType.Q4_0:
        out = dequant_q4_0(raw)
    elif t == GGMLType.Q4_1:
        out = dequant_q4_1(raw)
    elif t == GGMLType.Q5_0:
        out = dequant_q5_0(raw)
    elif t == GGMLType.Q5_1:
        out = dequant_q5_1(raw)
    elif t == GGMLType.Q2_K:
        out = dequant_q2_k(raw)
    elif t == GGMLType.Q3_K:
        out = dequant_q3_k(raw)
    elif t == GGMLType.Q4_K:
        out = dequant_q4_k(raw)
    elif t == GGMLType.Q5_K:
        out = dequant_q5_k(raw)
    elif t == GGMLType.Q6_K:
        out = dequant_q6_k(raw)
    elif t == GGMLType.IQ4_NL:
        out = dequant_iq4_nl(raw)
    elif t == GGMLType.IQ4_XS:
        out = dequant_iq4_xs(raw)
    elif t == GGMLType.TQ1_0:
        out = dequant_tq1_0(raw)
    elif t == GGMLType.TQ2_0:
        out = dequant_tq2_0(raw)
    elif t == GGMLType.MXFP4:
        out = dequant_mxfp4(raw)
    elif t in _IQ_DECODERS:
        out = _IQ_DECODERS[t](raw)
    else:
        raise NotImplementedError(f"dequantize: unsupported type {t!r}")
    return out.reshape(shape) if shape is not None else out


# --------------------------------------------------------------------------------------------
# quantise


def _nearest_int(x: np.ndarray) -> np.ndarray:
    return np.rint(x).astype(np.int32)


def quant_q8_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    amax = np.abs(x).max(axis=1)
    df = (amax / np.float32(127.0)).astype(np.float32)            # the inverse uses the unrounded scale (quantize_row_q8_0_ref)
    d = df.astype(np.float16)
    inv = np.where(df > 0, np.float32(1.0) / np.where(df > 0, df, np.float32(1.0)), np.float32(0.0)).astype(np.float32)
    q = np.clip(_nearest_int(x * inv[:, None]), -127, 127).astype(np.int8)
    out = np.empty((x.shape[0], 34), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q.view(np.uint8)
    return out.reshape(-1)


def quant_q4_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    idx = np.abs(x).argmax(axis=1)
    mx = x[np.arange(x.shape[0]), idx]
    df = (mx / np.float32(-8.0)).astype(np.float32)
    d = df.astype(np.float16)
    inv = np.where(df != 0, np.float32(1.0) / np.where(df != 0, df, np.float32(1.0)), np.float32(0.0)).astype(np.float32)
    q = np.clip(np.trunc(x * inv[:, None] + 8.5), 0, 15).astype(np.uint8)      # (int8)(x/d + 8.5), as quantize_row_q4_0_ref
    out = np.empty((x.shape[0], 18), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


def quant_q4_1(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    mn, mx = x.min(axis=1), x.max(axis=1)
    d = (mx - mn) / 15.0
    inv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 0.0).astype(np.float32)
    q = np.clip(np.trunc((x - mn[:, None]) * inv[:, None] + 0.5), 0, 15).astype(np.uint8)
    out = np.empty((x.shape[0], 20), np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = mn.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 4:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


def _pack_q5(q: np.ndarray):
    """5-bit values [nb, 32] -> (qh [nb, 4] bytes, qs [nb, 16] nibbles)."""
    hb = (q >> 4).astype(np.uint32)
    qh = (hb << np.arange(32, dtype=np.uint32)[None, :]).sum(axis=1, dtype=np.uint64).astype("<u4")
    lo = (q & 0x0F).astype(np.uint8)
    return qh.view(np.uint8).reshape(-1, 4), lo[:, :16] | (lo[:, 16:] << 4)


def quant_q5_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    idx = np.abs(x).argmax(axis=1)
    mx = x[np.arange(x.shape[0]), idx]
    d = mx / -16.0
    inv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 0.0).astype(np.float32)
    q = np.clip(np.trunc(x * inv[:, None] + 16.5), 0, 31).astype(np.uint8)
    out = np.empty((x.shape[0], 22), np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:6], out[:, 6:] = _pack_q5(q)
    return out.reshape(-1)


def quant_q5_1(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 32)
    mn, mx = x.min(axis=1), x.max(axis=1)
    d = (mx - mn) / 31.0
    inv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 0.0).astype(np.float32)
    q = np.clip(np.trunc((x - mn[:, None]) * inv[:, None] + 0.5), 0, 31).astype(np.uint8)
    out = np.empty((x.shape[0], 24), np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = mn.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 4:8], out[:, 8:] = _pack_q5(q)
    return out.reshape(-1)


def quant_q2_k(x: np.ndarray) -> np.ndarray:
    """Closed-form Q2_K encoder (per 16-element sub-block affine fit, 4-bit scale / min; no search)."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 16, 16)
    nb = x.shape[0]
    mn = np.minimum(x.min(axis=2), 0.0)
    scale = (x.max(axis=2) - mn) / 3.0
    d = scale.max(axis=1) / 15.0
    dmin = (-mn).max(axis=1) / 15.0
    d16, dmin16 = d.astype(np.float16).astype(np.float32), dmin.astype(np.float16).astype(np.float32)
    sc = np.clip(_nearest_int(scale / np.where(d16 > 0, d16, 1.0)[:, None]), 0, 15) * (d16 > 0)[:, None]
    m = np.clip(_nearest_int(-mn / np.where(dmin16 > 0, dmin16, 1.0)[:, None]), 0, 15) * (dmin16 > 0)[:, None]
    dl, ml = d16[:, None] * sc, dmin16[:, None] * m
    q = np.clip(_nearest_int((x + ml[:, :, None]) / np.where(dl > 0, dl, 1.0)[:, :, None]), 0, 3).astype(np.uint8) * (dl > 0)[:, :, None].astype(np.uint8)
    out = np.empty((nb, 84), np.uint8)
    out[:, 0:16] = (sc | (m << 4)).astype(np.uint8)
    q4 = q.reshape(nb, 2, 4, 32)                                          # [half, shift, 32]
    out[:, 16:80] = (q4[:, :, 0] | (q4[:, :, 1] << 2) | (q4[:, :, 2] << 4) | (q4[:, :, 3] << 6)).reshape(nb, 64)
    out[:, 80:82] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 82:84] = dmin.astype(np.float16).view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def quant_q3_k(x: np.ndarray) -> np.ndarray:
    """Closed-form Q3_K encoder (symmetric 3-bit per 16-element sub-block, 6-bit scales)."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 16, 16)
    nb = x.shape[0]
    idx = np.abs(x).argmax(axis=2)
    mx = np.take_along_axis(x, idx[:, :, None], axis=2)[:, :, 0]
    scale = mx / -4.0                                                     # value = scale * q, q in [-4, 3]
    amax_s = np.abs(scale).max(axis=1)
    d = amax_s / 31.0
    d16 = d.astype(np.float16).astype(np.float32)
    ls = np.clip(_nearest_int(scale / np.where(d16 > 0, d16, 1.0)[:, None]), -32, 31) * (d16 > 0)[:, None]
    eff = d16[:, None] * ls
    q = np.clip(_nearest_int(x / np.where(eff != 0, eff, 1.0)[:, :, None]), -4, 3) * (eff != 0)[:, :, None]
    u = (q + 4).astype(np.uint8)                                          # 0..7
    low, high = (u & 3).reshape(nb, 2, 4, 32), (u >> 2).reshape(nb, 2, 4, 32)
    out = np.zeros((nb, 110), np.uint8)
    bit = np.arange(8, dtype=np.uint8).reshape(1, 2, 4, 1)
    out[:, 0:32] = (high << bit).sum(axis=(1, 2)).astype(np.uint8)
    out[:, 32:96] = (low[:, :, 0] | (low[:, :, 1] << 2) | (low[:, :, 2] << 4) | (low[:, :, 3] << 6)).reshape(nb, 64)
    l6 = (ls + 32).astype(np.uint8)
    out[:, 96:104] = (l6[:, 0:8] & 0x0F) | ((l6[:, 8:16] & 0x0F) << 4)
    h2 = l6 >> 4
    out[:, 104:108] = h2[:, 0:4] | (h2[:, 4:8] << 2) | (h2[:, 8:12] << 4) | (h2[:, 12:16] << 6)
    out[:, 108:110] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def _kquant_affine_fit(x: np.ndarray, nmax: int):
    """Per-sub-block affine fit  x ~= scale*q - minv  with q in [0, nmax], minv >= 0.

    x: [nb, 8, 32].  Returns float scale/minv [nb, 8]."""
    mn = np.minimum(x.min(axis=2), 0.0)
    mx = x.max(axis=2)
    scale = (mx - mn) / float(nmax)
    return scale, -mn


def importance_weights(x: np.ndarray, imp: np.ndarray, block: int) -> np.ndarray:
    """Per-element error weights from an importance matrix row (mean squared activation per input column):
    w_j = imp_j * sqrt(sigma^2 + x_j^2) with sigma^2 = 2 * mean(x^2) over the block - the form the reference's weighted
    encoders use (`quantize_row_q4_K_impl`, ggml-quants.c): columns that see large activations count more, and within a
    block large weights count more than small ones."""
    xb = x.reshape(-1, block)
    sigma2 = 2.0 * (xb * xb).mean(axis=1, keepdims=True)
    reps = xb.shape[0] // max(1, imp.size // block) if imp.size >= block else 1
    cols = x.shape[-1]
    imp_rows = np.broadcast_to(imp.reshape(1, cols), (x.size // cols, cols)).reshape(-1, block)
    return (imp_rows * np.sqrt(sigma2 + xb * xb)).astype(np.float32)


def _kquant_affine_fit_weighted(x: np.ndarray, w: np.ndarray, nmax: int, n_try: int = 9):
    """Weighted version of `_kquant_affine_fit`: for a few shrunken ranges, round, then solve the weighted least-squares problem
    min sum w (x - s q - b)^2 for (s, b <= 0) in closed form and keep the candidate with the smallest weighted error.
    x, w: [nb, n_sub, sub].  Returns scale, minv (= -b) [nb, n_sub]."""
    mn = np.minimum(x.min(axis=2), 0.0)
    mx = x.max(axis=2)
    rng = np.maximum(mx - mn, 1e-30)
    best_s, best_m = rng / float(nmax), -mn
    best_err = None
    for k in range(n_try):
        s0 = rng / (float(nmax) + 0.15 * k - 0.3)          # from slightly wider than the range to ~25 % tighter
        q = np.clip(_nearest_int((x - mn[:, :, None]) / s0[:, :, None]), 0, nmax).astype(np.float32)
        Sw, Sq, Sqq = w.sum(2), (w * q).sum(2), (w * q * q).sum(2)
        Sx, Sxq = (w * x).sum(2), (w * x * q).sum(2)
        det = Sw * Sqq - Sq * Sq
        ok = det > 1e-30
        s = np.where(ok, (Sw * Sxq - Sx * Sq) / np.where(ok, det, 1.0), s0)
        b = np.where(ok, (Sqq * Sx - Sq * Sxq) / np.where(ok, det, 1.0), mn)
        pos = b > 0                                          # the format stores a non-negative minimum: refit through the origin
        s = np.where(pos, Sxq / np.maximum(Sqq, 1e-30), s)
        b = np.where(pos, 0.0, b)
        bad = ~(s > 0)
        s, b = np.where(bad, s0, s), np.where(bad, mn, b)
        q2 = np.clip(_nearest_int((x - b[:, :, None]) / s[:, :, None]), 0, nmax)
        err = (w * (x - (s[:, :, None] * q2 + b[:, :, None])) ** 2).sum(2)
        if best_err is None:
            best_err, best_s, best_m = err, s, -b
        else:
            better = err < best_err
            best_err = np.where(better, err, best_err)
            best_s, best_m = np.where(better, s, best_s), np.where(better, -b, best_m)
    return best_s.astype(np.float32), best_m.astype(np.float32)


def _kquant_pack_super(scale: np.ndarray, minv: np.ndarray):
    """6-bit quantise sub-block scales/mins against fp16 super-scales d / dmin."""
    smax = scale.max(axis=1)
    mmax = minv.max(axis=1)
    d = (smax / 63.0).astype(np.float16)
    dmin = (mmax / 63.0).astype(np.float16)
    df = d.astype(np.float32)
    dmf = dmin.astype(np.float32)
    inv_d = np.where(df > 0, 1.0 / np.where(df > 0, df, 1.0), 0.0)
    inv_m = np.where(dmf > 0, 1.0 / np.where(dmf > 0, dmf, 1.0), 0.0)
    sc = np.clip(_nearest_int(scale * inv_d[:, None]), 0, 63).astype(np.uint8)
    mq = np.clip(_nearest_int(minv * inv_m[:, None]), 0, 63).astype(np.uint8)
    return d, dmin, sc, mq


def quant_q4_k(x: np.ndarray, weights: np.ndarray | None = None) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 8, 32)
    scale, minv = _kquant_affine_fit(x, 15) if weights is None else _kquant_affine_fit_weighted(x, weights.reshape(x.shape), 15)
    d, dmin, sc, mq = _kquant_pack_super(scale, minv)
    eff_s = d.astype(np.float32)[:, None] * sc
    eff_m = dmin.astype(np.float32)[:, None] * mq
    inv = np.where(eff_s > 0, 1.0 / np.where(eff_s > 0, eff_s, 1.0), 0.0)
    q = np.clip(_nearest_int((x + eff_m[:, :, None]) * inv[:, :, None]), 0, 15).astype(np.uint8)
    out = np.empty((x.shape[0], 144), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = dmin.view(np.uint8).reshape(-1, 2)
    out[:, 4:16] = pack_scale_min_k4(sc, mq)
    q = q.reshape(-1, 4, 2, 32)
    out[:, 16:144] = (q[:, :, 0, :] | (q[:, :, 1, :] << 4)).reshape(-1, 128)
    return out.reshape(-1)


def quant_q5_k(x: np.ndarray, weights: np.ndarray | None = None) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 8, 32)
    scale, minv = _kquant_affine_fit(x, 31) if weights is None else _kquant_affine_fit_weighted(x, weights.reshape(x.shape), 31)
    d, dmin, sc, mq = _kquant_pack_super(scale, minv)
    eff_s = d.astype(np.float32)[:, None] * sc
    eff_m = dmin.astype(np.float32)[:, None] * mq
    inv = np.where(eff_s > 0, 1.0 / np.where(eff_s > 0, eff_s, 1.0), 0.0)
    q = np.clip(_nearest_int((x + eff_m[:, :, None]) * inv[:, :, None]), 0, 31).astype(np.uint8)
    out = np.empty((x.shape[0], 176), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = dmin.view(np.uint8).reshape(-1, 2)
    out[:, 4:16] = pack_scale_min_k4(sc, mq)
    hb = (q >> 4) & 1  # [nb, 8, 32]
    qh = np.zeros((x.shape[0], 32), np.uint8)
    for j in range(8):
        qh |= (hb[:, j, :] << j).astype(np.uint8)
    out[:, 16:48] = qh
    q4 = (q & 0x0F).reshape(-1, 4, 2, 32)
    out[:, 48:176] = (q4[:, :, 0, :] | (q4[:, :, 1, :] << 4)).reshape(-1, 128)
    return out.reshape(-1)


def _symmetric_scale_weighted(x: np.ndarray, w: np.ndarray, qmax: int, n_try: int = 9) -> np.ndarray:
    """Weighted search for a symmetric scale: candidates around amax / qmax, each refined by s = sum(w x q) / sum(w q q)."""
    amax = np.maximum(np.abs(x).max(axis=2), 1e-30)
    best, best_err = amax / qmax, None
    for k in range(n_try):
        s0 = amax / (qmax + 0.2 * k - 0.4)
        q = np.clip(_nearest_int(x / s0[:, :, None]), -qmax - 1, qmax).astype(np.float32)
        den = (w * q * q).sum(2)
        s = np.where(den > 0, (w * x * q).sum(2) / np.where(den > 0, den, 1.0), s0)
        s = np.where(s > 0, s, s0)
        q2 = np.clip(_nearest_int(x / s[:, :, None]), -qmax - 1, qmax)
        err = (w * (x - s[:, :, None] * q2) ** 2).sum(2)
        if best_err is None:
            best, best_err = s, err
        else:
            better = err < best_err
            best, best_err = np.where(better, s, best), np.where(better, err, best_err)
    return best.astype(np.float32)


def quant_q6_k(x: np.ndarray, weights: np.ndarray | None = None) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, 16, 16)  # 16 sub-blocks of 16
    amax = np.abs(x).max(axis=2)  # [nb, 16]
    sub = amax / 31.0 if weights is None else _symmetric_scale_weighted(x, weights.reshape(x.shape), 31)  # symmetric: q in [-31, 31]
    smax = sub.max(axis=1)
    d = (smax / 127.0).astype(np.float16)
    df = d.astype(np.float32)
    inv_d = np.where(df > 0, 1.0 / np.where(df > 0, df, 1.0), 0.0)
    sc = np.clip(_nearest_int(sub * inv_d[:, None]), 0, 127).astype(np.int8)
    eff = df[:, None] * sc.astype(np.float32)
    inv = np.where(eff > 0, 1.0 / np.where(eff > 0, eff, 1.0), 0.0)
    q = (np.clip(_nearest_int(x * inv[:, :, None]), -32, 31) + 32).astype(np.uint8)  # 0..63
    q = q.reshape(-1, 2, 4, 32)  # [nb, half, quarter k, l]; value index = half*128 + k*32 + l
    lo = q & 0x0F
    hi = q >> 4
    ql = np.empty((x.shape[0], 2, 64), np.uint8)
    ql[:, :, :32] = lo[:, :, 0, :] | (lo[:, :, 2, :] << 4)
    ql[:, :, 32:] = lo[:, :, 1, :] | (lo[:, :, 3, :] << 4)
    qh = hi[:, :, 0, :] | (hi[:, :, 1, :] << 2) | (hi[:, :, 2, :] << 4) | (hi[:, :, 3, :] << 6)
    out = np.empty((x.shape[0], 210), np.uint8)
    out[:, 0:128] = ql.reshape(-1, 128)
    out[:, 128:192] = qh.reshape(-1, 64)
    out[:, 192:208] = sc.view(np.uint8)
    out[:, 208:210] = d.view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def _nearest_code(v: np.ndarray, table: np.ndarray) -> np.ndarray:
    """Index of the table entry closest to each value (first one on ties)."""
    return np.abs(v[..., None] - table).argmin(-1).astype(np.uint8)


def _signed_max(x: np.ndarray) -> np.ndarray:
    """The element of largest magnitude along the last axis, with its sign."""
    i = np.abs(x).argmax(-1)
    return np.take_along_axis(x, i[..., None], -1)[..., 0]


def _pack_nibble_halves(codes: np.ndarray) -> np.ndarray:
    h = codes.shape[-1] // 2
    return (codes[..., :h] | (codes[..., h:] << 4)).astype(np.uint8)


def quant_iq4_nl(x: np.ndarray) -> np.ndarray:
    """The value table is asymmetric (-127 .. 113): the element of largest magnitude is mapped onto -127 by letting the block
    scale carry its sign; every other element takes the nearest table value."""
    xb = x.reshape(-1, 32)
    d = (_signed_max(xb) / -127.0).astype(np.float16)
    df = d.astype(np.float32)
    inv = np.where(df != 0, 1.0 / np.where(df != 0, df, 1), 0).astype(np.float32)
    out = np.empty((xb.shape[0], 18), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    out[:, 2:] = _pack_nibble_halves(_nearest_code(xb * inv[:, None], IQ4_VALUES))
    return out.reshape(-1)


def quant_iq4_xs(x: np.ndarray) -> np.ndarray:
    """Eight 32-value sub-blocks share an fp16 scale; each has a signed 6-bit multiplier (stored + 32)."""
    xb = x.reshape(-1, 8, 32)
    dl = _signed_max(xb) / -127.0                                            # ideal sub-block scales [nb, 8]
    big = _signed_max(dl)
    d = (big / -32.0).astype(np.float16)                                      # the largest one lands on multiplier -32
    df = d.astype(np.float32)
    inv_d = np.where(df != 0, 1.0 / np.where(df != 0, df, 1), 0).astype(np.float32)
    ls = np.clip(_nearest_int(dl * inv_d[:, None]), -32, 31).astype(np.int32)
    sub = df[:, None] * ls
    inv = np.where(sub != 0, 1.0 / np.where(sub != 0, sub, 1), 0).astype(np.float32)
    codes = _nearest_code(xb * inv[:, :, None], IQ4_VALUES)
    u = (ls + 32).astype(np.uint32)
    out = np.empty((xb.shape[0], 136), np.uint8)
    out[:, 0:2] = d.view(np.uint8).reshape(-1, 2)
    sh = ((u >> 4) << (2 * np.arange(8, dtype=np.uint32))[None, :]).sum(-1).astype("<u2")
    out[:, 2:4] = sh.view(np.uint8).reshape(-1, 2)
    lo = (u & 0x0F).astype(np.uint8).reshape(-1, 4, 2)
    out[:, 4:8] = lo[:, :, 0] | (lo[:, :, 1] << 4)
    out[:, 8:] = _pack_nibble_halves(codes).reshape(-1, 128)
    return out.reshape(-1)


def _ternary(x: np.ndarray):
    xb = x.reshape(-1, 256)
    d = np.abs(xb).max(-1).astype(np.float16)
    df = d.astype(np.float32)
    inv = np.where(df != 0, 1.0 / np.where(df != 0, df, 1), 0).astype(np.float32)
    return d, (np.clip(_nearest_int(xb * inv[:, None]), -1, 1) + 1).astype(np.uint8)


def quant_tq2_0(x: np.ndarray) -> np.ndarray:
    d, q = _ternary(x)
    planes = q.reshape(-1, 2, 4, 32)
    out = np.empty((q.shape[0], 66), np.uint8)
    out[:, 0:64] = (planes << (2 * np.arange(4, dtype=np.uint8)).reshape(1, 1, 4, 1)).sum(2).astype(np.uint8).reshape(-1, 64)
    out[:, 64:66] = d.view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def quant_tq1_0(x: np.ndarray) -> np.ndarray:
    d, q = _ternary(x)

    def pack(trits: np.ndarray) -> np.ndarray:                              # [nb, n, m] -> [nb, m]; trit 0 is the most significant
        n = trits.shape[1]
        v = (trits.astype(np.uint32) * (3 ** (4 - np.arange(n, dtype=np.uint32))).reshape(1, n, 1)).sum(1)
        return ((v * 256 + 242) // 243).astype(np.uint8)                    # ceil(v * 256 / 243): the byte as a base-3 fraction

    out = np.empty((q.shape[0], 54), np.uint8)
    out[:, 0:32] = pack(q[:, 0:160].reshape(-1, 5, 32))
    out[:, 32:48] = pack(q[:, 160:240].reshape(-1, 5, 16))
    out[:, 48:52] = pack(q[:, 240:256].reshape(-1, 4, 4))
    out[:, 52:54] = d.view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def quant_mxfp4(x: np.ndarray) -> np.ndarray:
    """Shared power-of-two scale per 32 values chosen so the largest magnitude falls in the top binade of E2M1 (4 .. 6)."""
    xb = x.reshape(-1, 32)
    amax = np.abs(xb).max(-1)
    e = np.where(amax > 0, np.floor(np.log2(np.where(amax > 0, amax, 1))) - 2 + 127, 0)
    e = np.clip(e, 0, 254).astype(np.uint8)
    d = e8m0_to_f32_half(e)
    out = np.empty((xb.shape[0], 17), np.uint8)
    out[:, 0] = e
    out[:, 1:] = _pack_nibble_halves(_nearest_code(xb / d[:, None], MXFP4_VALUES))
    return out.reshape(-1)



WEIGHTED_TYPES = (GGMLType.Q4_K, GGMLType.Q5_K, GGMLType.Q6_K)


def quantize(x: np.ndarray, ggml_type: GGMLType, importance: np.ndarray | None = None) -> np.ndarray:
    """Encode float32 data (last dim multiple of block size) -> raw uint8 bytes.  `importance` ([K], one value per input column,
    e.g. a row of an imatrix file) switches the K-quant encoders to their error-weighted fits; other types ignore it."""
    t = GGMLType(ggml_type)
    x = np.ascontiguousarray(x, dtype=np.float32)
    if t == GGMLType.F32:
        return x.astype("<f4").view(np.uint8).reshape(-1)
    if t == GGMLType.F16:
        return x.astype("<f2").view(np.uint8).reshape(-1)
    if t == GGMLType.BF16:
        return f32_to_bf16(x).astype("<u2").view(np.uint8).reshape(-1)
    per, _ = BLOCK_GEOMETRY[t]
    if x.shape[-1] % per:
        raise ValueError(f"last dim {x.shape[-1]} not a multiple of {per} for {t!r}")
    fn = {
        GGMLType.Q8_0: quant_q8_0,
        GGMLType.Q4_0: quant_q4_0,
        GGMLType.Q4_1: quant_q4_1,
        GGMLType.Q5_0: quant_q5_0,
        GGMLType.Q5_1: quant_q5_1,
        GGMLType.Q2_K: quant_q2_k,
        GGMLType.Q3_K: quant_q3_k,
        GGMLType.Q4_K: quant_q4_k,
        GGMLType.Q5_K: quant_q5_k,
        GGMLType.Q6_K: quant_q6_k,
        GGMLType.IQ4_NL: quant_iq4_nl,
        GGMLType.IQ4_XS: quant_iq4_xs,
        GGMLType.TQ1_0: quant_tq1_0,
        GGMLType.TQ2_0: quant_tq2_0,
        GGMLType.MXFP4: quant_mxfp4,
    }.get(t)
    if fn is None:
        raise NotImplementedError(f"quantize: unsupported type {t!r}")
    if importance is not None and t in WEIGHTED_TYPES:
        imp = np.maximum(np.asarray(importance, np.float32).reshape(-1), 0.0)
        if imp.size != x.shape[-1]:
            raise ValueError(f"importance has {imp.size} entries for rows of {x.shape[-1]}")
        if float(imp.sum()) > 0:
            return fn(x, importance_weights(x, imp, 256))
    return fn(x)


# --------------------------------------------------------------------------------------------
# direct random block synthesis (fast path for multi-GB synthetic checkpoints)


def random_blocks(ggml_type: GGMLType, n_elements: int, rng: np.random.Generator, std: float = 0.02) -> np.ndarray:
    """Produce *valid* random blocks of `ggml_type` whose decoded values have roughly N(0, std) spread,
    without ever materialising float weights (used for 8B/70B-shape random-init checkpoints)."""
    t = GGMLType(ggml_type)
    per, nbytes = BLOCK_GEOMETRY[t]
    if n_elements % per:
        raise ValueError("element count must be a multiple of the block size")
    nb = n_elements // per
    if t in (GGMLType.F32, GGMLType.F16, GGMLType.BF16):
        x = rng.standard_normal(n_elements, dtype=np.float32) * std
        return quantize(x, t)
    out = rng.integers(0, 256, size=(nb, nbytes), dtype=np.uint8)

    def put_f16(col: int, val: np.ndarray):
        out[:, col:col + 2] = val.astype(np.float16).view(np.uint8).reshape(-1, 2)

    jitter = rng.uniform(0.75, 1.25, size=nb).astype(np.float32)
    if t == GGMLType.Q8_0:
        put_f16(0, jitter * (3.0 * std / 127.0))
        # shape int8 payload ~ N(0, 42) so 3 sigma ~ 127
        q = np.clip(np.rint(rng.standard_normal((nb, 32), dtype=np.float32) * 42.0), -127, 127).astype(np.int8)
        out[:, 2:] = q.view(np.uint8)
    elif t == GGMLType.Q4_0:
        put_f16(0, jitter * (3.0 * std / 8.0))
    elif t in (GGMLType.Q4_K, GGMLType.Q5_K):
        nmax = 15 if t == GGMLType.Q4_K else 31
        # value = d*sc*q - dmin*m ; centre around zero: dmin*m ~ d*sc*nmax/2
        put_f16(0, jitter * (6.0 * std / nmax / 48.0))
        put_f16(2, jitter * (3.0 * std / 48.0))
        sc = rng.integers(32, 64, size=(nb, 8), dtype=np.uint8)
        mn = rng.integers(40, 56, size=(nb, 8), dtype=np.uint8)
        out[:, 4:16] = pack_scale_min_k4(sc, mn)
    elif t == GGMLType.Q6_K:
        out[:, 192:208] = rng.integers(32, 127, size=(nb, 16), dtype=np.int8).view(np.uint8)
        put_f16(208, jitter * (3.0 * std / 32.0 / 96.0))
    else:
        raise NotImplementedError(f"random_blocks: unsupported type {t!r}")
    return out.reshape(-1)
