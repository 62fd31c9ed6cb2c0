// Device-side decoding of GGUF block-quant weights into tensor-core / FMA friendly operands.
//
// Storage layouts the kernels consume (see ops/weights.py for the load-time repack):
//   W_F16 / W_BF16 : dense [N][K]
//   W_Q8_0         : SoA   int8 qs[N][K] | fp16 d[N][K/32]              (repacked from 34-byte blocks)
//   W_Q4_K         : native 144-byte super-blocks (16-byte aligned already)
//   W_Q5_K         : native 176-byte super-blocks (16-byte aligned already)
//   W_Q6_K         : SoA   u8 ql[N][K/2] | u8 qh[N][K/4] | i8 sc[N][K/16] | fp16 d[N][K/256]
// Bit layouts are the GGUF on-disk contract (reference structs: llama.cpp/ggml/src/ggml-common.h:170-336;
// reference decoders: ggml-quants.c:401,703,1352,1762).  The decode strategy is ours: a lane takes one
// 32-element "chunk" = two groups of 16 elements, each group = 8 half2 of *exact integer* values plus an
// fp32 (scale, bias) pair, so that  w = scale * q + bias.
#pragma once
#include "common.cuh"

namespace dlp {

struct QWeight {
    const uint8_t* base;   // W_F16/BF16: dense; Q4_K/Q5_K: blocks; Q8_0: qs; Q6_K: ql
    const uint8_t* aux0;   // Q8_0: d (fp16) ; Q6_K: qh
    const uint8_t* aux1;   // Q6_K: sc (int8)
    const uint8_t* aux2;   // Q6_K: d (fp16)
    int N, K;
};

struct QChunk {
    __half2 q[2][8];  // two groups of 16 exact-integer values
    float scale[2];
    float bias[2];
    int k[2];         // element offset of each group inside the row
};

#ifdef __CUDACC__

// 4 bytes (each 0..255) -> two half2 holding (b0,b1) and (b2,b3) minus `sub` (exact for |v| <= 1024)
__device__ __forceinline__ void bytes4_to_h2(uint32_t bytes, __half2 sub, __half2& lo, __half2& hi) {
    const uint32_t a = __byte_perm(bytes, 0x64646464u, 0x4140);  // [b0,0x64,b1,0x64] = half2(1024+b0, 1024+b1)
    const uint32_t b = __byte_perm(bytes, 0x64646464u, 0x4342);
    lo = __hsub2(*reinterpret_cast<const __half2*>(&a), sub);
    hi = __hsub2(*reinterpret_cast<const __half2*>(&b), sub);
}

__device__ __forceinline__ uint32_t k4_byte(uint32_t w0, uint32_t w1, uint32_t w2, int idx) {
    const uint32_t w = idx < 4 ? w0 : (idx < 8 ? w1 : w2);
    return (w >> ((idx & 3) * 8)) & 0xFFu;
}
// 6-bit scale / min of sub-block j from the 12 packed bytes (held in three registers, no local memory)
__device__ __forceinline__ void scale_min_k4(uint32_t w0, uint32_t w1, uint32_t w2, int j, int& sc, int& mn) {
    if (j < 4) {
        sc = k4_byte(w0, w1, w2, j) & 63;
        mn = k4_byte(w0, w1, w2, j + 4) & 63;
    } else {
        sc = (k4_byte(w0, w1, w2, j + 4) & 0x0F) | ((k4_byte(w0, w1, w2, j - 4) >> 6) << 4);
        mn = (k4_byte(w0, w1, w2, j + 4) >> 4) | ((k4_byte(w0, w1, w2, j) >> 6) << 4);
    }
}

template <int WT>
__device__ __forceinline__ void load_chunk(const QWeight& w, int row, int c, QChunk& out);

template <>
__device__ __forceinline__ void load_chunk<W_Q8_0>(const QWeight& w, int row, int c, QChunk& out) {
    const uint8_t* qs = w.base + (size_t)row * w.K + (size_t)c * 32;
    const uint4 v0 = ldg_nc_v4(qs);
    const uint4 v1 = ldg_nc_v4(qs + 16);
    const __half d = reinterpret_cast<const __half*>(w.aux0)[(size_t)row * (w.K >> 5) + c];
    const __half2 sub = __float2half2_rn(1152.f);  // 1024 + 128 (bytes are xor 0x80 -> q + 128)
    const uint32_t wd[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
        bytes4_to_h2(wd[i] ^ 0x80808080u, sub, out.q[i >> 2][(i & 3) * 2], out.q[i >> 2][(i & 3) * 2 + 1]);
    const float df = __half2float(d);
    out.scale[0] = out.scale[1] = df;
    out.bias[0] = out.bias[1] = 0.f;
    out.k[0] = c * 32;
    out.k[1] = c * 32 + 16;
}

template <>
__device__ __forceinline__ void load_chunk<W_Q4_K>(const QWeight& w, int row, int c, QChunk& out) {
    const int blk = c >> 3, j = (c & 7) >> 1, l0 = (c & 1) * 16;
    const uint8_t* b = w.base + ((size_t)row * (w.K >> 8) + blk) * 144;
    const uint4 hdr = __ldg(reinterpret_cast<const uint4*>(b));
    const uint4 v = ldg_nc_v4(b + 16 + j * 32 + l0);
    const float2 dm = h2_to_f2(hdr.x);
    int sc0, mn0, sc1, mn1;
    scale_min_k4(hdr.y, hdr.z, hdr.w, 2 * j, sc0, mn0);
    scale_min_k4(hdr.y, hdr.z, hdr.w, 2 * j + 1, sc1, mn1);
    const __half2 sub = __float2half2_rn(1024.f);
    const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        bytes4_to_h2(wd[i] & 0x0F0F0F0Fu, sub, out.q[0][i * 2], out.q[0][i * 2 + 1]);
        bytes4_to_h2((wd[i] >> 4) & 0x0F0F0F0Fu, sub, out.q[1][i * 2], out.q[1][i * 2 + 1]);
    }
    out.scale[0] = dm.x * sc0;
    out.scale[1] = dm.x * sc1;
    out.bias[0] = -dm.y * mn0;
    out.bias[1] = -dm.y * mn1;
    out.k[0] = blk * 256 + j * 64 + l0;
    out.k[1] = out.k[0] + 32;
}

template <>
__device__ __forceinline__ void load_chunk<W_Q5_K>(const QWeight& w, int row, int c, QChunk& out) {
    const int blk = c >> 3, j = (c & 7) >> 1, l0 = (c & 1) * 16;
    const uint8_t* b = w.base + ((size_t)row * (w.K >> 8) + blk) * 176;
    const uint4 hdr = __ldg(reinterpret_cast<const uint4*>(b));
    const uint4 h = __ldg(reinterpret_cast<const uint4*>(b + 16 + l0));
    const uint4 v = ldg_nc_v4(b + 48 + j * 32 + l0);
    const float2 dm = h2_to_f2(hdr.x);
    int sc0, mn0, sc1, mn1;
    scale_min_k4(hdr.y, hdr.z, hdr.w, 2 * j, sc0, mn0);
    scale_min_k4(hdr.y, hdr.z, hdr.w, 2 * j + 1, sc1, mn1);
    const __half2 sub = __float2half2_rn(1024.f);
    const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
    const uint32_t hd[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t lo = (wd[i] & 0x0F0F0F0Fu) | (((hd[i] >> (2 * j)) & 0x01010101u) << 4);
        const uint32_t hi = ((wd[i] >> 4) & 0x0F0F0F0Fu) | (((hd[i] >> (2 * j + 1)) & 0x01010101u) << 4);
        bytes4_to_h2(lo, sub, out.q[0][i * 2], out.q[0][i * 2 + 1]);
        bytes4_to_h2(hi, sub, out.q[1][i * 2], out.q[1][i * 2 + 1]);
    }
    out.scale[0] = dm.x * sc0;
    out.scale[1] = dm.x * sc1;
    out.bias[0] = -dm.y * mn0;
    out.bias[1] = -dm.y * mn1;
    out.k[0] = blk * 256 + j * 64 + l0;
    out.k[1] = out.k[0] + 32;
}

template <>
__device__ __forceinline__ void load_chunk<W_Q6_K>(const QWeight& w, int row, int c, QChunk& out) {
    const int blk = c >> 3, cc = c & 7, half = cc >> 2, l0 = (cc & 3) * 16;
    const int lq = l0 & 31;            // offset inside the 32-wide quarter
    const int upper = l0 >> 5;         // 0: (q1,q3) pair ; 1: (q2,q4) pair
    const size_t nblk = (size_t)row * (w.K >> 8) + blk;
    const uint4 v = ldg_nc_v4(w.base + nblk * 128 + half * 64 + l0);
    const uint4 h = ldg_nc_v4(w.aux0 + nblk * 64 + half * 32 + lq);
    const int8_t* sc = reinterpret_cast<const int8_t*>(w.aux1) + nblk * 16 + half * 8 + (lq >> 4) + upper * 2;
    const float d = __half2float(reinterpret_cast<const __half*>(w.aux2)[nblk]);
    const int sh = upper * 2;
    const __half2 sub = __float2half2_rn(1056.f);  // 1024 + 32
    const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
    const uint32_t hd[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t lo = (wd[i] & 0x0F0F0F0Fu) | (((hd[i] >> sh) & 0x03030303u) << 4);
        const uint32_t hi = ((wd[i] >> 4) & 0x0F0F0F0Fu) | (((hd[i] >> (sh + 4)) & 0x03030303u) << 4);
        bytes4_to_h2(lo, sub, out.q[0][i * 2], out.q[0][i * 2 + 1]);
        bytes4_to_h2(hi, sub, out.q[1][i * 2], out.q[1][i * 2 + 1]);
    }
    out.scale[0] = d * (float)sc[0];
    out.scale[1] = d * (float)sc[4];
    out.bias[0] = out.bias[1] = 0.f;
    out.k[0] = blk * 256 + half * 128 + upper * 32 + lq;
    out.k[1] = out.k[0] + 64;
}

#endif  // __CUDACC__
}  // namespace dlp
