// Small fused ops around the GEMMs: embedding gather (+dequant), RMSNorm (+residual), RoPE fused with
// the paged-KV write, SwiGLU, dequant-to-fp16, sampling.
//
// Reference counterparts (llama.cpp/ggml/src/ggml-cuda/): getrows.cu, norm.cu:107 (fused x weight + residual),
// rope.cu:44-111 with the set_rows fusion (:58-78), set-rows.cu, unary.cu (GLU), convert.cu, argmax.cu.
// The sampler chain order follows common/sampling.cpp (top-k -> top-p -> min-p -> temperature -> dist) but runs
// on the device so only the chosen token id crosses PCIe.
#include "kernels.cuh"

namespace dlp {

// ---------------------------------------------------------------------------------------------
// generic "decode 16 consecutive elements of a row" used by embedding gather and dequant-to-fp16
template <int WT>
__device__ __forceinline__ void decode_chunk_to_f32(const QWeight& w, int row, int c, float* out32, int* k0, int* k1) {
    QChunk ch;
    load_chunk<WT>(w, row, c, ch);
#pragma unroll
    for (int g = 0; g < 2; ++g)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float2 f = __half22float2(ch.q[g][i]);
            out32[g * 16 + i * 2] = fmaf(ch.scale[g], f.x, ch.bias[g]);
            out32[g * 16 + i * 2 + 1] = fmaf(ch.scale[g], f.y, ch.bias[g]);
        }
    *k0 = ch.k[0];
    *k1 = ch.k[1];
}

template <int WT, typename OutT>
__global__ void dequant_rows_kernel(const QWeight w, const int* __restrict__ row_ids, OutT* __restrict__ out, int ld_out) {
    // grid.x = output rows; each thread decodes 32-element chunks
    const int orow = blockIdx.x;
    pdl_wait();
    pdl_launch_dependents();
    const int row = row_ids ? row_ids[orow] : orow;
    const int nchunk = w.K / 32;
    for (int c = threadIdx.x; c < nchunk; c += blockDim.x) {
        float v[32];
        int k0, k1;
        decode_chunk_to_f32<WT>(w, row, c, v, &k0, &k1);
        OutT* o = out + (size_t)orow * ld_out;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            o[k0 + i] = (OutT)v[i];
            o[k1 + i] = (OutT)v[16 + i];
        }
    }
}

template <typename InT, typename OutT>
__global__ void dense_rows_kernel(const InT* __restrict__ w, int K, const int* __restrict__ row_ids, OutT* __restrict__ out, int ld_out) {
    const int orow = blockIdx.x;
    pdl_wait();
    pdl_launch_dependents();
    const int row = row_ids ? row_ids[orow] : orow;
    const int k = (blockIdx.y * blockDim.x + threadIdx.x) * 8;   // K is a multiple of 8 for every supported model
    if (k + 8 <= K) {
#pragma unroll
        for (int j = 0; j < 8; ++j) out[(size_t)orow * ld_out + k + j] = (OutT)(float)w[(size_t)row * K + k + j];
    } else {
        for (int j = k; j < K; ++j) out[(size_t)orow * ld_out + j] = (OutT)(float)w[(size_t)row * K + j];
    }
}

template <typename OutT>
static int dequant_rows_t(const QWeight& w, int wtype, const int* row_ids, int n_rows, OutT* out, int ld_out, cudaStream_t s) {
    DLP_REQUIRE(n_rows > 0, "dequant_rows: no rows");
    const int threads = 128;
    switch (wtype) {
        case W_F16: DLP_CHECK_CUDA(launch_pdl(dense_rows_kernel<__half, OutT>, dim3(n_rows, (w.K / 8 + 255) / 256 + ((w.K % 2048) ? 1 : 0)), dim3(256), 0, s, (const __half*)w.base, w.K, row_ids, out, ld_out)); break;
        case W_BF16: DLP_CHECK_CUDA(launch_pdl(dense_rows_kernel<__nv_bfloat16, OutT>, dim3(n_rows, (w.K / 8 + 255) / 256 + ((w.K % 2048) ? 1 : 0)), dim3(256), 0, s, (const __nv_bfloat16*)w.base, w.K, row_ids, out, ld_out)); break;
        case W_F32: DLP_CHECK_CUDA(launch_pdl(dense_rows_kernel<float, OutT>, dim3(n_rows, (w.K / 8 + 255) / 256 + ((w.K % 2048) ? 1 : 0)), dim3(256), 0, s, (const float*)w.base, w.K, row_ids, out, ld_out)); break;
        case W_Q8_0: DLP_CHECK_CUDA(launch_pdl(dequant_rows_kernel<W_Q8_0, OutT>, dim3(n_rows), dim3(threads), 0, s, w, row_ids, out, ld_out)); break;
        case W_Q4_K: DLP_CHECK_CUDA(launch_pdl(dequant_rows_kernel<W_Q4_K, OutT>, dim3(n_rows), dim3(threads), 0, s, w, row_ids, out, ld_out)); break;
        case W_Q5_K: DLP_CHECK_CUDA(launch_pdl(dequant_rows_kernel<W_Q5_K, OutT>, dim3(n_rows), dim3(threads), 0, s, w, row_ids, out, ld_out)); break;
        case W_Q6_K: DLP_CHECK_CUDA(launch_pdl(dequant_rows_kernel<W_Q6_K, OutT>, dim3(n_rows), dim3(threads), 0, s, w, row_ids, out, ld_out)); break;
        default: DLP_REQUIRE(false, "dequant_rows: unsupported weight type %d", wtype);
    }
    DLP_LAUNCH_CHECK();
    return 0;
}

int dequant_rows_f32(const QWeight& w, int wtype, const int* row_ids, int n_rows, float* out, int ld_out, cudaStream_t s) {
    return dequant_rows_t<float>(w, wtype, row_ids, n_rows, out, ld_out, s);
}
int dequant_rows_f16(const QWeight& w, int wtype, const int* row_ids, int n_rows, __half* out, int ld_out, cudaStream_t s) {
    return dequant_rows_t<__half>(w, wtype, row_ids, n_rows, out, ld_out, s);
}

// ---------------------------------------------------------------------------------------------
// RMSNorm:  h (fp32) [+= delta]  ->  out fp16 = h * rsqrt(mean(h^2) + eps) * w
__global__ void __launch_bounds__(256) rmsnorm_kernel(float* __restrict__ h, const float* __restrict__ delta,
                                                      const float* __restrict__ w, __half* __restrict__ out, int E, float eps) {
    __shared__ float red[32];
    const size_t base = (size_t)blockIdx.x * E;
    pdl_wait();
    pdl_launch_dependents();
    float ss = 0.f;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(h + base + k);
        if (delta) {
            const float4 d = *reinterpret_cast<const float4*>(delta + base + k);
            v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            *reinterpret_cast<float4*>(h + base + k) = v;
        }
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = block_sum(ss, red);
    const float rs = rsqrtf(ss / (float)E + eps);
    if (out == nullptr) return;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4*>(h + base + k);
        const float4 g = *reinterpret_cast<const float4*>(w + k);
        __half2 a = __floats2half2_rn(v.x * rs * g.x, v.y * rs * g.y);
        __half2 b = __floats2half2_rn(v.z * rs * g.z, v.w * rs * g.w);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&a);
        o.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(out + base + k) = o;
    }
}

int rmsnorm(float* h, const float* delta, const float* w, __half* out, int T, int E, float eps, cudaStream_t s) {
    DLP_REQUIRE(E % 4 == 0 && T > 0, "rmsnorm: bad shape T=%d E=%d", T, E);
    DLP_CHECK_CUDA(launch_pdl(rmsnorm_kernel, dim3(T), dim3(256), 0, s, h, delta, w, out, E, eps));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// LayerNorm (the LLM_NORM families: starcoder2, falcon, olmo, command-r, orion):
//   h (fp32) [+= delta]  ->  out fp16 = (h - mean) * rsqrt(var + eps) * w + b      (w, b optional: olmo has neither, command-r no b)
// Two passes over the row for the variance (mean first): the row is at most a few tens of KB and stays in L1/L2.
// ref: norm_f32 (ggml-cuda/norm.cu:5-45) followed by the mul / add nodes of build_norm.
__global__ void __launch_bounds__(256) layernorm_kernel(float* __restrict__ h, const float* __restrict__ delta, const float* __restrict__ w,
                                                        const float* __restrict__ b, __half* __restrict__ out, int E, float eps) {
    __shared__ float red[32];
    const size_t base = (size_t)blockIdx.x * E;
    pdl_wait();
    pdl_launch_dependents();
    float sum = 0.f;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(h + base + k);
        if (delta) {
            const float4 d = *reinterpret_cast<const float4*>(delta + base + k);
            v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            *reinterpret_cast<float4*>(h + base + k) = v;
        }
        sum += v.x + v.y + v.z + v.w;
    }
    const float mean = block_sum(sum, red) / (float)E;
    float ss = 0.f;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {      // each thread re-reads exactly the elements it wrote above
        const float4 v = *reinterpret_cast<const float4*>(h + base + k);
        const float a = v.x - mean, c = v.y - mean, d = v.z - mean, e = v.w - mean;
        ss += a * a + c * c + d * d + e * e;
    }
    const float rs = rsqrtf(block_sum(ss, red) / (float)E + eps);
    if (out == nullptr) return;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        const float4 v = *reinterpret_cast<const float4*>(h + base + k);
        float4 g = make_float4(1.f, 1.f, 1.f, 1.f), o4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (w) g = *reinterpret_cast<const float4*>(w + k);
        if (b) o4 = *reinterpret_cast<const float4*>(b + k);
        __half2 lo = __floats2half2_rn((v.x - mean) * rs * g.x + o4.x, (v.y - mean) * rs * g.y + o4.y);
        __half2 hi = __floats2half2_rn((v.z - mean) * rs * g.z + o4.z, (v.w - mean) * rs * g.w + o4.w);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&lo);
        o.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(out + base + k) = o;
    }
}

int layernorm(float* h, const float* delta, const float* w, const float* b, __half* out, int T, int E, float eps, cudaStream_t s) {
    DLP_REQUIRE(E % 4 == 0 && T > 0, "layernorm: bad shape T=%d E=%d", T, E);
    DLP_CHECK_CUDA(launch_pdl(layernorm_kernel, dim3(T), dim3(256), 0, s, h, delta, w, b, out, E, eps));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Bias + activation between the two mat-muls of a non-gated MLP:  out fp16 [T, F] = act(x fp32 [T, F] + bias[F]).
// act: 0 identity, 1 GELU (tanh form, what ggml's GELU computes), 2 ReLU^2, 3 SiLU, 4 GELU (erf form).
// ref: gelu / silu / relu + sqr (ggml-cuda/unary.cu), bias add (binbcast.cu).
__device__ __forceinline__ float act_apply(float v, int act) {
    switch (act) {
        case 1: { const float u = 0.7978845608028654f * v * (1.f + 0.044715f * v * v); return 0.5f * v * (1.f + tanhf(u)); }
        case 2: { const float r = fmaxf(v, 0.f); return r * r; }
        case 3: return v / (1.f + __expf(-v));
        case 4: return 0.5f * v * (1.f + erff(v * 0.7071067811865476f));
        default: return v;
    }
}

__global__ void __launch_bounds__(256) bias_act_kernel(const float* __restrict__ x, const float* __restrict__ bias, __half* __restrict__ out,
                                                       size_t n4, int F, int act) {
    pdl_wait();
    pdl_launch_dependents();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = *reinterpret_cast<const float4*>(x + i * 4);
        if (bias) {
            const float4 b = *reinterpret_cast<const float4*>(bias + (i * 4) % (size_t)F);
            v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
        }
        __half2 lo = __floats2half2_rn(act_apply(v.x, act), act_apply(v.y, act));
        __half2 hi = __floats2half2_rn(act_apply(v.z, act), act_apply(v.w, act));
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&lo);
        o.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(out + i * 4) = o;
    }
}

int bias_act(const float* x, const float* bias, __half* out, int T, int F, int act, cudaStream_t s) {
    DLP_REQUIRE(F % 4 == 0 && T > 0 && act >= 0 && act <= 4, "bias_act: bad arguments T=%d F=%d act=%d", T, F, act);
    const size_t n4 = (size_t)T * F / 4;
    const int blocks = (int)std::min<size_t>((n4 + 255) / 256, (size_t)148 * 8);
    DLP_CHECK_CUDA(launch_pdl(bias_act_kernel, dim3(blocks), dim3(256), 0, s, x, bias, out, n4, F, act));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// RoPE (adjacent-pair "norm" or half-split "neox") on Q and K, fused with the paged-KV write of K and V.
//   qkv   : [T, (Hq + 2*Hkv) * D]  fp32 or fp16 (row = q heads | k heads | v heads)
//   q_out : [T, Hq * D] fp16 (rotated, pre-scaled by q_scale)
//   k/v cache : [n_pages][Hkv][page_size][D] fp16 ; slot[t] = page * page_size + offset   (slot < 0 => skip write)
//   EX variant (other architecture families): bias [H*D] added to q/k/v before anything else (qwen2), and a per-head
//   RMSNorm over D with weights qn / kn applied to q / k heads before the rotation (qwen3).  A head's D/2 pairs sit in
//   D/64 consecutive warps of one CTA (256 % (D/2) == 0), so the sum of squares is a shuffle plus a few smem words.
template <typename InT, bool EX>
__global__ void rope_kv_kernel(const InT* __restrict__ qkv, const int* __restrict__ pos, const int* __restrict__ slot,
                               const float* __restrict__ inv_freq, __half* __restrict__ q_out, __half* __restrict__ kc,
                               __half* __restrict__ vc, int Hq, int Hkv, int D, int n_rot, int page_size, int neox, float q_scale,
                               const float* __restrict__ bias, const float* __restrict__ qn, const float* __restrict__ kn,
                               float eps, const char* __restrict__ pf0, unsigned pf0_bytes, const char* __restrict__ pf1,
                               unsigned pf1_bytes, unsigned long long* trace) {
    // one (head, pair) per thread: grid = (T, ceil(H * D/2 / 256)) so that even a single decode token spreads over
    // a dozen CTAs and every load is issued at once (a serial per-thread loop here is pure latency)
    const int t = blockIdx.x;
    const int H = Hq + 2 * Hkv;
    const int half_d = D / 2;
    const int idx = blockIdx.y * blockDim.x + threadIdx.x;
    if (threadIdx.x == 0) {   // weights of the next mat-vec (Wo) -> L2 while this kernel and attention run; no dependency
        prefetch_share_l2(pf0, pf0_bytes, blockIdx.x * gridDim.y + blockIdx.y, gridDim.x * gridDim.y);
        prefetch_share_l2(pf1, pf1_bytes, blockIdx.x * gridDim.y + blockIdx.y, gridDim.x * gridDim.y);
    }
    trace_event(trace, TR_ROPE | TR_BEGIN);
    pdl_wait();
    trace_event(trace, TR_ROPE | TR_WAITED);
    pdl_launch_dependents();
    const bool live = idx < H * half_d;
    if (!EX && !live) return;
    const float p = (float)pos[t];
    const int sl = slot[t];
    const int page = sl >= 0 ? sl / page_size : 0, off = sl >= 0 ? sl % page_size : 0;
    const InT* row = qkv + (size_t)t * H * D;
    const int hh = live ? idx / half_d : 0, i = idx % half_d;
    int d0, d1;
    if (neox) { d0 = i; d1 = i + n_rot / 2; } else { d0 = 2 * i; d1 = 2 * i + 1; }
    const bool rot = hh < Hq + Hkv && (neox ? i < n_rot / 2 : 2 * i < n_rot);
    if (neox && !rot && i >= n_rot / 2) {  // tail dims beyond the rotated span: map pair index onto them
        d0 = n_rot + 2 * (i - n_rot / 2);
        d1 = d0 + 1;
    }
    float x0 = live ? (float)row[hh * D + d0] : 0.f, x1 = live ? (float)row[hh * D + d1] : 0.f;
    if (EX) {
        if (bias != nullptr && live) { x0 += bias[hh * D + d0]; x1 += bias[hh * D + d1]; }
        if (qn != nullptr) {
            __shared__ float wsum[8];
            float ss = x0 * x0 + x1 * x1;
            if (half_d >= 32) {
                ss = warp_sum(ss);
                if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = ss;
                __syncthreads();
                const int wph = half_d / 32, w0 = (threadIdx.x >> 5) / wph * wph;
                ss = 0.f;
                for (int w = 0; w < wph; ++w) ss += wsum[w0 + w];
            } else {
                for (int o = half_d / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            }
            if (hh < Hq + Hkv) {
                const float r = rsqrtf(ss / (float)D + eps);
                const float* nw = hh < Hq ? qn : kn;
                x0 *= r * nw[d0];
                x1 *= r * nw[d1];
            }
        }
        if (!live) return;
    }
    if (rot) {
        float sn, cs;
        sincosf(p * inv_freq[i], &sn, &cs);
        const float r0 = x0 * cs - x1 * sn, r1 = x0 * sn + x1 * cs;
        x0 = r0; x1 = r1;
    }
    if (hh < Hq) {
        __half* o = q_out + ((size_t)t * Hq + hh) * D;
        o[d0] = __float2half_rn(x0 * q_scale);
        o[d1] = __float2half_rn(x1 * q_scale);
    } else if (sl >= 0) {
        const bool is_k = hh < Hq + Hkv;
        const int kvh = is_k ? hh - Hq : hh - Hq - Hkv;
        __half* o = (is_k ? kc : vc) + (((size_t)page * Hkv + kvh) * page_size + off) * D;
        o[d0] = __float2half_rn(x0);
        o[d1] = __float2half_rn(x1);
    }
}

int rope_kv(const void* qkv, int qkv_is_f16, const int* pos, const int* slot, const float* inv_freq, __half* q_out,
            __half* kc, __half* vc, int T, int Hq, int Hkv, int D, int n_rot, int page_size, int neox, float q_scale,
            cudaStream_t s, const float* bias, const float* qn, const float* kn, float eps, const void* pf0, size_t pf0_bytes,
            const void* pf1, size_t pf1_bytes) {
    DLP_REQUIRE(pf0_bytes < (1ull << 32) && pf1_bytes < (1ull << 32) && pf0_bytes % 16 == 0 && pf1_bytes % 16 == 0 &&
                ((uintptr_t)pf0 & 15) == 0 && ((uintptr_t)pf1 & 15) == 0, "rope_kv: bad prefetch range");
    DLP_REQUIRE(T > 0 && D % 2 == 0 && n_rot <= D && n_rot % 2 == 0, "rope_kv: bad shape");
    DLP_REQUIRE((qn == nullptr) == (kn == nullptr), "rope_kv: q/k norm weights come as a pair");
    const bool ex = bias != nullptr || qn != nullptr;
    if (qn != nullptr) DLP_REQUIRE(256 % (D / 2) == 0 && (D / 2 >= 32 || (D / 2 & (D / 2 - 1)) == 0), "rope_kv: qk-norm needs D/2 | 256 (D=%d)", D);
    const dim3 grid(T, ((Hq + 2 * Hkv) * (D / 2) + 255) / 256);
#define DLP_ROPE_LAUNCH(TY, EXF) DLP_CHECK_CUDA(launch_pdl(rope_kv_kernel<TY, EXF>, grid, dim3(256), 0, s, (const TY*)qkv, pos, slot, inv_freq, q_out, kc, vc, Hq, Hkv, D, n_rot, page_size, neox, q_scale, bias, qn, kn, eps, (const char*)pf0, (unsigned)pf0_bytes, (const char*)pf1, (unsigned)pf1_bytes, trace_buffer()))
    if (qkv_is_f16) { if (ex) DLP_ROPE_LAUNCH(__half, true); else DLP_ROPE_LAUNCH(__half, false); }
    else            { if (ex) DLP_ROPE_LAUNCH(float, true);  else DLP_ROPE_LAUNCH(float, false); }
#undef DLP_ROPE_LAUNCH
    return 0;
}

// ---------------------------------------------------------------------------------------------
// SwiGLU on a fused [T, 2F] (gate | up) fp16/fp32 buffer -> fp16 [T, F]
template <typename InT>
__global__ void swiglu_kernel(const InT* __restrict__ gu, __half* __restrict__ out, int F, int ld_in) {
    const int t = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < F) {
        const float g = (float)gu[(size_t)t * ld_in + i], u = (float)gu[(size_t)t * ld_in + F + i];
        out[(size_t)t * F + i] = __float2half_rn(g / (1.f + __expf(-g)) * u);
    }
}
int swiglu(const void* gu, int is_f16, __half* out, int T, int F, int ld_in, cudaStream_t s) {
    dim3 grid((F + 255) / 256, T);
    if (is_f16) swiglu_kernel<__half><<<grid, 256, 0, s>>>((const __half*)gu, out, F, ld_in);
    else swiglu_kernel<float><<<grid, 256, 0, s>>>((const float*)gu, out, F, ld_in);
    DLP_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// decode bookkeeping on the device (keeps the per-token loop free of host round trips).  For batch entry s
// (sequence slot sid = active[s], or s when active == NULL):
//   pos[s] = seq_len[sid]; slot[s] = block_table[sid][pos / page] * page + pos % page;
//   kv_len[s] = pos + 1; seq_len[sid] += 1
__global__ void advance_kernel(int* __restrict__ seq_len, const int* __restrict__ active, int* __restrict__ pos,
                               int* __restrict__ slot, int* __restrict__ kv_len, const int* __restrict__ block_table,
                               int max_pages, int page_size, int n) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    pdl_wait();
    pdl_launch_dependents();
    if (s >= n) return;
    const int sid = active ? active[s] : s;
    const int p = seq_len[sid];
    pos[s] = p;
    const int pg = p / page_size;
    slot[s] = (pg < max_pages) ? block_table[(size_t)sid * max_pages + pg] * page_size + p % page_size : -1;
    if (kv_len) kv_len[s] = p + 1;
    seq_len[sid] = p + 1;
}
int advance(int* seq_len, const int* active, int* pos, int* slot, int* kv_len, const int* block_table, int max_pages,
            int page_size, int n, cudaStream_t s) {
    DLP_CHECK_CUDA(launch_pdl(advance_kernel, dim3((n + 127) / 128), dim3(128), 0, s, seq_len, active, pos, slot, kv_len, block_table, max_pages, page_size, n));
    return 0;
}

// Same bookkeeping, one CTA per batch entry, plus the step's RoPE table: rope_cs[s][i] = (cos, sin)(pos[s] * inv_freq[i]).
// Every layer's attention kernel of this decode step reads the table instead of evaluating sincosf itself.
__global__ void __launch_bounds__(128) advance_rope_kernel(int* __restrict__ seq_len, const int* __restrict__ active,
                                                           int* __restrict__ pos, int* __restrict__ slot, int* __restrict__ kv_len,
                                                           const int* __restrict__ block_table, int max_pages, int page_size,
                                                           const float* __restrict__ inv_freq, int n_pairs,
                                                           float2* __restrict__ rope_cs) {
    const int s = blockIdx.x;
    pdl_wait();
    pdl_launch_dependents();
    const int sid = active ? active[s] : s;
    const int p = seq_len[sid];
    __syncthreads();                    // everybody has read the old length before thread 0 bumps it
    if (threadIdx.x == 0) {
        pos[s] = p;
        const int pg = p / page_size;
        slot[s] = (pg < max_pages) ? block_table[(size_t)sid * max_pages + pg] * page_size + p % page_size : -1;
        if (kv_len) kv_len[s] = p + 1;
        seq_len[sid] = p + 1;
    }
    for (int i = threadIdx.x; i < n_pairs; i += blockDim.x) {
        float sn, cs;
        sincosf((float)p * inv_freq[i], &sn, &cs);
        rope_cs[(size_t)s * n_pairs + i] = make_float2(cs, sn);
    }
}
int advance_rope(int* seq_len, const int* active, int* pos, int* slot, int* kv_len, const int* block_table, int max_pages,
                 int page_size, int n, const float* inv_freq, int n_pairs, float2* rope_cs, cudaStream_t s) {
    DLP_REQUIRE(n > 0 && n_pairs > 0, "advance_rope: bad shape");
    DLP_CHECK_CUDA(launch_pdl(advance_rope_kernel, dim3(n), dim3(128), 0, s, seq_len, active, pos, slot, kv_len, block_table, max_pages, page_size, inv_freq, n_pairs, rope_cs));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Sampling.  One CTA per row of logits [R, V] (fp32).
__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t f2ord(float f) {  // order-preserving float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void argmax_row(const float* __restrict__ row, int V, int* __restrict__ out_slot) {
    __shared__ float sv[32];
    __shared__ int si[32];
    float best = -INFINITY;
    int bi = 0;
    if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        const int V4 = V / 4;
#pragma unroll 8
        for (int i = threadIdx.x; i < V4; i += blockDim.x) {   // 16-byte loads, 8 in flight per thread
            const float4 v = *reinterpret_cast<const float4*>(row + 4 * i);
            if (v.x > best) { best = v.x; bi = 4 * i; }       // within a thread indices only grow: '>' keeps the lowest
            if (v.y > best) { best = v.y; bi = 4 * i + 1; }
            if (v.z > best) { best = v.z; bi = 4 * i + 2; }
            if (v.w > best) { best = v.w; bi = 4 * i + 3; }
        }
        for (int i = V4 * 4 + threadIdx.x; i < V; i += blockDim.x) {
            const float v = row[i];
            if (v > best) { best = v; bi = i; }
        }
    } else {
        for (int i = threadIdx.x; i < V; i += blockDim.x) {
            const float v = row[i];
            if (v > best) { best = v; bi = i; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sv[warp] = best; si[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        best = lane < nw ? sv[lane] : -INFINITY;
        bi = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) *out_slot = bi;
    }
}

__global__ void __launch_bounds__(1024) argmax_kernel(const float* __restrict__ logits, int V, int ld, int* __restrict__ out) {
    pdl_wait();
    pdl_launch_dependents();
    argmax_row(logits + (size_t)blockIdx.x * ld, V, out + blockIdx.x);
}

int argmax(const float* logits, int R, int V, int ld, int* out, cudaStream_t s) {
    DLP_REQUIRE(R > 0 && V > 0, "argmax: bad shape");
    DLP_CHECK_CUDA(launch_pdl(argmax_kernel, dim3(R), dim3(1024), 0, s, logits, V, ld, out));
    return 0;
}

constexpr int SAMPLE_MAX_CAND = 1024;

// top-k -> top-p -> min-p -> temperature -> multinomial draw
__device__ __forceinline__ void sample_row(const float* __restrict__ row, int V, int top_k, float top_p, float min_p, float temp,
                                           uint32_t seed, uint32_t ctr, int* __restrict__ out_slot) {
    __shared__ int s_count;
    __shared__ uint32_t s_lo, s_hi;
    __shared__ float cv[SAMPLE_MAX_CAND];
    __shared__ int ci[SAMPLE_MAX_CAND];
    __shared__ int s_n;
    const int k = max(1, min(top_k <= 0 ? SAMPLE_MAX_CAND : top_k, min(V, SAMPLE_MAX_CAND)));
    // radix select (4 passes of 8 bits, most significant first) on the order-preserving integer image of the logits:
    // after the passes `prefix` is the key of the k-th largest element
    __shared__ int hist[256];
    uint32_t prefix = 0, mask = 0;
    int want = k;   // rank (from the top) still to be located inside the current prefix bucket
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        for (int b = threadIdx.x; b < 256; b += blockDim.x) hist[b] = 0;
        __syncthreads();
#pragma unroll 4
        for (int i = threadIdx.x; i < V; i += blockDim.x) {
            const uint32_t key = f2ord(row[i]);
            if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255], 1);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int acc = 0, b = 255;
            for (; b > 0; --b) {
                if (acc + hist[b] >= want) break;
                acc += hist[b];
            }
            s_lo = (uint32_t)b;
            s_count = want - acc;
        }
        __syncthreads();
        prefix |= s_lo << shift;
        mask |= 0xFFu << shift;
        want = s_count;
        __syncthreads();
    }
    if (threadIdx.x == 0) s_lo = prefix;
    __syncthreads();
    const uint32_t thr = s_lo;  // largest threshold with count(>= thr) >= k
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
        const float v = row[i];
        if (f2ord(v) >= thr) {
            const int slot = atomicAdd(&s_n, 1);
            if (slot < SAMPLE_MAX_CAND) { cv[slot] = v; ci[slot] = i; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int n = min(s_n, SAMPLE_MAX_CAND);
        // small insertion sort, descending by (value, -index) for determinism
        for (int a = 1; a < n; ++a) {
            const float v = cv[a]; const int id = ci[a];
            int b = a - 1;
            while (b >= 0 && (cv[b] < v || (cv[b] == v && ci[b] > id))) { cv[b + 1] = cv[b]; ci[b + 1] = ci[b]; --b; }
            cv[b + 1] = v; ci[b + 1] = id;
        }
        if (n > k) n = k;  // ties at the threshold
        const float mx = cv[0];
        float sum = 0.f;
        for (int a = 0; a < n; ++a) sum += __expf(cv[a] - mx);
        // top-p (keep the smallest prefix with cumulative prob >= top_p), then min-p relative to the best
        if (top_p < 1.f) {
            float cum = 0.f;
            int keep = n;
            for (int a = 0; a < n; ++a) {
                cum += __expf(cv[a] - mx) / sum;
                if (cum >= top_p) { keep = a + 1; break; }
            }
            n = keep;
        }
        if (min_p > 0.f) {
            int keep = 1;
            for (int a = 1; a < n; ++a) if (__expf(cv[a] - mx) >= min_p) keep = a + 1; else break;
            n = keep;
        }
        int choice = 0;
        if (temp > 0.f && n > 1) {
            float z = 0.f;
            for (int a = 0; a < n; ++a) z += __expf((cv[a] - mx) / temp);
            const uint32_t r = hash_u32(seed ^ hash_u32(ctr + 0x85ebca6bu));
            const float u = (float)(r >> 8) * (1.0f / 16777216.0f) * z;
            float cum = 0.f;
            choice = n - 1;
            for (int a = 0; a < n; ++a) {
                cum += __expf((cv[a] - mx) / temp);
                if (u < cum) { choice = a; break; }
            }
        }
        *out_slot = ci[choice];
    }
}

__global__ void __launch_bounds__(1024) sample_kernel(const float* __restrict__ logits, int V, int ld, int top_k, float top_p,
                                                      float min_p, float temp, uint32_t seed, const int* __restrict__ step_ctr,
                                                      int* __restrict__ out) {
    pdl_wait();
    pdl_launch_dependents();
    sample_row(logits + (size_t)blockIdx.x * ld, V, top_k, top_p, min_p, temp, seed,
               (step_ctr ? (uint32_t)step_ctr[0] : 0u) * 0x9E3779B9u + blockIdx.x, out + blockIdx.x);
}

// Sampler parameters read from device memory: one record per sequence slot, picked through the batch-row -> slot map the
// decode step already keeps on the device.  One captured graph therefore serves every request (nothing about a request is
// baked into the launch), greedy and sampled rows can share a batch, and the random stream is per request: the record's own
// draw counter starts at zero when the host (re)writes the record and advances here, so the same seed and prompt give the
// same tokens whatever the server did before.  ref: per-slot `common_sampler` with its own RNG (common/sampling.cpp:179-311).
__global__ void __launch_bounds__(1024) sample_dev_kernel(const float* __restrict__ logits, int V, int ld, SampleParams* __restrict__ params,
                                                          const int* __restrict__ row_slot, int* __restrict__ out) {
    pdl_wait();
    pdl_launch_dependents();
    SampleParams* pp = params + (row_slot ? row_slot[blockIdx.x] : (int)blockIdx.x);
    const SampleParams p = *pp;
    const float* row = logits + (size_t)blockIdx.x * ld;
    if (p.temp <= 0.f) {
        argmax_row(row, V, out + blockIdx.x);
        return;
    }
    __syncthreads();   // every thread holds its copy of the record before thread 0 advances the counter
    if (threadIdx.x == 0) pp->ctr = p.ctr + 1;
    sample_row(row, V, p.top_k, p.top_p, p.min_p, p.temp, p.seed, p.ctr * 0x9E3779B9u, out + blockIdx.x);
}

int sample_dev(const float* logits, int R, int V, int ld, SampleParams* params, const int* row_slot, int* out, cudaStream_t s) {
    DLP_REQUIRE(R > 0 && V > 0 && params != nullptr, "sample_dev: bad argument");
    DLP_CHECK_CUDA(launch_pdl(sample_dev_kernel, dim3(R), dim3(1024), 0, s, logits, V, ld, params, row_slot, out));
    return 0;
}

int sample(const float* logits, int R, int V, int ld, int top_k, float top_p, float min_p, float temp, uint32_t seed,
           const int* step_ctr, int* out, cudaStream_t s) {
    DLP_REQUIRE(R > 0 && V > 0, "sample: bad shape");
    if (temp <= 0.f) return argmax(logits, R, V, ld, out, s);
    DLP_CHECK_CUDA(launch_pdl(sample_kernel, dim3(R), dim3(1024), 0, s, logits, V, ld, top_k, top_p, min_p, temp, seed, step_ctr, out));
    return 0;
}

// cast helpers -------------------------------------------------------------------------------
__global__ void f32_to_f16_kernel(const float* __restrict__ in, __half* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2half_rn(in[i]);
}
int f32_to_f16(const float* in, __half* out, size_t n, cudaStream_t s) {
    f32_to_f16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, out, n);
    DLP_LAUNCH_CHECK();
    return 0;
}

}  // namespace dlp
