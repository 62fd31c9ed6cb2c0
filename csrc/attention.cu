// Paged-KV attention for decode / small query counts: split-KV flash-decoding + LSE combine.
//
// Reference counterparts: `flash_attn_ext_vec` (llama.cpp/ggml/src/ggml-cuda/fattn-vec.cuh) with
// `parallel_blocks` split-KV and `flash_attn_combine_results` (fattn-common.cuh:721-772).  The reference
// cache is a contiguous ring per stream; here K/V live in fixed-size pages
// [n_pages][Hkv][page_size][D] addressed through a block table, so sequences grow without copies and
// TP ranks own disjoint KV heads.  One CTA = (split, kv head [x q-head group], query); the G <= 4 query
// heads that share a KV head are processed together so each K/V byte is read once per group.
#include "kernels.cuh"
#include <cooperative_groups.h>

namespace dlp {

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
    const uint32_t s = smem_u32(smem);
    const int sz = pred ? 16 : 0;  // src-size 0 => zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, int G>
__global__ void __launch_bounds__(128) attn_decode_kernel(const AttnParams p) {
    constexpr int DL = D / 8;  // dims per lane (8 lanes share one token)
    static_assert(DL % 8 == 0, "head dim must be a multiple of 64");
    const int split = blockIdx.x;
    const int groups_per_kv = (p.Hq / p.Hkv) / G;
    const int kvh = blockIdx.y / groups_per_kv;
    const int hq0 = kvh * (p.Hq / p.Hkv) + (blockIdx.y % groups_per_kv) * G;
    const int qi = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane >> 3, sub = lane & 7;

    trace_event(p.trace, TR_ATTN | TR_BEGIN);
    pdl_wait();
    trace_event(p.trace, TR_ATTN | TR_WAITED);
    const int len = p.kv_len[qi];
    // effective split count grows with the context (one split per 64 keys): a short context is handled by a single CTA
    // that writes the final result itself; the launch shape (n_splits) stays fixed so the step can live in a CUDA graph
    const int eff = min(p.n_splits, max(1, (len + 63) / 64));
    if (split >= eff) return;
    const int per = (len + eff - 1) / eff;
    const int begin = split * per, end = min(len, begin + per);
    const int* bt = p.block_table + (size_t)(p.bt_row ? p.bt_row[qi] : qi) * p.max_pages;

    float q[G][DL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const __half* qp = p.q + ((size_t)qi * p.Hq + hq0 + g) * D + sub * DL;
#pragma unroll
        for (int i = 0; i < DL; i += 8) {
            const uint4 v = *reinterpret_cast<const uint4*>(qp + i);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = h2_to_f2(w[j]);
                q[g][i + 2 * j] = f.x;
                q[g][i + 2 * j + 1] = f.y;
            }
        }
    }
    float m[G], l[G], o[G][DL];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m[g] = -INFINITY;
        l[g] = 0.f;
#pragma unroll
        for (int i = 0; i < DL; ++i) o[g][i] = 0.f;
    }

    // warp-uniform loop bound: all 32 lanes take part in every shuffle; groups past `end` re-read the last valid
    // token and discard the result.  The raw K/V rows of the *next* iteration are requested before the current one is
    // reduced (the page lookup -> K/V load chain is pure latency otherwise).
    constexpr int NV = DL / 8;   // 16-byte loads per lane and tensor
    uint4 kr[NV], vr[NV], kn[NV], vn[NV];
    auto fetch = [&](int tok0, uint4 (&kd)[NV], uint4 (&vd)[NV]) {
        const int tok = min(tok0 + grp, end - 1);
        const int page = bt[tok / p.page_size], off = tok % p.page_size;
        const size_t base = (((size_t)page * p.Hkv + kvh) * p.page_size + off) * D + sub * DL;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            kd[i] = ldg_nc_v4(p.kc + base + i * 8);
            vd[i] = ldg_nc_v4(p.vc + base + i * 8);
        }
    };
    const int t_first = begin + warp * 4;
    if (t_first < end) fetch(t_first, kr, vr);
    for (int tok0 = t_first; tok0 < end; tok0 += 16) {
        const bool valid = tok0 + grp < end;
        const bool more = tok0 + 16 < end;
        if (more) fetch(tok0 + 16, kn, vn);
        float kf[DL], vf[DL];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const uint32_t kw[4] = {kr[i].x, kr[i].y, kr[i].z, kr[i].w}, vw[4] = {vr[i].x, vr[i].y, vr[i].z, vr[i].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 a = h2_to_f2(kw[j]), b = h2_to_f2(vw[j]);
                kf[i * 8 + 2 * j] = a.x; kf[i * 8 + 2 * j + 1] = a.y;
                vf[i * 8 + 2 * j] = b.x; vf[i * 8 + 2 * j + 1] = b.y;
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < DL; ++i) s = fmaf(q[g][i], kf[i], s);
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            if (!valid) continue;
            const float mn = fmaxf(m[g], s);
            const float alpha = __expf(m[g] - mn), pr = __expf(s - mn);
            m[g] = mn;
            l[g] = l[g] * alpha + pr;
#pragma unroll
            for (int i = 0; i < DL; ++i) o[g][i] = fmaf(pr, vf[i], o[g][i] * alpha);
        }
        if (more) {
#pragma unroll
            for (int i = 0; i < NV; ++i) { kr[i] = kn[i]; vr[i] = vn[i]; }
        }
    }
    trace_event(p.trace, TR_ATTN | TR_MAIN_DONE);
    pdl_launch_dependents();
    // merge the 4 lane groups of the warp, then the 4 warps of the CTA
    __shared__ float sm_m[4][G], sm_l[4][G];
    __shared__ float sm_o[4][G][D];
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
        for (int off = 8; off <= 16; off <<= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[g], off);
            const float l2 = __shfl_xor_sync(0xffffffffu, l[g], off);
            const float mn = fmaxf(m[g], m2);
            const float a1 = (m[g] == -INFINITY) ? 0.f : __expf(m[g] - mn);
            const float a2 = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
#pragma unroll
            for (int i = 0; i < DL; ++i) {
                const float o2 = __shfl_xor_sync(0xffffffffu, o[g][i], off);
                o[g][i] = o[g][i] * a1 + o2 * a2;
            }
            l[g] = l[g] * a1 + l2 * a2;
            m[g] = mn;
        }
        if (grp == 0) {
#pragma unroll
            for (int i = 0; i < DL; ++i) sm_o[warp][g][sub * DL + i] = o[g][i];
            if (sub == 0) { sm_m[warp][g] = m[g]; sm_l[warp][g] = l[g]; }
        }
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
        const int g = idx / D, d = idx % D;
        float mn = -INFINITY;
#pragma unroll
        for (int w = 0; w < 4; ++w) mn = fmaxf(mn, sm_m[w][g]);
        float acc = 0.f, ls = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const float a = (sm_m[w][g] == -INFINITY) ? 0.f : __expf(sm_m[w][g] - mn);
            acc += sm_o[w][g][d] * a;
            ls += sm_l[w][g] * a;
        }
        const int hq = hq0 + g;
        if (eff == 1) {
            p.out[((size_t)qi * p.Hq + hq) * D + d] = __float2half_rn(ls > 0.f ? acc / ls : 0.f);
        } else {
            const size_t pi = ((size_t)qi * p.Hq + hq) * p.n_splits + split;
            p.part_o[pi * D + d] = acc;
            if (d == 0) { p.part_ml[pi * 2] = mn; p.part_ml[pi * 2 + 1] = ls; }
        }
    }
    trace_event(p.trace, TR_ATTN | TR_END);       // (CTA 0 is rarely the combining CTA: its END marks the partial write)
    if (eff == 1 || p.counters == nullptr) return;
    // fused LSE combine: the last split CTA of this (query, kv-head group) to finish merges all partials - no second launch
    __shared__ int is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        int* ctr = p.counters + (size_t)qi * gridDim.y + blockIdx.y;
        const int prev = atomicAdd(ctr, 1);
        is_last = (prev == eff - 1);
        if (is_last) *ctr = 0;   // ready for the next launch
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
        const int g = idx / D, d = idx % D;
        const size_t pi0 = ((size_t)qi * p.Hq + hq0 + g) * p.n_splits;
        float mn = -INFINITY;
        for (int sp = 0; sp < eff; ++sp) mn = fmaxf(mn, __ldcg(p.part_ml + (pi0 + sp) * 2));
        float acc = 0.f, ls = 0.f;
        for (int sp = 0; sp < eff; ++sp) {
            const float ms = __ldcg(p.part_ml + (pi0 + sp) * 2);
            const float a = (ms == -INFINITY) ? 0.f : __expf(ms - mn);
            acc += __ldcg(p.part_o + (pi0 + sp) * D + d) * a;
            ls += __ldcg(p.part_ml + (pi0 + sp) * 2 + 1) * a;
        }
        p.out[((size_t)qi * p.Hq + hq0 + g) * D + d] = __float2half_rn(ls > 0.f ? acc / ls : 0.f);
    }
}

// ---------------------------------------------------------------------------------------------
// Decode step in ONE launch: RoPE on Q and on the new token's K, paged write of that token's K/V, split-KV attention over the
// cache, and the split combine through distributed shared memory.
//
//   * The splits of one (query, KV-head group) form a thread-block CLUSTER (grid.x = cluster size = n_splits <= 16): partial
//     (o, m, l) triples are stored straight into the leader CTA's shared memory (st.shared::cluster through
//     cluster.map_shared_rank) and merged after one cluster barrier - no global workspace, no __threadfence + atomic ticket,
//     no L2 round trips, no second kernel.
//   * Q arrives un-rotated as the fp32 output row of the fused QKV mat-vec; every lane rotates the 16 dims it owns (the
//     half-split "neox" partner dim sits 4 lanes away: one shuffle).  The new token (position kv_len-1) is not in the cache
//     yet: the CTA that owns the tail of the key range rotates K, writes K/V to the page and folds the token in from
//     registers, so the separate RoPE + KV-write launch of the unfused path (rope_kv_kernel) disappears from the decode step.
// Reference counterparts: ggml_rope_ext + ggml_set_rows (K/V cache write) + flash_attn_ext_vec + flash_attn_combine_results -
// four graph nodes / launches per layer in `llama.cpp/src/models/llama.cpp:60-110`, `ggml-cuda/fattn-vec.cuh`.
template <int D, int G>
__global__ void __launch_bounds__(128) attn_decode_rope_kernel(const AttnParams p, const float* __restrict__ qkv,
                                                               const float* __restrict__ inv_freq,
                                                               const float2* __restrict__ rope_cs, __half* __restrict__ kc_w,
                                                               __half* __restrict__ vc_w, int neox, float q_scale) {
    namespace cg = cooperative_groups;
    constexpr int DL = D / 8;     // dims per lane (8 lanes share one token)
    constexpr int NV = DL / 8;    // 16-byte loads per lane and tensor
    constexpr int MAXS = 16;
    static_assert(DL % 8 == 0, "head dim must be a multiple of 64");
    cg::cluster_group cluster = cg::this_cluster();
    const int split = blockIdx.x;                      // == rank in the cluster (cluster dims = (gridDim.x, 1, 1))
    const int gqa = p.Hq / p.Hkv;
    const int groups_per_kv = gqa / G;
    const int kvh = blockIdx.y / groups_per_kv;
    const int hq0 = kvh * gqa + (blockIdx.y % groups_per_kv) * G;
    const int qi = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane >> 3, sub = lane & 7;
    const int H = p.Hq + 2 * p.Hkv;

    // (a variant that staged 64-key K/V tiles in 64 KB of dynamic shared memory with cp.async measured slower: the large
    // carve-out keeps the kernel from sharing an SM with the neighbouring mat-vecs and costs the PDL overlap)
    __shared__ float sm_m[4][G], sm_l[4][G];
    __shared__ float sm_o[4][G][D];
    __shared__ float cl_o[MAXS][G][D];                 // leader only: landing zone of the splits' partial outputs
    __shared__ float cl_ml[MAXS][G][2];

    trace_event(p.trace, TR_ATTN | TR_BEGIN);
    // Everything the previous kernel (the QKV mat-vec) did NOT write is fetched before the dependency wait: sequence length,
    // block table and the step's rotation table come from the bookkeeping kernel at the start of the step (at least two launches
    // back, hence complete once this grid may start), the cached K/V rows from earlier steps.  Only `qkv` needs the wait, so the
    // dependent chain length -> page id -> K/V row runs under the predecessor's tail instead of after it.
    const int len = p.kv_len[qi];                      // keys visible to this query, the new token included
    const int n_old = len - 1;                         // ... of which this many are already in the cache
    const int eff = min((int)gridDim.x, max(1, (len + p.keys_per_split - 1) / p.keys_per_split));
    if (eff == 1 && split > 0) return;                 // uniform over the cluster: nobody waits on a barrier in that case
    const bool active = split < eff;
    const int per = (len + eff - 1) / eff;
    const int begin = split * per, end = min(len, begin + per);
    const int end_c = min(end, n_old);                 // cached part of this CTA's key range
    const bool owns_new = active && end == len;
    const int* bt = p.block_table + (size_t)(p.bt_row ? p.bt_row[qi] : qi) * p.max_pages;
    const float* row = qkv + (size_t)qi * H * D;
    uint4 kr[NV], vr[NV], kn[NV], vn[NV];
    auto fetch = [&](int tok0, uint4 (&kd)[NV], uint4 (&vd)[NV]) {
        const int tok = min(tok0 + grp, end_c - 1);
        const int page = bt[tok / p.page_size], off = tok % p.page_size;
        const size_t base = (((size_t)page * p.Hkv + kvh) * p.page_size + off) * D + sub * DL;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            kd[i] = ldg_nc_v4(p.kc + base + i * 8);
            vd[i] = ldg_nc_v4(p.vc + base + i * 8);
        }
    };
    const int t_first = begin + warp * 4;
    if (active && t_first < end_c) fetch(t_first, kr, vr);
    pdl_wait();
    if (p.pdl_early) pdl_launch_dependents();
    trace_event(p.trace, TR_ATTN | TR_WAITED);

    if (active) {

        // rotation terms of the DL dims this lane owns (shared by the G query heads and the new key): the per-step table
        // [nq][D/2] of (cos, sin) written by advance_rope_kernel once for all layers, or computed here
        float cs[DL], sn[DL];
        if (rope_cs != nullptr) {
            const float2* tab = rope_cs + (size_t)qi * (D / 2);
            if (neox) {
#pragma unroll
                for (int i = 0; i < DL; i += 2) {
                    const float4 v = *reinterpret_cast<const float4*>(tab + (sub & 3) * DL + i);
                    cs[i] = v.x; sn[i] = v.y; cs[i + 1] = v.z; sn[i + 1] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < DL; i += 4) {
                    const float4 v = *reinterpret_cast<const float4*>(tab + sub * (DL / 2) + i / 2);
                    cs[i] = cs[i + 1] = v.x; sn[i] = sn[i + 1] = v.y;
                    cs[i + 2] = cs[i + 3] = v.z; sn[i + 2] = sn[i + 3] = v.w;
                }
            }
        } else {
            const float pos = (float)n_old;
#pragma unroll
            for (int i = 0; i < DL; ++i) {
                const int d = sub * DL + i;
                const int fi = neox ? (d % (D / 2)) : (d >> 1);
                sincosf(pos * inv_freq[fi], &sn[i], &cs[i]);
            }
        }
        auto rotate = [&](float (&x)[DL]) {
            if (neox) {
#pragma unroll
                for (int i = 0; i < DL; ++i) {
                    const float other = __shfl_xor_sync(0xffffffffu, x[i], 4);     // dim d +- D/2 lives in lane sub ^ 4
                    x[i] = sub < 4 ? x[i] * cs[i] - other * sn[i] : x[i] * cs[i] + other * sn[i];
                }
            } else {
#pragma unroll
                for (int i = 0; i < DL; i += 2) {
                    const float a = x[i], b = x[i + 1];
                    x[i] = a * cs[i] - b * sn[i];
                    x[i + 1] = a * sn[i] + b * cs[i];
                }
            }
        };
        auto load_row = [&](const float* src, float (&x)[DL]) {
#pragma unroll
            for (int i = 0; i < DL; i += 4) {
                const float4 v = *reinterpret_cast<const float4*>(src + sub * DL + i);
                x[i] = v.x; x[i + 1] = v.y; x[i + 2] = v.z; x[i + 3] = v.w;
            }
        };
        float q[G][DL];
        float kf_new[DL], vf_new[DL];
        const bool do_new = owns_new && warp == 0;     // the new token's K/V rows travel with the Q rows: one L2 round trip
        if (do_new) {
            load_row(row + (size_t)(p.Hq + kvh) * D, kf_new);
            load_row(row + (size_t)(p.Hq + p.Hkv + kvh) * D, vf_new);
        }
#pragma unroll
        for (int g = 0; g < G; ++g) load_row(row + (size_t)(hq0 + g) * D, q[g]);
#pragma unroll
        for (int g = 0; g < G; ++g) {
            rotate(q[g]);
#pragma unroll
            for (int i = 0; i < DL; ++i) q[g][i] *= q_scale;
        }
        float m[G], l[G], o[G][DL];
#pragma unroll
        for (int g = 0; g < G; ++g) {
            m[g] = -INFINITY;
            l[g] = 0.f;
#pragma unroll
            for (int i = 0; i < DL; ++i) o[g][i] = 0.f;
        }
        auto accumulate = [&](const float (&kf)[DL], const float (&vf)[DL], bool valid) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float sc = 0.f;
#pragma unroll
                for (int i = 0; i < DL; ++i) sc = fmaf(q[g][i], kf[i], sc);
                sc += __shfl_xor_sync(0xffffffffu, sc, 1);
                sc += __shfl_xor_sync(0xffffffffu, sc, 2);
                sc += __shfl_xor_sync(0xffffffffu, sc, 4);
                if (!valid) continue;
                const float mn = fmaxf(m[g], sc);
                const float alpha = __expf(m[g] - mn), pr = __expf(sc - mn);
                m[g] = mn;
                l[g] = l[g] * alpha + pr;
#pragma unroll
                for (int i = 0; i < DL; ++i) o[g][i] = fmaf(pr, vf[i], o[g][i] * alpha);
            }
        };

        if (do_new) {
            // the new token first (online softmax does not care about order): rotate K, round K/V to the cache precision,
            // write the page slot, fold it in (lane group 0)
            rotate(kf_new);
            uint32_t kh[DL / 2], vh[DL / 2];
#pragma unroll
            for (int i = 0; i < DL; i += 2) {
                const __half2 a = __floats2half2_rn(kf_new[i], kf_new[i + 1]), b = __floats2half2_rn(vf_new[i], vf_new[i + 1]);
                kh[i / 2] = *reinterpret_cast<const uint32_t*>(&a);
                vh[i / 2] = *reinterpret_cast<const uint32_t*>(&b);
                const float2 af = __half22float2(a), bf = __half22float2(b);
                kf_new[i] = af.x; kf_new[i + 1] = af.y; vf_new[i] = bf.x; vf_new[i + 1] = bf.y;
            }
            if (grp == 0 && blockIdx.y % groups_per_kv == 0) {
                const int page = bt[n_old / p.page_size], off = n_old % p.page_size;
                const size_t base = (((size_t)page * p.Hkv + kvh) * p.page_size + off) * D + sub * DL;
#pragma unroll
                for (int i = 0; i < NV; ++i) {
                    *reinterpret_cast<uint4*>(kc_w + base + i * 8) = make_uint4(kh[4 * i], kh[4 * i + 1], kh[4 * i + 2], kh[4 * i + 3]);
                    *reinterpret_cast<uint4*>(vc_w + base + i * 8) = make_uint4(vh[4 * i], vh[4 * i + 1], vh[4 * i + 2], vh[4 * i + 3]);
                }
            }
            accumulate(kf_new, vf_new, grp == 0);
        }
        for (int tok0 = t_first; tok0 < end_c; tok0 += 16) {
            const bool valid = tok0 + grp < end_c;
            const bool more = tok0 + 16 < end_c;
            if (more) fetch(tok0 + 16, kn, vn);
            float kf[DL], vf[DL];
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const uint32_t kw[4] = {kr[i].x, kr[i].y, kr[i].z, kr[i].w}, vw[4] = {vr[i].x, vr[i].y, vr[i].z, vr[i].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 a = h2_to_f2(kw[j]), b = h2_to_f2(vw[j]);
                    kf[i * 8 + 2 * j] = a.x; kf[i * 8 + 2 * j + 1] = a.y;
                    vf[i * 8 + 2 * j] = b.x; vf[i * 8 + 2 * j + 1] = b.y;
                }
            }
            accumulate(kf, vf, valid);
            if (more) {
#pragma unroll
                for (int i = 0; i < NV; ++i) { kr[i] = kn[i]; vr[i] = vn[i]; }
            }
        }
        trace_event(p.trace, TR_ATTN | TR_MAIN_DONE);
        pdl_launch_dependents();
        // merge the 4 lane groups of the warp, then the 4 warps of the CTA
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int off = 8; off <= 16; off <<= 1) {
                const float m2 = __shfl_xor_sync(0xffffffffu, m[g], off);
                const float l2 = __shfl_xor_sync(0xffffffffu, l[g], off);
                const float mn = fmaxf(m[g], m2);
                const float a1 = (m[g] == -INFINITY) ? 0.f : __expf(m[g] - mn);
                const float a2 = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
#pragma unroll
                for (int i = 0; i < DL; ++i) {
                    const float o2 = __shfl_xor_sync(0xffffffffu, o[g][i], off);
                    o[g][i] = o[g][i] * a1 + o2 * a2;
                }
                l[g] = l[g] * a1 + l2 * a2;
                m[g] = mn;
            }
            if (grp == 0) {
#pragma unroll
                for (int i = 0; i < DL; ++i) sm_o[warp][g][sub * DL + i] = o[g][i];
                if (sub == 0) { sm_m[warp][g] = m[g]; sm_l[warp][g] = l[g]; }
            }
        }
        __syncthreads();
        float* dst_o = eff == 1 ? nullptr : cluster.map_shared_rank(&cl_o[0][0][0], 0);
        float* dst_ml = eff == 1 ? nullptr : cluster.map_shared_rank(&cl_ml[0][0][0], 0);
        for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
            const int g = idx / D, d = idx % D;
            float mn = -INFINITY;
#pragma unroll
            for (int w = 0; w < 4; ++w) mn = fmaxf(mn, sm_m[w][g]);
            float acc = 0.f, ls = 0.f;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const float a = (sm_m[w][g] == -INFINITY) ? 0.f : __expf(sm_m[w][g] - mn);
                acc += sm_o[w][g][d] * a;
                ls += sm_l[w][g] * a;
            }
            if (eff == 1) {
                p.out[((size_t)qi * p.Hq + hq0 + g) * D + d] = __float2half_rn(ls > 0.f ? acc / ls : 0.f);
            } else {
                dst_o[((size_t)split * G + g) * D + d] = acc;
                if (d == 0) { dst_ml[(split * G + g) * 2] = mn; dst_ml[(split * G + g) * 2 + 1] = ls; }
            }
        }
    } else {
        pdl_launch_dependents();
    }
    if (eff == 1) { trace_event(p.trace, TR_ATTN | TR_END); return; }
    cluster.sync();                                    // release/acquire at cluster scope: the partials are visible in rank 0
    if (split != 0) return;
    for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
        const int g = idx / D, d = idx % D;
        float mn = -INFINITY;
        for (int sp = 0; sp < eff; ++sp) mn = fmaxf(mn, cl_ml[sp][g][0]);
        float acc = 0.f, ls = 0.f;
        for (int sp = 0; sp < eff; ++sp) {
            const float ms = cl_ml[sp][g][0];
            const float a = (ms == -INFINITY) ? 0.f : __expf(ms - mn);
            acc += cl_o[sp][g][d] * a;
            ls += cl_ml[sp][g][1] * a;
        }
        p.out[((size_t)qi * p.Hq + hq0 + g) * D + d] = __float2half_rn(ls > 0.f ? acc / ls : 0.f);
    }
    trace_event(p.trace, TR_ATTN | TR_END);
}

template <int D, int G>
static int launch_attn_rope_g(const AttnParams& p, const float* qkv, const float* inv_freq, const float2* rope_cs, __half* kc_w,
                              __half* vc_w, int neox, float q_scale, cudaStream_t s) {
    auto kern = attn_decode_rope_kernel<D, G>;
    if (first_use_on_device((const void*)kern))   // cluster sizes above 8 are "non-portable": opt in once per device
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    const int gqa = p.Hq / p.Hkv;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.Hkv * (gqa / G), p.nq);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = p.n_splits;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    DLP_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale));
    return 0;
}

template <int D>
static int launch_attn_rope(const AttnParams& p, const float* qkv, const float* inv_freq, const float2* rope_cs, __half* kc_w,
                            __half* vc_w, int neox, float q_scale, cudaStream_t s) {
    const int gqa = p.Hq / p.Hkv;
    if (gqa % 4 == 0) return launch_attn_rope_g<D, 4>(p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale, s);
    if (gqa % 2 == 0) return launch_attn_rope_g<D, 2>(p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale, s);
    return launch_attn_rope_g<D, 1>(p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale, s);
}

int attn_decode_rope(const AttnParams& p_in, const float* qkv, const float* inv_freq, const float2* rope_cs, __half* kc_w,
                     __half* vc_w, int D, int neox, float q_scale, cudaStream_t s) {
    AttnParams p = p_in;
    p.trace = trace_buffer();
    static int early = -1, kps = 0;
    if (early < 0) {
        const char* e = getenv("DLP_PDL_EARLY");
        early = (e && e[0] == '0') ? 0 : 1;
        const char* k = getenv("DLP_ATTN_KEYS_PER_SPLIT");
        kps = k ? atoi(k) : 32;
        if (kps < 16) kps = 16;
    }
    p.pdl_early = early;
    p.keys_per_split = kps;
    DLP_REQUIRE(inv_freq != nullptr || rope_cs != nullptr, "attn_decode_rope: need inv_freq or a (cos, sin) table");
    DLP_REQUIRE(p.nq > 0 && p.Hq % p.Hkv == 0, "attn_decode_rope: bad head configuration Hq=%d Hkv=%d", p.Hq, p.Hkv);
    DLP_REQUIRE(p.n_splits >= 1 && p.n_splits <= 16, "attn_decode_rope: cluster size %d outside 1..16", p.n_splits);
    if (D == 64) return launch_attn_rope<64>(p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale, s);
    if (D == 128) return launch_attn_rope<128>(p, qkv, inv_freq, rope_cs, kc_w, vc_w, neox, q_scale, s);
    DLP_REQUIRE(false, "attn_decode_rope: head dim %d unsupported (64, 128)", D);
}

template <int D>
__global__ void attn_combine_kernel(const AttnParams p) {
    const int hq = blockIdx.x, qi = blockIdx.y;
    const size_t pi0 = ((size_t)qi * p.Hq + hq) * p.n_splits;
    pdl_wait();
    pdl_launch_dependents();
    const int len = p.kv_len[qi];
    const int eff = min(p.n_splits, max(1, (len + 63) / 64));
    if (eff == 1) return;   // the single split already wrote the final output
    float mn = -INFINITY;
    for (int s = 0; s < eff; ++s) mn = fmaxf(mn, p.part_ml[(pi0 + s) * 2]);
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float acc = 0.f, ls = 0.f;
        for (int s = 0; s < eff; ++s) {
            const float ms = p.part_ml[(pi0 + s) * 2];
            const float a = (ms == -INFINITY) ? 0.f : __expf(ms - mn);
            acc += p.part_o[(pi0 + s) * D + d] * a;
            ls += p.part_ml[(pi0 + s) * 2 + 1] * a;
        }
        p.out[((size_t)qi * p.Hq + hq) * D + d] = __float2half_rn(ls > 0.f ? acc / ls : 0.f);
    }
}

template <int D>
static int launch_attn(const AttnParams& p, cudaStream_t s) {
    const int gqa = p.Hq / p.Hkv;
    const int G = gqa % 4 == 0 ? 4 : (gqa % 2 == 0 ? 2 : 1);
    dim3 grid(p.n_splits, p.Hkv * (gqa / G), p.nq);
    if (G == 4) DLP_CHECK_CUDA(launch_pdl(attn_decode_kernel<D, 4>, grid, dim3(128), 0, s, p));
    else if (G == 2) DLP_CHECK_CUDA(launch_pdl(attn_decode_kernel<D, 2>, grid, dim3(128), 0, s, p));
    else DLP_CHECK_CUDA(launch_pdl(attn_decode_kernel<D, 1>, grid, dim3(128), 0, s, p));
    if (p.n_splits > 1 && p.counters == nullptr)
        DLP_CHECK_CUDA(launch_pdl(attn_combine_kernel<D>, dim3(p.Hq, p.nq), dim3(D), 0, s, p));
    return 0;
}

int attn_paged(const AttnParams& p_in, int D, cudaStream_t s) {
    AttnParams p = p_in;
    p.trace = trace_buffer();
    DLP_REQUIRE(p.nq > 0 && p.Hq % p.Hkv == 0, "attn: bad head configuration Hq=%d Hkv=%d", p.Hq, p.Hkv);
    DLP_REQUIRE(p.n_splits >= 1 && (p.n_splits == 1 || (p.part_o && p.part_ml)), "attn: split-KV needs a workspace");
    if (D == 64) return launch_attn<64>(p, s);
    if (D == 128) return launch_attn<128>(p, s);
    DLP_REQUIRE(false, "attn: head dim %d unsupported (64, 128)", D);
}

}  // namespace dlp

// =================================================================================================
// Causal flash-prefill over the paged cache (one sequence chunk).
//
// Reference counterpart: `flash_attn_ext_f16` (ggml-cuda/fattn-mma-f16.cuh).  FlashAttention-2 schedule:
// one CTA = 64 query rows of one head (4 warps x 16 rows), K/V streamed page-by-page through shared
// memory with cp.async, scores and P.V on the legacy tensor-core path (mma.sync m16n8k16, fp32
// accumulate), online softmax in registers.  Attention is ~1% of prefill FLOPs at the benchmark shapes
// (the linear layers run on tcgen05), so this kernel favours simplicity; a TMEM-resident variant is the
// planned upgrade.
namespace dlp {


__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(smem)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(smem)));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

template <int D>
__global__ void __launch_bounds__(128) attn_prefill_kernel(const __half* __restrict__ q, const __half* __restrict__ kc,
                                                           const __half* __restrict__ vc, const int* __restrict__ bt,
                                                           __half* __restrict__ out, int T, int pos0, int Hq, int Hkv, int PS) {
    constexpr int LD = D + 8;      // padded row (halves) -> conflict-free ldmatrix
    constexpr int BQ = 64, BKV = 64;
    extern __shared__ __align__(16) uint8_t att_smem[];
    __half* sQ = reinterpret_cast<__half*>(att_smem);
    __half* sKV = sQ + BQ * LD;              // two stages of (K tile | V tile): the next tile lands while this one is consumed
    // heaviest query tiles first: under the causal mask the last tile has the most key tiles to walk
    const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * BQ, hq = blockIdx.y, kvh = hq / (Hq / Hkv);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    constexpr int CPR = D / 8;  // 16-byte chunks per row

    for (int i = tid; i < BQ * CPR; i += 128) {
        const int r = i / CPR, c = i % CPR;
        const bool ok = q0 + r < T;
        cp_async16(sQ + r * LD + c * 8, q + ((size_t)(ok ? q0 + r : 0) * Hq + hq) * D + c * 8, ok);
    }
    const int n_kv = pos0 + min(q0 + BQ, T);
    const int n_tiles = (n_kv + BKV - 1) / BKV;

    uint32_t qf[D / 16][4];
    float o[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    constexpr float LOG2E = 1.4426950408889634f;

    auto load_kv = [&](int kt) {
        __half* dK = sKV + (kt & 1) * 2 * BKV * LD;
        __half* dV = dK + BKV * LD;
        for (int i = tid; i < BKV * CPR; i += 128) {
            const int r = i / CPR, c = i % CPR;
            const int tok = kt * BKV + r;
            const bool ok = tok < n_kv;
            const int tk = ok ? tok : 0;
            const size_t src = (((size_t)bt[tk / PS] * Hkv + kvh) * PS + tk % PS) * D + c * 8;
            cp_async16(dK + r * LD + c * 8, kc + src, ok);
            cp_async16(dV + r * LD + c * 8, vc + src, ok);
        }
        cp_async_commit();
    };
    load_kv(0);                                // (commits the Q rows requested above with it)
    for (int kt = 0; kt < n_tiles; ++kt) {
        __syncthreads();                       // tile kt-1 fully consumed: its stage may be refilled
        if (kt + 1 < n_tiles) { load_kv(kt + 1); cp_async_wait_group<1>(); } else { cp_async_wait_group<0>(); }
        __syncthreads();                       // tile kt (and Q) visible to every warp
        const __half* sK = sKV + (kt & 1) * 2 * BKV * LD;
        const __half* sV = sK + BKV * LD;
        if (kt == 0) {
#pragma unroll
            for (int ks = 0; ks < D / 16; ++ks)
                ldmatrix_x4(qf[ks], sQ + (warp * 16 + (lane & 15)) * LD + ks * 16 + (lane >> 4) * 8);
        }
        // ---- S = Q K^T (16 x 64 per warp) ----
        float s[BKV / 8][4];
#pragma unroll
        for (int i = 0; i < BKV / 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
            for (int np = 0; np < BKV / 16; ++np) {
                uint32_t kb[4];
                const int mat = lane >> 3, r = lane & 7;
                ldmatrix_x4(kb, sK + (np * 16 + (mat >> 1) * 8 + r) * LD + ks * 16 + (mat & 1) * 8);
                mma_16816(s[np * 2], qf[ks], kb[0], kb[1]);
                mma_16816(s[np * 2 + 1], qf[ks], kb[2], kb[3]);
            }
        }
        // ---- causal mask + online softmax (rows g and g+8 of this warp's 16) ----
        const int qpos0 = pos0 + q0 + warp * 16 + g;
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < BKV / 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int kpos = kt * BKV + nt * 8 + t4 * 2 + (e & 1);
                const int qpos = qpos0 + (e >> 1) * 8;
                if (kpos > qpos) s[nt][e] = -INFINITY;
                mx[e >> 1] = fmaxf(mx[e >> 1], s[nt][e]);
            }
        }
        float alpha[2], msc[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
            const float mn = fmaxf(m_run[h], mx[h]);
            alpha[h] = (m_run[h] == -INFINITY) ? 0.f : exp2f((m_run[h] - mn) * LOG2E);
            msc[h] = (mn == -INFINITY) ? 0.f : mn * LOG2E;
            m_run[h] = mn;
        }
        float rs[2] = {0.f, 0.f};
#pragma unroll
        for (int nt = 0; nt < BKV / 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float pv = exp2f(s[nt][e] * LOG2E - msc[e >> 1]);  // exp2(-inf) = 0 for masked keys
                s[nt][e] = pv;
                rs[e >> 1] += pv;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) l_run[h] = l_run[h] * alpha[h] + rs[h];
#pragma unroll
        for (int i = 0; i < D / 8; ++i) {
            o[i][0] *= alpha[0]; o[i][1] *= alpha[0];
            o[i][2] *= alpha[1]; o[i][3] *= alpha[1];
        }
        // ---- O += P V ----
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
            uint32_t pa[4];
            pa[0] = pack_h2(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack_h2(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack_h2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack_h2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int dp = 0; dp < D / 16; ++dp) {
                uint32_t vb[4];
                const int mat = lane >> 3, r = lane & 7;
                ldmatrix_x4_trans(vb, sV + (kk * 16 + (mat & 1) * 8 + r) * LD + dp * 16 + (mat >> 1) * 8);
                mma_16816(o[dp * 2], pa, vb[0], vb[1]);
                mma_16816(o[dp * 2 + 1], pa, vb[2], vb[3]);
            }
        }
    }
    // ---- finalise: row sums live per-quad; reduce then normalise and store ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
        l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int row = q0 + warp * 16 + g + h * 8;
        if (row < T) {
            const float inv = l_run[h] > 0.f ? 1.f / l_run[h] : 0.f;
            __half* op = out + ((size_t)row * Hq + hq) * D;
#pragma unroll
            for (int i = 0; i < D / 8; ++i) {
                const uint32_t v = pack_h2(o[i][h * 2] * inv, o[i][h * 2 + 1] * inv);
                *reinterpret_cast<uint32_t*>(op + i * 8 + t4 * 2) = v;
            }
        }
    }
}

int attn_prefill(const __half* q, const __half* kc, const __half* vc, const int* bt_row, __half* out, int T, int pos0,
                 int Hq, int Hkv, int D, int page_size, cudaStream_t s) {
    DLP_REQUIRE(T > 0 && Hq % Hkv == 0, "attn_prefill: bad shape");
    dim3 grid((T + 63) / 64, Hq);
    if (D == 128) {
        const int smem = 5 * 64 * (128 + 8) * 2;   // Q + 2 stages x (K, V)
        if (first_use_on_device((const void*)attn_prefill_kernel<128>))
            DLP_CHECK_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attn_prefill_kernel<128><<<grid, 128, smem, s>>>(q, kc, vc, bt_row, out, T, pos0, Hq, Hkv, page_size);
    } else if (D == 64) {
        const int smem = 5 * 64 * (64 + 8) * 2;
        if (first_use_on_device((const void*)attn_prefill_kernel<64>))
            DLP_CHECK_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attn_prefill_kernel<64><<<grid, 128, smem, s>>>(q, kc, vc, bt_row, out, T, pos0, Hq, Hkv, page_size);
    } else {
        DLP_REQUIRE(false, "attn_prefill: head dim %d unsupported (64, 128)", D);
    }
    DLP_LAUNCH_CHECK();
    return 0;
}

}  // namespace dlp

// =================================================================================================
// Split-KV paged attention without a caller-provided workspace.
//
// Some hosts of this library are not in a position to hand us scratch memory: a serving process that carved
// the whole HBM into weights + KV pages at start-up and then froze its allocator, a plug-in running under a
// "no device allocation after init" rule, or code sitting between cudaStreamBeginCapture / EndCapture where
// cudaMalloc is illegal.  For them the split-KV partial results (the only scratch `attn_paged` needs) live in
// a fixed array inside this module's device image.  Costs and rules, all of which the header repeats:
//   * the array is ATTN_STATIC_SCRATCH_FLOATS floats (4 MiB) on every GPU that loads the library, until exit;
//   * a launch needs nq * Hq * n_splits * (D + 2) floats and is refused with an error beyond that;
//   * every call on a device shares the same array, so calls that overlap on the GPU (different streams, or
//     graphs replayed concurrently) race - serialise them;
//   * nothing has to start at zero: the partial kernel writes every (query, head, split) slot - empty splits
//     write m = -inf, l = 0 - before the combine kernel reads it, so stale contents of an earlier call are
//     never observed.
// In host code the array's name is only a handle: the device address is obtained once per device with
// cudaGetSymbolAddress (device current, return code checked) on the first call there, which therefore must
// happen outside stream capture.
namespace dlp {

constexpr size_t ATTN_STATIC_SCRATCH_FLOATS = size_t(1) << 20;  // keep in sync with DLP_ATTN_STATIC_SCRATCH_FLOATS
static __device__ __align__(256) float g_attn_static_scratch[ATTN_STATIC_SCRATCH_FLOATS];

int attn_paged_static(AttnParams p, int D, cudaStream_t s) {
    DLP_REQUIRE(p.nq > 0 && p.Hq > 0 && p.n_splits >= 1 && D > 0, "attn_paged_static: bad shape");
    const size_t n_part = (size_t)p.nq * p.Hq * p.n_splits;
    const size_t need = p.n_splits > 1 ? n_part * (size_t)(D + 2) : 0;
    DLP_REQUIRE(need <= ATTN_STATIC_SCRATCH_FLOATS,
                "attn_paged_static: nq*Hq*n_splits*(D+2) = %zu floats exceeds the built-in scratch of %zu floats; "
                "lower n_splits / batch or use dlp_attn_paged with your own workspace",
                need, (size_t)ATTN_STATIC_SCRATCH_FLOATS);
    constexpr int MAX_DEV = 64;
    static float* dev_ptr[MAX_DEV] = {nullptr};
    int dev = -1;
    DLP_CHECK_CUDA(cudaGetDevice(&dev));
    DLP_REQUIRE(dev >= 0 && dev < MAX_DEV, "attn_paged_static: device ordinal %d out of range", dev);
    float* ptr = dev_ptr[dev];
    if (ptr == nullptr) {  // once per device, with that device current (racing first calls store the same value)
        DLP_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_attn_static_scratch));
        DLP_REQUIRE(ptr != nullptr, "attn_paged_static: cudaGetSymbolAddress returned a null address");
        dev_ptr[dev] = ptr;
    }
    p.counters = nullptr;                // this entry point keeps the two-kernel combine
    p.part_o = ptr;                      // [nq, Hq, n_splits, D]
    p.part_ml = ptr + n_part * (size_t)D;  // [nq, Hq, n_splits, 2]
    return attn_paged(p, D, s);
}

}  // namespace dlp
