// Blackwell-native causal flash-prefill over the paged KV cache: tcgen05 MMAs for Q.K^T and P.V with S and O in tensor memory,
// K / V pages brought in by TMA.
//
// Reference counterpart: flash_attn_ext_f16 (llama.cpp/ggml/src/ggml-cuda/fattn-mma-f16.cuh:38-104), an FA-2 schedule on
// mma.sync + cp.async with the accumulators in registers - which is also what attention.cu:attn_prefill_kernel is.  This kernel
// is the sm_100a form of the same algorithm:
//
//   CTA = one tile of 128 queries of one query head.  192 threads:
//     warp 0      producer   TMA: the Q tile once, then per 128-key tile the K and V rows page by page (the block table gives the
//                            page ids; a page is a contiguous [page_size x D] block, box {64, page_size} of a 2-D map over the cache)
//     warp 1      MMA        one thread issues  S = Q.K^T  (M = 128 queries, N = 128 keys, K = D) into one of two TMEM S buffers,
//                            and  O += P.V  (M = 128, N = D, K = 128 keys) into the TMEM O accumulator.  V stays in its cache
//                            layout [key][D]: it is the *MN-major* B operand (descriptor: 128-byte swizzle atoms of 8 keys x 64
//                            dims, SBO = 1024 B between key groups, LBO = distance between the two 64-dim halves).
//     warps 2-5   softmax    thread r owns query row r: tcgen05.ld of its S row, causal mask on the diagonal tiles, running
//                            max / sum, P = exp2(..) written as fp16 into shared memory in the K-major swizzled layout the
//                            second MMA reads as its A operand.  O is rescaled in TMEM only when a row maximum actually grew
//                            (warp-uniform test), and normalised + stored as fp16 at the end.
//   QK^T of tile i+1 is issued before P.V of tile i, so the tensor pipe works on the next scores while the softmax warps are
//   busy; K/V stages form a 2-deep ring.
#include <string.h>

#include "kernels.cuh"

namespace dlp {

int make_tmap_2d(CUtensorMap* out, const void* base, int is_bf16, uint64_t rows, uint64_t K, uint64_t ld, uint32_t box_rows);

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
        "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
          "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
          "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
          "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// mbarrier wait that turns a protocol error into a trapped kernel (an error the host sees) instead of a GPU that spins forever
__device__ __forceinline__ void mbar_wait_g(uint64_t* bar, uint32_t parity) {
    for (unsigned spins = 0; !mbar_try_wait(bar, parity); ++spins)
        if (spins > (1u << 24)) __trap();
}

// MN-major operand in the 128-byte-swizzled layout TMA produces for rows of 64 fp16: atoms of 8 rows (1024 B); `lbo_bytes` =
// distance between consecutive 64-element chunks along the MN dimension, SBO = 1024 B between groups of 8 rows along K.
__device__ __forceinline__ uint64_t make_smem_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16 instruction descriptor with an MN-major B operand (bit 16)
__host__ __device__ constexpr uint32_t make_idesc_f16_bmn(int m, int n) { return make_idesc_f16(m, n, 0) | (1u << 16); }

struct FaParams {
    const int* bt_row;      // block-table row of the sequence
    __half* out;            // [T, Hq*D]
    int T, pos0, Hq, Hkv, page_size, max_pages;
};

template <int D>
struct FaCfg {
    static constexpr int BM = 128, BN = 128;
    static constexpr int KB = D / 64;
    static constexpr int Q_BYTES = BM * D * 2;
    static constexpr int KV_BYTES = BN * D * 2;
    static constexpr int P_BYTES = BM * BN * 2;
    static constexpr int STAGES = 2;
    static constexpr int SMEM_BYTES = Q_BYTES + STAGES * 2 * KV_BYTES + P_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;            // S0 [0,128)  S1 [128,256)  O [256, 256 + D)
    static constexpr int O_COL = 256;
};

template <int D>
__global__ void __launch_bounds__(192, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const FaParams p) {
    using Cfg = FaCfg<D>;
    constexpr int BM = Cfg::BM, BN = Cfg::BN, KB = Cfg::KB;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sm_q = smem;                                        // [KB][128 rows][128 B]
    uint8_t* sm_k = sm_q + Cfg::Q_BYTES;                         // [stage][KB][BN rows][128 B]
    uint8_t* sm_v = sm_k + Cfg::STAGES * Cfg::KV_BYTES;          // [stage][KB (dim halves)][BN keys][128 B]
    uint8_t* sm_p = sm_v + Cfg::STAGES * Cfg::KV_BYTES;          // [BN / 64][128 rows][128 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm_p + Cfg::P_BYTES);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;      // [2]
    uint64_t* v_full = bars + 3;      // [2]
    uint64_t* k_empty = bars + 5;     // [2]
    uint64_t* v_empty = bars + 7;     // [2]
    uint64_t* s_full = bars + 9;      // [2]
    uint64_t* s_free = bars + 11;     // [2]
    uint64_t* p_full = bars + 13;
    uint64_t* pv_done = bars + 14;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = gridDim.x - 1 - blockIdx.x;        // heaviest (latest) query tiles first
    const int t0 = qt * BM;
    const int hq = blockIdx.y;
    const int kvh = hq / (p.Hq / p.Hkv);
    const int q_hi = p.pos0 + min(p.T, t0 + BM);      // keys visible to the last query of this tile
    const int n_tiles = (q_hi + BN - 1) / BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_q);
        tma_prefetch_desc(&tm_k);
        tma_prefetch_desc(&tm_v);
        mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&k_full[s], 1);
            mbar_init(&v_full[s], 1);
            mbar_init(&k_empty[s], 1);
            mbar_init(&v_empty[s], 1);
            mbar_init(&s_full[s], 1);
            mbar_init(&s_free[s], 4);
        }
        mbar_init(p_full, 4);
        mbar_init(pv_done, 1);
        mbar_fence_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            mbar_expect_tx(q_full, Cfg::Q_BYTES);
            for (int kb = 0; kb < KB; ++kb) tma_load_2d(sm_q + kb * (BM * 128), &tm_q, q_full, hq * D + kb * 64, t0);
            const int ps = p.page_size, chunks = BN / ps;
            for (int i = 0; i < n_tiles; ++i) {
                const int s = i & 1;
                const uint32_t ph = (uint32_t)(i >> 1) & 1u;
                const int n0 = i * BN;
                mbar_wait_g(&k_empty[s], ph ^ 1);
                mbar_expect_tx(&k_full[s], Cfg::KV_BYTES);
                for (int pc = 0; pc < chunks; ++pc) {
                    const int pi = min((n0 + pc * ps) / ps, p.max_pages - 1);
                    const int row = (p.bt_row[pi] * p.Hkv + kvh) * ps;
                    for (int kb = 0; kb < KB; ++kb)
                        tma_load_2d(sm_k + s * Cfg::KV_BYTES + kb * (BN * 128) + pc * (ps * 128), &tm_k, &k_full[s], kb * 64, row);
                }
                mbar_wait_g(&v_empty[s], ph ^ 1);
                mbar_expect_tx(&v_full[s], Cfg::KV_BYTES);
                for (int pc = 0; pc < chunks; ++pc) {
                    const int pi = min((n0 + pc * ps) / ps, p.max_pages - 1);
                    const int row = (p.bt_row[pi] * p.Hkv + kvh) * ps;
                    for (int kb = 0; kb < KB; ++kb)
                        tma_load_2d(sm_v + s * Cfg::KV_BYTES + kb * (BN * 128) + pc * (ps * 128), &tm_v, &v_full[s], kb * 64, row);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc_s = make_idesc_f16(BM, BN, 0);
            constexpr uint32_t idesc_o = make_idesc_f16_bmn(BM, D);
            auto issue_pv = [&](int j) {
                const int s = j & 1;
                mbar_wait_g(p_full, (uint32_t)j & 1u);
                mbar_wait_g(&v_full[s], (uint32_t)(j >> 1) & 1u);
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < BN / 16; ++kk) {
                    const uint64_t da = make_smem_desc_sw128(smem_u32(sm_p + (kk >> 2) * (BM * 128))) + (uint64_t)(2 * (kk & 3));
                    const uint64_t db = make_smem_desc_sw128_mn(smem_u32(sm_v + s * Cfg::KV_BYTES + kk * 16 * 128), BN * 128);
                    tc_mma_f16(tmem_base + Cfg::O_COL, da, db, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
                }
                tc_commit(&v_empty[s]);
                tc_commit(pv_done);
            };
            mbar_wait_g(q_full, 0);
            for (int i = 0; i < n_tiles; ++i) {
                const int s = i & 1;
                const uint32_t ph = (uint32_t)(i >> 1) & 1u;
                mbar_wait_g(&k_full[s], ph);
                mbar_wait_g(&s_free[s], ph ^ 1);
                tc_fence_after();
#pragma unroll
                for (int kb = 0; kb < KB; ++kb) {
                    const uint64_t da = make_smem_desc_sw128(smem_u32(sm_q + kb * (BM * 128)));
                    const uint64_t db = make_smem_desc_sw128(smem_u32(sm_k + s * Cfg::KV_BYTES + kb * (BN * 128)));
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        tc_mma_f16(tmem_base + s * BN, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc_s, (kb > 0 || k > 0) ? 1u : 0u);
                }
                tc_commit(&k_empty[s]);
                tc_commit(&s_full[s]);
                if (i > 0) issue_pv(i - 1);       // scores of tile i are on their way while P.V of tile i-1 runs
            }
            issue_pv(n_tiles - 1);
        }
    } else {
        // ===== softmax / correction / epilogue: thread r <-> query row r =====
        const int q4 = warp & 3;                       // TMEM lane quarter this warp may access
        const int r = q4 * 32 + lane;
        const int q_pos = p.pos0 + t0 + r;             // this query attends to keys 0 .. q_pos
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q4 * 32) << 16);
        constexpr float LOG2E = 1.4426950408889634f;
        float m_run = -INFINITY, l_run = 0.f;
        uint8_t* prow = sm_p + r * 128;
        const int sw = r & 7;
        for (int i = 0; i < n_tiles; ++i) {
            const int s = i & 1;
            const int n0 = i * BN;
            mbar_wait_g(&s_full[s], (uint32_t)(i >> 1) & 1u);
            tc_fence_after();
            uint32_t sv[BN];
#pragma unroll
            for (int c = 0; c < BN / 32; ++c) tmem_ld_32x32(lane_addr + s * BN + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&sv[c * 32]));
            tc_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_free[s]);    // the next-but-one score tile may overwrite this buffer
            float mx = m_run;
            const bool diag = n0 + BN - 1 > p.pos0 + t0;          // some (query, key) pairs of this tile are masked
#pragma unroll
            for (int j = 0; j < BN; ++j) {
                float v = __uint_as_float(sv[j]);
                if (diag && n0 + j > q_pos) v = -INFINITY;
                sv[j] = __float_as_uint(v);
                mx = fmaxf(mx, v);
            }
            const float m_new = mx;
            const float ms = (m_new == -INFINITY) ? 0.f : m_new * LOG2E;
            const float alpha = (m_run == -INFINITY) ? 0.f : exp2f(m_run * LOG2E - ms);
            float sum = 0.f;
            uint32_t pk[BN / 2];
#pragma unroll
            for (int j = 0; j < BN; j += 2) {
                const float p0 = exp2f(__uint_as_float(sv[j]) * LOG2E - ms);
                const float p1 = exp2f(__uint_as_float(sv[j + 1]) * LOG2E - ms);
                const __half2 h = __floats2half2_rn(p0, p1);
                const float2 hf = __half22float2(h);             // the sum uses the rounded values the tensor core will see
                sum += hf.x + hf.y;
                pk[j / 2] = *reinterpret_cast<const uint32_t*>(&h);
            }
            l_run = l_run * alpha + sum;
            m_run = m_new;
            if (i > 0) {
                mbar_wait_g(pv_done, (uint32_t)(i - 1) & 1u);       // P.V of the previous tile has read P and updated O
                tc_fence_after();
                if (__any_sync(0xffffffffu, alpha != 1.f)) {      // a row maximum grew: rescale this warp's 32 rows of O
#pragma unroll 1
                    for (int c = 0; c < D / 32; ++c) {
                        uint32_t o[32];
                        tmem_ld_32x32(lane_addr + Cfg::O_COL + c * 32, o);
                        tc_wait_ld();
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
                        tmem_st_32x32(lane_addr + Cfg::O_COL + c * 32, o);
                    }
                    tc_wait_st();
                }
            }
            // P row -> shared memory, K-major swizzled (16-byte chunk c of 64-key block b lands at chunk c ^ (r & 7))
#pragma unroll
            for (int b = 0; b < BN / 64; ++b)
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    *reinterpret_cast<uint4*>(prow + b * (BM * 128) + ((c ^ sw) << 4)) =
                        make_uint4(pk[b * 32 + c * 4], pk[b * 32 + c * 4 + 1], pk[b * 32 + c * 4 + 2], pk[b * 32 + c * 4 + 3]);
            fence_proxy_async();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        // epilogue: O / l -> fp16
        mbar_wait_g(pv_done, (uint32_t)(n_tiles - 1) & 1u);
        tc_fence_after();
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
        const int t = t0 + r;
        __half* orow = p.out + ((size_t)t * p.Hq + hq) * D;
#pragma unroll 1
        for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(lane_addr + Cfg::O_COL + c * 32, o);
            tc_wait_ld();
            if (t < p.T) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    uint32_t w[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const __half2 h = __floats2half2_rn(__uint_as_float(o[j + 2 * e]) * inv, __uint_as_float(o[j + 2 * e + 1]) * inv);
                        w[e] = *reinterpret_cast<const uint32_t*>(&h);
                    }
                    *reinterpret_cast<uint4*>(orow + c * 32 + j) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        __syncwarp();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

template <int D>
static int launch_fa(const __half* q, const __half* kc, const __half* vc, const FaParams& p, int n_pages, cudaStream_t s) {
    using Cfg = FaCfg<D>;
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap_2d(&tq, q, 0, (uint64_t)p.T, (uint64_t)p.Hq * D, (uint64_t)p.Hq * D, Cfg::BM))) return rc;
    const uint64_t rows = (uint64_t)n_pages * p.Hkv * p.page_size;
    if ((rc = make_tmap_2d(&tk, kc, 0, rows, D, D, (uint32_t)p.page_size))) return rc;
    if ((rc = make_tmap_2d(&tv, vc, 0, rows, D, D, (uint32_t)p.page_size))) return rc;
    auto kern = attn_prefill_tc_kernel<D>;
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    dim3 grid((p.T + Cfg::BM - 1) / Cfg::BM, p.Hq);
    kern<<<grid, 192, Cfg::SMEM_BYTES, s>>>(tq, tk, tv, p);
    DLP_LAUNCH_CHECK();
    return 0;
}

int attn_prefill_tc(const __half* q, const __half* kc, const __half* vc, const int* bt_row, __half* out, int T, int pos0, int Hq, int Hkv,
                    int D, int page_size, int n_pages, int max_pages, cudaStream_t s) {
    DLP_REQUIRE(T > 0 && Hq % Hkv == 0 && n_pages > 0 && max_pages > 0, "attn_prefill_tc: bad shape");
    DLP_REQUIRE(page_size % 8 == 0 && page_size <= 128 && 128 % page_size == 0, "attn_prefill_tc: page size %d (multiple of 8 that divides 128)", page_size);
    FaParams p;
    p.bt_row = bt_row; p.out = out; p.T = T; p.pos0 = pos0; p.Hq = Hq; p.Hkv = Hkv; p.page_size = page_size; p.max_pages = max_pages;
    if (D == 128) return launch_fa<128>(q, kc, vc, p, n_pages, s);
    if (D == 64) return launch_fa<64>(q, kc, vc, p, n_pages, s);
    DLP_REQUIRE(false, "attn_prefill_tc: head dim %d unsupported (64, 128)", D);
}

}  // namespace dlp
