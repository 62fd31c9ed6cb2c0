// tcgen05 / TMEM / TMA GEMM for sm_100a:   Y[T, N] (op)= X[T, K] . W[N, K]^T
//
// Role in the engine: every prefill / batched-decode linear layer (QKV, attn-out, gate/up, down, lm_head).
// The reference runs these through cuBLAS GemmEx or its mma.sync MMQ kernels
// (llama.cpp/ggml/src/ggml-cuda/ggml-cuda.cu:1227-1371, mmq.cuh:3459); this is the Blackwell-native
// replacement: weights are the UMMA "A" operand (M = 128 output features per CTA), tokens are the "B"
// operand (N = BT tokens, 16..256), so small decode batches still fill the 128-lane datapath
// ("swap-AB"), accumulators live in TMEM and the epilogue (fp16/fp32 store, fp32 residual
// accumulate / split-K reduction, SwiGLU over a gate+up accumulator pair) reads them back with
// tcgen05.ld.  Operands arrive through TMA with the 128-byte swizzle; a 3..8 stage mbarrier ring
// decouples the TMA warp from the single MMA-issuing thread.
#include <string.h>

#include "kernels.cuh"

namespace dlp {

constexpr int GEMM_BM = 128;  // output features per CTA (UMMA M)
constexpr int GEMM_BK = 64;   // K elements per pipeline stage (= 128 B of fp16: one swizzle atom)

enum GemmEpilogue : int { EPI_STORE_F16 = 0, EPI_STORE_F32 = 1, EPI_ATOMIC_ADD_F32 = 2, EPI_SWIGLU_F16 = 3,
                          EPI_STORE_BF16 = 4, EPI_RS_PUSH_F32 = 5 };

struct GemmParams {
    void* y;       // output [T, ldy]
    int T, N, K;   // tokens, output features, reduction
    int ldy;       // row stride of y in elements
    int epilogue;
    int k_blocks_per_split;
    GemmCollective coll;   // fused all-gather (X side) / reduce-scatter (Y side); zero-initialised = none
    GemmGroup grp;         // grouped (MoE) launch: per-tile expert / row range table built on the device
    // split-K with a deterministic fix-up (any epilogue): every split stores its fp32 partial tile into `ws`, the split that
    // arrives last at the tile's counter adds the partials in split order and runs the epilogue.  nullptr: not used.
    float* ws;
    unsigned* ws_counters;
};

template <int BT, int NA>
struct GemmCfg {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = BT * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = NA * A_BYTES + B_BYTES;
    static constexpr int STAGES_RAW = (200 * 1024) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int TMEM_COLS_RAW = NA * BT;
    static constexpr int TMEM_COLS = TMEM_COLS_RAW <= 32 ? 32 : TMEM_COLS_RAW <= 64 ? 64 : TMEM_COLS_RAW <= 128 ? 128
                                     : TMEM_COLS_RAW <= 256 ? 256 : 512;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float silu_f(float x) { return x / (1.f + __expf(-x)); }

// ---- dequant-fused A operand ---------------------------------------------------------------------------
// For block-quantised weights the four epilogue warps (idle while the accumulator is being built) act as the
// A-operand producer: thread r owns row r of the 128-row tile, decodes that row's 64-element K slice from
// the GGUF blocks straight into shared memory as fp16 - in exactly the layout TMA would have produced
// (128-byte rows, 16-byte chunks XOR-swizzled by row % 8) - then fences the generic-proxy writes towards the
// async proxy and arrives on the stage's "full" barrier.  The tensor core never sees the difference; no
// dequantised copy of the weights ever exists in global memory.
template <int WT>
__device__ __forceinline__ void dequant_row_slice(const QWeight& w, int row, int kc, uint8_t* tile, int r) {
    uint8_t* dst = tile + r * 128;
    const int sw = r & 7;
    if (row >= w.N) {
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(dst + c * 16) = make_uint4(0, 0, 0, 0);
        return;
    }
    constexpr int NCH = (WT == W_Q6_K) ? 4 : 2;
    int cbase;
    if (WT == W_Q8_0) cbase = kc >> 5;
    else if (WT == W_Q6_K) cbase = (kc >> 8) * 8 + ((kc & 255) >> 7) * 4;
    else cbase = (kc >> 8) * 8 + ((kc & 255) >> 6) * 2;
    QChunk ch[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) load_chunk<WT>(w, row, cbase + i, ch[i]);   // all loads in flight before any math
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int e0 = ch[i].k[g] - kc;   // 16-aligned element offset inside the 64-wide slice (or outside it)
            if (e0 < 0 || e0 >= 64) continue;
            uint32_t o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float2 f = __half22float2(ch[i].q[g][j]);
                const __half2 h = __floats2half2_rn(fmaf(ch[i].scale[g], f.x, ch[i].bias[g]), fmaf(ch[i].scale[g], f.y, ch[i].bias[g]));
                o[j] = *reinterpret_cast<const uint32_t*>(&h);
            }
            const int c0 = e0 >> 3;
            *reinterpret_cast<uint4*>(dst + ((c0 ^ sw) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<uint4*>(dst + (((c0 + 1) ^ sw) << 4)) = make_uint4(o[4], o[5], o[6], o[7]);
        }
    }
}

template <int WT>
__device__ __forceinline__ void prefetch_row_slice(const QWeight& w, int row, int kc) {
    if (row >= w.N || kc >= w.K) return;
    if (WT == W_Q8_0) prefetch_l2(w.base + (size_t)row * w.K + kc);
    else if (WT == W_Q4_K) prefetch_l2(w.base + ((size_t)row * (w.K >> 8) + (kc >> 8)) * 144 + 16 + ((kc & 255) >> 6) * 32);
    else if (WT == W_Q5_K) prefetch_l2(w.base + ((size_t)row * (w.K >> 8) + (kc >> 8)) * 176 + 48 + ((kc & 255) >> 6) * 32);
    else if (WT == W_Q6_K) prefetch_l2(w.base + ((size_t)row * (w.K >> 8) + (kc >> 8)) * 128 + ((kc & 255) >> 7) * 64);
}

template <int MODE, int CH>
__device__ __forceinline__ void epi_loop(const GemmParams& p, const uint32_t (&r0)[CH], const uint32_t (&r1)[CH], int tb, int n, int t_end) {
    if (MODE == EPI_RS_PUSH_F32) {
        // GEMM -> reduce-scatter: token t is owned by rank t / rs_rows; our partial for it goes straight into that rank's
        // reduce buffer, slot [our rank][t % rs_rows][n] (NVLink store).  Owner / local row advance incrementally.
        int owner = tb / p.coll.rs_rows;
        int local = tb - owner * p.coll.rs_rows;
        float* dst = p.coll.rs_peer_buf[owner] + ((size_t)p.coll.rs_rank * p.coll.rs_rows + local) * p.N + n;
#pragma unroll
        for (int j = 0; j < CH; ++j) {
            if (tb + j < t_end) *dst = __uint_as_float(r0[j]);
            dst += p.N;
            if (++local == p.coll.rs_rows) {
                local = 0;
                ++owner;
                if (owner < p.coll.rs_tp) dst = p.coll.rs_peer_buf[owner] + (size_t)p.coll.rs_rank * p.coll.rs_rows * p.N + n;
            }
        }
        return;
    }
    size_t o = (size_t)tb * p.ldy + n;
#pragma unroll
    for (int j = 0; j < CH; ++j, o += p.ldy) {
        if (tb + j >= t_end) break;
        const float v = __uint_as_float(r0[j]);
        if (MODE == EPI_STORE_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(v);
        else if (MODE == EPI_STORE_BF16) reinterpret_cast<__nv_bfloat16*>(p.y)[o] = __float2bfloat16_rn(v);
        else if (MODE == EPI_STORE_F32) reinterpret_cast<float*>(p.y)[o] = v;
        else if (MODE == EPI_ATOMIC_ADD_F32) atomicAdd(reinterpret_cast<float*>(p.y) + o, v);
        else if (MODE == EPI_SWIGLU_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(silu_f(v) * __uint_as_float(r1[j]));
    }
}

template <int BT, int NA, int AB_FMT, int WT>
__global__ void __launch_bounds__(192, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tm_w0, const __grid_constant__ CUtensorMap tm_w1,
                    const __grid_constant__ CUtensorMap tm_x, const GemmParams p, const QWeight qw0, const QWeight qw1) {
    constexpr bool QUANT = (WT != W_F16);
    using Cfg = GemmCfg<BT, NA>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment is required by the 128B swizzle atoms
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a0 = smem;
    uint8_t* smem_a1 = smem + STAGES * Cfg::A_BYTES;  // only when NA == 2
    uint8_t* smem_b = smem + NA * STAGES * Cfg::A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES;
    uint64_t* tmem_full_bar = bars + 2 * STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    int n0 = blockIdx.x * GEMM_BM;      // first weight row of this tile (A operand)
    int t0 = blockIdx.y * BT;           // first token row (B operand, output row)
    int t_end = p.T;                    // output rows >= t_end are not stored
    const int n_out0 = n0;              // first output column
    if (p.grp.tile_expert != nullptr) {
        // grouped launch: this token tile belongs to one expert; its weight rows sit at expert * rows_per_expert in the stack
        if ((int)blockIdx.y >= *p.grp.n_tiles) return;          // (uniform over the CTA, before any barrier / TMEM allocation)
        n0 += p.grp.tile_expert[blockIdx.y] * p.grp.rows_per_expert;
        t0 = p.grp.tile_row0[blockIdx.y];
        t_end = t0 + p.grp.tile_rows[blockIdx.y];
    }
    const int total_kb = p.K / GEMM_BK;
    const int kb_begin = blockIdx.z * p.k_blocks_per_split;
    const int kb_end = min(total_kb, kb_begin + p.k_blocks_per_split);
    const int num_kb = kb_end - kb_begin;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_w0);
        if (NA == 2) tma_prefetch_desc(&tm_w1);
        tma_prefetch_desc(&tm_x);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], QUANT ? 5 : 1);   // TMA thread (+ one lane of each of the 4 dequant warps)
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
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
        // ===== TMA producer =====
        if (lane == 0) {
            if (p.coll.x_flags != nullptr) {
                // all-gather -> GEMM: the token rows of this tile were pushed into our gather buffer by their owner
                // ranks; wait for exactly those owners, then order the acquired generic-proxy view before the TMA reads
                const int s0 = t0 / p.coll.x_rows_per_src, s1 = min(p.T - 1, t0 + BT - 1) / p.coll.x_rows_per_src;
                const long long c0 = clock64();
                const uint32_t x_epoch = sp_epoch(p.coll.step, p.coll.x_epoch);
                for (int src = s0; src <= s1; ++src) {
                    uint32_t v;
                    do {
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p.coll.x_flags + src) : "memory");
                        if (clock64() - c0 > (1ll << 34)) { if (p.coll.error) atomicExch(p.coll.error, 1); break; }
                    } while ((int32_t)(v - x_epoch) < 0);
                }
                fence_proxy_async_global();
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                mbar_expect_tx(&full_bar[stage], QUANT ? Cfg::B_BYTES : Cfg::STAGE_BYTES);
                const int kc = (kb_begin + kb) * GEMM_BK;
                if (!QUANT) {
                    tma_load_2d(smem_a0 + stage * Cfg::A_BYTES, &tm_w0, &full_bar[stage], kc, n0);
                    if (NA == 2) tma_load_2d(smem_a1 + stage * Cfg::A_BYTES, &tm_w1, &full_bar[stage], kc, n0);
                }
                tma_load_2d(smem_b + stage * Cfg::B_BYTES, &tm_x, &full_bar[stage], kc, t0);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(GEMM_BM, BT, AB_FMT);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint64_t da0 = make_smem_desc_sw128(smem_u32(smem_a0 + stage * Cfg::A_BYTES));
                const uint64_t da1 = make_smem_desc_sw128(smem_u32(smem_a1 + stage * Cfg::A_BYTES));
                const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + stage * Cfg::B_BYTES));
#pragma unroll
                for (int k = 0; k < GEMM_BK / 16; ++k) {
                    const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
                    // +32 bytes (>>4 = 2) per 16-element K step inside the swizzle atom
                    tc_mma_f16(tmem_base, da0 + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, acc);
                    if (NA == 2) tc_mma_f16(tmem_base + BT, da1 + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, acc);
                }
                tc_commit(&empty_bar[stage]);  // frees this smem slot once the MMAs above have read it
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            tc_commit(tmem_full_bar);  // accumulator complete
        }
    } else {
        if constexpr (QUANT) {
            // ===== dequant producers (the epilogue warps, before there is anything to drain) =====
            const int r = (warp - 2) * 32 + lane;
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < num_kb; ++kb) {
                const int kc = (kb_begin + kb) * GEMM_BK;
                prefetch_row_slice<WT>(qw0, n0 + r, kc + 4 * GEMM_BK);
                if (NA == 2) prefetch_row_slice<WT>(qw1, n0 + r, kc + 4 * GEMM_BK);
                mbar_wait(&empty_bar[stage], phase ^ 1);
                dequant_row_slice<WT>(qw0, n0 + r, kc, smem_a0 + stage * Cfg::A_BYTES, r);
                if (NA == 2) dequant_row_slice<WT>(qw1, n0 + r, kc, smem_a1 + stage * Cfg::A_BYTES, r);
                fence_proxy_async();   // generic-proxy smem writes -> visible to tcgen05 (async proxy)
                __syncwarp();
                if (lane == 0) mbar_arrive(&full_bar[stage]);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
        // ===== epilogue: 4 warps, warp%4 selects the TMEM lane quarter =====
        const int q = warp & 3;
        const int n_local = q * 32 + lane;
        const int n = n_out0 + n_local;
        if (num_kb > 0) {
            mbar_wait(tmem_full_bar, 0);
            tc_fence_after();
        }
        const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
        constexpr int CH = BT >= 32 ? 32 : 16;
        const int S = gridDim.z;
        const bool split_ws = p.ws != nullptr && S > 1;
        const size_t tile = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
        auto load_acc = [&](int c, uint32_t (&r0)[CH], uint32_t (&r1)[CH]) {
            if (num_kb > 0) {
                if constexpr (CH == 32) {
                    tmem_ld_32x32(lane_addr + c * CH, r0);
                    if (NA == 2) tmem_ld_32x32(lane_addr + BT + c * CH, r1);
                } else {
                    tmem_ld_32x16(lane_addr + c * CH, r0);
                    if (NA == 2) tmem_ld_32x16(lane_addr + BT + c * CH, r1);
                }
                tc_wait_ld();
            } else {
#pragma unroll
                for (int j = 0; j < CH; ++j) { r0[j] = 0; r1[j] = 0; }
            }
        };
        auto run_epilogue = [&](int c, const uint32_t (&r0)[CH], const uint32_t (&r1)[CH]) {
            if (n < p.N) {
                // one specialised store loop per epilogue (dispatch hoisted out of the per-element loop: a switch with the
                // reduce-scatter address arithmetic inside it cost 35-50 % of the whole GEMM, see profiles/r1_gemm_ab.txt)
                const int tb = t0 + c * CH;
                switch (p.epilogue) {
                    case EPI_STORE_F16: epi_loop<EPI_STORE_F16, CH>(p, r0, r1, tb, n, t_end); break;
                    case EPI_STORE_BF16: epi_loop<EPI_STORE_BF16, CH>(p, r0, r1, tb, n, t_end); break;
                    case EPI_STORE_F32: epi_loop<EPI_STORE_F32, CH>(p, r0, r1, tb, n, t_end); break;
                    case EPI_ATOMIC_ADD_F32: epi_loop<EPI_ATOMIC_ADD_F32, CH>(p, r0, r1, tb, n, t_end); break;
                    case EPI_SWIGLU_F16: if (NA == 2) epi_loop<EPI_SWIGLU_F16, CH>(p, r0, r1, tb, n, t_end); break;
                    case EPI_RS_PUSH_F32: epi_loop<EPI_RS_PUSH_F32, CH>(p, r0, r1, tb, n, t_end); break;
                }
            }
        };
        if (!split_ws) {
#pragma unroll 1
            for (int c = 0; c < BT / CH; ++c) {
                if (t0 + c * CH >= t_end) break;  // warp-uniform
                uint32_t r0[CH], r1[CH];
                load_acc(c, r0, r1);
                run_epilogue(c, r0, r1);
            }
        } else {
            // partial tile -> workspace [tile][split][accumulator][token][128 features] (128 consecutive floats per store round)
            __shared__ int s_last;
            float* wsp = p.ws + ((tile * S + blockIdx.z) * NA) * (size_t)BT * GEMM_BM;
#pragma unroll 1
            for (int c = 0; c < BT / CH; ++c) {
                if (t0 + c * CH >= t_end) break;
                uint32_t r0[CH], r1[CH];
                load_acc(c, r0, r1);
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    __stcg(wsp + (size_t)(c * CH + j) * GEMM_BM + n_local, __uint_as_float(r0[j]));
                    if (NA == 2) __stcg(wsp + (size_t)(BT + c * CH + j) * GEMM_BM + n_local, __uint_as_float(r1[j]));
                }
            }
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 64) s_last = (atomicAdd(p.ws_counters + tile, 1u) == (unsigned)(S - 1)) ? 1 : 0;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (s_last) {
                __threadfence();
                const float* base = p.ws + (tile * S * NA) * (size_t)BT * GEMM_BM;
#pragma unroll 1
                for (int c = 0; c < BT / CH; ++c) {
                    if (t0 + c * CH >= t_end) break;
                    uint32_t r0[CH], r1[CH];
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        float a0 = 0.f, a1 = 0.f;
                        for (int z = 0; z < S; ++z) {      // fixed order: the result does not depend on which split finished last
                            const float* pz = base + (size_t)z * NA * BT * GEMM_BM;
                            a0 += __ldcg(pz + (size_t)(c * CH + j) * GEMM_BM + n_local);
                            if (NA == 2) a1 += __ldcg(pz + (size_t)(BT + c * CH + j) * GEMM_BM + n_local);
                        }
                        r0[j] = __float_as_uint(a0);
                        r1[j] = __float_as_uint(a1);
                    }
                    run_epilogue(c, r0, r1);
                }
                if (threadIdx.x == 64) p.ws_counters[tile] = 0;    // ready for the next launch
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
    if (p.epilogue == EPI_RS_PUSH_F32 && threadIdx.x == 64) {
        // every CTA's stores are ordered before its counter bump; the last CTA publishes this rank's epoch to all peers
        __threadfence_system();
        const unsigned total = gridDim.x * gridDim.y * gridDim.z;
        if (atomicAdd(p.coll.rs_done, 1u) == total - 1) {
            *p.coll.rs_done = 0;
            __threadfence_system();
            for (int pr = 0; pr < p.coll.rs_tp; ++pr)
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.coll.rs_peer_flags[pr] + p.coll.rs_rank), "r"(sp_epoch(p.coll.step, p.coll.rs_epoch)) : "memory");
        }
    }
}

// ---------------------------------------------------------------------------------------------
// host side

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// 2-D K-major tensor map: dims {K, rows}, row stride ld elements, box {64, box_rows}, 128B swizzle.
int make_tmap_2d(CUtensorMap* out, const void* base, int is_bf16, uint64_t rows, uint64_t K, uint64_t ld, uint32_t box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    DLP_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    DLP_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer must be 16-byte aligned");
    DLP_REQUIRE((ld * 2) % 16 == 0, "TMA row stride must be a multiple of 16 bytes (ld=%llu)", (unsigned long long)ld);
    cuuint64_t dims[2] = {K, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {GEMM_BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    DLP_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

template <int BT, int NA, int AB_FMT, int WT = W_F16>
static int launch_gemm(const CUtensorMap& tw0, const CUtensorMap& tw1, const CUtensorMap& tx, const GemmParams& p,
                       int splits, cudaStream_t stream, const QWeight& qw0 = QWeight(), const QWeight& qw1 = QWeight()) {
    using Cfg = GemmCfg<BT, NA>;
    auto kern = gemm_tcgen05_kernel<BT, NA, AB_FMT, WT>;
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    dim3 grid((p.N + GEMM_BM - 1) / GEMM_BM, p.grp.tile_expert ? p.grp.max_tiles : (p.T + BT - 1) / BT, splits);
    kern<<<grid, 192, Cfg::SMEM_BYTES, stream>>>(tw0, tw1, tx, p, qw0, qw1);
    DLP_LAUNCH_CHECK();
    return 0;
}

// Split-K through the workspace for the epilogues that cannot reduce across CTAs themselves: skinny problems (decode batches,
// narrow tensor-parallel shards) have fewer 128-row tiles than the machine has SMs, and a CTA per tile streams its weight rows at
// a fraction of the HBM rate.  Workspace layout: 4096 tile counters (zero-initialised by the owner, self-resetting), then
// [tile][split][accumulator][BT][128] fp32 partials.
constexpr size_t GEMM_WS_COUNTERS = 4096;
static int auto_split_ws(int epilogue, int split_k, long tiles, int total_kb, int na, int bt, void* ws, size_t ws_bytes, GemmParams& p) {
    // Measured (profiles/r2_gemm_splitk_ab.txt): the fix-up pays for token tiles up to 32 (decode batches: +19 % at B = 8); for
    // prefill-sized tiles the serial reload of 128 x 256 partials costs more than the idle SMs (54 k -> 29 k tok/s at pp512), and
    // the residual-add epilogue reduces across CTAs with atomics at no extra cost.
    if (split_k > 0 || ws == nullptr || tiles <= 0 || tiles > (long)GEMM_WS_COUNTERS || bt > 32 || epilogue == EPI_ATOMIC_ADD_F32) return 1;
    static int enabled = -1;
    if (enabled < 0) { const char* e = getenv("DLP_GEMM_SPLITK"); enabled = (e && e[0] == '0') ? 0 : 1; }
    if (!enabled) return 1;
    int s = sm_count() / (int)tiles;
    if (s > 8) s = 8;
    while (s > 1 && total_kb / s < 4) --s;
    const size_t per = (size_t)na * bt * GEMM_BM * sizeof(float);
    while (s > 1 && GEMM_WS_COUNTERS * 4 + (size_t)tiles * s * per > ws_bytes) --s;
    if (s <= 1) return 1;
    p.ws_counters = reinterpret_cast<unsigned*>(ws);
    p.ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + GEMM_WS_COUNTERS * 4);
    return s;
}

// Y[T,N] (op)= X[T,K] . W0[N,K]^T  (and W1 for SwiGLU).  X/W are fp16 (is_bf16=0) or bf16 (is_bf16=1).
int gemm_f16(const void* w0, const void* w1, int ldw, const void* x, int ldx, void* y, int ldy, int T, int N, int K,
             int epilogue, int is_bf16, int split_k, cudaStream_t stream, const GemmCollective* coll, const GemmGroup* grp, void* ws,
             size_t ws_bytes) {
    // grouped: N = output features of ONE expert, w0/w1 = stacks of all experts (rows_total = what the tensor map must cover)
    DLP_REQUIRE(T > 0 && N > 0 && K > 0, "gemm: empty problem T=%d N=%d K=%d", T, N, K);
    DLP_REQUIRE(grp == nullptr || (grp->tile_expert && grp->tile_row0 && grp->tile_rows && grp->n_tiles && grp->max_tiles > 0 && split_k <= 1 &&
                                   epilogue != EPI_RS_PUSH_F32 && epilogue != EPI_ATOMIC_ADD_F32), "gemm: bad grouped descriptor");
    if (grp && split_k == 1) { ws = nullptr; ws_bytes = 0; }
    DLP_REQUIRE(K % GEMM_BK == 0, "gemm: K=%d must be a multiple of %d", K, GEMM_BK);
    const int na = (epilogue == EPI_SWIGLU_F16) ? 2 : 1;
    DLP_REQUIRE(na == 1 || w1 != nullptr, "gemm: SwiGLU epilogue needs the up-projection weight");
    const int total_kb = K / GEMM_BK;
    int splits = split_k;
    if (splits <= 0) {  // auto: fill the machine when the epilogue can reduce across CTAs
        splits = 1;
        if (epilogue == EPI_ATOMIC_ADD_F32) {
            const int bt_guess = T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256;
            const long tiles = (long)((N + GEMM_BM - 1) / GEMM_BM) * ((T + bt_guess - 1) / bt_guess);
            const int sms = sm_count();
            while (tiles * splits * 2 <= sms && total_kb / (splits * 2) >= 8) splits *= 2;
        }
    }
    DLP_REQUIRE(splits == 1 || epilogue == EPI_ATOMIC_ADD_F32, "gemm: split-K needs the atomic-add epilogue or a workspace");
    if (splits > total_kb) splits = total_kb;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    if (coll) p.coll = *coll;
    DLP_REQUIRE(epilogue != EPI_RS_PUSH_F32 || (coll && coll->rs_rows > 0 && coll->rs_done && coll->rs_tp >= 1 && coll->rs_tp <= 8 && splits == 1),
                "gemm: the reduce-scatter epilogue needs a collective descriptor and no atomic split-K");
    p.y = y; p.T = T; p.N = N; p.K = K; p.ldy = ldy; p.epilogue = epilogue;
    if (grp) p.grp = *grp;
    const int bt = grp ? grp->bt : (T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256);
    if (splits == 1)
        splits = auto_split_ws(epilogue, split_k, (long)((N + GEMM_BM - 1) / GEMM_BM) * (grp ? grp->max_tiles : (T + bt - 1) / bt), total_kb, na, bt,
                               ws, ws_bytes, p);
    p.k_blocks_per_split = (total_kb + splits - 1) / splits;
    splits = (total_kb + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
    if (splits == 1) { p.ws = nullptr; p.ws_counters = nullptr; }

    DLP_REQUIRE(bt == 16 || bt == 32 || bt == 64 || bt == 128 || bt == 256, "gemm: token tile %d", bt);
    const uint64_t w_rows = grp ? (uint64_t)grp->rows_per_expert * grp->n_expert : (uint64_t)N;   // grouped: the map spans the whole stack
    CUtensorMap tw0, tw1, tx;
    int rc;
    if ((rc = make_tmap_2d(&tw0, w0, is_bf16, w_rows, K, ldw, GEMM_BM))) return rc;
    tw1 = tw0;
    if (na == 2 && (rc = make_tmap_2d(&tw1, w1, is_bf16, w_rows, K, ldw, GEMM_BM))) return rc;
    if ((rc = make_tmap_2d(&tx, x, is_bf16, T, K, ldx, bt))) return rc;

#define DLP_GEMM_CASE(BT_)                                                                         \
    case BT_:                                                                                      \
        if (na == 2) return is_bf16 ? launch_gemm<BT_, 2, 1>(tw0, tw1, tx, p, splits, stream)      \
                                    : launch_gemm<BT_, 2, 0>(tw0, tw1, tx, p, splits, stream);     \
        return is_bf16 ? launch_gemm<BT_, 1, 1>(tw0, tw1, tx, p, splits, stream)                   \
                       : launch_gemm<BT_, 1, 0>(tw0, tw1, tx, p, splits, stream);
    switch (bt) {
        DLP_GEMM_CASE(16)
        DLP_GEMM_CASE(32)
        DLP_GEMM_CASE(64)
        DLP_GEMM_CASE(128)
        DLP_GEMM_CASE(256)
    }
#undef DLP_GEMM_CASE
    return -1;
}

// Y[T,N] (op)= X[T,K] . dequant(W)[N,K]^T with the block-quant decode fused into the A-operand producer.
int gemm_quant(const QWeight& w0, const QWeight* w1, int wtype, const void* x, int ldx, void* y, int ldy, int T, int epilogue,
               int split_k, cudaStream_t stream, const GemmCollective* coll, const GemmGroup* grp, void* ws, size_t ws_bytes) {
    // grouped: w0 / w1 describe the stack of all experts (w0.N = n_expert * rows_per_expert); one expert's width tiles the grid
    const int N = grp ? grp->rows_per_expert : w0.N, K = w0.K;
    DLP_REQUIRE(T > 0 && N > 0 && K > 0 && K % 256 == 0, "gemm_quant: bad problem T=%d N=%d K=%d", T, N, K);
    DLP_REQUIRE(grp == nullptr || (grp->tile_expert && grp->tile_row0 && grp->tile_rows && grp->n_tiles && grp->max_tiles > 0 && split_k <= 1 &&
                                   epilogue != EPI_RS_PUSH_F32 && epilogue != EPI_ATOMIC_ADD_F32 && w0.N == grp->rows_per_expert * grp->n_expert),
                "gemm_quant: bad grouped descriptor");
    const int na = (epilogue == EPI_SWIGLU_F16) ? 2 : 1;
    DLP_REQUIRE(na == 1 || w1 != nullptr, "gemm_quant: SwiGLU epilogue needs the up-projection weight");
    const int total_kb = K / GEMM_BK;
    int splits = split_k;
    if (splits <= 0) {
        splits = 1;
        if (epilogue == EPI_ATOMIC_ADD_F32) {
            const int bt_guess = T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256;
            const long tiles = (long)((N + GEMM_BM - 1) / GEMM_BM) * ((T + bt_guess - 1) / bt_guess);
            const int sms = sm_count();
            while (tiles * splits * 2 <= sms && total_kb / (splits * 2) >= 8) splits *= 2;
        }
    }
    DLP_REQUIRE(splits == 1 || epilogue == EPI_ATOMIC_ADD_F32, "gemm_quant: split-K needs the atomic-add epilogue or a workspace");
    if (splits > total_kb) splits = total_kb;
    GemmParams p;
    memset(&p, 0, sizeof(p));
    if (coll) p.coll = *coll;
    DLP_REQUIRE(epilogue != EPI_RS_PUSH_F32 || (coll && coll->rs_rows > 0 && coll->rs_done && coll->rs_tp >= 1 && coll->rs_tp <= 8 && splits == 1),
                "gemm: the reduce-scatter epilogue needs a collective descriptor and no atomic split-K");
    p.y = y; p.T = T; p.N = N; p.K = K; p.ldy = ldy; p.epilogue = epilogue;
    if (grp) p.grp = *grp;
    const int bt = grp ? grp->bt : (T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256);
    if (splits == 1)
        splits = auto_split_ws(epilogue, split_k, (long)((N + GEMM_BM - 1) / GEMM_BM) * (grp ? grp->max_tiles : (T + bt - 1) / bt), total_kb, na, bt,
                               ws, ws_bytes, p);
    p.k_blocks_per_split = (total_kb + splits - 1) / splits;
    splits = (total_kb + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
    if (splits == 1) { p.ws = nullptr; p.ws_counters = nullptr; }
    CUtensorMap tx;
    int rc;
    if ((rc = make_tmap_2d(&tx, x, 0, T, K, ldx, bt))) return rc;
    const QWeight up = w1 ? *w1 : QWeight();
#define DLP_QG_BT(WT_, BT_)                                                                                   \
    case BT_:                                                                                                 \
        return na == 2 ? launch_gemm<BT_, 2, 0, WT_>(tx, tx, tx, p, splits, stream, w0, up)                   \
                       : launch_gemm<BT_, 1, 0, WT_>(tx, tx, tx, p, splits, stream, w0, up);
#define DLP_QG_TYPE(WT_)                                                                                      \
    case WT_:                                                                                                 \
        switch (bt) { DLP_QG_BT(WT_, 16) DLP_QG_BT(WT_, 32) DLP_QG_BT(WT_, 64) DLP_QG_BT(WT_, 128) DLP_QG_BT(WT_, 256) } \
        break;
    switch (wtype) {
        DLP_QG_TYPE(W_Q8_0)
        DLP_QG_TYPE(W_Q4_K)
        DLP_QG_TYPE(W_Q5_K)
        DLP_QG_TYPE(W_Q6_K)
    }
#undef DLP_QG_TYPE
#undef DLP_QG_BT
    DLP_REQUIRE(false, "gemm_quant: unsupported weight type %d", wtype);
}

}  // namespace dlp
