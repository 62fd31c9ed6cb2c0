// Bandwidth-bound decode linear layers:  y[T, N] (op)= x[T, K] . W[N, K]^T   for T in {1, 2, 4}
//
// Role: single-stream / tiny-batch decode, where every weight byte is read exactly once per token and
// HBM bandwidth is the roofline.  Reference counterparts: `mul_mat_vec_f` (llama.cpp/ggml/src/ggml-cuda/
// mmvf.cu) and `mul_mat_vec_q` (mmvq.cu:143) incl. their gate+up+GLU fusion (mmvq.cu:181-189,297-330).
// Differences by design: activations stay fp16/fp32 (no q8_1 activation quantise pass), RMSNorm is fused
// into the prologue (each CTA normalises x once into shared memory), the residual add is fused into the
// epilogue, and every lane keeps 8 x 16-byte weight loads in flight (two row streams x 4-deep unroll).
#include <cooperative_groups.h>
#include <stdlib.h>

#include "kernels.cuh"
#include "tpcomm.cuh"

namespace dlp {

__device__ __forceinline__ float gv_silu(float x) { return x / (1.f + __expf(-x)); }

// ---- fused all-reduce, consumer side -------------------------------------------------------------
// Row-parallel partial results of the previous mat-vec arrive in this rank's buffer of the site ([parity][tp][AR_MAX_T][K] fp32),
// stored by every rank (ourselves included) over NVLink.  The consumer spins on the values themselves (tpcomm.cuh), forms
// h_out = h_in + sum_p buf[p]  in rank order (bitwise identical on every rank) and re-arms the other parity.
// fp32 staging of the folded residual row(s) behind the fp16 activations: one row, or two (ping-pong) when NT > 1
__host__ __device__ __forceinline__ size_t ar_stage_bytes(bool ar, int NT, int K) { return ar ? (size_t)(NT > 1 ? 2 : 1) * K * 4 : 0; }

template <int NT>
__device__ __forceinline__ void stage_x(const GemvParams& p, int K, __half* xs, float* red) {
    // xs[t][k] <- fp16 activations (optionally RMS-normalised from the fp32 residual stream)
    const int tid = threadIdx.x, nth = blockDim.x;
    if (p.norm_w == nullptr) {
        const uint4* src = reinterpret_cast<const uint4*>(p.x);
        uint4* dst = reinterpret_cast<uint4*>(xs);
        const int n16 = NT * K / 8;
        for (int i = tid; i < n16; i += nth) dst[i] = src[i];
        __syncthreads();
        return;
    }
    const float* h = reinterpret_cast<const float*>(p.x);
    auto normalise = [&](const float* src, int t) {   // src: fp32 row (global h or the staged, folded row)
        float ss = 0.f;
        for (int k = tid * 4; k < K; k += nth * 4) {
            const float4 v = *reinterpret_cast<const float4*>(src + k);
            ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
        ss = block_sum(ss, red);
        const float rs = rsqrtf(ss / (float)K + p.eps);
        for (int k = tid * 4; k < K; k += nth * 4) {
            const float4 v = *reinterpret_cast<const float4*>(src + k);
            const float4 g = *reinterpret_cast<const float4*>(p.norm_w + k);
            __half2 a = __floats2half2_rn(v.x * rs * g.x, v.y * rs * g.y);
            __half2 b = __floats2half2_rn(v.z * rs * g.z, v.w * rs * g.w);
            uint2 o;
            o.x = *reinterpret_cast<uint32_t*>(&a);
            o.y = *reinterpret_cast<uint32_t*>(&b);
            *reinterpret_cast<uint2*>(xs + (size_t)t * K + k) = o;
        }
    };
    if (p.ar_in.buf == nullptr) {
#pragma unroll
        for (int t = 0; t < NT; ++t) normalise(h + (size_t)t * K, t);
        __syncthreads();
        return;
    }
    // Fold the all-reduce.  Every CTA needs the whole folded row, but reading tp x K partials in each of ~300 CTAs costs more
    // than the mat-vec itself at tp = 8 (measured: 11 us per site).  The CTAs of a thread-block cluster therefore split the row:
    // each folds K / C columns (all tp loads of a thread are issued before the first is checked) and stores its slice into the
    // shared memory of all C cluster members through DSMEM; one cluster barrier later everybody holds the full row.
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    const unsigned C = cl.num_blocks(), cr = cl.block_rank();
    const int tp = p.ar_in.tp;
    const float* arb = p.ar_in.buf + ar_parity_off(p.ar_in.epoch, tp, K);
    float* hs = reinterpret_cast<float*>(xs + (size_t)NT * K);
    const int kslice = K / (int)C, k_lo = (int)cr * kslice;
    float* peer_hs[8];
#pragma unroll
    for (unsigned c = 0; c < 8; ++c) peer_hs[c] = c < C ? cl.map_shared_rank(hs, c) : nullptr;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        const int boff = NT > 1 ? (t & 1) * K : 0;
        for (int k = k_lo + tid * 4; k < k_lo + kslice; k += nth * 4) {
            float4 v = *reinterpret_cast<const float4*>(h + (size_t)t * K + k);
            float4 d[8];
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < tp) d[r] = ld_volatile_f4(arb + ((size_t)r * AR_MAX_T + t) * K + k);
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < tp) {
                    if (ar_has_empty(d[r])) d[r] = ar_take4(arb + ((size_t)r * AR_MAX_T + t) * K + k, p.ar_in.error);
                    v.x += d[r].x; v.y += d[r].y; v.z += d[r].z; v.w += d[r].w;
                }
            if (blockIdx.x < C) *reinterpret_cast<float4*>(p.ar_in.h_out + (size_t)t * K + k) = v;
#pragma unroll
            for (unsigned c = 0; c < 8; ++c)
                if (c < C) *reinterpret_cast<float4*>(peer_hs[c] + boff + k) = v;
        }
        cl.sync();     // slices of token t are in everybody's staging row; rows of token t - 2 are no longer read by anyone
        normalise(hs + boff, t);
    }
    trace_event(p.trace, TR_GEMV | TR_AR_DONE);
    ar_rearm(p.ar_in.buf, p.ar_in.epoch, tp, K);   // the other parity: next used two steps from now
    __syncthreads();
}

template <int NT>
__device__ __forceinline__ void gv_write(const GemvParams& p, int row, float (&acc)[2][NT], bool dual, int lane) {
    // acc[s][t] is warp-uniform after the reductions; lanes split the stores
    if (dual) {
        if (lane < NT) {
            float g = 0.f, u = 0.f;
#pragma unroll
            for (int t = 0; t < NT; ++t) if (t == lane) { g = acc[0][t]; u = acc[1][t]; }
            reinterpret_cast<__half*>(p.y)[(size_t)lane * p.ldy + row] = __float2half_rn(gv_silu(g) * u);
        }
        return;
    }
    if (p.out_mode == GV_PUSH_F32) {   // partial sums of this rank's K shard -> slot [rank][t][row, row+1] of every rank's buffer
        if (lane < NT) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int t = 0; t < NT; ++t) if (t == lane) { a0 = acc[0][t]; a1 = acc[1][t]; }
            ar_push2(p.ar_out, ar_parity_off(p.ar_out.epoch, p.ar_out.tp, p.w0.N), lane, p.w0.N, row, a0, a1);
        }
        return;
    }
    if (lane < 2 * NT) {
        const int s = lane & 1, t = lane >> 1;
        float v = 0.f;
#pragma unroll
        for (int tt = 0; tt < NT; ++tt) if (tt == t) v = s ? acc[1][tt] : acc[0][tt];
        const size_t o = (size_t)t * p.ldy + row + s;
        if (row + s < p.w0.N) {
            if (p.out_mode == GV_STORE_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(v);
            else if (p.out_mode == GV_STORE_F32) reinterpret_cast<float*>(p.y)[o] = v;
            else reinterpret_cast<float*>(p.y)[o] += v;  // GV_ADD_F32: each output has exactly one writer
        }
    }
}

// ---- dense fp16 / bf16 weights -----------------------------------------------------------------
template <int NT, bool BF16>
__global__ void __launch_bounds__(256) gemv_dense_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t gv_smem[];
    const int K = p.w0.K, N = p.w0.N;
    __half* xs = reinterpret_cast<__half*>(gv_smem);
    __shared__ float red[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rows_per_it = dual ? 1 : 2;
    const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
    constexpr int U = 4;
    {   // weights do not depend on the previous kernel: pull this warp's first rows towards L2 while we wait for it
        const int row = gw * rows_per_it;
        if (row < N) {
            const uint8_t* r0 = p.w0.base + (size_t)row * K * 2;
            const uint8_t* r1 = dual ? p.w1.base + (size_t)row * K * 2 : p.w0.base + (size_t)min(row + 1, N - 1) * K * 2;
            for (int b = lane * 128; b < K * 2; b += 32 * 128) { prefetch_l2(r0 + b); prefetch_l2(r1 + b); }
        }
    }
    pdl_wait();
    stage_x<NT>(p, K, xs, red);

    for (int row = gw * rows_per_it; row < N; row += nw * rows_per_it) {
        const uint8_t* wp[2];
        wp[0] = p.w0.base + (size_t)row * K * 2;
        wp[1] = dual ? p.w1.base + (size_t)row * K * 2 : p.w0.base + (size_t)min(row + 1, N - 1) * K * 2;
        float acc[2][NT];
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[s][t] = 0.f;
        for (int k = lane * 8; k < K; k += 256 * U) {
            uint4 w[2][U];
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int s = 0; s < 2; ++s)
                    w[s][u] = (k + u * 256 < K) ? ldg_nc_v4(wp[s] + (size_t)(k + u * 256) * 2) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (k + u * 256 < K) {
                    float2 wf[2][4];
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                        const uint32_t ww[4] = {w[s][u].x, w[s][u].y, w[s][u].z, w[s][u].w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) wf[s][i] = BF16 ? bf2_to_f2(ww[i]) : h2_to_f2(ww[i]);
                    }
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)t * K + k + u * 256);
                        const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float2 xf = h2_to_f2(xw[i]);
#pragma unroll
                            for (int s = 0; s < 2; ++s) {
                                acc[s][t] = fmaf(wf[s][i].x, xf.x, acc[s][t]);
                                acc[s][t] = fmaf(wf[s][i].y, xf.y, acc[s][t]);
                            }
                        }
                    }
                }
            }
        }
        if (row + nw * rows_per_it >= N) pdl_launch_dependents();  // last row group of this warp: let the next kernel ramp up
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[s][t] = warp_sum(acc[s][t]);
        gv_write<NT>(p, row, acc, dual, lane);
    }
    pdl_launch_dependents();
}



// ---- dense weights, v3: balanced row ranges + in-block dynamic units + register double buffering ---------------------
// ncu on v1 (profiles/r1_ncu_*): the small matrices of a decode step (QKV 50 MB, attn-out 34 MB, down 118 MB) ran at
// 29-49 % of DRAM throughput because (a) the grid used 0.43-0.65 of the resident CTA slots, (b) each warp walked its rows
// in 4-14 dependent load rounds and (c) nothing was requested until the RMSNorm prologue had finished.  v3:
//   * every CTA owns a contiguous, equally sized row range (+-1 row), so the machine is filled whatever N is;
//   * the range is cut into units of 2 rows x (256*U) columns = one "round" of U 16-byte loads per row and lane; warps
//     take units from a shared-memory counter, always holding the *next* unit's loads in a second register set while
//     they reduce the current one - the load stream never waits for the math;
//   * the first unit is requested before griddepcontrol.wait and before the prologue;
//   * per-unit partial sums go to shared memory and are combined in a fixed order (deterministic) after one barrier.
template <int NT, bool BF16, int U, int DEPTH>
__global__ void __launch_bounds__(256, 2) gemv_dense_v3_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t gv_smem[];
    const int K = p.w0.K, N = p.w0.N;
    constexpr int KCW = 256 * U;               // columns per unit
    const int KC = K / KCW;                    // K chunks per row (host guarantees divisibility)
    const unsigned kc_inv = (unsigned)((0x100000000ull + (unsigned)KC - 1) / (unsigned)KC);
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rpu = dual ? 1 : 2;              // rows per unit (dual: one gate row + the matching up row)
    __half* xs = reinterpret_cast<__half*>(gv_smem);
    size_t off = (size_t)NT * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, NT, K);
    off = (off + 15) & ~size_t(15);
    float* partial = reinterpret_cast<float*>(gv_smem + off);   // [n_groups][KC][2][NT]
    __shared__ float red[32];
    __shared__ int next_unit;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // balanced contiguous row range of this CTA (in units of rpu rows)
    const int groups_total = (N + rpu - 1) / rpu;
    const int g_lo = (int)(((long long)groups_total * blockIdx.x) / gridDim.x);
    const int g_hi = (int)(((long long)groups_total * (blockIdx.x + 1)) / gridDim.x);
    const int n_units = (g_hi - g_lo) * KC;

    constexpr bool PRE2 = DEPTH == 3;          // experiment: both register sets are requested before the dependency wait
    uint4 buf[2][2][U];   // [register set][stream][load]
    int roff = 0;         // expert row offset (MoE), known only after the router kernel has finished
    auto issue = [&](int u, int b) {
        const int gq = KC == 1 ? u : (int)__umulhi((unsigned)u, kc_inv);   // u / KC without the integer-division sequence
        const int g = g_lo + gq, c = u - gq * KC;
        const int row = g * rpu;
        const uint8_t* s0 = p.w0.base + ((size_t)(row + roff) * K + (size_t)c * KCW + lane * 8) * 2;
        const uint8_t* s1 = dual ? p.w1.base + ((size_t)(row + roff) * K + (size_t)c * KCW + lane * 8) * 2
                                 : p.w0.base + ((size_t)(min(row + 1, N - 1) + roff) * K + (size_t)c * KCW + lane * 8) * 2;
#pragma unroll
        for (int i = 0; i < U; ++i) {
            buf[b][0][i] = ldg_nc_v4(s0 + (size_t)i * 512);
            buf[b][1][i] = ldg_nc_v4(s1 + (size_t)i * 512);
        }
    };
    auto reduce = [&](int u, int b) {
        const int g = KC == 1 ? u : (int)__umulhi((unsigned)u, kc_inv), c = u - g * KC;
        float acc[2][NT];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[m][t] = 0.f;
#pragma unroll
        for (int i = 0; i < U; ++i) {
            float2 wf[2][4];
#pragma unroll
            for (int m = 0; m < 2; ++m) {
                const uint32_t ww[4] = {buf[b][m][i].x, buf[b][m][i].y, buf[b][m][i].z, buf[b][m][i].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) wf[m][j] = BF16 ? bf2_to_f2(ww[j]) : h2_to_f2(ww[j]);
            }
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)t * K + (size_t)c * KCW + i * 256 + lane * 8);
                const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 xf = h2_to_f2(xw[j]);
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        acc[m][t] = fmaf(wf[m][j].x, xf.x, acc[m][t]);
                        acc[m][t] = fmaf(wf[m][j].y, xf.y, acc[m][t]);
                    }
                }
            }
        }
        float* dst = partial + ((size_t)g * KC + c) * 2 * NT;
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                const float v = warp_sum(acc[m][t]);
                if (lane == 0) dst[m * NT + t] = v;
            }
    };

    // Two register sets, hand-unrolled (DEPTH == 2): the next unit is requested *before* the current one is reduced.  A generic
    // ring that refills a slot after reducing it measured 8 % slower on the fp16 step (316 vs 345 tok/s), see profiles/.
    static_assert(DEPTH == 2 || DEPTH == 3, "dense v3 is tuned for two register sets");
    int u = warp;                              // the first round is assigned statically so that it can start right now
    if (threadIdx.x == 0) next_unit = PRE2 ? 16 : 8;
    const bool dyn_rows = p.row_off_dev != nullptr;
    if (!dyn_rows && u < n_units) issue(u, 0); // weights are static: requested before the predecessor has finished
    if (PRE2 && !dyn_rows && u + 8 < n_units) issue(u + 8, 1);
    trace_event(p.trace, TR_GEMV | TR_BEGIN);
    trace_event_cta(p.trace, TR_BEGIN);
    pdl_wait();
    if (p.pdl_early) pdl_launch_dependents();   // right after the wait: the successor's launch latency hides under our main loop,
                                                // and "my predecessor is complete when my successor starts" still holds
    trace_event(p.trace, TR_GEMV | TR_WAITED);
    trace_event_cta(p.trace, TR_WAITED);
    if (dyn_rows) {
        roff = (*p.row_off_dev) * p.row_off_mul;
        if (u < n_units) issue(u, 0);
    }
    const float oscale = p.out_scale_dev ? *p.out_scale_dev : 1.f;
    stage_x<NT>(p, K, xs, red);                // (ends with __syncthreads: next_unit is visible)

    if (PRE2 && u < n_units) {                 // peeled first round: set 1 (unit u + 8) is already in flight
        const int un = u + 8;
        if (un < n_units && dyn_rows) issue(un, 1);
        reduce(u, 0);
        if (un < n_units) {
            int u2 = 0;
            if (lane == 0) u2 = atomicAdd(&next_unit, 1);
            u2 = __shfl_sync(0xffffffffu, u2, 0);
            if (u2 < n_units) issue(u2, 0);
            reduce(un, 1);
            u = u2;
        } else {
            u = n_units;
        }
    }
    while (u < n_units) {                      // software pipeline, unrolled by two for static register indexing
        int un = 0;
        if (lane == 0) un = atomicAdd(&next_unit, 1);
        un = __shfl_sync(0xffffffffu, un, 0);
        if (un < n_units) issue(un, 1);
        reduce(u, 0);
        if (un >= n_units) break;
        int u2 = 0;
        if (lane == 0) u2 = atomicAdd(&next_unit, 1);
        u2 = __shfl_sync(0xffffffffu, u2, 0);
        if (u2 < n_units) issue(u2, 0);
        reduce(un, 1);
        u = u2;
    }
    trace_event(p.trace, TR_GEMV | TR_MAIN_DONE);
    trace_event_cta(p.trace, TR_MAIN_DONE);
    pdl_launch_dependents();
    __syncthreads();
    // fixed-order combination of the K-chunk partials + fused epilogue
    const int n_rows = (g_hi - g_lo) * rpu;
    if (p.out_mode == GV_PUSH_F32) {   // one thread per (row pair, token): 8-byte self-signalling stores into every rank's buffer
        const size_t poff = ar_parity_off(p.ar_out.epoch, p.ar_out.tp, N);
        for (int idx = threadIdx.x; idx < (g_hi - g_lo) * NT; idx += blockDim.x) {
            const int g = idx / NT, t = idx % NT;
            float v0 = 0.f, v1 = 0.f;
            for (int c = 0; c < KC; ++c) {
                const float* q = partial + ((size_t)g * KC + c) * 2 * NT;
                v0 += q[t];
                v1 += q[NT + t];
            }
            ar_push2(p.ar_out, poff, t, N, (g_lo + g) * 2, v0 * oscale, v1 * oscale);
        }
        return;
    }
    for (int idx = threadIdx.x; idx < n_rows * NT; idx += blockDim.x) {
        const int rl = idx / NT, t = idx % NT;
        const int g = rl / rpu, m = dual ? 0 : (rl % rpu);
        const int row = (g_lo + g) * rpu + (dual ? 0 : m);
        if (row >= N) continue;
        float v0 = 0.f, v1 = 0.f;
        for (int c = 0; c < KC; ++c) {
            const float* q = partial + ((size_t)g * KC + c) * 2 * NT;
            v0 += q[m * NT + t];
            if (dual) v1 += q[NT + t];
        }
        if (!dual) v0 *= oscale;
        const size_t o = (size_t)t * p.ldy + row;
        if (p.out_mode == GV_STORE_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(v0);
        else if (p.out_mode == GV_STORE_F32) reinterpret_cast<float*>(p.y)[o] = v0;
        else if (p.out_mode == GV_ADD_F32) reinterpret_cast<float*>(p.y)[o] += v0;
        else reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(gv_silu(v0) * v1);
    }
    trace_event(p.trace, TR_GEMV | TR_END);
    trace_event_cta(p.trace, TR_END);
}

// CTAs per cluster for kernels that fold an all-reduce in their prologue (stage_x): the cluster splits the fold.
// Pairs by default: 148 SMs hold 74 pairs exactly, while clusters of 4 strand 16 SMs (measured: 66 of the 74 clusters of a
// 296-CTA grid were resident, the rest ran as a second wave and doubled the kernel time).  DLP_AR_CLUSTER=n overrides.
static unsigned ar_cluster(const GemvParams& p, int blocks) {
    if (p.ar_in.buf == nullptr) return 1;
    static int forced = -1;
    if (forced < 0) { const char* e = getenv("DLP_AR_CLUSTER"); forced = e ? atoi(e) : 0; }
    unsigned c = forced > 0 ? (unsigned)forced : 2u;
    while (c > 1 && (p.w0.K % (int)(4 * c) != 0 || blocks < (int)c)) c >>= 1;
    return c;
}
// largest grid (in CTAs) of `cl`-CTA clusters that is resident at once; the balanced row ranges of the kernels take any grid size
template <typename Kern>
static int resident_cluster_blocks(Kern kern, unsigned cl, size_t smem, int threads, int want) {
    if (cl <= 1) return want;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)want);
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cl;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); return want; }
    const int cap = n * (int)cl;
    return want < cap ? want : cap;
}

template <typename Kern>
static int launch_gemv_v3(Kern kern, const GemvParams& p, int U, size_t x_smem, cudaStream_t stream) {
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rpu = dual ? 1 : 2;
    const int groups = (p.w0.N + rpu - 1) / rpu;
    int blocks = sm_count() * 2;
    if (blocks > groups) blocks = groups;
    blocks -= blocks % (int)ar_cluster(p, blocks);
    const int KC = p.w0.K / (256 * U);
    const int max_groups = (groups + blocks - 1) / blocks + 1;
    const size_t smem = ((x_smem + 15) & ~size_t(15)) + (size_t)max_groups * KC * 2 * p.T * sizeof(float);
    DLP_REQUIRE(smem <= 100 * 1024, "gemv v3: %zu bytes of shared memory needed", smem);
    const unsigned cl = ar_cluster(p, blocks);
    blocks = resident_cluster_blocks(kern, cl, smem, 256, blocks);   // (fewer CTAs only lengthen the row ranges: smem bound still holds
    blocks -= blocks % (int)cl;                                        //  when the cap is >= 1/2 of the request; checked below)
    DLP_REQUIRE(((size_t)((groups + blocks - 1) / blocks + 1) * KC * 2 * p.T * sizeof(float)) + ((x_smem + 15) & ~size_t(15)) <= 100 * 1024,
                "gemv v3: row range too long for shared memory");
    const size_t smem2 = ((x_smem + 15) & ~size_t(15)) + (size_t)((groups + blocks - 1) / blocks + 1) * KC * 2 * p.T * sizeof(float);
    DLP_CHECK_CUDA(launch_pdl_cluster(kern, dim3(blocks), dim3(256), smem2, stream, cl, p));
    return 0;
}

// ---- dense weights, bulk-async pipelined variant (experimental, DLP_GEMV_BULK=1) -----------------------------------------
// Memory-bound mat-vecs live or die by how many bytes are in flight and how early the first byte is requested.
// Here one elected thread streams whole groups of weight rows into a shared-memory ring with cp.async.bulk (TMA,
// mbarrier completion): no registers are tied up by loads, 160-200 KB per SM are in flight, and the first S stages
// are requested *before* griddepcontrol.wait - i.e. while the previous kernel is still draining - and before the
// RMSNorm prologue.  Eight consumer warps take the rows of a stage out of shared memory (R rows per stage, 8/R warps
// per row splitting K), reduce across warps through a small per-stage partial array and apply the fused epilogue.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct BulkCfg { int R; int S; int n_groups; };   // rows per stage, ring depth, total row groups

template <int NT, bool BF16>
__global__ void __launch_bounds__(288, 1) gemv_bulk_kernel(const GemvParams p, const BulkCfg cfg) {
    extern __shared__ __align__(128) uint8_t gv_smem[];
    const int K = p.w0.K, N = p.w0.N, R = cfg.R, S = cfg.S;
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int nmat = dual ? 2 : 1;
    const uint32_t row_bytes = (uint32_t)K * 2;
    const uint32_t stage_bytes = (uint32_t)R * row_bytes * nmat;
    __half* xs = reinterpret_cast<__half*>(gv_smem);
    size_t off = (size_t)NT * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, NT, K);
    off = (off + 127) & ~size_t(127);
    uint8_t* ring = gv_smem + off;
    off += (size_t)S * stage_bytes;
    float* partials = reinterpret_cast<float*>(gv_smem + off);          // [S][8 warps][2][NT]
    off += (size_t)S * 8 * 2 * NT * sizeof(float);
    uint64_t* full = reinterpret_cast<uint64_t*>(gv_smem + off);
    uint64_t* empty = full + S;
    __shared__ float red[32];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_local = (cfg.n_groups - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // groups of this CTA
    auto issue = [&](int it) {   // producer lane only: request row group `it` of this CTA into its ring slot
        const int s = it % S;
        const int row0 = ((int)blockIdx.x + it * (int)gridDim.x) * R;
        const int nrows = min(R, N - row0);
        const uint32_t bytes = (uint32_t)nrows * row_bytes;
        mbar_expect_tx(&full[s], bytes * nmat);
        uint8_t* dst = ring + (size_t)s * stage_bytes;
        bulk_g2s(dst, p.w0.base + (size_t)row0 * row_bytes, bytes, &full[s]);
        if (dual) bulk_g2s(dst + (size_t)R * row_bytes, p.w1.base + (size_t)row0 * row_bytes, bytes, &full[s]);
    };
    if (tid == 256) {
        for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
        mbar_fence_init();
        for (int it = 0; it < min(S, n_local); ++it) issue(it);   // weights are static: no need to wait for the predecessor
    }
    pdl_wait();
    stage_x<NT>(p, K, xs, red);   // all 288 threads (ends with __syncthreads: barrier inits are visible to everyone)

    if (warp == 8) {
        // ===== producer: refill ring slots as the consumers release them =====
        if (lane == 0) {
            for (int it = S; it < n_local; ++it) {
                mbar_wait(&empty[it % S], ((it / S) - 1) & 1);
                issue(it);
            }
        }
        pdl_launch_dependents();
    } else {
        // ===== consumers =====
        const int wpr = 8 / R;                       // warps per row
        const int r = warp / wpr, sub = warp % wpr;
        const int kspan = K / wpr, k_begin = sub * kspan;
        for (int it = 0; it < n_local; ++it) {
            const int s = it % S;
            mbar_wait(&full[s], (it / S) & 1);
            const uint8_t* st = ring + (size_t)s * stage_bytes;
            const uint8_t* w0r = st + (size_t)r * row_bytes;
            const uint8_t* w1r = st + (size_t)(R + r) * row_bytes;
            float acc[2][NT];
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int t = 0; t < NT; ++t) acc[m][t] = 0.f;
#pragma unroll 2
            for (int k = k_begin + lane * 8; k < k_begin + kspan; k += 256) {
                const uint4 a = *reinterpret_cast<const uint4*>(w0r + (size_t)k * 2);
                uint4 b = make_uint4(0, 0, 0, 0);
                if (dual) b = *reinterpret_cast<const uint4*>(w1r + (size_t)k * 2);
                const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
                float2 af[4], bf[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    af[i] = BF16 ? bf2_to_f2(aw[i]) : h2_to_f2(aw[i]);
                    bf[i] = BF16 ? bf2_to_f2(bw[i]) : h2_to_f2(bw[i]);
                }
#pragma unroll
                for (int t = 0; t < NT; ++t) {
                    const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)t * K + k);
                    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float2 xf = h2_to_f2(xw[i]);
                        acc[0][t] = fmaf(af[i].x, xf.x, acc[0][t]);
                        acc[0][t] = fmaf(af[i].y, xf.y, acc[0][t]);
                        if (dual) {
                            acc[1][t] = fmaf(bf[i].x, xf.x, acc[1][t]);
                            acc[1][t] = fmaf(bf[i].y, xf.y, acc[1][t]);
                        }
                    }
                }
            }
            float* pw = partials + ((size_t)s * 8 + warp) * 2 * NT;
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int t = 0; t < NT; ++t) {
                    const float v = warp_sum(acc[m][t]);
                    if (lane == 0) pw[m * NT + t] = v;
                }
            asm volatile("bar.sync 1, 256;" ::: "memory");   // the 8 consumer warps: partials of this stage complete
            if (tid < R * NT) {
                const int rr = tid / NT, t = tid % NT;
                const int row = ((int)blockIdx.x + it * (int)gridDim.x) * R + rr;
                if (row < N) {
                    float v0 = 0.f, v1 = 0.f;
                    for (int wv = rr * wpr; wv < (rr + 1) * wpr; ++wv) {
                        const float* q = partials + ((size_t)s * 8 + wv) * 2 * NT;
                        v0 += q[t];
                        v1 += q[NT + t];
                    }
                    const size_t o = (size_t)t * p.ldy + row;
                    if (p.out_mode == GV_PUSH_F32) ar_push1(p.ar_out, ar_parity_off(p.ar_out.epoch, p.ar_out.tp, N), t, N, row, v0);
                    else if (p.out_mode == GV_STORE_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(v0);
                    else if (p.out_mode == GV_STORE_F32) reinterpret_cast<float*>(p.y)[o] = v0;
                    else if (p.out_mode == GV_ADD_F32) reinterpret_cast<float*>(p.y)[o] += v0;
                    else reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(gv_silu(v0) * v1);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    }
}

// picks (R, S) for the bulk kernel; returns false when the shape does not fit its constraints
static bool bulk_config(const GemvParams& p, int T, size_t fixed_smem, BulkCfg& cfg, size_t& smem) {
    const int K = p.w0.K, N = p.w0.N;
    const int nmat = p.out_mode == GV_SWIGLU_F16 ? 2 : 1;
    const size_t budget = 212 * 1024;
    for (int R = 8; R >= 1; R >>= 1) {
        const int wpr = 8 / R;
        if ((K / wpr) % 256 != 0 || K % wpr != 0) continue;
        const size_t stage = (size_t)R * K * 2 * nmat;
        if (stage > 40 * 1024 && R > 1) continue;
        const size_t fixed = ((fixed_smem + 127) & ~size_t(127)) + 256;
        int S = 8;
        while (S >= 2 && fixed + (size_t)S * (stage + 8 * 2 * T * 4 + 16) > budget) --S;
        if (S < 2) continue;
        cfg.R = R; cfg.S = S; cfg.n_groups = (N + R - 1) / R;
        smem = fixed + (size_t)S * (stage + 8 * 2 * T * 4 + 16);
        return true;
    }
    return false;
}

template <typename Kern>
static int launch_gemv_bulk(Kern kern, const GemvParams& p, const BulkCfg& cfg, size_t smem, cudaStream_t stream) {
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    int blocks = sm_count();
    if (blocks > cfg.n_groups) blocks = cfg.n_groups;
    const unsigned cl = ar_cluster(p, blocks);
    blocks -= blocks % (int)cl;
    DLP_CHECK_CUDA(launch_pdl_cluster(kern, dim3(blocks), dim3(288), smem, stream, cl, p, cfg));
    return 0;
}

// ---- block-quantised weights ---------------------------------------------------------------------
template <int WT, int NT>
__global__ void __launch_bounds__(256) gemv_quant_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t gv_smem[];
    const int K = p.w0.K, N = p.w0.N;
    __half* xs = reinterpret_cast<__half*>(gv_smem);
    // [NT][K/16] sums of x per 16 elements (placed after the optional fp32 AR staging row)
    float* xsum = reinterpret_cast<float*>(gv_smem + (size_t)NT * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, NT, K));
    __shared__ float red[32];
    pdl_wait();
    stage_x<NT>(p, K, xs, red);
    constexpr bool HAS_BIAS = (WT == W_Q4_K || WT == W_Q5_K);
    if (HAS_BIAS) {
        for (int i = threadIdx.x; i < NT * (K / 16); i += blockDim.x) {
            const __half2* xp = reinterpret_cast<const __half2*>(xs + (size_t)i * 16);
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float2 f = __half22float2(xp[j]); s += f.x + f.y; }
            xsum[i] = s;
        }
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rows_per_it = dual ? 1 : 2;
    const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
    const int nchunk = K / 32;
    for (int row = gw * rows_per_it; row < N; row += nw * rows_per_it) {
        float acc[2][NT];
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[s][t] = 0.f;
        const int row1 = dual ? row : min(row + 1, N - 1);
#pragma unroll 2
        for (int c = lane; c < nchunk; c += 32) {
            QChunk ch[2];
            load_chunk<WT>(p.w0, row, c, ch[0]);
            load_chunk<WT>(dual ? p.w1 : p.w0, row1, c, ch[1]);
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                const int kk = ch[0].k[g];  // identical for both streams (same chunk index)
#pragma unroll
                for (int t = 0; t < NT; ++t) {
                    const uint4 xa = *reinterpret_cast<const uint4*>(xs + (size_t)t * K + kk);
                    const uint4 xb = *reinterpret_cast<const uint4*>(xs + (size_t)t * K + kk + 8);
                    const uint32_t xw[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
                    float d0 = 0.f, d1 = 0.f;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float2 xf = h2_to_f2(xw[i]);
                        const float2 q0 = __half22float2(ch[0].q[g][i]);
                        const float2 q1 = __half22float2(ch[1].q[g][i]);
                        d0 = fmaf(q0.x, xf.x, d0); d0 = fmaf(q0.y, xf.y, d0);
                        d1 = fmaf(q1.x, xf.x, d1); d1 = fmaf(q1.y, xf.y, d1);
                    }
                    acc[0][t] = fmaf(ch[0].scale[g], d0, acc[0][t]);
                    acc[1][t] = fmaf(ch[1].scale[g], d1, acc[1][t]);
                    if (HAS_BIAS) {
                        const float sx = xsum[t * (K / 16) + (kk >> 4)];
                        acc[0][t] = fmaf(ch[0].bias[g], sx, acc[0][t]);
                        acc[1][t] = fmaf(ch[1].bias[g], sx, acc[1][t]);
                    }
                }
            }
        }
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[s][t] = warp_sum(acc[s][t]);
        gv_write<NT>(p, row, acc, dual, lane);
    }
    pdl_launch_dependents();
}


// ---- block-quantised weights, v3: int8 activations + dp4a, balanced / dynamic / double-buffered like the dense v3 ------
// The fp32-decode kernel below (v1) needs ~150 instructions per 32 weights and is issue-bound (Q8_0 decoded at 43 % of
// HBM bandwidth, no faster than fp16).  v3 quantises the (already normalised) activations once per CTA to int8 with one
// fp32 scale per 32 values - the numerics class of the reference's q8_1 path (mmvq.cu / quantize.cu) - and reduces each
// 32-weight chunk with 8 dp4a, so all formats stream at memory speed.  Per chunk:  w_i = scale*q_i + bias  gives
//   sum_i w_i x_i  ~=  scale * dx * dot_int(q, xq) + bias * sum_i x_i     (the last sum is kept in fp32, exact)
// For Q6_K the stored 6-bit value is q+32: dot_int(u, xq) - 32 * sum_int(xq) recovers the signed product exactly.
// Raw register images of one 32-weight chunk.  qraw_load() only *issues* loads (no instruction may depend on their
// results, or the warp would stall for a full memory latency per unit); all decoding happens in qraw_dot().
template <int WT> struct QRaw;
template <> struct QRaw<W_Q8_0> { uint4 a, b, c, e; unsigned d2; };     // 64 int8 weights + their two fp16 block scales
// K-quants 4/5: one lane owns a whole 64-weight group (32 bytes of nibbles = sub-blocks 2j | 2j+1 of a super-block), so the two
// 6-bit scales / mins are decoded once per 64 weights and the per-unit overhead (warp sums, queue, partials) is halved
template <> struct QRaw<W_Q4_K> { uint4 hdr, qs0, qs1; };
template <> struct QRaw<W_Q5_K> { uint4 hdr, qs0, qs1, qh0, qh1; };
template <int WT> struct QChunkW { static constexpr int value = (WT == W_Q4_K || WT == W_Q5_K || WT == W_Q8_0) ? 64 : 32; };   // weights per lane-chunk
template <> struct QRaw<W_Q6_K> { uint4 ql, qh; unsigned short d; signed char s0, s1; };

template <int WT> __device__ __forceinline__ void qraw_load(const QWeight& w, int row, int c, QRaw<WT>& r);
template <> __device__ __forceinline__ void qraw_load<W_Q8_0>(const QWeight& w, int row, int c, QRaw<W_Q8_0>& r) {
    const uint8_t* qs = w.base + (size_t)row * w.K + (size_t)c * 64;          // c counts 64-weight groups (two Q8_0 blocks)
    r.a = ldg_nc_v4(qs);
    r.b = ldg_nc_v4(qs + 16);
    r.c = ldg_nc_v4(qs + 32);
    r.e = ldg_nc_v4(qs + 48);
    r.d2 = __ldg(reinterpret_cast<const unsigned*>(w.aux0) + (((size_t)row * (w.K >> 5)) >> 1) + c);
}
template <> __device__ __forceinline__ void qraw_load<W_Q4_K>(const QWeight& w, int row, int c, QRaw<W_Q4_K>& r) {
    const int blk = c >> 2, j = c & 3;           // c counts 64-weight groups
    const uint8_t* b = w.base + ((size_t)row * (w.K >> 8) + blk) * 144;
    r.hdr = __ldg(reinterpret_cast<const uint4*>(b));
    r.qs0 = ldg_nc_v4(b + 16 + j * 32);
    r.qs1 = ldg_nc_v4(b + 16 + j * 32 + 16);
}
template <> __device__ __forceinline__ void qraw_load<W_Q5_K>(const QWeight& w, int row, int c, QRaw<W_Q5_K>& r) {
    const int blk = c >> 2, j = c & 3;
    const uint8_t* b = w.base + ((size_t)row * (w.K >> 8) + blk) * 176;
    r.hdr = __ldg(reinterpret_cast<const uint4*>(b));
    r.qh0 = __ldg(reinterpret_cast<const uint4*>(b + 16));
    r.qh1 = __ldg(reinterpret_cast<const uint4*>(b + 32));
    r.qs0 = ldg_nc_v4(b + 48 + j * 32);
    r.qs1 = ldg_nc_v4(b + 48 + j * 32 + 16);
}
template <> __device__ __forceinline__ void qraw_load<W_Q6_K>(const QWeight& w, int row, int c, QRaw<W_Q6_K>& r) {
    const int blk = c >> 3, cc = c & 7, half = cc >> 2, l0 = (cc & 3) * 16;
    const int lq = l0 & 31, upper = l0 >> 5;
    const size_t nblk = (size_t)row * (w.K >> 8) + blk;
    r.ql = ldg_nc_v4(w.base + nblk * 128 + half * 64 + l0);
    r.qh = ldg_nc_v4(w.aux0 + nblk * 64 + half * 32 + lq);
    const signed char* sc = reinterpret_cast<const signed char*>(w.aux1) + nblk * 16 + half * 8 + (lq >> 4) + upper * 2;
    r.s0 = __ldg(sc);
    r.s1 = __ldg(sc + 4);
    r.d = __ldg(reinterpret_cast<const unsigned short*>(w.aux2) + nblk);
}

// element offsets of the two 16-element groups of chunk c (same mapping as load_chunk in quant.cuh)
template <int WT> __device__ __forceinline__ void chunk_offsets(int c, int& k0, int& k1) {
    if (WT == W_Q8_0) { k0 = c * 64; k1 = k0 + 32; }
    else if (WT == W_Q6_K) {
        const int blk = c >> 3, cc = c & 7, half = cc >> 2, l0 = (cc & 3) * 16;
        k0 = blk * 256 + half * 128 + (l0 >> 5) * 32 + (l0 & 31); k1 = k0 + 64;
    } else { k0 = (c >> 2) * 256 + (c & 3) * 64; k1 = k0 + 32; }   // 64-weight group: low-nibble sub-block at k0, high at k0 + 32
}

__device__ __forceinline__ int dp4a4(const uint32_t (&q)[4], const uint4& x) {
    int s = __dp4a((int)q[0], (int)x.x, 0);
    s = __dp4a((int)q[1], (int)x.y, s);
    s = __dp4a((int)q[2], (int)x.z, s);
    return __dp4a((int)q[3], (int)x.w, s);
}

__device__ __forceinline__ float us2f(unsigned short h) { return __half2float(__ushort_as_half(h)); }

// dot of one raw chunk with token t's int8 activations; xq int8 [K], xd fp32 [K/32] scales, xs16 fp32 [K/16] exact sums,
// xqs16 int [K/16] sums of the int8 values.  `c` is the chunk index (needed to pick the sub-block scales).  Unsigned quant
// values are centred ( sum (q-z) xq = dp4a(q, xq) - z * sum xq ), which keeps the activation-rounding noise off the large offset
// term; the sum of xq per 16 values comes from the prologue instead of a second set of dp4a per chunk.
template <int WT>
__device__ __forceinline__ float qraw_dot(const QRaw<WT>& r, const int8_t* xq, const float* xd, const float* xs16, const int* xqs16,
                                          int c, int k0, int k1) {
    if constexpr (WT == W_Q8_0) {
        const int rot = (c >> 1) & 3;                      // granule rotation of this 64-byte group (see the quantise prologue)
        const int8_t* xg = xq + k0;
        const uint4 x0 = *reinterpret_cast<const uint4*>(xg + ((rot & 3) << 4)), x1 = *reinterpret_cast<const uint4*>(xg + (((1 + rot) & 3) << 4));
        const uint4 x2 = *reinterpret_cast<const uint4*>(xg + (((2 + rot) & 3) << 4)), x3 = *reinterpret_cast<const uint4*>(xg + (((3 + rot) & 3) << 4));
        const uint32_t qa[4] = {r.a.x, r.a.y, r.a.z, r.a.w}, qb[4] = {r.b.x, r.b.y, r.b.z, r.b.w};
        const uint32_t qc[4] = {r.c.x, r.c.y, r.c.z, r.c.w}, qe[4] = {r.e.x, r.e.y, r.e.z, r.e.w};
        const int i0 = dp4a4(qa, x0) + dp4a4(qb, x1), i1 = dp4a4(qc, x2) + dp4a4(qe, x3);
        const float2 dd = h2_to_f2(r.d2);
        return dd.x * xd[k0 >> 5] * (float)i0 + dd.y * xd[k1 >> 5] * (float)i1;
    } else if constexpr (WT == W_Q4_K || WT == W_Q5_K) {
        const int j = c & 3;
        const int rot = (c >> 1) & 3;                      // granule rotation of this 64-byte group (see the quantise prologue)
        const int8_t* xg = xq + k0;
        const uint4 xa0 = *reinterpret_cast<const uint4*>(xg + ((rot & 3) << 4)), xa1 = *reinterpret_cast<const uint4*>(xg + (((1 + rot) & 3) << 4));
        const uint4 xb0 = *reinterpret_cast<const uint4*>(xg + (((2 + rot) & 3) << 4)), xb1 = *reinterpret_cast<const uint4*>(xg + (((3 + rot) & 3) << 4));
        const float2 dm = h2_to_f2(r.hdr.x);
        int sc0, mn0, sc1, mn1;
        scale_min_k4(r.hdr.y, r.hdr.z, r.hdr.w, 2 * j, sc0, mn0);
        scale_min_k4(r.hdr.y, r.hdr.z, r.hdr.w, 2 * j + 1, sc1, mn1);
        const float s0 = dm.x * sc0, s1 = dm.x * sc1;
        const uint32_t w0[4] = {r.qs0.x, r.qs0.y, r.qs0.z, r.qs0.w}, w1[4] = {r.qs1.x, r.qs1.y, r.qs1.z, r.qs1.w};
        uint32_t lo0[4], hi0[4], lo1[4], hi1[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            lo0[i] = w0[i] & 0x0F0F0F0Fu;
            hi0[i] = (w0[i] >> 4) & 0x0F0F0F0Fu;
            lo1[i] = w1[i] & 0x0F0F0F0Fu;
            hi1[i] = (w1[i] >> 4) & 0x0F0F0F0Fu;
        }
        constexpr int Z = (WT == W_Q5_K) ? 16 : 8;
        if constexpr (WT == W_Q5_K) {
            const uint32_t h0[4] = {r.qh0.x, r.qh0.y, r.qh0.z, r.qh0.w}, h1[4] = {r.qh1.x, r.qh1.y, r.qh1.z, r.qh1.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                lo0[i] |= ((h0[i] >> (2 * j)) & 0x01010101u) << 4;
                hi0[i] |= ((h0[i] >> (2 * j + 1)) & 0x01010101u) << 4;
                lo1[i] |= ((h1[i] >> (2 * j)) & 0x01010101u) << 4;
                hi1[i] |= ((h1[i] >> (2 * j + 1)) & 0x01010101u) << 4;
            }
        }
        const int ka = k0 >> 4, kb = k1 >> 4;
        const int ia = dp4a4(lo0, xa0) + dp4a4(lo1, xa1) - Z * (xqs16[ka] + xqs16[ka + 1]);
        const int ib = dp4a4(hi0, xb0) + dp4a4(hi1, xb1) - Z * (xqs16[kb] + xqs16[kb + 1]);
        const float b0 = (float)Z * s0 - dm.y * mn0, b1 = (float)Z * s1 - dm.y * mn1;
        return s0 * xd[k0 >> 5] * (float)ia + s1 * xd[k1 >> 5] * (float)ib + b0 * (xs16[ka] + xs16[ka + 1]) + b1 * (xs16[kb] + xs16[kb + 1]);
    } else {   // Q6_K: stored value = q + 32
        const uint4 xa = *reinterpret_cast<const uint4*>(xq + k0);
        const uint4 xb = *reinterpret_cast<const uint4*>(xq + k1);
        const int sh = (((c & 7) & 3) * 16 >> 5) * 2;
        const uint32_t l4[4] = {r.ql.x, r.ql.y, r.ql.z, r.ql.w}, h4[4] = {r.qh.x, r.qh.y, r.qh.z, r.qh.w};
        uint32_t lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            lo[i] = (l4[i] & 0x0F0F0F0Fu) | (((h4[i] >> sh) & 0x03030303u) << 4);
            hi[i] = ((l4[i] >> 4) & 0x0F0F0F0Fu) | (((h4[i] >> (sh + 4)) & 0x03030303u) << 4);
        }
        const int ia = dp4a4(lo, xa) - 32 * xqs16[k0 >> 4];
        const int ib = dp4a4(hi, xb) - 32 * xqs16[k1 >> 4];
        const float d = us2f(r.d);
        return d * (float)r.s0 * xd[k0 >> 5] * (float)ia + d * (float)r.s1 * xd[k1 >> 5] * (float)ib;
    }
}

template <int WT, int NT>
__global__ void __launch_bounds__(256, 2) gemv_quant_v3_kernel(const GemvParams p) {
    extern __shared__ __align__(16) uint8_t gv_smem[];
    const int K = p.w0.K, N = p.w0.N;
    constexpr int CW = QChunkW<WT>::value;       // weights per lane-chunk (32, or 64 for the 4/5-bit K-quants)
    const int nchunk = K / CW;
    const int nb32 = K >> 5;                    // 32-value activation blocks per token
    const int KC = (nchunk + 31) >> 5;          // units of 32 chunks (1024 or 2048 columns) per row; the last may be partial
    const unsigned kc_inv = (unsigned)((0x100000000ull + (unsigned)KC - 1) / (unsigned)KC);   // exact for u < 2^32 / KC (KC == 1 is special-cased: the reciprocal would not fit)
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rpu = dual ? 1 : 2;
    // shared memory: fp16 x [NT][K] (stage_x output, AR staging behind it) | int8 xq [NT][K] | xd [NT][K/32] | xs16 [NT][K/16] | xqs16 [NT][K/16] | partials
    __half* xs = reinterpret_cast<__half*>(gv_smem);
    size_t off = (size_t)NT * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, NT, K);
    off = (off + 15) & ~size_t(15);
    int8_t* xq = reinterpret_cast<int8_t*>(gv_smem + off);
    off += (size_t)NT * K;
    float* xd = reinterpret_cast<float*>(gv_smem + off);
    off += (size_t)NT * nb32 * 4;
    float* xs16 = reinterpret_cast<float*>(gv_smem + off);
    off += (size_t)NT * (K >> 4) * 4;
    int* xqs16 = reinterpret_cast<int*>(gv_smem + off);
    off += (size_t)NT * (K >> 4) * 4;
    float* partial = reinterpret_cast<float*>(gv_smem + off);
    __shared__ float red[32];
    __shared__ int next_unit;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int groups_total = (N + rpu - 1) / rpu;
    const int g_lo = (int)(((long long)groups_total * blockIdx.x) / gridDim.x);
    const int g_hi = (int)(((long long)groups_total * (blockIdx.x + 1)) / gridDim.x);
    const int n_units = (g_hi - g_lo) * KC;

    constexpr int DEPTH = sizeof(QRaw<WT>) >= 64 ? 2 : sizeof(QRaw<WT>) >= 36 ? (NT > 1 ? 2 : 3) : 4;   // units in flight per warp (ncu: two were not enough to cover HBM latency, 1.5-3.3 TB/s)
    QRaw<WT> buf[DEPTH][2];   // [ring slot][row stream]
    int roff = 0;             // expert row offset (MoE)
    auto issue = [&](int u, int b) {
        const int gq = KC == 1 ? u : (int)__umulhi((unsigned)u, kc_inv);   // u / KC without the ~20-instruction integer division
        const int g = g_lo + gq, c = (u - gq * KC) * 32 + lane;
        if (c >= nchunk) return;
        const int row = g * rpu;
        qraw_load<WT>(p.w0, row + roff, c, buf[b][0]);
        if (dual) qraw_load<WT>(p.w1, row + roff, c, buf[b][1]);
        else qraw_load<WT>(p.w0, min(row + 1, N - 1) + roff, c, buf[b][1]);
    };
    auto reduce = [&](int u, int b) {
        const int g = KC == 1 ? u : (int)__umulhi((unsigned)u, kc_inv), cu = u - g * KC, c = cu * 32 + lane;
        float acc[2][NT];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[m][t] = 0.f;
        if (c < nchunk) {
            int k0, k1;
            chunk_offsets<WT>(c, k0, k1);
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                acc[0][t] = qraw_dot<WT>(buf[b][0], xq + (size_t)t * K, xd + (size_t)t * nb32, xs16 + (size_t)t * (K >> 4), xqs16 + (size_t)t * (K >> 4), c, k0, k1);
                acc[1][t] = qraw_dot<WT>(buf[b][1], xq + (size_t)t * K, xd + (size_t)t * nb32, xs16 + (size_t)t * (K >> 4), xqs16 + (size_t)t * (K >> 4), c, k0, k1);
            }
        }
        float* dst = partial + ((size_t)g * KC + cu) * 2 * NT;
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int t = 0; t < NT; ++t) {
                const float v = warp_sum(acc[m][t]);
                if (lane == 0) dst[m * NT + t] = v;
            }
    };

    int uu[DEPTH];
    const bool dyn_rows = p.row_off_dev != nullptr;
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) {          // the first DEPTH rounds are assigned statically so they can start right now,
        uu[d] = warp + 8 * d;                  // before the dependency wait: weights are static
        if (!dyn_rows && uu[d] < n_units) issue(uu[d], d);
    }
    if (threadIdx.x == 0) next_unit = 8 * DEPTH;
    pdl_wait();
    if (p.pdl_early) pdl_launch_dependents();
    if (dyn_rows) {
        roff = (*p.row_off_dev) * p.row_off_mul;
#pragma unroll
        for (int d = 0; d < DEPTH; ++d)
            if (uu[d] < n_units) issue(uu[d], d);
    }
    const float oscale = p.out_scale_dev ? *p.out_scale_dev : 1.f;
    stage_x<NT>(p, K, xs, red);
    // quantise the staged activations: one 32-block per 8 threads (4 values each)
    for (int i = threadIdx.x; i < NT * (K >> 2); i += blockDim.x) {
        const __half2* hp = reinterpret_cast<const __half2*>(xs) + (size_t)i * 2;
        const float2 a = __half22float2(hp[0]), b = __half22float2(hp[1]);
        float amax = fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(b.x), fabsf(b.y)));
        float s4 = a.x + a.y + b.x + b.y;
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 4));
        s4 += __shfl_xor_sync(0xffffffffu, s4, 1);
        s4 += __shfl_xor_sync(0xffffffffu, s4, 2);           // sum over 16 values (4 threads)
        const float d = amax / 127.f, inv = amax > 0.f ? 127.f / amax : 0.f;
        const int q0 = __float2int_rn(a.x * inv), q1 = __float2int_rn(a.y * inv), q2 = __float2int_rn(b.x * inv), q3 = __float2int_rn(b.y * inv);
        // (4/5-bit K-quants read 64 contiguous bytes per lane: the 16-byte granules of every 64-byte group are rotated by
        //  (group >> 1) & 3 so that the 32 lanes of one load instruction spread over all banks instead of two)
        int wi = i;
        if (CW == 64) {
            const int t = i / (K >> 2), e = i - t * (K >> 2);           // token, word index inside the token
            const int grp = e >> 4, gran = (e >> 2) & 3, w = e & 3;     // 64-byte group, 16-byte granule, word
            wi = t * (K >> 2) + (grp << 4) + (((gran + (grp >> 1)) & 3) << 2) + w;
        }
        reinterpret_cast<uint32_t*>(xq)[wi] = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
        int qs4 = q0 + q1 + q2 + q3;
        qs4 += __shfl_xor_sync(0xffffffffu, qs4, 1);
        qs4 += __shfl_xor_sync(0xffffffffu, qs4, 2);         // integer sum over the same 16 values
        if ((i & 7) == 0) xd[i >> 3] = d;
        if ((i & 3) == 0) { xs16[i >> 2] = s4; xqs16[i >> 2] = qs4; }
    }
    __syncthreads();

    for (;;) {                                 // ring of DEPTH prefetched units per warp, refilled from the shared counter
        bool did = false;
#pragma unroll
        for (int d = 0; d < DEPTH; ++d) {
            if (uu[d] < n_units) {
                reduce(uu[d], d);
                int un = 0;
                if (lane == 0) un = atomicAdd(&next_unit, 1);
                un = __shfl_sync(0xffffffffu, un, 0);
                uu[d] = un;
                if (un < n_units) issue(un, d);
                did = true;
            }
        }
        if (!did) break;
    }
    pdl_launch_dependents();
    __syncthreads();
    const int n_rows = (g_hi - g_lo) * rpu;
    if (p.out_mode == GV_PUSH_F32) {   // one thread per (row pair, token): 8-byte self-signalling stores into every rank's buffer
        const size_t poff = ar_parity_off(p.ar_out.epoch, p.ar_out.tp, N);
        for (int idx = threadIdx.x; idx < (g_hi - g_lo) * NT; idx += blockDim.x) {
            const int g = idx / NT, t = idx % NT;
            float v0 = 0.f, v1 = 0.f;
            for (int c = 0; c < KC; ++c) {
                const float* q = partial + ((size_t)g * KC + c) * 2 * NT;
                v0 += q[t];
                v1 += q[NT + t];
            }
            ar_push2(p.ar_out, poff, t, N, (g_lo + g) * 2, v0 * oscale, v1 * oscale);
        }
        return;
    }
    for (int idx = threadIdx.x; idx < n_rows * NT; idx += blockDim.x) {
        const int rl = idx / NT, t = idx % NT;
        const int g = rl / rpu, m = dual ? 0 : (rl % rpu);
        const int row = (g_lo + g) * rpu + (dual ? 0 : m);
        if (row >= N) continue;
        float v0 = 0.f, v1 = 0.f;
        for (int c = 0; c < KC; ++c) {
            const float* q = partial + ((size_t)g * KC + c) * 2 * NT;
            v0 += q[m * NT + t];
            if (dual) v1 += q[NT + t];
        }
        if (!dual) v0 *= oscale;
        const size_t o = (size_t)t * p.ldy + row;
        if (p.out_mode == GV_STORE_F16) reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(v0);
        else if (p.out_mode == GV_STORE_F32) reinterpret_cast<float*>(p.y)[o] = v0;
        else if (p.out_mode == GV_ADD_F32) reinterpret_cast<float*>(p.y)[o] += v0;
        else reinterpret_cast<__half*>(p.y)[o] = __float2half_rn(gv_silu(v0) * v1);
    }
}

template <typename Kern>
static int launch_gemv_qv3(Kern kern, const GemvParams& p, size_t x_smem, cudaStream_t stream, int chunk_w = 32) {
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rpu = dual ? 1 : 2;
    const int groups = (p.w0.N + rpu - 1) / rpu;
    int blocks = sm_count() * 2;
    if (blocks > groups) blocks = groups;
    const unsigned cl = ar_cluster(p, blocks);
    const int K = p.w0.K, KC = ((K / chunk_w) + 31) >> 5;
    const size_t fixed = ((x_smem + 15) & ~size_t(15)) + (size_t)p.T * (K + (K >> 5) * 4 + (K >> 4) * 8);
    blocks = resident_cluster_blocks(kern, cl, fixed + 4096, 256, blocks);
    blocks -= blocks % (int)cl;
    const int max_groups = (groups + blocks - 1) / blocks + 1;
    const size_t smem = fixed + (size_t)max_groups * KC * 2 * p.T * sizeof(float);
    if (smem > 100 * 1024) return -2;   // caller falls back to v1
    DLP_CHECK_CUDA(launch_pdl_cluster(kern, dim3(blocks), dim3(256), smem, stream, cl, p));
    return 0;
}

// ---- MoE router: softmax over all experts -> top-k -> renormalise (llama-graph.cpp:1039-1139; topk-moe.cu) ---------------
__global__ void moe_route_kernel(const float* __restrict__ logits, int n_expert, int top_k, int* __restrict__ ids, float* __restrict__ w) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x, lane = threadIdx.x;
    float v = lane < n_expert ? logits[(size_t)t * n_expert + lane] : -INFINITY;
    const float mx = warp_max(v);
    float pr = lane < n_expert ? __expf(v - mx) : 0.f;
    const float sum = warp_sum(pr);
    pr /= sum;
    float wsum = 0.f, mine_w = 0.f;
    int mine_slot = -1;
    for (int j = 0; j < top_k; ++j) {          // iterative arg-max, ties to the lower expert index
        float best = pr;
        int bi = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == bi) { mine_slot = j; mine_w = pr; pr = -1.f; }
        wsum += best;
    }
    if (mine_slot >= 0) {
        ids[t * top_k + mine_slot] = lane;
        w[t * top_k + mine_slot] = mine_w / wsum;
    }
}

int moe_route(const float* logits, int T, int n_expert, int top_k, int* ids, float* w, cudaStream_t s) {
    DLP_REQUIRE(n_expert >= 1 && n_expert <= 32 && top_k >= 1 && top_k <= n_expert, "moe_route: n_expert must be <= 32 (got %d)", n_expert);
    DLP_CHECK_CUDA(launch_pdl(moe_route_kernel, dim3(T), dim3(32), 0, s, logits, n_expert, top_k, ids, w));
    return 0;
}

// ---------------------------------------------------------------------------------------------
template <typename Kern>
static int launch_gemv(Kern kern, const GemvParams& p, size_t smem, cudaStream_t stream) {
    if (first_use_on_device((const void*)kern))
        DLP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    const bool dual = p.out_mode == GV_SWIGLU_F16;
    const int rows_per_block = 8 * (dual ? 1 : 2);
    int blocks = (p.w0.N + rows_per_block - 1) / rows_per_block;
    const int max_blocks = sm_count() * 4;
    if (blocks > max_blocks) blocks = max_blocks;
    const unsigned cl = ar_cluster(p, blocks);
    blocks -= blocks % (int)cl;      // (grid-stride rows: any block count covers N)
    DLP_CHECK_CUDA(launch_pdl_cluster(kern, dim3(blocks), dim3(256), smem, stream, cl, p));
    return 0;
}

int gemv(const GemvParams& p_in, int wtype, cudaStream_t stream) {
    GemvParams p = p_in;
    p.trace = trace_buffer();
    static int early = -1;
    if (early < 0) { const char* e = getenv("DLP_PDL_EARLY"); early = (e && e[0] == '0') ? 0 : 1; }
    p.pdl_early = early && p.ar_in.buf == nullptr;   // (measured: no gain, slightly slower, on the cluster launches that fold an all-reduce)
    const int K = p.w0.K, N = p.w0.N, T = p.T;
    DLP_REQUIRE(T == 1 || T == 2 || T == 4, "gemv: T=%d unsupported (1, 2 or 4; use the tcgen05 GEMM above that)", T);
    DLP_REQUIRE(K % 256 == 0, "gemv: K=%d must be a multiple of 256", K);
    DLP_REQUIRE(N % 2 == 0, "gemv: N=%d must be even", N);
    DLP_REQUIRE(p.out_mode != GV_SWIGLU_F16 || p.w1.base != nullptr, "gemv: SwiGLU needs both gate and up weights");
    size_t smem = (size_t)T * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, T, K);
    DLP_REQUIRE(p.ar_in.buf == nullptr || p.norm_w != nullptr, "gemv: the fused all-reduce prologue needs the RMSNorm fusion");
    DLP_REQUIRE(p.out_mode != GV_PUSH_F32 || (p.ar_out.tp >= 1 && p.ar_out.tp <= 8), "gemv: bad tensor-parallel size");
    const bool quant_bias = (wtype == W_Q4_K || wtype == W_Q5_K);
    if (quant_bias) smem += (size_t)T * (K / 16) * 4;
    DLP_REQUIRE(smem <= 200 * 1024, "gemv: activations (%zu B) do not fit in shared memory", smem);
#define DLP_GV_T(KERN, ...)                                                            \
    switch (T) {                                                                       \
        case 1: return launch_gemv(KERN<__VA_ARGS__ 1>, p, smem, stream);              \
        case 2: return launch_gemv(KERN<__VA_ARGS__ 2>, p, smem, stream);              \
        default: return launch_gemv(KERN<__VA_ARGS__ 4>, p, smem, stream);             \
    }
    static int use_bulk = -1;
    // experimental: measured slower than the register-pipelined kernel in r1 (consumer side is the limit)
    if (use_bulk < 0) { const char* e = getenv("DLP_GEMV_BULK"); use_bulk = (e && e[0] == '1') ? 1 : 0; }
    if (use_bulk && (wtype == W_F16 || wtype == W_BF16)) {
        BulkCfg cfg;
        size_t bsm = 0;
        if (bulk_config(p, T, smem, cfg, bsm)) {
            const bool bf = wtype == W_BF16;
            switch (T) {
                case 1: return bf ? launch_gemv_bulk(gemv_bulk_kernel<1, true>, p, cfg, bsm, stream) : launch_gemv_bulk(gemv_bulk_kernel<1, false>, p, cfg, bsm, stream);
                case 2: return bf ? launch_gemv_bulk(gemv_bulk_kernel<2, true>, p, cfg, bsm, stream) : launch_gemv_bulk(gemv_bulk_kernel<2, false>, p, cfg, bsm, stream);
                default: return bf ? launch_gemv_bulk(gemv_bulk_kernel<4, true>, p, cfg, bsm, stream) : launch_gemv_bulk(gemv_bulk_kernel<4, false>, p, cfg, bsm, stream);
            }
        }
    }
    DLP_REQUIRE(p.row_off_dev == nullptr || (T == 1 && K % 256 == 0), "gemv: expert-indexed weights need T == 1");
    static int use_v3 = -1, v3_cfg = 0;
    if (use_v3 < 0) {
        const char* e = getenv("DLP_GEMV_V3");
        use_v3 = (e && e[0] == '0') ? 0 : 1;
        const char* c = getenv("DLP_GEMV_UD");   // "42" = 4 loads/row/unit x 2 units in flight (default), "24" = 2 x 4
        v3_cfg = (c && c[0] == '2') ? 1 : 0;
    }
    if (use_v3 && (wtype == W_F16 || wtype == W_BF16)) {
        int U = (K % 1024 == 0) ? 4 : (K % 512 == 0) ? 2 : 1;
        if (v3_cfg == 1 && U == 4) U = 2;
        const bool bf = wtype == W_BF16;
        const size_t xsm = smem;
        static int pre2 = -1;
        if (pre2 < 0) { const char* e = getenv("DLP_GEMV_PRE2"); pre2 = (e && e[0] == '1') ? 1 : 0; }
#define DLP_V3(NT_, U_, D_)                                                                                       \
    return bf ? launch_gemv_v3(gemv_dense_v3_kernel<NT_, true, U_, D_>, p, U_, xsm, stream)                        \
              : launch_gemv_v3(gemv_dense_v3_kernel<NT_, false, U_, D_>, p, U_, xsm, stream);
#define DLP_V3_T(NT_)                                                                                             \
    if (U == 4) { if (pre2) { DLP_V3(NT_, 4, 3) } else { DLP_V3(NT_, 4, 2) } } else if (U == 2) { DLP_V3(NT_, 2, 2) } else { DLP_V3(NT_, 1, 2) }
        if (xsm + 8192 <= 100 * 1024) {   // x (+ partials) must fit next to a second resident CTA
            if (T == 1) { DLP_V3_T(1) }
            if (T == 2) { DLP_V3_T(2) }
            if (T == 4) { DLP_V3_T(4) }
        }
#undef DLP_V3_T
#undef DLP_V3
    }
    static int use_qv3 = -1;
    if (use_qv3 < 0) { const char* e = getenv("DLP_GEMV_QV3"); use_qv3 = (e && e[0] == '0') ? 0 : 1; }
    if (use_qv3 && (wtype == W_Q8_0 || wtype == W_Q4_K || wtype == W_Q5_K || wtype == W_Q6_K)) {
        const size_t xsm = (size_t)T * K * 2 + ar_stage_bytes(p.ar_in.buf != nullptr, T, K);
        int rc = -2;
#define DLP_QV3(WT_)                                                                                       \
    case WT_:                                                                                              \
        rc = T == 1 ? launch_gemv_qv3(gemv_quant_v3_kernel<WT_, 1>, p, xsm, stream, QChunkW<WT_>::value)   \
           : T == 2 ? launch_gemv_qv3(gemv_quant_v3_kernel<WT_, 2>, p, xsm, stream, QChunkW<WT_>::value)   \
                    : launch_gemv_qv3(gemv_quant_v3_kernel<WT_, 4>, p, xsm, stream, QChunkW<WT_>::value);  \
        break;
        switch (wtype) { DLP_QV3(W_Q8_0) DLP_QV3(W_Q4_K) DLP_QV3(W_Q5_K) DLP_QV3(W_Q6_K) }
#undef DLP_QV3
        if (rc != -2) return rc;
    }
    switch (wtype) {
        case W_F16:
            switch (T) {
                case 1: return launch_gemv(gemv_dense_kernel<1, false>, p, smem, stream);
                case 2: return launch_gemv(gemv_dense_kernel<2, false>, p, smem, stream);
                default: return launch_gemv(gemv_dense_kernel<4, false>, p, smem, stream);
            }
        case W_BF16:
            switch (T) {
                case 1: return launch_gemv(gemv_dense_kernel<1, true>, p, smem, stream);
                case 2: return launch_gemv(gemv_dense_kernel<2, true>, p, smem, stream);
                default: return launch_gemv(gemv_dense_kernel<4, true>, p, smem, stream);
            }
        case W_Q8_0: DLP_GV_T(gemv_quant_kernel, W_Q8_0, )
        case W_Q4_K: DLP_GV_T(gemv_quant_kernel, W_Q4_K, )
        case W_Q5_K: DLP_GV_T(gemv_quant_kernel, W_Q5_K, )
        case W_Q6_K: DLP_GV_T(gemv_quant_kernel, W_Q6_K, )
    }
#undef DLP_GV_T
    DLP_REQUIRE(false, "gemv: unsupported weight type %d", wtype);
}

}  // namespace dlp
