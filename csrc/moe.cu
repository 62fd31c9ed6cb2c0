// Mixture-of-experts routing plan and data movement around the grouped tcgen05 GEMM - all on the device, so a MoE layer is a
// fixed sequence of launches (graph-capturable) whatever the router decides.
//
//   router logits [T, nE] --moe_route_ex--> ids / weights [T, k]
//   ids --moe_plan--> expert-sorted order: pos[t*k+j] (sorted row of each assignment), src[row] (assignment of each sorted row),
//                     bounds[nE+1], and the token-tile table of the grouped GEMMs (expert, first row, rows) + tile count
//   xn [T, E] --moe_gather--> xs [T*k, E] (expert-sorted activations: the B operand of the grouped gate/up GEMM)
//   grouped GEMM gate/up + SwiGLU -> act [T*k, F];  grouped GEMM down -> y [T*k, E] fp32
//   y --moe_combine--> h[t] += sum_j w[t,j] * y[pos[t,j]]     (fixed order: deterministic, no atomics)
//
// Reference counterparts: topk_moe_cuda (ggml-cuda/topk-moe.cu), mm_ids_helper (mmid.cu: expert-sorted token lists +
// expert_bounds), mul_mat_q with ids_dst / expert_bounds (mmq.cuh:3459), build_moe_ffn (llama.cpp/src/llama-graph.cpp:999-1266).
#include "kernels.cuh"

namespace dlp {

constexpr int MOE_MAX_EXPERTS = 256;

// softmax over all experts -> top-k -> (optionally) renormalise.  One warp per token, up to 256 experts (8 per lane).
__global__ void moe_route_ex_kernel(const float* __restrict__ logits, int ld, int n_expert, int top_k, int renorm, int* __restrict__ ids,
                                    float* __restrict__ w) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x, lane = threadIdx.x;
    float v[8];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int e = lane + 32 * i;
        v[i] = e < n_expert ? logits[(size_t)t * ld + e] : -INFINITY;
        mx = fmaxf(mx, v[i]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        v[i] = (lane + 32 * i) < n_expert ? __expf(v[i] - mx) : -1.f;   // -1: not a candidate
        if (v[i] > 0.f) sum += v[i];
    }
    sum = warp_sum(sum);
    float wsum = 0.f;
    for (int j = 0; j < top_k; ++j) {          // iterative arg-max over (probability, lower expert index)
        float best = -1.f;
        int bi = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (v[i] > best) { best = v[i]; bi = lane + 32 * i; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if ((bi & 31) == lane) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i == (bi >> 5)) v[i] = -1.f;
        }
        if (lane == 0) { ids[t * top_k + j] = bi; w[t * top_k + j] = best / sum; }
        wsum += best / sum;
    }
    if (renorm) {
        __syncwarp();
        if (lane < top_k) w[t * top_k + lane] /= wsum;
    }
}

int moe_route_ex(const float* logits, int ld, int T, int n_expert, int top_k, int renorm, int* ids, float* w, cudaStream_t s) {
    DLP_REQUIRE(n_expert >= 1 && n_expert <= MOE_MAX_EXPERTS && top_k >= 1 && top_k <= n_expert && top_k <= 32,
                "moe_route: n_expert %d (<= %d), top_k %d (<= 32)", n_expert, MOE_MAX_EXPERTS, top_k);
    DLP_CHECK_CUDA(launch_pdl(moe_route_ex_kernel, dim3(T), dim3(32), 0, s, logits, ld, n_expert, top_k, renorm, ids, w));
    return 0;
}

// One CTA builds the whole plan.  Warp e' handles experts e', e'+nw, ...: it walks the assignment list in order (ballot + popc),
// so rows of one expert keep the token order (stable, deterministic).
__global__ void __launch_bounds__(1024) moe_plan_kernel(const int* __restrict__ ids, int n_assign, int n_expert, int bt, int* __restrict__ pos,
                                                        int* __restrict__ src, int* __restrict__ bounds, int* __restrict__ tile_expert,
                                                        int* __restrict__ tile_row0, int* __restrict__ tile_rows, int* __restrict__ n_tiles) {
    __shared__ int cnt[MOE_MAX_EXPERTS + 1];
    __shared__ int start[MOE_MAX_EXPERTS + 1];
    __shared__ int tile0[MOE_MAX_EXPERTS + 1];
    pdl_wait();
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int e = warp; e < n_expert; e += nw) {           // pass 1: how many assignments per expert
        int c = 0;
        for (int i0 = 0; i0 < n_assign; i0 += 32) {
            const int i = i0 + lane;
            c += __popc(__ballot_sync(0xffffffffu, i < n_assign && ids[i] == e));
        }
        if (lane == 0) cnt[e] = c;
    }
    __syncthreads();
    if (threadIdx.x == 0) {                               // exclusive scans: sorted-row bounds and first tile of each expert
        int acc = 0, tacc = 0;
        for (int e = 0; e < n_expert; ++e) {
            start[e] = acc;
            tile0[e] = tacc;
            bounds[e] = acc;
            acc += cnt[e];
            tacc += (cnt[e] + bt - 1) / bt;
        }
        start[n_expert] = acc;
        bounds[n_expert] = acc;
        tile0[n_expert] = tacc;
        *n_tiles = tacc;
    }
    __syncthreads();
    for (int e = warp; e < n_expert; e += nw) {           // pass 2: sorted position of every assignment, in token order
        int base = start[e];
        for (int i0 = 0; i0 < n_assign; i0 += 32) {
            const int i = i0 + lane;
            const bool mine = i < n_assign && ids[i] == e;
            const unsigned m = __ballot_sync(0xffffffffu, mine);
            if (mine) {
                const int r = base + __popc(m & ((1u << lane) - 1u));
                pos[i] = r;
                src[r] = i;
            }
            base += __popc(m);
        }
        const int nt = tile0[e + 1] - tile0[e];           // this expert's token tiles
        for (int j = lane; j < nt; j += 32) {
            tile_expert[tile0[e] + j] = e;
            tile_row0[tile0[e] + j] = start[e] + j * bt;
            tile_rows[tile0[e] + j] = min(bt, cnt[e] - j * bt);
        }
    }
}

int moe_plan(const int* ids, int n_assign, int n_expert, int bt, int* pos, int* src, int* bounds, int* tile_expert, int* tile_row0,
             int* tile_rows, int* n_tiles, cudaStream_t s) {
    DLP_REQUIRE(n_assign > 0 && n_expert >= 1 && n_expert <= MOE_MAX_EXPERTS && bt >= 16, "moe_plan: bad shape");
    const int threads = n_expert >= 32 ? 1024 : 32 * n_expert;
    DLP_CHECK_CUDA(launch_pdl(moe_plan_kernel, dim3(1), dim3(threads), 0, s, ids, n_assign, n_expert, bt, pos, src, bounds, tile_expert,
                              tile_row0, tile_rows, n_tiles));
    return 0;
}

// xs[row] = x[src[row] / top_k]: one CTA per sorted row, 16-byte copies
__global__ void __launch_bounds__(128) moe_gather_kernel(const __half* __restrict__ x, int ldx, const int* __restrict__ src, int top_k, int K,
                                                         __half* __restrict__ xs) {
    pdl_wait();
    pdl_launch_dependents();
    const int row = blockIdx.x;
    const uint4* s = reinterpret_cast<const uint4*>(x + (size_t)(src[row] / top_k) * ldx);
    uint4* d = reinterpret_cast<uint4*>(xs + (size_t)row * K);
    for (int i = threadIdx.x; i < K / 8; i += blockDim.x) d[i] = s[i];
}

int moe_gather(const __half* x, int ldx, const int* src, int top_k, int n_assign, int K, __half* xs, cudaStream_t s) {
    DLP_REQUIRE(K % 8 == 0 && ldx % 8 == 0, "moe_gather: K and ldx must be multiples of 8");
    DLP_CHECK_CUDA(launch_pdl(moe_gather_kernel, dim3(n_assign), dim3(128), 0, s, x, ldx, src, top_k, K, xs));
    return 0;
}

// h[t] += sum_j w[t,j] * y[pos[t*k+j]]   (one CTA per token; experts are added in slot order j = 0..k-1)
__global__ void __launch_bounds__(256) moe_combine_kernel(const float* __restrict__ y, const int* __restrict__ pos, const float* __restrict__ w,
                                                          int top_k, int E, float* __restrict__ h) {
    pdl_wait();
    pdl_launch_dependents();
    const int t = blockIdx.x;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        float4 acc = *reinterpret_cast<const float4*>(h + (size_t)t * E + k);
        for (int j = 0; j < top_k; ++j) {
            const float wj = w[t * top_k + j];
            const float4 v = *reinterpret_cast<const float4*>(y + (size_t)pos[t * top_k + j] * E + k);
            acc.x = fmaf(wj, v.x, acc.x); acc.y = fmaf(wj, v.y, acc.y); acc.z = fmaf(wj, v.z, acc.z); acc.w = fmaf(wj, v.w, acc.w);
        }
        *reinterpret_cast<float4*>(h + (size_t)t * E + k) = acc;
    }
}

int moe_combine(const float* y, const int* pos, const float* w, int T, int top_k, int E, float* h, cudaStream_t s) {
    DLP_REQUIRE(E % 4 == 0 && T > 0, "moe_combine: bad shape");
    DLP_CHECK_CUDA(launch_pdl(moe_combine_kernel, dim3(T), dim3(256), 0, s, y, pos, w, top_k, E, h));
    return 0;
}

}  // namespace dlp
