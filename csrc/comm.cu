// Symmetric (peer-mapped) device memory + the stand-alone pieces of the fused tensor-parallel all-reduce.
//
// The reference's tensor parallelism moves every activation through one hub GPU with cudaMemcpyPeerAsync
// (llama.cpp/ggml/src/ggml-cuda/ggml-cuda.cu:1680-1744) and has no reduction at all.  Here every rank owns
// a "reduce buffer" that all peers map through CUDA IPC; the row-parallel mat-vec kernels push their partial
// rows straight into the peers' buffers over NVLink (gemv.cu, GV_PUSH_F32) and the next kernel's prologue
// sums them (ArIn).  This file provides the allocation / handle exchange primitives, the epoch counter bump
// and a stand-alone consumer for places where no mat-vec follows (pipeline-stage boundary).
#include <string.h>

#include "kernels.cuh"
#include "tpcomm.cuh"

namespace dlp {

__global__ void bump_kernel(int* ctr) { pdl_wait(); pdl_launch_dependents(); *ctr += 1; }

__global__ void __launch_bounds__(256) ar_consume_kernel(const float* __restrict__ h_in, float* __restrict__ h_out, const ArIn a, int K, int T) {
    pdl_wait();
    pdl_launch_dependents();
    const float* arb = a.buf + ar_parity_off(a.epoch, a.tp, K);
    for (int t = blockIdx.x; t < T; t += gridDim.x) {
        for (int k = threadIdx.x * 4; k < K; k += blockDim.x * 4) {
            float4 v = *reinterpret_cast<const float4*>(h_in + (size_t)t * K + k);
            for (int r = 0; r < a.tp; ++r) {
                const float4 d = ar_take4(arb + ((size_t)r * AR_MAX_T + t) * K + k, a.error);
                v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            }
            *reinterpret_cast<float4*>(h_out + (size_t)t * K + k) = v;
        }
    }
    ar_rearm(a.buf, a.epoch, a.tp, K);
}

__global__ void fill_u32_kernel(uint32_t* p, size_t n, uint32_t v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
int fill_u32(uint32_t* p, size_t n, uint32_t v, cudaStream_t s) {
    fill_u32_kernel<<<(unsigned)((n + 1023) / 1024 < 1184 ? (n + 1023) / 1024 : 1184), 256, 0, s>>>(p, n, v);
    DLP_LAUNCH_CHECK();
    return 0;
}

// Sharded greedy sampling under tensor parallelism: every rank holds the logits of its vocabulary shard only.  One CTA per
// batch row finds the shard's best (value, index), stores it as one 8-byte word into slot [rank][row] of every rank (a value is
// its own arrival signal: 0 = empty), waits for the tp words and keeps the global winner - the same on every rank (ties go to
// the lower token id).  No logits all-gather, no NCCL call inside the decode step.
// ref: the reference gathers full logits on the hub GPU and samples on the CPU (llama-context.cpp:1591, server-context.cpp:2742).
__device__ __forceinline__ uint32_t ord_f32(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__global__ void __launch_bounds__(1024) argmax_tp_kernel(const float* __restrict__ logits, int V, int ld, int v0, const TokX x, int* __restrict__ out) {
    __shared__ float sv[32];
    __shared__ int si[32];
    pdl_wait();
    pdl_launch_dependents();
    const int b = blockIdx.x;
    const float* row = logits + (size_t)b * ld;
    float best = -INFINITY;
    int bi = 0;
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
        const float v = row[i];
        if (v > best) { best = v; bi = i; }
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
    if (warp != 0) return;
    const int nw = blockDim.x >> 5;
    best = lane < nw ? sv[lane] : -INFINITY;
    bi = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int par = *x.epoch & 1;
    const size_t slot = ((size_t)par * x.tp + x.rank) * TOKX_MAX_R + b;
    if (lane == 0) {   // larger word = better candidate: value first, then the lower global token id
        const unsigned long long w = ((unsigned long long)ord_f32(best) << 32) | (0xFFFFFFFFu - (uint32_t)(v0 + bi));
        if (x.mc_slots != nullptr) {
            asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.mc_slots + slot), "l"(w) : "memory");
        } else {
            for (int pr = 0; pr < x.tp; ++pr) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.peer_slots[pr] + slot), "l"(w) : "memory");
        }
    }
    unsigned long long w = 0;
    if (lane < x.tp) {
        const unsigned long long* p = x.slots + ((size_t)par * x.tp + lane) * TOKX_MAX_R + b;
        const long long t0 = clock64();
        while (true) {
            asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
            if (w != 0) break;
            if (x.error && *reinterpret_cast<volatile int*>(x.error)) break;
            if (clock64() - t0 > (1ll << 34)) { if (x.error) atomicExch(x.error, 1); break; }
        }
        x.slots[((size_t)(par ^ 1) * x.tp + lane) * TOKX_MAX_R + b] = 0;   // re-arm the other parity
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long ow = __shfl_xor_sync(0xffffffffu, w, o);
        w = ow > w ? ow : w;
    }
    if (lane == 0) out[b] = (int)(0xFFFFFFFFu - (uint32_t)(w & 0xFFFFFFFFull));
}

int argmax_tp(const float* logits, int R, int V, int ld, int v0, const TokX& x, int* out, cudaStream_t s) {
    DLP_REQUIRE(R >= 1 && R <= TOKX_MAX_R && V > 0 && x.tp >= 1 && x.tp <= 8 && x.slots && x.epoch, "argmax_tp: bad argument");
    DLP_CHECK_CUDA(launch_pdl(argmax_tp_kernel, dim3(R), dim3(1024), 0, s, logits, V, ld, v0, x, out));
    return 0;
}

int ar_consume(const float* h_in, float* h_out, const ArIn& a, int K, int T, cudaStream_t s) {
    DLP_REQUIRE(a.buf && a.epoch && a.tp >= 1 && a.tp <= 8 && T >= 1 && T <= AR_MAX_T && K % 4 == 0, "ar_consume: bad descriptor");
    DLP_CHECK_CUDA(launch_pdl(ar_consume_kernel, dim3(T), dim3(256), 0, s, h_in, h_out, a, K, T));
    return 0;
}

// Reduce-scatter consumer + RMSNorm + all-gather producer for the sequence-parallel prefill path.
// One CTA per owned token row: wait for the tp partial pushes of the preceding row-parallel GEMM, fold them into the
// fp32 residual shard, normalise, and store the fp16 row into EVERY peer's gather buffer (the next column-parallel
// GEMM's X operand); the last CTA publishes the gather epoch.
__global__ void __launch_bounds__(256) rs_norm_ag_kernel(const RsNormAg a) {
    __shared__ float red[32];
    extern __shared__ __align__(16) uint8_t rsn_smem[];
    float* hrow = reinterpret_cast<float*>(rsn_smem);   // [E]
    const int row = blockIdx.x, E = a.E;
    if (a.red_buf != nullptr) {
        if ((int)threadIdx.x < a.tp && !(a.error && *reinterpret_cast<volatile int*>(a.error))) {
            const long long t0 = clock64();
            const uint32_t red_epoch = sp_epoch(a.step, a.red_epoch);
            uint32_t v;
            do {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.red_flags + threadIdx.x) : "memory");
                if (clock64() - t0 > (1ll << 34)) { if (a.error) atomicExch(a.error, 1); break; }
            } while ((int32_t)(v - red_epoch) < 0);
        }
        __syncthreads();
    }
    float* h = a.h_shard + (size_t)row * E;
    float ss = 0.f;
    for (int k = threadIdx.x * 4; k < E; k += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(h + k);
        if (a.red_buf != nullptr) {
            for (int r = 0; r < a.tp; ++r) {
                const float4 d = __ldcg(reinterpret_cast<const float4*>(a.red_buf + ((size_t)r * a.rows + row) * E + k));
                v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
            }
            *reinterpret_cast<float4*>(h + k) = v;
        }
        *reinterpret_cast<float4*>(hrow + k) = v;
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    if (a.norm_w == nullptr) return;
    ss = block_sum(ss, red);
    const float rs = rsqrtf(ss / (float)E + a.eps);
    const size_t grow = (size_t)a.rank * a.rows + row;   // row index in the gathered [T, E] operand
    for (int k = threadIdx.x * 8; k < E; k += blockDim.x * 8) {
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __half2 hh = __floats2half2_rn(hrow[k + 2 * j] * rs * a.norm_w[k + 2 * j], hrow[k + 2 * j + 1] * rs * a.norm_w[k + 2 * j + 1]);
            o[j] = *reinterpret_cast<const uint32_t*>(&hh);
        }
        const uint4 v = make_uint4(o[0], o[1], o[2], o[3]);
        for (int pr = 0; pr < a.tp; ++pr) *reinterpret_cast<uint4*>(a.ag_peer_buf[pr] + grow * E + k) = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (atomicAdd(a.done, 1u) == gridDim.x - 1) {
            *a.done = 0;
            __threadfence_system();
            for (int pr = 0; pr < a.tp; ++pr)
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.ag_peer_flags[pr] + a.rank), "r"(sp_epoch(a.step, a.ag_epoch)) : "memory");
        }
    }
}

int rs_norm_ag(const RsNormAg& a, cudaStream_t s) {
    DLP_REQUIRE(a.h_shard && a.rows > 0 && a.E % 8 == 0 && a.tp >= 1 && a.tp <= 8, "rs_norm_ag: bad descriptor");
    DLP_REQUIRE(a.norm_w == nullptr || a.done != nullptr, "rs_norm_ag: the gather needs a completion counter");
    const size_t smem = (size_t)a.E * 4;
    DLP_REQUIRE(smem <= 200 * 1024, "rs_norm_ag: n_embd %d does not fit the row staging buffer", a.E);
    if (first_use_on_device((const void*)rs_norm_ag_kernel))   // rows wider than 12288 need more than the default 48 KB
        DLP_CHECK_CUDA(cudaFuncSetAttribute(rs_norm_ag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    rs_norm_ag_kernel<<<a.rows, 256, smem, s>>>(a);
    DLP_LAUNCH_CHECK();
    return 0;
}

int bump(int* ctr, cudaStream_t s) {
    DLP_CHECK_CUDA(launch_pdl(bump_kernel, dim3(1), dim3(1), 0, s, ctr));
    return 0;
}

}  // namespace dlp

using namespace dlp;

extern "C" {

int dlp_symm_alloc(size_t bytes, void** ptr) {
    DLP_REQUIRE(ptr && bytes > 0, "dlp_symm_alloc: bad argument");
    DLP_CHECK_CUDA(cudaMalloc(ptr, bytes));
    DLP_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
    DLP_CHECK_CUDA(cudaDeviceSynchronize());
    return 0;
}
int dlp_symm_free(void* ptr) {
    DLP_CHECK_CUDA(cudaFree(ptr));
    return 0;
}
int dlp_ipc_export(void* ptr, void* handle64) {
    DLP_REQUIRE(ptr && handle64, "dlp_ipc_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
    DLP_CHECK_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), ptr));
    return 0;
}
int dlp_ipc_import(const void* handle64, void** ptr) {
    DLP_REQUIRE(ptr && handle64, "dlp_ipc_import: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    DLP_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
int dlp_ipc_close(void* ptr) {
    DLP_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}
int dlp_bump(int32_t* ctr, void* stream) {
    DLP_REQUIRE(ctr, "dlp_bump: null argument");
    return bump(ctr, (cudaStream_t)stream);
}
int dlp_fill_u32(void* p, size_t n_words, uint32_t value, void* stream) {
    DLP_REQUIRE(p && n_words > 0, "dlp_fill_u32: bad argument");
    return fill_u32(reinterpret_cast<uint32_t*>(p), n_words, value, (cudaStream_t)stream);
}
int dlp_memcpy_d2h(void* dst, const void* src, size_t bytes) {
    DLP_CHECK_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"
