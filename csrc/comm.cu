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

namespace dlp {

__global__ void bump_kernel(int* ctr) { pdl_wait(); pdl_launch_dependents(); *ctr += 1; }

__global__ void __launch_bounds__(256) ar_consume_kernel(const float* __restrict__ h_in, float* __restrict__ h_out, const ArIn a, int K, int T) {
    pdl_wait();
    pdl_launch_dependents();
    if ((int)threadIdx.x < a.tp && !(a.error && *reinterpret_cast<volatile int*>(a.error))) {
        const uint32_t want = (uint32_t)*a.epoch;
        const long long t0 = clock64();
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flags + threadIdx.x) : "memory");
            if (clock64() - t0 > (1ll << 34)) { if (a.error) atomicExch(a.error, 1); break; }
        } while ((int32_t)(v - want) < 0);
    }
    __syncthreads();
    const int t = blockIdx.x;
    for (int k = threadIdx.x * 4; k < K; k += blockDim.x * 4) {
        float4 v = *reinterpret_cast<const float4*>(h_in + (size_t)t * K + k);
        for (int r = 0; r < a.tp; ++r) {
            const float4 d = __ldcg(reinterpret_cast<const float4*>(a.buf + ((size_t)r * T + t) * K + k));
            v.x += d.x; v.y += d.y; v.z += d.z; v.w += d.w;
        }
        *reinterpret_cast<float4*>(h_out + (size_t)t * K + k) = v;
    }
}

int ar_consume(const float* h_in, float* h_out, const ArIn& a, int K, int T, cudaStream_t s) {
    DLP_REQUIRE(a.buf && a.flags && a.epoch && a.tp >= 1 && a.tp <= 8, "ar_consume: bad descriptor");
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
            uint32_t v;
            do {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.red_flags + threadIdx.x) : "memory");
                if (clock64() - t0 > (1ll << 34)) { if (a.error) atomicExch(a.error, 1); break; }
            } while ((int32_t)(v - a.red_epoch) < 0);
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
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.ag_peer_flags[pr] + a.rank), "r"(a.ag_epoch) : "memory");
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
int dlp_memcpy_d2h(void* dst, const void* src, size_t bytes) {
    DLP_CHECK_CUDA(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"
