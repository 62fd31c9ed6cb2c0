// Device helpers of the in-kernel tensor-parallel collectives for decode (T <= 4): self-signalling stores over NVLink.
//
// The reference moves every split matmul's operands through a hub GPU with cudaMemcpyPeerAsync and has no reduction
// (llama.cpp/ggml/src/ggml-cuda/ggml-cuda.cu:1680-1744).  Here a row-parallel mat-vec stores its partial rows straight into
// all ranks' buffers from its epilogue - through the NVLS multicast alias when the fabric offers one (`multimem.st`: one
// store, replicated by the switch) or with one peer store per rank - and the next mat-vec's prologue spins on the data itself.
#pragma once
#include "kernels.cuh"

namespace dlp {

__device__ __forceinline__ float ar_clean(float v) { return __float_as_uint(v) == AR_EMPTY ? 0.f : v; }   // -0.0 means "not here yet"

__device__ __forceinline__ size_t ar_parity_off(const int* epoch, int tp, int K) {
    return (size_t)(*epoch & 1) * tp * AR_MAX_T * K;
}

// two adjacent rows of token t -> slot [rank][t][row, row+1] of every rank's buffer (row even: 8-byte aligned)
__device__ __forceinline__ void ar_push2(const ArOut& a, size_t poff, int t, int K, int row, float v0, float v1) {
    v0 = ar_clean(v0);
    v1 = ar_clean(v1);
    const size_t off = poff + ((size_t)a.rank * AR_MAX_T + t) * K + row;
    if (a.mc_buf != nullptr) {
        asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(a.mc_buf + off), "f"(v0), "f"(v1) : "memory");
    } else {
        for (int pr = 0; pr < a.tp; ++pr)
            asm volatile("st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(a.peer_buf[pr] + off), "f"(v0), "f"(v1) : "memory");
    }
}
__device__ __forceinline__ void ar_push1(const ArOut& a, size_t poff, int t, int K, int row, float v) {
    v = ar_clean(v);
    const size_t off = poff + ((size_t)a.rank * AR_MAX_T + t) * K + row;
    if (a.mc_buf != nullptr) {
        asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(a.mc_buf + off), "f"(v) : "memory");
    } else {
        for (int pr = 0; pr < a.tp; ++pr)
            asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(a.peer_buf[pr] + off), "f"(v) : "memory");
    }
}

__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool ar_has_empty(const float4& v) {
    return __float_as_uint(v.x) == AR_EMPTY || __float_as_uint(v.y) == AR_EMPTY || __float_as_uint(v.z) == AR_EMPTY ||
           __float_as_uint(v.w) == AR_EMPTY;
}
// four consecutive values of one source rank; spins until all four have arrived (a dead peer costs ~9 s once, then the error
// flag short-cuts every later wait and the step falls through with whatever is there: the host checks the flag)
__device__ __forceinline__ float4 ar_take4(const float* p, int* error) {
    float4 v = ld_volatile_f4(p);
    if (ar_has_empty(v)) {
        const long long t0 = clock64();
        while (true) {
            v = ld_volatile_f4(p);
            if (!ar_has_empty(v)) break;
            if (error && *reinterpret_cast<volatile int*>(error)) break;
            if (clock64() - t0 > (1ll << 34)) { if (error) atomicExch(error, 1); break; }
        }
    }
    return v;
}

// re-arm the other parity of a site: every CTA of the consuming kernel takes a slice
__device__ __forceinline__ void ar_rearm(float* buf, const int* epoch, int tp, int K) {
    const size_t n4 = (size_t)tp * AR_MAX_T * K / 4;
    uint4* dst = reinterpret_cast<uint4*>(buf + (size_t)((*epoch & 1) ^ 1) * tp * AR_MAX_T * K);
    const uint4 e = make_uint4(AR_EMPTY, AR_EMPTY, AR_EMPTY, AR_EMPTY);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = e;
}

}  // namespace dlp
