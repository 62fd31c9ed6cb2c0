// dlp-worker: native CPU pipeline-stage worker (BASELINE config 1: layer-split pipeline over localhost TCP workers).
//
// Reference counterpart: `rpc-server` + the ggml-RPC backend (llama.cpp/tools/rpc/rpc-server.cpp, ggml/src/ggml-rpc/
// ggml-rpc.cpp).  There the client owns the graph and the weights: it uploads every tensor (SET_TENSOR), serialises
// the compute graph per step (GRAPH_COMPUTE) and moves activations worker -> client -> worker (two TCP hops).  This
// worker is a different design with the same capability:
//   * it memory-maps the GGUF file itself and owns layers [l0, l1) (+ embedding on the first, norm + lm_head on the
//     last stage) - no weights ever cross the socket;
//   * the wire carries only token ids / hidden states in and hidden states / logits out, per call:
//       request : u8 cmd | u32 seq | u32 n_tokens | u32 pos0 | u32 flags | u64 nbytes | payload
//       response: u32 status | u64 nbytes | payload
//     cmds: 1 HELLO (-> info), 2 RUN, 3 RESET seq, 4 TRUNCATE seq to pos0, 5 SHUTDOWN;
//   * it keeps a per-sequence KV cache for its layers and validates every size it is given.
// Kernels: fp32 accumulate over on-the-fly decoded F32/F16/Q8_0/Q4_K/Q5_K/Q6_K rows (bit layouts: the GGUF contract,
// ggml-common.h:170-336), OpenMP over output rows.
//
// Build: g++ -O3 -std=c++17 -fopenmp -mavx2 -mfma -mf16c worker.cpp -o dlp-worker
#include <arpa/inet.h>
#include <fcntl.h>
#include <immintrin.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include <omp.h>

namespace {

constexpr uint32_t PROTO_VERSION = 1;
enum Cmd : uint8_t { HELLO = 1, RUN = 2, RESET = 3, TRUNCATE = 4, SHUTDOWN = 5 };
}  // namespace
#include "gguf_lite.h"
namespace {
using namespace dlp_gguf;

// ---- Q8_0 fast path: int8 activations (one fp32 scale per 32 values) x int8 weights with AVX2 integer dot products.
// Same numerics class as the reference's CPU vec-dot for Q8_0 (q8_0 x q8_0, ggml-cpu/quants.c); ~4x the throughput of
// decoding to fp32 first, which matters because the CPU configuration is the only one where compute, not HBM, limits us.
struct QAct { std::vector<int8_t> q; std::vector<float> d, c; };   // c[t][b] = -128 * d * sum(q): offset term of the VNNI path

// DLP_WORKER_PROF=1: seconds per phase of every RUN on stderr (where a prefill goes: mat-muls, attention, the rest)
struct Prof {
    bool on = getenv("DLP_WORKER_PROF") != nullptr;
    double t0 = 0, acc[6] = {0, 0, 0, 0, 0, 0};
    void start() { if (on) t0 = omp_get_wtime(); }
    void lap(int i) { if (on) { const double t = omp_get_wtime(); acc[i] += t - t0; t0 = t; } }
    void report(int T) {
        if (!on) return;
        fprintf(stderr, "prof T=%d: matmul %.1f ms, attention %.1f ms, norm %.1f ms, rope+kv %.1f ms, act %.1f ms, head %.1f ms\n", T,
                acc[0] * 1e3, acc[1] * 1e3, acc[2] * 1e3, acc[3] * 1e3, acc[4] * 1e3, acc[5] * 1e3);
        for (double& a : acc) a = 0;
    }
};

void quantize_act(const float* x, int T, uint64_t K, QAct& a) {
    a.q.resize((size_t)T * K);
    a.d.resize((size_t)T * (K / 32));
    a.c.resize((size_t)T * (K / 32));
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)(T * (K / 32)); ++b) {
        const float* p = x + (size_t)b * 32;
        float amax = 0.f;
        for (int i = 0; i < 32; ++i) amax = std::max(amax, std::fabs(p[i]));
        const float d = amax / 127.f, inv = amax > 0 ? 127.f / amax : 0.f;
        a.d[b] = d;
        int sum = 0;
        for (int i = 0; i < 32; ++i) { const int8_t v = (int8_t)std::lrintf(p[i] * inv); a.q[(size_t)b * 32 + i] = v; sum += v; }
        a.c[b] = -128.f * d * (float)sum;
    }
}

inline float dot_q8_0_row(const uint8_t* wrow, const int8_t* xq, const float* xd, uint64_t nblk) {
    __m256 acc = _mm256_setzero_ps();
    const __m256i ones = _mm256_set1_epi16(1);
    for (uint64_t b = 0; b < nblk; ++b) {
        const uint8_t* p = wrow + b * 34;
        uint16_t dh; memcpy(&dh, p, 2);
        const __m256i wv = _mm256_loadu_si256((const __m256i*)(p + 2));
        const __m256i xv = _mm256_loadu_si256((const __m256i*)(xq + b * 32));
        const __m256i aw = _mm256_sign_epi8(wv, wv);            // |w|
        const __m256i sx = _mm256_sign_epi8(xv, wv);            // x * sign(w)
        const __m256i p16 = _mm256_maddubs_epi16(aw, sx);       // 16 x int16 pair sums
        const __m256i p32 = _mm256_madd_epi16(p16, ones);       // 8 x int32
        acc = _mm256_fmadd_ps(_mm256_cvtepi32_ps(p32), _mm256_set1_ps(f16(dh) * xd[b]), acc);
    }
    __m128 s = _mm_add_ps(_mm256_castps256_ps128(acc), _mm256_extractf128_ps(acc, 1));
    s = _mm_hadd_ps(s, s); s = _mm_hadd_ps(s, s);
    return _mm_cvtss_f32(s);
}

inline float hsum256(__m256 v) {
    __m128 s = _mm_add_ps(_mm256_castps256_ps128(v), _mm256_extractf128_ps(v, 1));
    s = _mm_hadd_ps(s, s); s = _mm_hadd_ps(s, s);
    return _mm_cvtss_f32(s);
}

// R weight rows x 4 tokens per pass over K: each weight block is decoded (|w|, sign source, scale) once for four tokens and
// each activation block is loaded once for R rows.
template <int R>
inline void tile_q8_0(const uint8_t* wr, size_t rb, const int8_t* xq, const float* xd, uint64_t K, uint64_t nblk, float* out /*[R][4]*/) {
    const __m256i ones = _mm256_set1_epi16(1);
    __m256 acc[R][4];
    for (int r = 0; r < R; ++r) for (int j = 0; j < 4; ++j) acc[r][j] = _mm256_setzero_ps();
    for (uint64_t b = 0; b < nblk; ++b) {
        __m256i wv[R], aw[R];
        float dw[R];
        for (int r = 0; r < R; ++r) {
            const uint8_t* p = wr + r * rb + b * 34;
            uint16_t dh; memcpy(&dh, p, 2);
            dw[r] = f16(dh);
            wv[r] = _mm256_loadu_si256((const __m256i*)(p + 2));
            aw[r] = _mm256_sign_epi8(wv[r], wv[r]);
        }
#pragma GCC unroll 4
        for (int j = 0; j < 4; ++j) {
            const __m256i xv = _mm256_loadu_si256((const __m256i*)(xq + (size_t)j * K + b * 32));
            const float dx = xd[(size_t)j * nblk + b];
            for (int r = 0; r < R; ++r) {
                const __m256i p32 = _mm256_madd_epi16(_mm256_maddubs_epi16(aw[r], _mm256_sign_epi8(xv, wv[r])), ones);
                acc[r][j] = _mm256_fmadd_ps(_mm256_cvtepi32_ps(p32), _mm256_set1_ps(dw[r] * dx), acc[r][j]);
            }
        }
    }
    for (int r = 0; r < R; ++r) for (int j = 0; j < 4; ++j) out[r * 4 + j] = hsum256(acc[r][j]);
}

// Cache blocking for prefill: a panel of weight rows (L2-resident) is swept by groups of four tokens whose int8 activations
// (4 x K bytes) stay in L1 across the panel, instead of re-streaming the whole [T][K] activation matrix for every weight row.
void matmul_q8_0(const Tensor& w, const QAct& a, int T, float* y, bool accumulate) {
    const uint64_t K = w.cols(), N = w.rows(), nblk = K / 32;
    const size_t rb = row_bytes(w.type, K);
    constexpr int64_t PANEL = 32;
    const int64_t n_panels = ((int64_t)N + PANEL - 1) / PANEL;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t pi = 0; pi < n_panels; ++pi) {
        const int64_t n0 = pi * PANEL, n1 = std::min<int64_t>(N, n0 + PANEL);
        int t = 0;
        for (; t + 4 <= T; t += 4) {
            const int8_t* xq = a.q.data() + (size_t)t * K;
            const float* xd = a.d.data() + (size_t)t * nblk;
            float out[8];
            int64_t n = n0;
            for (; n + 2 <= n1; n += 2) {
                tile_q8_0<2>(w.data + n * rb, rb, xq, xd, K, nblk, out);
                for (int r = 0; r < 2; ++r)
                    for (int j = 0; j < 4; ++j) {
                        float& d = y[(size_t)(t + j) * N + n + r];
                        d = accumulate ? d + out[r * 4 + j] : out[r * 4 + j];
                    }
            }
            for (; n < n1; ++n) {
                tile_q8_0<1>(w.data + n * rb, rb, xq, xd, K, nblk, out);
                for (int j = 0; j < 4; ++j) {
                    float& d = y[(size_t)(t + j) * N + n];
                    d = accumulate ? d + out[j] : out[j];
                }
            }
        }
        for (; t < T; ++t)
            for (int64_t n = n0; n < n1; ++n) {
                const float v = dot_q8_0_row(w.data + n * rb, a.q.data() + (size_t)t * K, a.d.data() + (size_t)t * nblk, nblk);
                if (accumulate) y[(size_t)t * N + n] += v; else y[(size_t)t * N + n] = v;
            }
    }
}

// ---- AVX-512 VNNI variant (chosen at run time).  vpdpbusd multiplies unsigned by signed bytes into int32 lanes without the
// int16 saturation of vpmaddubsw, so the weights go in as w + 128 (one xor per weight block, shared by the token group) and the
// offset 128 * sum(x) leaves again through a per-(token, block) term prepared by quantize_act.  One 512-bit operation covers two
// 32-value blocks; their two scales sit in the low and high halves of the scale vector.  Per 64 multiply-adds: vpdpbusd, convert,
// scale product, fma - against six AVX2 operations per 32.
#define DLP_VNNI __attribute__((target("avx512f,avx512bw,avx512vl,avx512dq,avx512vnni,avx2,fma,f16c")))

DLP_VNNI inline __m512 pair_scale(const float* p, __m512i half_sel) {   // (p[0] x 8 | p[1] x 8)
    double two; memcpy(&two, p, 8);
    return _mm512_permutexvar_ps(half_sel, _mm512_castpd_ps(_mm512_set1_pd(two)));
}

template <int R>
DLP_VNNI inline void tile_q8_0_vnni(const uint8_t* wr, size_t rb, const float* dws, const int8_t* xq, const float* xd, const float* xc,
                                    uint64_t K, uint64_t nblk, float* out /*[R][4]*/) {
    const __m512i half_sel = _mm512_set_epi32(1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0);
    const __m512i bias = _mm512_set1_epi8((char)0x80);
    __m512 acc[R][4];
    for (int r = 0; r < R; ++r) for (int j = 0; j < 4; ++j) acc[r][j] = _mm512_setzero_ps();
    for (uint64_t b = 0; b + 2 <= nblk; b += 2) {
        __m512i wu[R];
        __m512 dwv[R];
        for (int r = 0; r < R; ++r) {
            const uint8_t* p = wr + r * rb + b * 34;
            const __m256i lo = _mm256_loadu_si256((const __m256i*)(p + 2)), hi = _mm256_loadu_si256((const __m256i*)(p + 36));
            wu[r] = _mm512_xor_si512(_mm512_inserti64x4(_mm512_castsi256_si512(lo), hi, 1), bias);
            dwv[r] = pair_scale(dws + r * nblk + b, half_sel);
        }
#pragma GCC unroll 4
        for (int j = 0; j < 4; ++j) {
            const __m512i xv = _mm512_loadu_si512((const void*)(xq + (size_t)j * K + b * 32));
            const __m512 dxv = pair_scale(xd + (size_t)j * nblk + b, half_sel);
            for (int r = 0; r < R; ++r) {
                const __m512i dp = _mm512_dpbusd_epi32(_mm512_setzero_si512(), wu[r], xv);
                acc[r][j] = _mm512_fmadd_ps(_mm512_cvtepi32_ps(dp), _mm512_mul_ps(dwv[r], dxv), acc[r][j]);
            }
        }
    }
    for (int r = 0; r < R; ++r)
        for (int j = 0; j < 4; ++j) {
            const float* dw = dws + r * nblk;
            const float* c = xc + (size_t)j * nblk;
            __m512 corr = _mm512_setzero_ps();
            uint64_t b = 0;
            for (; b + 16 <= nblk; b += 16) corr = _mm512_fmadd_ps(_mm512_loadu_ps(dw + b), _mm512_loadu_ps(c + b), corr);
            float tail = 0.f;
            for (; b < nblk; ++b) tail += dw[b] * c[b];
            out[r * 4 + j] = _mm512_reduce_add_ps(_mm512_add_ps(acc[r][j], corr)) + tail;
        }
}

DLP_VNNI void matmul_q8_0_vnni(const Tensor& w, const QAct& a, int T, float* y, bool accumulate) {
    const uint64_t K = w.cols(), N = w.rows(), nblk = K / 32;
    const size_t rb = row_bytes(w.type, K);
    constexpr int64_t PANEL = 32;
    const int64_t n_panels = ((int64_t)N + PANEL - 1) / PANEL;
#pragma omp parallel
    {
        std::vector<float> dws((size_t)PANEL * nblk);
#pragma omp for schedule(dynamic, 1)
        for (int64_t pi = 0; pi < n_panels; ++pi) {
            const int64_t n0 = pi * PANEL, n1 = std::min<int64_t>(N, n0 + PANEL);
            int t = 0;
            if (T >= 4)      // the panel's block scales as fp32, once for all token groups
                for (int64_t n = n0; n < n1; ++n)
                    for (uint64_t b = 0; b < nblk; ++b) { uint16_t dh; memcpy(&dh, w.data + n * rb + b * 34, 2); dws[(size_t)(n - n0) * nblk + b] = f16(dh); }
            for (; t + 4 <= T; t += 4) {
                const int8_t* xq = a.q.data() + (size_t)t * K;
                const float* xd = a.d.data() + (size_t)t * nblk;
                const float* xc = a.c.data() + (size_t)t * nblk;
                float out[16];
                int64_t n = n0;
                for (; n + 4 <= n1; n += 4) {
                    tile_q8_0_vnni<4>(w.data + n * rb, rb, dws.data() + (size_t)(n - n0) * nblk, xq, xd, xc, K, nblk, out);
                    for (int r = 0; r < 4; ++r)
                        for (int j = 0; j < 4; ++j) {
                            float& d = y[(size_t)(t + j) * N + n + r];
                            d = accumulate ? d + out[r * 4 + j] : out[r * 4 + j];
                        }
                }
                for (; n < n1; ++n) {
                    tile_q8_0_vnni<1>(w.data + n * rb, rb, dws.data() + (size_t)(n - n0) * nblk, xq, xd, xc, K, nblk, out);
                    for (int j = 0; j < 4; ++j) {
                        float& d = y[(size_t)(t + j) * N + n];
                        d = accumulate ? d + out[j] : out[j];
                    }
                }
            }
            for (; t < T; ++t)
                for (int64_t n = n0; n < n1; ++n) {
                    const float v = dot_q8_0_row(w.data + n * rb, a.q.data() + (size_t)t * K, a.d.data() + (size_t)t * nblk, nblk);
                    if (accumulate) y[(size_t)t * N + n] += v; else y[(size_t)t * N + n] = v;
                }
        }
    }
}

bool have_vnni() {
    static const bool ok = __builtin_cpu_supports("avx512vnni") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl") &&
                           __builtin_cpu_supports("avx512dq") && getenv("DLP_WORKER_NO_VNNI") == nullptr;
    return ok;
}

// An activation matrix used by several mat-muls (q / k / v, gate / up): its int8 form is made once, on first need.
struct Act {
    const float* x; int T; uint64_t K; QAct qa; bool have = false;
    Act(const float* x_, int T_, uint64_t K_) : x(x_), T(T_), K(K_) {}
    const QAct& q8() { if (!have) { quantize_act(x, T, K, qa); have = true; } return qa; }
};

// y[T][N] = x[T][K] . W[N][K]^T     (W decoded row by row, shared across the T tokens)
void matmul(const Tensor& w, Act& act, float* y, bool accumulate = false) {
    const uint64_t K = w.cols(), N = w.rows();
    const float* x = act.x;
    const int T = act.T;
    if (w.type == T_Q8_0 && K % 32 == 0) {
        if (T >= 4 && K % 64 == 0 && have_vnni()) matmul_q8_0_vnni(w, act.q8(), T, y, accumulate);
        else matmul_q8_0(w, act.q8(), T, y, accumulate);
        return;
    }
    const size_t rb = row_bytes(w.type, K);
#pragma omp parallel
    {
        std::vector<float> row(K);
#pragma omp for schedule(static)
        for (int64_t n = 0; n < (int64_t)N; ++n) {
            const float* r;
            if (w.type == T_F32) r = (const float*)(w.data + n * rb);
            else { decode_row(w.type, w.data + n * rb, K, row.data()); r = row.data(); }
            for (int t = 0; t < T; ++t) {
                const float v = dot(r, x + (size_t)t * K, K);
                if (accumulate) y[(size_t)t * N + n] += v; else y[(size_t)t * N + n] = v;
            }
        }
    }
}

void rmsnorm(const float* x, const float* w, float eps, int T, int E, float* out) {
#pragma omp parallel for schedule(static) if (T >= 8)
    for (int t = 0; t < T; ++t) {
        const float* r = x + (size_t)t * E;
        double ss = 0;
        for (int i = 0; i < E; ++i) ss += (double)r[i] * r[i];
        const float s = 1.0f / std::sqrt((float)(ss / E) + eps);
        for (int i = 0; i < E; ++i) out[(size_t)t * E + i] = r[i] * s * w[i];
    }
}

// ---- model stage ------------------------------------------------------------------------------------------------------
struct Hparams { int n_embd, n_layer, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_rot, n_ctx; float eps, theta;
                 bool neox = false, qkv_bias = false, qk_norm = false, fused_qkv = false, fused_gate_up = false;
                 // granite / minicpm scalars (1 = off): h = embed * emb_scale; h += res_scale * branch; logits /= logit_scale; scores * attn_scale
                 float emb_scale = 1.f, res_scale = 1.f, logit_scale = 1.f, attn_scale = 0.f;
                 std::string ffn_norm = "ffn_norm.weight"; };

struct SeqCache { std::vector<std::vector<float>> k, v; int len = 0; };  // per local layer: [n_ctx][Hkv*D]

struct Stage {
    Gguf g;
    Hparams hp{};
    int l0 = 0, l1 = 0;
    bool first = false, last = false;
    std::vector<float> inv_freq;
    std::map<uint32_t, SeqCache> seqs;
    Prof prof;
    int max_seqs = 64;           // --max-seqs: cap on live sequences (each holds an n_ctx KV cache)

    void load(const std::string& path, int a, int b, int n_ctx) {
        g.open(path);
        const std::string ar = g.str.count("general.architecture") ? g.str["general.architecture"] : "llama";
        // dense RMSNorm + SwiGLU graphs (the fused-path families of models/config.py without routed experts)
        static const char* norm_rope[] = {"llama", "granite", "minicpm", "internlm2", "xverse", "ernie4_5", "baichuan", "mistral3"};
        static const char* neox_rope[] = {"qwen2", "qwen3", "phi3", "exaone", "seed_oss", "qwen"};
        bool known = false;
        for (const char* n : norm_rope) if (ar == n) { known = true; hp.neox = false; }
        for (const char* n : neox_rope) if (ar == n) { known = true; hp.neox = true; }
        if (!known) throw std::runtime_error("the CPU worker runs dense RMSNorm / SwiGLU checkpoints (llama, granite, minicpm, internlm2, xverse, "
                                             "ernie4_5, baichuan, mistral3, qwen, qwen2, qwen3, phi3, exaone, seed_oss); got " + ar);
        hp.qkv_bias = g.has("blk.0.attn_q.bias") || g.has("blk.0.attn_qkv.bias");   // qwen2, seed_oss; qwen: on the fused tensor
        hp.qk_norm = g.has("blk.0.attn_q_norm.weight");            // qwen3
        hp.fused_qkv = g.has("blk.0.attn_qkv.weight");             // phi3, qwen
        hp.fused_gate_up = ar == "phi3";
        if (ar == "seed_oss") hp.ffn_norm = "post_attention_norm.weight";
        auto N = [&](const char* k, double d) { return g.n(ar + "." + k, d); };
        hp.n_embd = (int)N("embedding_length", 0); hp.n_layer = (int)N("block_count", 0);
        hp.n_head = (int)N("attention.head_count", 0); hp.n_head_kv = (int)N("attention.head_count_kv", hp.n_head);
        hp.head_dim = (int)N("attention.key_length", hp.n_embd / std::max(1, hp.n_head));
        hp.n_ff = (int)N("feed_forward_length", 0) / (ar == "qwen" ? 2 : 1); hp.eps = (float)N("attention.layer_norm_rms_epsilon", 1e-5);
        hp.theta = (float)N("rope.freq_base", 10000.0); hp.n_rot = (int)N("rope.dimension_count", hp.head_dim);
        hp.n_vocab = (int)g.get("token_embd.weight").rows();
        hp.n_ctx = n_ctx;
        if (ar == "minicpm") { hp.emb_scale = 12.f; hp.res_scale = 1.4f / std::sqrt((float)hp.n_layer); hp.logit_scale = 256.f / hp.n_embd; }
        if (ar == "granite" || ar == "minicpm") {
            hp.emb_scale = (float)N("embedding_scale", hp.emb_scale); hp.res_scale = (float)N("residual_scale", hp.res_scale);
            hp.logit_scale = (float)N("logit_scale", hp.logit_scale); hp.attn_scale = (float)N("attention.scale", 0.0);
        }
        if (N("expert_count", 0) > 0) throw std::runtime_error("MoE checkpoints are not supported by the CPU worker");
        l0 = a; l1 = b < 0 ? hp.n_layer : b;
        if (l0 < 0 || l1 > hp.n_layer || l0 >= l1) throw std::runtime_error("bad layer range");
        first = l0 == 0; last = l1 == hp.n_layer;
        inv_freq.resize(hp.n_rot / 2);
        for (int i = 0; i < hp.n_rot / 2; ++i) inv_freq[i] = std::pow(hp.theta, -2.0f * i / hp.n_rot);
        if (g.has("rope_freqs.weight")) { const float* ff = (const float*)g.get("rope_freqs.weight").data; for (size_t i = 0; i < inv_freq.size(); ++i) inv_freq[i] /= ff[i]; }
    }

    SeqCache& seq(uint32_t id) {
        auto it = seqs.find(id);
        if (it != seqs.end()) return it->second;
        // every live sequence owns a full n_ctx KV cache: an unauthenticated peer must not be able to allocate without bound
        if (seqs.size() >= (size_t)max_seqs) throw std::runtime_error("too many live sequences (limit " + std::to_string(max_seqs) + "); RESET one first");
        SeqCache& c = seqs[id];
        const size_t kv = (size_t)hp.n_head_kv * hp.head_dim;
        c.k.assign(l1 - l0, std::vector<float>((size_t)hp.n_ctx * kv));
        c.v.assign(l1 - l0, std::vector<float>((size_t)hp.n_ctx * kv));
        return c;
    }

    void rope(float* x, int n_heads, const float* cs, const float* sn) const {
        const int D = hp.head_dim, R2 = hp.n_rot / 2;
        for (int h = 0; h < n_heads; ++h)
            for (int i = 0; i < R2; ++i) {
                float* p0 = x + (size_t)h * D + (hp.neox ? i : 2 * i);
                float* p1 = hp.neox ? p0 + R2 : p0 + 1;
                const float x0 = *p0, x1 = *p1;
                *p0 = x0 * cs[i] - x1 * sn[i]; *p1 = x0 * sn[i] + x1 * cs[i];
            }
    }

    // causal attention of T new queries (positions pos0 ..) over the fp32 cache of one layer; one (token, head) per task
    void attention(const float* q, const float* kc, const float* vc, float* att, int T, int pos0, float scale) const {
        const int D = hp.head_dim, Hq = hp.n_head, Hkv = hp.n_head_kv, qd = Hq * D, kvd = Hkv * D, gqa = Hq / Hkv;
#pragma omp parallel
        {
            std::vector<float> s((size_t)pos0 + T);
#pragma omp for collapse(2) schedule(dynamic, 8)
            for (int t = 0; t < T; ++t)
                for (int hq = 0; hq < Hq; ++hq) {
                    const int n_keys = pos0 + t + 1, kvh = hq / gqa;
                    const float* qv = q + (size_t)t * qd + (size_t)hq * D;
                    float mx = -INFINITY;
                    for (int j = 0; j < n_keys; ++j) { s[j] = scale * dot(qv, kc + (size_t)j * kvd + (size_t)kvh * D, D); mx = std::max(mx, s[j]); }
                    float sum = 0;
                    for (int j = 0; j < n_keys; ++j) { s[j] = std::exp(s[j] - mx); sum += s[j]; }
                    float* o = att + (size_t)t * qd + (size_t)hq * D;
                    for (int d = 0; d < D; ++d) o[d] = 0;
                    const float inv = 1.0f / sum;
                    for (int j = 0; j < n_keys; ++j) {
                        const float w = s[j] * inv;
                        const float* vv = vc + (size_t)j * kvd + (size_t)kvh * D;
#pragma omp simd
                        for (int d = 0; d < D; ++d) o[d] += w * vv[d];
                    }
                }
        }
    }

    // in: token ids (first stage) or hidden [T][E]; out: hidden [T][E] or logits ([1|T][vocab]) on the last stage
    std::vector<float> run(uint32_t sid, int T, int pos0, uint32_t flags, const uint8_t* payload, size_t nbytes) {
        const int E = hp.n_embd, D = hp.head_dim, Hq = hp.n_head, Hkv = hp.n_head_kv, F = hp.n_ff;
        if (T <= 0 || T > hp.n_ctx) throw std::runtime_error("bad token count");
        if (pos0 < 0 || pos0 > hp.n_ctx - T) throw std::runtime_error("context overflow");   // also rejects wire values >= 2^31
        SeqCache& c = seq(sid);
        if (pos0 != c.len) throw std::runtime_error("position " + std::to_string(pos0) + " does not continue sequence of length " + std::to_string(c.len));
        if (pos0 + T > hp.n_ctx) throw std::runtime_error("context overflow");
        std::vector<float> h((size_t)T * E);
        if (first) {
            if (nbytes != (size_t)T * 4) throw std::runtime_error("payload size mismatch (token ids)");
            const Tensor& emb = g.get("token_embd.weight");
            const size_t rb = row_bytes(emb.type, E);
            for (int t = 0; t < T; ++t) {
                int32_t id; memcpy(&id, payload + (size_t)t * 4, 4);
                if (id < 0 || id >= hp.n_vocab) throw std::runtime_error("token id out of range");
                decode_row(emb.type, emb.data + (size_t)id * rb, E, h.data() + (size_t)t * E);
            }
            if (hp.emb_scale != 1.f) for (float& x : h) x *= hp.emb_scale;
        } else {
            if (nbytes != (size_t)T * E * 4) throw std::runtime_error("payload size mismatch (hidden states)");
            memcpy(h.data(), payload, nbytes);
        }
        const int qd = Hq * D, kvd = Hkv * D, gqa = Hq / Hkv;
        std::vector<float> xn((size_t)T * E), q((size_t)T * qd), k((size_t)T * kvd), v((size_t)T * kvd), att((size_t)T * qd),
            gate((size_t)T * F), up((size_t)T * F);
        const float scale = hp.attn_scale > 0.f ? hp.attn_scale : 1.0f / std::sqrt((float)D);
        std::vector<float> branch(hp.res_scale != 1.f ? (size_t)T * E : 0);
        auto add_branch = [&](const Tensor& w, Act& a) {      // h += res_scale * (a . W^T)
            if (hp.res_scale == 1.f) { matmul(w, a, h.data(), true); return; }
            matmul(w, a, branch.data());
            for (size_t i = 0; i < branch.size(); ++i) h[i] += hp.res_scale * branch[i];
        };
        // rotation angles depend on (position, pair) only: one table per call instead of cos/sin per layer and head
        const int R2 = hp.n_rot / 2;
        std::vector<float> rc((size_t)T * R2), rs((size_t)T * R2);
#pragma omp parallel for schedule(static) if (T >= 8)
        for (int t = 0; t < T; ++t)
            for (int i = 0; i < R2; ++i) { const float a = (pos0 + t) * inv_freq[i]; rc[(size_t)t * R2 + i] = std::cos(a); rs[(size_t)t * R2 + i] = std::sin(a); }
        prof.start();
        for (int il = l0; il < l1; ++il) {
            const std::string p = "blk." + std::to_string(il) + ".";
            rmsnorm(h.data(), (const float*)g.get(p + "attn_norm.weight").data, hp.eps, T, E, xn.data());
            prof.lap(2);
            {
                Act a(xn.data(), T, E);
                if (hp.fused_qkv) {
                    const Tensor& w = g.get(p + "attn_qkv.weight");
                    matmul(slice_rows(w, 0, qd), a, q.data());
                    matmul(slice_rows(w, qd, qd + kvd), a, k.data());
                    matmul(slice_rows(w, qd + kvd, qd + 2 * kvd), a, v.data());
                } else {
                    matmul(g.get(p + "attn_q.weight"), a, q.data());
                    matmul(g.get(p + "attn_k.weight"), a, k.data());
                    matmul(g.get(p + "attn_v.weight"), a, v.data());
                }
            }
            prof.lap(0);
            if (hp.qkv_bias) {
                const bool fb = g.has(p + "attn_qkv.bias");       // qwen: one bias laid out like the fused matrix
                const float* bq = (const float*)g.get(p + (fb ? "attn_qkv.bias" : "attn_q.bias")).data;
                const float* bk = fb ? bq + qd : (const float*)g.get(p + "attn_k.bias").data;
                const float* bv = fb ? bq + qd + kvd : (const float*)g.get(p + "attn_v.bias").data;
                for (int t = 0; t < T; ++t) {
                    for (int i = 0; i < qd; ++i) q[(size_t)t * qd + i] += bq[i];
                    for (int i = 0; i < kvd; ++i) { k[(size_t)t * kvd + i] += bk[i]; v[(size_t)t * kvd + i] += bv[i]; }
                }
            }
            if (hp.qk_norm) {   // per-head RMSNorm over head_dim, before the rotation
                std::vector<float> tmp(std::max(qd, kvd));
                for (int t = 0; t < T; ++t) {
                    rmsnorm(q.data() + (size_t)t * qd, (const float*)g.get(p + "attn_q_norm.weight").data, hp.eps, Hq, D, tmp.data());
                    memcpy(q.data() + (size_t)t * qd, tmp.data(), (size_t)qd * 4);
                    rmsnorm(k.data() + (size_t)t * kvd, (const float*)g.get(p + "attn_k_norm.weight").data, hp.eps, Hkv, D, tmp.data());
                    memcpy(k.data() + (size_t)t * kvd, tmp.data(), (size_t)kvd * 4);
                }
            }
            std::vector<float>& kc = c.k[il - l0];
            std::vector<float>& vc = c.v[il - l0];
#pragma omp parallel for schedule(static) if (T >= 8)
            for (int t = 0; t < T; ++t) {
                rope(q.data() + (size_t)t * qd, Hq, rc.data() + (size_t)t * R2, rs.data() + (size_t)t * R2);
                rope(k.data() + (size_t)t * kvd, Hkv, rc.data() + (size_t)t * R2, rs.data() + (size_t)t * R2);
                memcpy(kc.data() + (size_t)(pos0 + t) * kvd, k.data() + (size_t)t * kvd, kvd * 4);
                memcpy(vc.data() + (size_t)(pos0 + t) * kvd, v.data() + (size_t)t * kvd, kvd * 4);
            }
            prof.lap(3);
            attention(q.data(), kc.data(), vc.data(), att.data(), T, pos0, scale);
            prof.lap(1);
            {
                Act a(att.data(), T, qd);
                add_branch(g.get(p + "attn_output.weight"), a);
            }
            prof.lap(0);
            rmsnorm(h.data(), (const float*)g.get(p + hp.ffn_norm).data, hp.eps, T, E, xn.data());
            prof.lap(2);
            {
                Act a(xn.data(), T, E);
                if (hp.fused_gate_up) {   // phi3: ffn_up = [gate | up]
                    const Tensor& w = g.get(p + "ffn_up.weight");
                    matmul(slice_rows(w, 0, F), a, gate.data());
                    matmul(slice_rows(w, F, 2 * (uint64_t)F), a, up.data());
                } else {
                    matmul(g.get(p + "ffn_gate.weight"), a, gate.data());
                    matmul(g.get(p + "ffn_up.weight"), a, up.data());
                }
            }
            prof.lap(0);
            const int64_t nact = (int64_t)T * F;
#pragma omp parallel for simd schedule(static) if (nact >= 16384)
            for (int64_t i = 0; i < nact; ++i) gate[i] = gate[i] / (1.0f + std::exp(-gate[i])) * up[i];
            prof.lap(4);
            {
                Act a(gate.data(), T, F);
                add_branch(g.get(p + "ffn_down.weight"), a);
            }
            prof.lap(0);
        }
        c.len = pos0 + T;
        if (!last) { prof.report(T); return h; }
        const bool all = flags & 1u;
        const int R = all ? T : 1;
        const float* src = all ? h.data() : h.data() + (size_t)(T - 1) * E;
        std::vector<float> hn((size_t)R * E), logits((size_t)R * hp.n_vocab);
        rmsnorm(src, (const float*)g.get("output_norm.weight").data, hp.eps, R, E, hn.data());
        Act a(hn.data(), R, E);
        matmul(g.has("output.weight") ? g.get("output.weight") : g.get("token_embd.weight"), a, logits.data());
        if (hp.logit_scale != 1.f) for (float& x : logits) x /= hp.logit_scale;
        prof.lap(5);
        prof.report(T);
        return logits;
    }
};

// ---- wire ----------------------------------------------------------------------------------------------------------------
bool recv_all(int fd, void* p, size_t n) { uint8_t* b = (uint8_t*)p; while (n) { ssize_t r = ::recv(fd, b, n, 0); if (r <= 0) return false; b += r; n -= (size_t)r; } return true; }
bool send_all(int fd, const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; while (n) { ssize_t r = ::send(fd, b, n, MSG_NOSIGNAL); if (r <= 0) return false; b += r; n -= (size_t)r; } return true; }
bool reply(int fd, uint32_t status, const void* p, uint64_t n) { return send_all(fd, &status, 4) && send_all(fd, &n, 8) && (n == 0 || send_all(fd, p, n)); }

#pragma pack(push, 1)
struct ReqHeader { uint8_t cmd; uint32_t seq, n_tokens, pos0, flags; uint64_t nbytes; };
struct Info { uint32_t version, n_embd, n_vocab, n_layer, l0, l1, first, last, n_ctx; };
#pragma pack(pop)

bool serve_client(int fd, Stage& st, bool verbose) {
    std::vector<uint8_t> payload;
    bool hello = false;
    for (;;) {
        ReqHeader rq;
        if (!recv_all(fd, &rq, sizeof rq)) return true;
        if (rq.nbytes > (uint64_t)st.hp.n_ctx * st.hp.n_embd * 4 + 64) { reply(fd, 2, "payload too large", 17); return true; }
        payload.resize(rq.nbytes);
        if (rq.nbytes && !recv_all(fd, payload.data(), rq.nbytes)) return true;
        if (!hello && rq.cmd != HELLO) { reply(fd, 3, "HELLO first", 11); return true; }   // same rule as the reference (ggml-rpc.cpp:1584)
        const uint32_t seq_id = rq.seq;   // the wire struct is packed: never bind references to its (misaligned) members
        try {
            switch (rq.cmd) {
                case HELLO: {
                    uint32_t ver = 0;
                    if (rq.nbytes >= 4) memcpy(&ver, payload.data(), 4);
                    if (ver != PROTO_VERSION) { reply(fd, 4, "protocol version mismatch", 25); return true; }
                    hello = true;
                    Info in{PROTO_VERSION, (uint32_t)st.hp.n_embd, (uint32_t)st.hp.n_vocab, (uint32_t)st.hp.n_layer, (uint32_t)st.l0, (uint32_t)st.l1,
                            st.first, st.last, (uint32_t)st.hp.n_ctx};
                    reply(fd, 0, &in, sizeof in);
                    break;
                }
                case RUN: {
                    std::vector<float> out = st.run(seq_id, (int)rq.n_tokens, (int)rq.pos0, rq.flags, payload.data(), payload.size());
                    if (verbose) fprintf(stderr, "RPC worker[%d:%d) seq %u: %u tokens @%u -> %zu floats\n", st.l0, st.l1, seq_id, (unsigned)rq.n_tokens, (unsigned)rq.pos0, out.size());
                    if (!reply(fd, 0, out.data(), out.size() * 4)) return true;
                    break;
                }
                case RESET: st.seqs.erase(seq_id); reply(fd, 0, nullptr, 0); break;
                case TRUNCATE: {
                    const uint32_t new_len = rq.pos0;   // unsigned on the wire: compare before any narrowing
                    if (new_len > (uint32_t)st.hp.n_ctx) throw std::runtime_error("TRUNCATE: length beyond the context window");
                    auto it = st.seqs.find(seq_id);
                    if (it != st.seqs.end()) it->second.len = std::max(0, std::min<int>(it->second.len, (int)new_len));
                    reply(fd, 0, nullptr, 0);
                    break;
                }
                case SHUTDOWN: reply(fd, 0, nullptr, 0); return false;
                default: reply(fd, 5, "unknown command", 15); return true;
            }
        } catch (const std::exception& e) {
            const std::string m = e.what();
            if (!reply(fd, 1, m.data(), m.size())) return true;
        }
    }
}

}  // namespace

int main(int argc, char** argv) {
    signal(SIGPIPE, SIG_IGN);
    std::string model, host = "127.0.0.1";
    int port = 50052, l0 = 0, l1 = -1, n_ctx = 2048, max_seqs = 64, bench_T = 0, bench_reps = 3;
    bool verbose = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> std::string { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); } return argv[++i]; };
        if (a == "-m" || a == "--model") model = next();
        else if (a == "-H" || a == "--host") host = next();
        else if (a == "-p" || a == "--port") port = atoi(next().c_str());
        else if (a == "--layers") { std::string v = next(); size_t c = v.find(':'); l0 = atoi(v.substr(0, c).c_str()); l1 = c == std::string::npos ? -1 : atoi(v.substr(c + 1).c_str()); }
        else if (a == "-c" || a == "--ctx-size") n_ctx = atoi(next().c_str());
        else if (a == "-v" || a == "--verbose") verbose = true;
        else if (a == "--max-seqs") max_seqs = std::max(1, atoi(next().c_str()));
        else if (a == "--bench") bench_T = atoi(next().c_str());
        else if (a == "--bench-reps") bench_reps = std::max(1, atoi(next().c_str()));
        else { fprintf(stderr, "dlp-worker -m model.gguf [--layers a:b] [-H host] [-p port] [-c n_ctx] [--max-seqs n] [--bench T [--bench-reps r]] [-v]\n"); return a == "-h" ? 0 : 2; }
    }
    if (model.empty()) { fprintf(stderr, "dlp-worker: -m model.gguf is required\n"); return 2; }
    Stage st;
    st.max_seqs = max_seqs;
    try { st.load(model, l0, l1, n_ctx); } catch (const std::exception& e) { fprintf(stderr, "dlp-worker: %s\n", e.what()); return 1; }
    if (bench_T > 0) {   // no sockets: time this stage on T synthetic tokens (prefill) and on one more token (decode)
        st.prof.on = true;
        std::vector<uint8_t> in;
        if (st.first) { in.resize((size_t)bench_T * 4); for (int t = 0; t < bench_T; ++t) { const int32_t id = 3 + (t * 7919) % (st.hp.n_vocab - 3); memcpy(in.data() + (size_t)t * 4, &id, 4); } }
        else { std::vector<float> hx((size_t)bench_T * st.hp.n_embd); for (size_t i = 0; i < hx.size(); ++i) hx[i] = 0.01f * (float)((int)(i % 201) - 100); in.resize(hx.size() * 4); memcpy(in.data(), hx.data(), in.size()); }
        const size_t per_tok = in.size() / bench_T;
        for (int r = 0; r < bench_reps; ++r) {
            try {
                double t0 = omp_get_wtime();
                st.run(1, bench_T, 0, 0, in.data(), in.size());
                const double tp = omp_get_wtime() - t0;
                t0 = omp_get_wtime();
                if (bench_T < n_ctx) st.run(1, 1, bench_T, 0, in.data(), per_tok);
                const double tg = omp_get_wtime() - t0;
                fprintf(stderr, "bench: prefill %d tokens %.1f ms (%.1f tok/s), one decode step %.2f ms\n", bench_T, tp * 1e3, bench_T / tp, tg * 1e3);
                st.seqs.clear();
            } catch (const std::exception& e) { fprintf(stderr, "dlp-worker: %s\n", e.what()); return 1; }
        }
        return 0;
    }
    if (host != "127.0.0.1" && host != "localhost")
        fprintf(stderr, "WARNING: listening on %s - the worker protocol is unauthenticated; never expose it to an untrusted network\n", host.c_str());
    int ls = ::socket(AF_INET, SOCK_STREAM, 0), one = 1;
    setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr{};
    addr.sin_family = AF_INET; addr.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host == "localhost" ? "127.0.0.1" : host.c_str(), &addr.sin_addr) != 1) addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    if (bind(ls, (sockaddr*)&addr, sizeof addr) != 0 || listen(ls, 4) != 0) { perror("bind/listen"); return 1; }
    fprintf(stderr, "RPC worker ready on %s:%d | layers [%d, %d) of %d%s%s | n_ctx %d\n", host.c_str(), port, st.l0, st.l1, st.hp.n_layer,
            st.first ? " +embedding" : "", st.last ? " +lm_head" : "", n_ctx);
    for (;;) {
        int fd = ::accept(ls, nullptr, nullptr);
        if (fd < 0) continue;
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        const bool keep = serve_client(fd, st, verbose);
        ::close(fd);
        st.seqs.clear();  // a client's buffers do not outlive its connection (as the reference frees per client)
        if (!keep) break;
    }
    return 0;
}
