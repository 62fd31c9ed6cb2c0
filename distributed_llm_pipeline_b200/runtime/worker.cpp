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
struct QAct { std::vector<int8_t> q; std::vector<float> d; };

void quantize_act(const float* x, int T, uint64_t K, QAct& a) {
    a.q.resize((size_t)T * K);
    a.d.resize((size_t)T * (K / 32));
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)(T * (K / 32)); ++b) {
        const float* p = x + (size_t)b * 32;
        float amax = 0.f;
        for (int i = 0; i < 32; ++i) amax = std::max(amax, std::fabs(p[i]));
        const float d = amax / 127.f, inv = amax > 0 ? 127.f / amax : 0.f;
        a.d[b] = d;
        for (int i = 0; i < 32; ++i) a.q[(size_t)b * 32 + i] = (int8_t)std::lrintf(p[i] * inv);
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

void matmul_q8_0(const Tensor& w, const QAct& a, int T, float* y, bool accumulate) {
    const uint64_t K = w.cols(), N = w.rows(), nblk = K / 32;
    const size_t rb = row_bytes(w.type, K);
    const __m256i ones = _mm256_set1_epi16(1);
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        const uint8_t* wr = w.data + n * rb;
        int t = 0;
        for (; t + 4 <= T; t += 4) {   // register blocking: one pass over the weight row feeds four tokens
            __m256 acc[4] = {_mm256_setzero_ps(), _mm256_setzero_ps(), _mm256_setzero_ps(), _mm256_setzero_ps()};
            for (uint64_t b = 0; b < nblk; ++b) {
                const uint8_t* p = wr + b * 34;
                uint16_t dh; memcpy(&dh, p, 2);
                const float dw = f16(dh);
                const __m256i wv = _mm256_loadu_si256((const __m256i*)(p + 2));
                const __m256i aw = _mm256_sign_epi8(wv, wv);
#pragma GCC unroll 4
                for (int j = 0; j < 4; ++j) {
                    const __m256i xv = _mm256_loadu_si256((const __m256i*)(a.q.data() + (size_t)(t + j) * K + b * 32));
                    const __m256i p32 = _mm256_madd_epi16(_mm256_maddubs_epi16(aw, _mm256_sign_epi8(xv, wv)), ones);
                    acc[j] = _mm256_fmadd_ps(_mm256_cvtepi32_ps(p32), _mm256_set1_ps(dw * a.d[(size_t)(t + j) * nblk + b]), acc[j]);
                }
            }
            for (int j = 0; j < 4; ++j) {
                const float v = hsum256(acc[j]);
                if (accumulate) y[(size_t)(t + j) * N + n] += v; else y[(size_t)(t + j) * N + n] = v;
            }
        }
        for (; t < T; ++t) {
            const float v = dot_q8_0_row(wr, a.q.data() + (size_t)t * K, a.d.data() + (size_t)t * nblk, nblk);
            if (accumulate) y[(size_t)t * N + n] += v; else y[(size_t)t * N + n] = v;
        }
    }
}

// y[T][N] = x[T][K] . W[N][K]^T     (W decoded row by row, shared across the T tokens)
void matmul(const Tensor& w, const float* x, int T, float* y, bool accumulate = false) {
    const uint64_t K = w.cols(), N = w.rows();
    if (w.type == T_Q8_0 && K % 32 == 0) {
        QAct a;
        quantize_act(x, T, K, a);
        matmul_q8_0(w, a, T, y, accumulate);
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
                 bool neox = false, qkv_bias = false, qk_norm = false, fused_qkv = false; };

struct SeqCache { std::vector<std::vector<float>> k, v; int len = 0; };  // per local layer: [n_ctx][Hkv*D]

struct Stage {
    Gguf g;
    Hparams hp{};
    int l0 = 0, l1 = 0;
    bool first = false, last = false;
    std::vector<float> inv_freq;
    std::map<uint32_t, SeqCache> seqs;
    int max_seqs = 64;           // --max-seqs: cap on live sequences (each holds an n_ctx KV cache)

    void load(const std::string& path, int a, int b, int n_ctx) {
        g.open(path);
        const std::string ar = g.str.count("general.architecture") ? g.str["general.architecture"] : "llama";
        if (ar != "llama" && ar != "qwen2" && ar != "qwen3" && ar != "phi3")
            throw std::runtime_error("the CPU worker runs llama / qwen2 / qwen3 / phi3 checkpoints (got " + ar + ")");
        hp.neox = ar != "llama";                                   // half-split rotation for the other families
        hp.qkv_bias = g.has("blk.0.attn_q.bias");                  // qwen2
        hp.qk_norm = g.has("blk.0.attn_q_norm.weight");            // qwen3
        hp.fused_qkv = g.has("blk.0.attn_qkv.weight");             // phi3
        auto N = [&](const char* k, double d) { return g.n(ar + "." + k, d); };
        hp.n_embd = (int)N("embedding_length", 0); hp.n_layer = (int)N("block_count", 0);
        hp.n_head = (int)N("attention.head_count", 0); hp.n_head_kv = (int)N("attention.head_count_kv", hp.n_head);
        hp.head_dim = (int)N("attention.key_length", hp.n_embd / std::max(1, hp.n_head));
        hp.n_ff = (int)N("feed_forward_length", 0); hp.eps = (float)N("attention.layer_norm_rms_epsilon", 1e-5);
        hp.theta = (float)N("rope.freq_base", 10000.0); hp.n_rot = (int)N("rope.dimension_count", hp.head_dim);
        hp.n_vocab = (int)g.get("token_embd.weight").rows();
        hp.n_ctx = n_ctx;
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

    void rope(float* x, int n_heads, int pos) const {
        const int D = hp.head_dim;
        for (int h = 0; h < n_heads; ++h)
            for (int i = 0; i < hp.n_rot / 2; ++i) {
                const float a = pos * inv_freq[i], c = std::cos(a), s = std::sin(a);
                float* p0 = x + (size_t)h * D + (hp.neox ? i : 2 * i);
                float* p1 = hp.neox ? p0 + hp.n_rot / 2 : p0 + 1;
                const float x0 = *p0, x1 = *p1;
                *p0 = x0 * c - x1 * s; *p1 = x0 * s + x1 * c;
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
        } else {
            if (nbytes != (size_t)T * E * 4) throw std::runtime_error("payload size mismatch (hidden states)");
            memcpy(h.data(), payload, nbytes);
        }
        const int qd = Hq * D, kvd = Hkv * D, gqa = Hq / Hkv;
        std::vector<float> xn((size_t)T * E), q((size_t)T * qd), k((size_t)T * kvd), v((size_t)T * kvd), att((size_t)T * qd),
            gate((size_t)T * F), up((size_t)T * F);
        const float scale = 1.0f / std::sqrt((float)D);
        for (int il = l0; il < l1; ++il) {
            const std::string p = "blk." + std::to_string(il) + ".";
            rmsnorm(h.data(), (const float*)g.get(p + "attn_norm.weight").data, hp.eps, T, E, xn.data());
            if (hp.fused_qkv) {
                const Tensor& w = g.get(p + "attn_qkv.weight");
                matmul(slice_rows(w, 0, qd), xn.data(), T, q.data());
                matmul(slice_rows(w, qd, qd + kvd), xn.data(), T, k.data());
                matmul(slice_rows(w, qd + kvd, qd + 2 * kvd), xn.data(), T, v.data());
            } else {
                matmul(g.get(p + "attn_q.weight"), xn.data(), T, q.data());
                matmul(g.get(p + "attn_k.weight"), xn.data(), T, k.data());
                matmul(g.get(p + "attn_v.weight"), xn.data(), T, v.data());
            }
            if (hp.qkv_bias) {
                const float* bq = (const float*)g.get(p + "attn_q.bias").data;
                const float* bk = (const float*)g.get(p + "attn_k.bias").data;
                const float* bv = (const float*)g.get(p + "attn_v.bias").data;
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
            for (int t = 0; t < T; ++t) {
                rope(q.data() + (size_t)t * qd, Hq, pos0 + t);
                rope(k.data() + (size_t)t * kvd, Hkv, pos0 + t);
                memcpy(kc.data() + (size_t)(pos0 + t) * kvd, k.data() + (size_t)t * kvd, kvd * 4);
                memcpy(vc.data() + (size_t)(pos0 + t) * kvd, v.data() + (size_t)t * kvd, kvd * 4);
            }
#pragma omp parallel for collapse(2) schedule(static)
            for (int t = 0; t < T; ++t)
                for (int hq = 0; hq < Hq; ++hq) {
                    const int n_keys = pos0 + t + 1, kvh = hq / gqa;
                    const float* qv = q.data() + (size_t)t * qd + (size_t)hq * D;
                    std::vector<float> s(n_keys);
                    float mx = -INFINITY;
                    for (int j = 0; j < n_keys; ++j) { s[j] = scale * dot(qv, kc.data() + (size_t)j * kvd + (size_t)kvh * D, D); mx = std::max(mx, s[j]); }
                    float sum = 0;
                    for (int j = 0; j < n_keys; ++j) { s[j] = std::exp(s[j] - mx); sum += s[j]; }
                    float* o = att.data() + (size_t)t * qd + (size_t)hq * D;
                    for (int d = 0; d < D; ++d) o[d] = 0;
                    for (int j = 0; j < n_keys; ++j) { const float w = s[j] / sum; const float* vv = vc.data() + (size_t)j * kvd + (size_t)kvh * D; for (int d = 0; d < D; ++d) o[d] += w * vv[d]; }
                }
            matmul(g.get(p + "attn_output.weight"), att.data(), T, h.data(), true);
            rmsnorm(h.data(), (const float*)g.get(p + "ffn_norm.weight").data, hp.eps, T, E, xn.data());
            if (hp.fused_qkv) {   // phi3: ffn_up = [gate | up]
                const Tensor& w = g.get(p + "ffn_up.weight");
                matmul(slice_rows(w, 0, F), xn.data(), T, gate.data());
                matmul(slice_rows(w, F, 2 * (uint64_t)F), xn.data(), T, up.data());
            } else {
                matmul(g.get(p + "ffn_gate.weight"), xn.data(), T, gate.data());
                matmul(g.get(p + "ffn_up.weight"), xn.data(), T, up.data());
            }
            for (size_t i = 0; i < (size_t)T * F; ++i) gate[i] = gate[i] / (1.0f + std::exp(-gate[i])) * up[i];
            matmul(g.get(p + "ffn_down.weight"), gate.data(), T, h.data(), true);
        }
        c.len = pos0 + T;
        if (!last) return h;
        const bool all = flags & 1u;
        const int R = all ? T : 1;
        const float* src = all ? h.data() : h.data() + (size_t)(T - 1) * E;
        std::vector<float> hn((size_t)R * E), logits((size_t)R * hp.n_vocab);
        rmsnorm(src, (const float*)g.get("output_norm.weight").data, hp.eps, R, E, hn.data());
        matmul(g.has("output.weight") ? g.get("output.weight") : g.get("token_embd.weight"), hn.data(), R, logits.data());
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
    int port = 50052, l0 = 0, l1 = -1, n_ctx = 2048, max_seqs = 64;
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
        else { fprintf(stderr, "dlp-worker -m model.gguf [--layers a:b] [-H host] [-p port] [-c n_ctx] [--max-seqs n] [-v]\n"); return a == "-h" ? 0 : 2; }
    }
    if (model.empty()) { fprintf(stderr, "dlp-worker: -m model.gguf is required\n"); return 2; }
    Stage st;
    st.max_seqs = max_seqs;
    try { st.load(model, l0, l1, n_ctx); } catch (const std::exception& e) { fprintf(stderr, "dlp-worker: %s\n", e.what()); return 1; }
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
