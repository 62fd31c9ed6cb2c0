// dlp-run: native single-GPU decode runner - GGUF in, token ids out, no Python and no PyTorch in the process.
//
// What it is for: the reference's inference runtime is native (libllama + llama-cli, llama.cpp/src/llama-context.cpp,
// tools/cli/cli.cpp); this binary is the same layer of this framework - it maps the GGUF file, repacks and uploads the weights
// (the rules of ops/weights.py), owns the paged KV cache and drives the sm_100a kernels of libdlp_b200.so through the C ABI
// (include/dlp_b200.h).  The decode step (bookkeeping + RoPE table -> embedding -> per layer: fused-norm QKV mat-vec ->
// one-launch RoPE / KV write / split-KV attention with cluster combine -> Wo (+residual) -> gate|up SwiGLU -> down (+residual)
// -> fused-norm lm_head -> arg-max) is captured once into a CUDA graph and replayed; the sampled token never leaves the GPU.
// Scope: llama-architecture dense checkpoints (F16 / BF16 / Q8_0 / Q4_K / Q5_K / Q6_K tensors), one sequence, greedy.  The Python
// engine (models/llama.py) is the full-featured path; tests/test_native_runner.py checks the two produce the same tokens.
//
//   dlp-run -m model.gguf -t 1,5,9,200 -n 32 [-c 2048] [--bench]        prints one JSON line
// Build: see runtime/native.py (g++ + libcudart + libdlp_b200.so).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <sstream>

#include "../../include/dlp_b200.h"
#include "gguf_lite.h"

using namespace dlp_gguf;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "dlp-run: %s failed: %s\n", #x, cudaGetErrorString(e_)); exit(3); } } while (0)
#define DK(x) do { int r_ = (x); if (r_ != 0) { fprintf(stderr, "dlp-run: %s failed (%d): %s\n", #x, r_, dlp_last_error()); exit(4); } } while (0)

namespace {

template <typename T> T* dev_alloc(size_t n) { void* p = nullptr; CK(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); CK(cudaMemset(p, 0, std::max<size_t>(n, 1) * sizeof(T))); return (T*)p; }
void* upload(const void* host, size_t bytes) { void* p = nullptr; CK(cudaMalloc(&p, bytes)); CK(cudaMemcpy(p, host, bytes, cudaMemcpyHostToDevice)); return p; }

inline uint16_t f32_to_f16(float f) {   // round-to-nearest-even, enough for norm-free fp32 tensors stored as fp16
    uint32_t x; memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    int32_t exp = (int32_t)((x >> 23) & 0xFF) - 127 + 15;
    uint32_t man = x & 0x7FFFFFu;
    if (exp <= 0) { if (exp < -10) return (uint16_t)sign; man |= 0x800000u; const uint32_t sh = (uint32_t)(14 - exp); uint32_t h = man >> sh; if ((man >> (sh - 1)) & 1u) h += 1; return (uint16_t)(sign | h); }
    if (exp >= 31) return (uint16_t)(sign | 0x7C00u);
    uint32_t h = sign | ((uint32_t)exp << 10) | (man >> 13);
    if ((man & 0x1000u) && ((man & 0x2FFFu) != 0)) h += 1;
    return (uint16_t)h;
}

// One or more row-stacked matrices of the same type -> kernel layout on the device (planes: see csrc/quant.cuh)
dlp_weight make_weight(const std::vector<Tensor>& parts) {
    const uint32_t type = parts[0].type;
    const uint64_t K = parts[0].cols();
    uint64_t N = 0;
    for (const auto& t : parts) { if (t.type != type || t.cols() != K) throw std::runtime_error("stacked tensors differ in type or width"); N += t.rows(); }
    if (K % 256) throw std::runtime_error("K must be a multiple of 256 for the device kernels");
    dlp_weight w{};
    w.N = (int32_t)N; w.K = (int32_t)K; w.wtype = (int32_t)type;
    const size_t rb = row_bytes(type, K);
    std::vector<uint8_t> raw(N * rb);
    size_t off = 0;
    for (const auto& t : parts) { memcpy(raw.data() + off, t.data, t.rows() * rb); off += t.rows() * rb; }
    if (type == T_F16 || type == T_BF16 || type == T_Q4_K || type == T_Q5_K) {
        w.base = upload(raw.data(), raw.size());
    } else if (type == T_F32) {
        std::vector<uint16_t> h(N * K);
        const float* f = (const float*)raw.data();
        for (size_t i = 0; i < N * K; ++i) h[i] = f32_to_f16(f[i]);
        w.wtype = T_F16;
        w.base = upload(h.data(), h.size() * 2);
    } else if (type == T_Q8_0) {          // [N][K/32] blocks of (fp16 d | 32 x int8) -> qs [N][K] + d [N][K/32]
        const size_t nb = N * (K / 32);
        std::vector<int8_t> qs(N * K);
        std::vector<uint16_t> d(nb);
        for (size_t b = 0; b < nb; ++b) { memcpy(&d[b], raw.data() + b * 34, 2); memcpy(qs.data() + b * 32, raw.data() + b * 34 + 2, 32); }
        w.base = upload(qs.data(), qs.size());
        w.aux0 = upload(d.data(), d.size() * 2);
    } else if (type == T_Q6_K) {          // 210-byte super-blocks -> ql [N][K/2] | qh [N][K/4] | scales [N][K/16] | d [N][K/256]
        const size_t nb = N * (K / 256);
        std::vector<uint8_t> ql(nb * 128), qh(nb * 64);
        std::vector<int8_t> sc(nb * 16);
        std::vector<uint16_t> d(nb);
        for (size_t b = 0; b < nb; ++b) {
            const uint8_t* s = raw.data() + b * 210;
            memcpy(ql.data() + b * 128, s, 128); memcpy(qh.data() + b * 64, s + 128, 64); memcpy(sc.data() + b * 16, s + 192, 16); memcpy(&d[b], s + 208, 2);
        }
        w.base = upload(ql.data(), ql.size()); w.aux0 = upload(qh.data(), qh.size());
        w.aux1 = upload(sc.data(), sc.size()); w.aux2 = upload(d.data(), d.size() * 2);
    } else {
        throw std::runtime_error("tensor type " + std::to_string(type) + " is not supported by dlp-run");
    }
    return w;
}

// Mixed-type Q|K|V (Q4_K_M keeps attn_v at more bits in some layers): decoded once on the host and stacked as fp16 - the
// same rule as models/weights.py, so both run-times stream identical bytes through the identical kernel
dlp_weight make_weight_f16_decoded(const std::vector<Tensor>& parts) {
    const uint64_t K = parts[0].cols();
    uint64_t N = 0;
    for (const auto& t : parts) { if (t.cols() != K) throw std::runtime_error("stacked tensors differ in width"); N += t.rows(); }
    std::vector<uint16_t> h(N * K);
    std::vector<float> row(K);
    size_t r = 0;
    for (const auto& t : parts) {
        const size_t rb = row_bytes(t.type, K);
        for (uint64_t i = 0; i < t.rows(); ++i, ++r) {
            decode_row(t.type, t.data + i * rb, K, row.data());
            for (uint64_t k = 0; k < K; ++k) h[r * K + k] = f32_to_f16(row[k]);
        }
    }
    dlp_weight w{};
    w.N = (int32_t)N; w.K = (int32_t)K; w.wtype = T_F16;
    w.base = upload(h.data(), h.size() * 2);
    return w;
}

float* upload_f32(const Tensor& t) {
    if (t.type != T_F32) throw std::runtime_error("expected an fp32 tensor");
    return (float*)upload(t.data, t.cols() * t.rows() * 4);
}

struct Layer { dlp_weight wqkv, wq, wk, wv, wo, gate, up, down; bool stacked; float *attn_norm, *ffn_norm; void *kc, *vc; };

}  // namespace

int main(int argc, char** argv) {
    std::string model, tok_csv = "1";
    int n_predict = 16, n_ctx = 2048;
    bool bench = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() -> std::string { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); } return argv[++i]; };
        if (a == "-m" || a == "--model") model = next();
        else if (a == "-t" || a == "--tokens") tok_csv = next();
        else if (a == "-n" || a == "--predict") n_predict = atoi(next().c_str());
        else if (a == "-c" || a == "--ctx-size") n_ctx = atoi(next().c_str());
        else if (a == "--bench") bench = true;
        else { fprintf(stderr, "usage: dlp-run -m model.gguf [-t id,id,...] [-n N] [-c CTX] [--bench]\n"); return 2; }
    }
    if (model.empty()) { fprintf(stderr, "dlp-run: -m model.gguf is required\n"); return 2; }
    std::vector<int32_t> prompt;
    { std::stringstream ss(tok_csv); std::string item; while (std::getline(ss, item, ',')) if (!item.empty()) prompt.push_back(atoi(item.c_str())); }
    if (prompt.empty()) prompt.push_back(1);
    try {
        Gguf g;
        g.open(model);
        const std::string ar = g.str.count("general.architecture") ? g.str["general.architecture"] : "llama";
        if (ar != "llama") throw std::runtime_error("dlp-run handles llama-architecture checkpoints (got " + ar + "); use the Python engine");
        auto N = [&](const char* k, double d) { return g.n(ar + "." + k, d); };
        const int E = (int)N("embedding_length", 0), L = (int)N("block_count", 0), Hq = (int)N("attention.head_count", 0);
        const int Hkv = (int)N("attention.head_count_kv", Hq), D = (int)N("attention.key_length", E / std::max(1, Hq));
        const int F = (int)N("feed_forward_length", 0), n_rot = (int)N("rope.dimension_count", D);
        const float eps = (float)N("attention.layer_norm_rms_epsilon", 1e-5), theta = (float)N("rope.freq_base", 10000.0);
        if (N("expert_count", 0) > 0) throw std::runtime_error("MoE checkpoints are not supported by dlp-run");
        if (n_rot != D || (D != 64 && D != 128)) throw std::runtime_error("dlp-run needs full-width RoPE and head_dim 64 or 128");
        const int V = (int)g.get("token_embd.weight").rows();
        const int page = 64, max_pages = (n_ctx + page - 1) / page, qkv_dim = (Hq + 2 * Hkv) * D;
        if ((int)prompt.size() + n_predict + 1 > n_ctx) throw std::runtime_error("prompt + n_predict exceed the context size");

        auto t_load = std::chrono::steady_clock::now();
        dlp_weight tok_embd = make_weight({g.get("token_embd.weight")});
        dlp_weight lm_head = g.has("output.weight") ? make_weight({g.get("output.weight")}) : tok_embd;
        float* out_norm = upload_f32(g.get("output_norm.weight"));
        std::vector<Layer> layers(L);
        for (int il = 0; il < L; ++il) {
            const std::string p = "blk." + std::to_string(il) + ".";
            Layer& l = layers[il];
            const Tensor &tq = g.get(p + "attn_q.weight"), &tk = g.get(p + "attn_k.weight"), &tv = g.get(p + "attn_v.weight");
            l.stacked = true;
            l.wqkv = (tq.type == tk.type && tk.type == tv.type) ? make_weight({tq, tk, tv}) : make_weight_f16_decoded({tq, tk, tv});
            l.wo = make_weight({g.get(p + "attn_output.weight")});
            l.gate = make_weight({g.get(p + "ffn_gate.weight")});
            l.up = make_weight({g.get(p + "ffn_up.weight")});
            l.down = make_weight({g.get(p + "ffn_down.weight")});
            l.attn_norm = upload_f32(g.get(p + "attn_norm.weight"));
            l.ffn_norm = upload_f32(g.get(p + "ffn_norm.weight"));
            l.kc = dev_alloc<uint16_t>((size_t)max_pages * Hkv * page * D);
            l.vc = dev_alloc<uint16_t>((size_t)max_pages * Hkv * page * D);
        }
        std::vector<float> inv(D / 2);
        for (int i = 0; i < D / 2; ++i) inv[i] = std::pow(theta, -2.0f * i / D);
        if (g.has("rope_freqs.weight")) { const float* ff = (const float*)g.get("rope_freqs.weight").data; for (int i = 0; i < D / 2; ++i) inv[i] /= ff[i]; }
        float* inv_freq = (float*)upload(inv.data(), inv.size() * 4);
        std::vector<int32_t> bt(max_pages);
        for (int i = 0; i < max_pages; ++i) bt[i] = i;
        int32_t* block_table = (int32_t*)upload(bt.data(), bt.size() * 4);
        const double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_load).count();

        // ---- static activation buffers (one sequence, one token per step) ----
        float* h = dev_alloc<float>(E);
        float* qkv = dev_alloc<float>(qkv_dim);
        uint16_t* att = dev_alloc<uint16_t>((size_t)Hq * D);
        uint16_t* act = dev_alloc<uint16_t>(F);
        float* logits = dev_alloc<float>(V);
        float* rope_cs = dev_alloc<float>(D);              // (cos, sin) per rotation pair
        int32_t *seq_len = dev_alloc<int32_t>(1), *pos = dev_alloc<int32_t>(1), *slot = dev_alloc<int32_t>(1), *kv_len = dev_alloc<int32_t>(1);
        int32_t* next_tok = dev_alloc<int32_t>(1);
        cudaStream_t st;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        const float q_scale = 1.0f / std::sqrt((float)D);
        int launches = 0;

        auto step = [&]() {      // one token: everything stays on the device; `next_tok` is both input and output
            launches = 0;
            DK(dlp_advance_rope(seq_len, nullptr, pos, slot, kv_len, block_table, max_pages, page, 1, inv_freq, D / 2, rope_cs, st)); ++launches;
            DK(dlp_dequant_rows_f32(&tok_embd, next_tok, 1, h, E, st)); ++launches;
            for (int il = 0; il < L; ++il) {
                Layer& l = layers[il];
                if (l.stacked) { DK(dlp_gemv(&l.wqkv, nullptr, h, l.attn_norm, eps, qkv, qkv_dim, DLP_OUT_F32, 1, st)); ++launches; }
                else {
                    DK(dlp_gemv(&l.wq, nullptr, h, l.attn_norm, eps, qkv, qkv_dim, DLP_OUT_F32, 1, st));
                    DK(dlp_gemv(&l.wk, nullptr, h, l.attn_norm, eps, qkv + (size_t)Hq * D, qkv_dim, DLP_OUT_F32, 1, st));
                    DK(dlp_gemv(&l.wv, nullptr, h, l.attn_norm, eps, qkv + (size_t)(Hq + Hkv) * D, qkv_dim, DLP_OUT_F32, 1, st));
                    launches += 3;
                }
                DK(dlp_attn_decode_rope(qkv, inv_freq, rope_cs, l.kc, l.vc, block_table, nullptr, kv_len, att, 1, Hq, Hkv, D, max_pages, page, 16, 0,
                                        q_scale, st)); ++launches;
                DK(dlp_gemv(&l.wo, nullptr, att, nullptr, 0.f, h, E, DLP_OUT_ADD_F32, 1, st)); ++launches;
                DK(dlp_gemv(&l.gate, &l.up, h, l.ffn_norm, eps, act, F, DLP_OUT_SWIGLU_F16, 1, st)); ++launches;
                DK(dlp_gemv(&l.down, nullptr, act, nullptr, 0.f, h, E, DLP_OUT_ADD_F32, 1, st)); ++launches;
            }
            DK(dlp_gemv(&lm_head, nullptr, h, out_norm, eps, logits, V, DLP_OUT_F32, 1, st)); ++launches;
            DK(dlp_argmax(logits, 1, V, V, next_tok, st)); ++launches;
        };

        // prompt: token by token through the same step (the batched tcgen05 prefill lives in the Python engine); the first call
        // runs eagerly so every kernel configures its attributes, then the step is captured once and replayed
        std::vector<int32_t> out;
        int32_t tok = prompt[0];
        CK(cudaMemcpyAsync(next_tok, &tok, 4, cudaMemcpyHostToDevice, st));
        step();
        CK(cudaStreamSynchronize(st));
        cudaGraph_t graph; cudaGraphExec_t exec;
        {   // capture: roll the bookkeeping back afterwards (capturing does not execute, so nothing to undo on the device)
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            step();
            CK(cudaStreamEndCapture(st, &graph));
            CK(cudaGraphInstantiate(&exec, graph, 0));
        }
        for (size_t i = 1; i < prompt.size(); ++i) {
            CK(cudaMemcpyAsync(next_tok, &prompt[i], 4, cudaMemcpyHostToDevice, st));
            CK(cudaGraphLaunch(exec, st));
        }
        CK(cudaMemcpyAsync(&tok, next_tok, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        out.push_back(tok);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, st));
        std::vector<int32_t> ring((size_t)std::max(1, n_predict - 1));
        int32_t* ring_d = dev_alloc<int32_t>(ring.size());
        for (int i = 1; i < n_predict; ++i) {            // device-resident loop: the token feeds back without a host round trip
            CK(cudaGraphLaunch(exec, st));
            CK(cudaMemcpyAsync(ring_d + (i - 1), next_tok, 4, cudaMemcpyDeviceToDevice, st));
        }
        CK(cudaEventRecord(e1, st));
        CK(cudaStreamSynchronize(st));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (n_predict > 1) { CK(cudaMemcpy(ring.data(), ring_d, (size_t)(n_predict - 1) * 4, cudaMemcpyDeviceToHost)); out.insert(out.end(), ring.begin(), ring.begin() + (n_predict - 1)); }

        printf("{\"tokens\": [");
        for (size_t i = 0; i < out.size(); ++i) printf("%s%d", i ? ", " : "", out[i]);
        const double per = n_predict > 1 ? ms / (n_predict - 1) : 0.0;
        printf("], \"n_prompt\": %zu, \"n_predict\": %d, \"ms_per_token\": %.4f, \"decode_tok_s\": %.2f, \"launches_per_step\": %d, \"load_s\": %.2f, "
               "\"n_layer\": %d, \"bench\": %s}\n", prompt.size(), n_predict, per, per > 0 ? 1e3 / per : 0.0, launches, load_s, L, bench ? "true" : "false");
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "dlp-run: %s\n", e.what());
        return 1;
    }
}
