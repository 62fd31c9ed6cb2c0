// extern "C" surface of libdlp_b200.so (declarations: include/dlp_b200.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "../include/dlp_b200.h"
#include "kernels.cuh"
#include <mutex>
#include <set>
#include <utility>

namespace dlp {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* get_error() { return g_err; }

static int g_pdl = -1;
bool pdl_enabled() {
    if (g_pdl < 0) {
        const char* e = getenv("DLP_PDL");
        g_pdl = (e && e[0] == '0') ? 0 : 1;
    }
    return g_pdl != 0;
}
void set_pdl(bool on) { g_pdl = on ? 1 : 0; }
bool first_use_on_device(const void* kernel) {
    static std::mutex mu;
    static std::set<std::pair<int, const void*>> seen;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lock(mu);
    return seen.insert({dev, kernel}).second;
}
static unsigned long long* g_trace = nullptr;
unsigned long long* trace_buffer() { return g_trace; }
void set_trace_buffer(unsigned long long* buf) { g_trace = buf; }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static QWeight to_qw(const dlp_weight* w) {
    QWeight q;
    q.base = (const uint8_t*)w->base;
    q.aux0 = (const uint8_t*)w->aux0;
    q.aux1 = (const uint8_t*)w->aux1;
    q.aux2 = (const uint8_t*)w->aux2;
    q.N = w->N;
    q.K = w->K;
    return q;
}

}  // namespace dlp

using namespace dlp;

extern "C" {

const char* dlp_last_error(void) { return get_error(); }
int dlp_version(void) { return 100; }
int dlp_sm_count(void) { return sm_count(); }
int dlp_set_pdl(int on) { set_pdl(on != 0); return 0; }
int dlp_set_trace(void* buf) { set_trace_buffer((unsigned long long*)buf); return 0; }

int dlp_gemv(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
             int ldy, int out_mode, int T, void* stream) {
    DLP_REQUIRE(w0 && w0->base && x && y, "dlp_gemv: null argument");
    DLP_REQUIRE(!w1 || (w1->wtype == w0->wtype && w1->N == w0->N && w1->K == w0->K), "dlp_gemv: gate/up mismatch");
    GemvParams p;
    memset(&p, 0, sizeof(p));
    p.w0 = to_qw(w0);
    if (w1) p.w1 = to_qw(w1);
    p.x = x; p.norm_w = norm_w; p.eps = eps; p.y = y; p.ldy = ldy; p.out_mode = out_mode; p.T = T;
    return gemv(p, w0->wtype, (cudaStream_t)stream);
}

static void fill_ar(GemvParams& p, const dlp_ar_in* in, const dlp_ar_out* out) {
    if (in && in->buf) {
        p.ar_in.buf = in->buf; p.ar_in.epoch = in->epoch; p.ar_in.h_out = in->h_out;
        p.ar_in.error = in->error; p.ar_in.tp = in->tp;
    }
    if (out) {
        for (int i = 0; i < 8; ++i) p.ar_out.peer_buf[i] = out->peer_buf[i];
        p.ar_out.mc_buf = out->mc_buf; p.ar_out.epoch = out->epoch; p.ar_out.rank = out->rank; p.ar_out.tp = out->tp;
    }
}

int dlp_gemv_tp(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
                int ldy, int out_mode, int T, const dlp_ar_in* ar_in, const dlp_ar_out* ar_out, void* stream) {
    DLP_REQUIRE(w0 && w0->base && x, "dlp_gemv_tp: null argument");
    DLP_REQUIRE(out_mode != DLP_OUT_PUSH_F32 || (ar_out && ar_out->epoch && T <= AR_MAX_T), "dlp_gemv_tp: push needs ar_out");
    DLP_REQUIRE(out_mode == DLP_OUT_PUSH_F32 || y, "dlp_gemv_tp: null output");
    DLP_REQUIRE(!(ar_in && ar_in->buf) || (ar_in->epoch && ar_in->h_out && norm_w), "dlp_gemv_tp: incomplete ar_in");
    GemvParams p;
    memset(&p, 0, sizeof(p));
    p.w0 = to_qw(w0);
    if (w1) p.w1 = to_qw(w1);
    p.x = x; p.norm_w = norm_w; p.eps = eps; p.y = y; p.ldy = ldy; p.out_mode = out_mode; p.T = T;
    fill_ar(p, ar_in, out_mode == DLP_OUT_PUSH_F32 ? ar_out : nullptr);
    return gemv(p, w0->wtype, (cudaStream_t)stream);
}

int dlp_gemv_moe(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y, int ldy,
                 int out_mode, const int32_t* expert_id_dev, int rows_per_expert, const float* out_scale_dev, void* stream) {
    DLP_REQUIRE(w0 && w0->base && x && y && expert_id_dev && rows_per_expert > 0, "dlp_gemv_moe: bad argument");
    GemvParams p;
    memset(&p, 0, sizeof(p));
    p.w0 = to_qw(w0);
    if (w1) p.w1 = to_qw(w1);
    p.w0.N = rows_per_expert;              // the kernel walks one expert's rows; the offset selects which
    if (w1) p.w1.N = rows_per_expert;
    p.x = x; p.norm_w = norm_w; p.eps = eps; p.y = y; p.ldy = ldy; p.out_mode = out_mode; p.T = 1;
    p.row_off_dev = expert_id_dev; p.row_off_mul = rows_per_expert; p.out_scale_dev = out_scale_dev;
    return gemv(p, w0->wtype, (cudaStream_t)stream);
}

int dlp_moe_route(const float* logits, int T, int n_expert, int top_k, int32_t* ids, float* weights, void* stream) {
    DLP_REQUIRE(logits && ids && weights && T > 0, "dlp_moe_route: bad argument");
    return moe_route(logits, T, n_expert, top_k, ids, weights, (cudaStream_t)stream);
}

int dlp_ar_consume(const float* h_in, float* h_out, const dlp_ar_in* ar_in, int E, int T, void* stream) {
    DLP_REQUIRE(h_in && h_out && ar_in && ar_in->buf, "dlp_ar_consume: null argument");
    ArIn a;
    a.buf = ar_in->buf; a.epoch = ar_in->epoch; a.h_out = h_out; a.error = ar_in->error; a.tp = ar_in->tp;
    return ar_consume(h_in, h_out, a, E, T, (cudaStream_t)stream);
}

int dlp_argmax_tp(const float* logits, int R, int V, int ld, int v0, const dlp_tokx* x, int32_t* out, void* stream) {
    DLP_REQUIRE(logits && x && out, "dlp_argmax_tp: null argument");
    TokX t;
    for (int i = 0; i < 8; ++i) t.peer_slots[i] = reinterpret_cast<unsigned long long*>(x->peer_slots[i]);
    t.mc_slots = reinterpret_cast<unsigned long long*>(x->mc_slots);
    t.slots = reinterpret_cast<unsigned long long*>(x->slots);
    t.epoch = x->epoch; t.error = x->error; t.rank = x->rank; t.tp = x->tp;
    return argmax_tp(logits, R, V, ld, v0, t, out, (cudaStream_t)stream);
}

int dlp_gemm_sp(const dlp_weight* w0, const dlp_weight* w1, const void* x, int ldx, void* y, int ldy, int epilogue, int T,
                const dlp_gemm_coll* coll, void* stream) {
    DLP_REQUIRE(w0 && w0->base && x, "dlp_gemm_sp: null argument");
    GemmCollective c;
    memset(&c, 0, sizeof(c));
    if (coll) {
        c.x_flags = coll->x_flags; c.x_rows_per_src = coll->x_rows_per_src; c.x_epoch = coll->x_epoch; c.error = coll->error;
        for (int i = 0; i < 8; ++i) { c.rs_peer_buf[i] = coll->rs_peer_buf[i]; c.rs_peer_flags[i] = coll->rs_peer_flags[i]; }
        c.rs_done = coll->rs_done; c.rs_rows = coll->rs_rows; c.rs_rank = coll->rs_rank; c.rs_tp = coll->rs_tp; c.rs_epoch = coll->rs_epoch; c.step = coll->step;
        DLP_REQUIRE(c.x_flags == nullptr || c.x_rows_per_src > 0, "dlp_gemm_sp: x_rows_per_src must be positive");
    }
    const int wt = w0->wtype;
    cudaStream_t s = (cudaStream_t)stream;
    void* ws = coll ? coll->ws : nullptr;
    const size_t ws_bytes = coll ? coll->ws_bytes : 0;
    const int sk = ws ? 0 : 1;     // with a workspace: automatic deterministic split-K for skinny shards
    if (wt == DLP_W_F16 || wt == DLP_W_BF16)
        return gemm_f16(w0->base, w1 ? w1->base : nullptr, w0->K, x, ldx, y, ldy, T, w0->N, w0->K, epilogue, wt == DLP_W_BF16, sk, s,
                        coll ? &c : nullptr, nullptr, ws, ws_bytes);
    const QWeight q0 = to_qw(w0);
    QWeight q1;
    if (w1) q1 = to_qw(w1);
    return gemm_quant(q0, w1 ? &q1 : nullptr, wt, x, ldx, y, ldy, T, epilogue, sk, s, coll ? &c : nullptr, nullptr, ws, ws_bytes);
}

int dlp_rs_norm_ag(const dlp_rs_norm_ag_desc* d, void* stream) {
    DLP_REQUIRE(d && d->h_shard, "dlp_rs_norm_ag: null argument");
    RsNormAg a;
    memset(&a, 0, sizeof(a));
    a.h_shard = d->h_shard; a.red_buf = d->red_buf; a.red_flags = d->red_flags; a.red_epoch = d->red_epoch;
    a.norm_w = d->norm_w; a.eps = d->eps;
    for (int i = 0; i < 8; ++i) { a.ag_peer_buf[i] = (__half*)d->ag_peer_buf[i]; a.ag_peer_flags[i] = d->ag_peer_flags[i]; }
    a.ag_epoch = d->ag_epoch; a.done = d->done; a.error = d->error; a.rank = d->rank; a.tp = d->tp; a.rows = d->rows; a.E = d->E; a.step = d->step;
    return rs_norm_ag(a, (cudaStream_t)stream);
}

int dlp_gemm_f16(const void* w0, const void* w1, int ldw, const void* x, int ldx, void* y, int ldy, int T, int N,
                 int K, int epilogue, int is_bf16, int split_k, void* stream) {
    DLP_REQUIRE(w0 && x && y, "dlp_gemm_f16: null argument");
    return gemm_f16(w0, w1, ldw, x, ldx, y, ldy, T, N, K, epilogue, is_bf16, split_k, (cudaStream_t)stream);
}

static int linear_impl(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
                       int ldy, int out_mode, int T, void* scratch, size_t scratch_bytes, void* ws, size_t ws_bytes, void* stream) {
    DLP_REQUIRE(w0 && w0->base && x && y, "dlp_linear: null argument");
    if (T == 1 || T == 2 || T == 4) return dlp_gemv(w0, w1, x, norm_w, eps, y, ldy, out_mode, T, stream);
    DLP_REQUIRE(norm_w == nullptr, "dlp_linear: RMSNorm fusion is only available on the T<=4 mat-vec path");
    const int epi = out_mode == DLP_OUT_F16 ? 0 : out_mode == DLP_OUT_F32 ? 1 : out_mode == DLP_OUT_ADD_F32 ? 2 : 3;
    const int wt = w0->wtype;
    cudaStream_t s = (cudaStream_t)stream;
    if (wt == DLP_W_F16 || wt == DLP_W_BF16)
        return gemm_f16(w0->base, w1 ? w1->base : nullptr, w0->K, x, w0->K, y, ldy, T, w0->N, w0->K, epi,
                        wt == DLP_W_BF16, 0, s, nullptr, nullptr, ws, ws_bytes);
    // block-quantised: the decode is fused into the GEMM's A-operand producer (no scratch).  DLP_GEMM_DEQUANT_SCRATCH=1
    // selects the older two-pass path (decode to fp16 scratch, then the dense GEMM) for A/B runs.
    static int two_pass = -1;
    if (two_pass < 0) { const char* e = getenv("DLP_GEMM_DEQUANT_SCRATCH"); two_pass = (e && e[0] == '1') ? 1 : 0; }
    if (!two_pass) {
        const QWeight q0 = to_qw(w0);
        QWeight q1;
        if (w1) q1 = to_qw(w1);
        return gemm_quant(q0, w1 ? &q1 : nullptr, wt, x, w0->K, y, ldy, T, epi, 0, s, nullptr, nullptr, ws, ws_bytes);
    }
    const size_t per = (size_t)w0->N * w0->K * 2;
    const size_t need = per * (w1 ? 2 : 1);
    DLP_REQUIRE(scratch != nullptr && scratch_bytes >= need, "dlp_linear: scratch too small (%zu < %zu bytes)", scratch_bytes, need);
    int rc = dequant_rows_f16(to_qw(w0), wt, nullptr, w0->N, (__half*)scratch, w0->K, s);
    if (rc) return rc;
    const void* up = nullptr;
    if (w1) {
        up = (const uint8_t*)scratch + per;
        rc = dequant_rows_f16(to_qw(w1), wt, nullptr, w1->N, (__half*)up, w1->K, s);
        if (rc) return rc;
    }
    return gemm_f16(scratch, up, w0->K, x, w0->K, y, ldy, T, w0->N, w0->K, epi, 0, 0, s, nullptr, nullptr, ws, ws_bytes);
}

int dlp_linear(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
               int ldy, int out_mode, int T, void* scratch, size_t scratch_bytes, void* stream) {
    return linear_impl(w0, w1, x, norm_w, eps, y, ldy, out_mode, T, scratch, scratch_bytes, nullptr, 0, stream);
}
int dlp_linear_ws(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
                  int ldy, int out_mode, int T, void* scratch, size_t scratch_bytes, void* ws, size_t ws_bytes, void* stream) {
    return linear_impl(w0, w1, x, norm_w, eps, y, ldy, out_mode, T, scratch, scratch_bytes, ws, ws_bytes, stream);
}

int dlp_moe_route_ex(const float* logits, int ld, int T, int n_expert, int top_k, int renorm, int32_t* ids, float* weights, void* stream) {
    DLP_REQUIRE(logits && ids && weights && T > 0 && ld >= n_expert, "dlp_moe_route_ex: bad argument");
    return moe_route_ex(logits, ld, T, n_expert, top_k, renorm, ids, weights, (cudaStream_t)stream);
}
int dlp_moe_plan(const int32_t* ids, int n_assign, int n_expert, int token_tile, dlp_moe_plan_bufs* b, void* stream) {
    DLP_REQUIRE(ids && b && b->pos && b->src && b->bounds && b->tile_expert && b->tile_row0 && b->tile_rows && b->n_tiles, "dlp_moe_plan: null argument");
    return moe_plan(ids, n_assign, n_expert, token_tile, b->pos, b->src, b->bounds, b->tile_expert, b->tile_row0, b->tile_rows, b->n_tiles,
                    (cudaStream_t)stream);
}
int dlp_moe_gather(const void* x, int ldx, const int32_t* src, int top_k, int n_assign, int K, void* xs, void* stream) {
    DLP_REQUIRE(x && src && xs && n_assign > 0, "dlp_moe_gather: bad argument");
    return moe_gather((const __half*)x, ldx, src, top_k, n_assign, K, (__half*)xs, (cudaStream_t)stream);
}
int dlp_moe_combine(const float* y, const int32_t* pos, const float* weights, int T, int top_k, int E, float* h, void* stream) {
    DLP_REQUIRE(y && pos && weights && h, "dlp_moe_combine: null argument");
    return moe_combine(y, pos, weights, T, top_k, E, h, (cudaStream_t)stream);
}
int dlp_linear_grouped(const dlp_weight* w0, const dlp_weight* w1, const void* xs, void* y, int ldy, int out_mode, int n_assign, int n_expert,
                       int rows_per_expert, int token_tile, const dlp_moe_plan_bufs* b, void* ws, size_t ws_bytes, void* stream) {
    DLP_REQUIRE(w0 && w0->base && xs && y && b && b->tile_expert && n_assign > 0 && n_expert > 0, "dlp_linear_grouped: bad argument");
    DLP_REQUIRE(out_mode == DLP_OUT_F16 || out_mode == DLP_OUT_F32 || out_mode == DLP_OUT_SWIGLU_F16, "dlp_linear_grouped: output mode %d", out_mode);
    DLP_REQUIRE(w0->N == n_expert * rows_per_expert, "dlp_linear_grouped: weight has %d rows, expected %d x %d", w0->N, n_expert, rows_per_expert);
    GemmGroup g;
    g.tile_expert = b->tile_expert; g.tile_row0 = b->tile_row0; g.tile_rows = b->tile_rows; g.n_tiles = b->n_tiles;
    g.rows_per_expert = rows_per_expert; g.n_expert = n_expert; g.bt = token_tile;
    g.max_tiles = (n_assign + token_tile - 1) / token_tile + n_expert;     // every expert may end with one partial tile
    const int epi = out_mode == DLP_OUT_F16 ? 0 : out_mode == DLP_OUT_F32 ? 1 : 3;
    const int wt = w0->wtype;
    cudaStream_t s = (cudaStream_t)stream;
    if (wt == DLP_W_F16 || wt == DLP_W_BF16)
        return gemm_f16(w0->base, w1 ? w1->base : nullptr, w0->K, xs, w0->K, y, ldy, n_assign, rows_per_expert, w0->K, epi, wt == DLP_W_BF16,
                        ws ? 0 : 1, s, nullptr, &g, ws, ws_bytes);
    const QWeight q0 = to_qw(w0);
    QWeight q1;
    if (w1) q1 = to_qw(w1);
    return gemm_quant(q0, w1 ? &q1 : nullptr, wt, xs, w0->K, y, ldy, n_assign, epi, ws ? 0 : 1, s, nullptr, &g, ws, ws_bytes);
}

int dlp_dequant_rows_f32(const dlp_weight* w, const int32_t* row_ids, int n_rows, float* out, int ld_out, void* stream) {
    DLP_REQUIRE(w && w->base && out, "dlp_dequant_rows_f32: null argument");
    return dequant_rows_f32(to_qw(w), w->wtype, row_ids, n_rows, out, ld_out, (cudaStream_t)stream);
}
int dlp_dequant_rows_f16(const dlp_weight* w, const int32_t* row_ids, int n_rows, void* out, int ld_out, void* stream) {
    DLP_REQUIRE(w && w->base && out, "dlp_dequant_rows_f16: null argument");
    return dequant_rows_f16(to_qw(w), w->wtype, row_ids, n_rows, (__half*)out, ld_out, (cudaStream_t)stream);
}

int dlp_rmsnorm(float* h, const float* delta, const float* w, void* out_f16, int T, int E, float eps, void* stream) {
    DLP_REQUIRE(h && (w || !out_f16), "dlp_rmsnorm: null argument");
    return rmsnorm(h, delta, w, (__half*)out_f16, T, E, eps, (cudaStream_t)stream);
}

int dlp_layernorm(float* h, const float* delta, const float* w, const float* b, void* out_f16, int T, int E, float eps, void* stream) {
    DLP_REQUIRE(h, "dlp_layernorm: null argument");
    return layernorm(h, delta, w, b, (__half*)out_f16, T, E, eps, (cudaStream_t)stream);
}
int dlp_bias_act(const float* x, const float* bias, void* out_f16, int T, int F, int act, void* stream) {
    DLP_REQUIRE(x && out_f16, "dlp_bias_act: null argument");
    return bias_act(x, bias, (__half*)out_f16, T, F, act, (cudaStream_t)stream);
}

int dlp_rope_kv(const void* qkv, int qkv_is_f16, const int32_t* pos, const int32_t* slot, const float* inv_freq,
                void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D, int n_rot, int page_size,
                int neox, float q_scale, void* stream) {
    DLP_REQUIRE(qkv && pos && slot && inv_freq && q_out && k_cache && v_cache, "dlp_rope_kv: null argument");
    return rope_kv(qkv, qkv_is_f16, pos, slot, inv_freq, (__half*)q_out, (__half*)k_cache, (__half*)v_cache, T, Hq, Hkv,
                   D, n_rot, page_size, neox, q_scale, (cudaStream_t)stream);
}

int dlp_rope_kv_ex(const void* qkv, int qkv_is_f16, const int32_t* pos, const int32_t* slot, const float* inv_freq,
                   void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D, int n_rot, int page_size,
                   int neox, float q_scale, const float* bias, const float* q_norm, const float* k_norm, float eps,
                   const void* pf0, size_t pf0_bytes, const void* pf1, size_t pf1_bytes, void* stream) {
    DLP_REQUIRE(qkv && pos && slot && inv_freq && q_out && k_cache && v_cache, "dlp_rope_kv_ex: null argument");
    return rope_kv(qkv, qkv_is_f16, pos, slot, inv_freq, (__half*)q_out, (__half*)k_cache, (__half*)v_cache, T, Hq, Hkv,
                   D, n_rot, page_size, neox, q_scale, (cudaStream_t)stream, bias, q_norm, k_norm, eps, pf0, pf0_bytes, pf1,
                   pf1_bytes);
}

int dlp_swiglu(const void* gate_up, int is_f16, void* out_f16, int T, int F, int ld_in, void* stream) {
    DLP_REQUIRE(gate_up && out_f16, "dlp_swiglu: null argument");
    return swiglu(gate_up, is_f16, (__half*)out_f16, T, F, ld_in, (cudaStream_t)stream);
}

int dlp_advance(int32_t* seq_len, const int32_t* active, int32_t* pos, int32_t* slot, int32_t* kv_len,
                const int32_t* block_table, int max_pages, int page_size, int n, void* stream) {
    DLP_REQUIRE(seq_len && pos && slot && block_table && n > 0, "dlp_advance: bad argument");
    return advance(seq_len, active, pos, slot, kv_len, block_table, max_pages, page_size, n, (cudaStream_t)stream);
}

int dlp_attn_paged(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                   const int32_t* bt_row, const int32_t* kv_len, void* out, float* part_o, float* part_ml, int nq,
                   int Hq, int Hkv, int D, int max_pages, int page_size, int n_splits, void* stream) {
    DLP_REQUIRE(q && k_cache && v_cache && block_table && kv_len && out, "dlp_attn_paged: null argument");
    AttnParams p;
    p.q = (const __half*)q; p.kc = (const __half*)k_cache; p.vc = (const __half*)v_cache;
    p.block_table = block_table; p.bt_row = bt_row; p.kv_len = kv_len; p.out = (__half*)out;
    p.part_o = part_o; p.part_ml = part_ml; p.counters = nullptr;
    p.nq = nq; p.Hq = Hq; p.Hkv = Hkv; p.max_pages = max_pages; p.page_size = page_size; p.n_splits = n_splits;
    return attn_paged(p, D, (cudaStream_t)stream);
}

int dlp_attn_paged_fused(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                         const int32_t* bt_row, const int32_t* kv_len, void* out, float* part_o, float* part_ml,
                         int32_t* counters, int nq, int Hq, int Hkv, int D, int max_pages, int page_size, int n_splits,
                         void* stream) {
    DLP_REQUIRE(q && k_cache && v_cache && block_table && kv_len && out && counters, "dlp_attn_paged_fused: null argument");
    AttnParams p;
    p.q = (const __half*)q; p.kc = (const __half*)k_cache; p.vc = (const __half*)v_cache;
    p.block_table = block_table; p.bt_row = bt_row; p.kv_len = kv_len; p.out = (__half*)out;
    p.part_o = part_o; p.part_ml = part_ml; p.counters = counters;
    p.nq = nq; p.Hq = Hq; p.Hkv = Hkv; p.max_pages = max_pages; p.page_size = page_size; p.n_splits = n_splits;
    return attn_paged(p, D, (cudaStream_t)stream);
}

int dlp_advance_rope(int32_t* seq_len, const int32_t* active, int32_t* pos, int32_t* slot, int32_t* kv_len,
                     const int32_t* block_table, int max_pages, int page_size, int n, const float* inv_freq, int n_pairs,
                     float* rope_cs, void* stream) {
    DLP_REQUIRE(seq_len && pos && slot && block_table && inv_freq && rope_cs, "dlp_advance_rope: null argument");
    return advance_rope(seq_len, active, pos, slot, kv_len, block_table, max_pages, page_size, n, inv_freq, n_pairs,
                        (float2*)rope_cs, (cudaStream_t)stream);
}

int dlp_attn_decode_rope(const float* qkv, const float* inv_freq, const float* rope_cs, void* k_cache, void* v_cache, const int32_t* block_table,
                         const int32_t* bt_row, const int32_t* kv_len, void* out, int nq, int Hq, int Hkv, int D, int max_pages,
                         int page_size, int n_splits, int neox, float q_scale, void* stream) {
    DLP_REQUIRE(qkv && (inv_freq || rope_cs) && k_cache && v_cache && block_table && kv_len && out, "dlp_attn_decode_rope: null argument");
    DLP_REQUIRE(((uintptr_t)rope_cs & 15) == 0 && ((uintptr_t)qkv & 15) == 0, "dlp_attn_decode_rope: qkv / rope_cs must be 16-byte aligned");
    AttnParams p;
    p.q = nullptr; p.kc = (const __half*)k_cache; p.vc = (const __half*)v_cache;
    p.block_table = block_table; p.bt_row = bt_row; p.kv_len = kv_len; p.out = (__half*)out;
    p.part_o = nullptr; p.part_ml = nullptr; p.counters = nullptr;
    p.nq = nq; p.Hq = Hq; p.Hkv = Hkv; p.max_pages = max_pages; p.page_size = page_size; p.n_splits = n_splits;
    return attn_decode_rope(p, qkv, inv_freq, (const float2*)rope_cs, (__half*)k_cache, (__half*)v_cache, D, neox, q_scale,
                            (cudaStream_t)stream);
}

int dlp_attn_paged_static(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                          const int32_t* bt_row, const int32_t* kv_len, void* out, int nq, int Hq, int Hkv, int D,
                          int max_pages, int page_size, int n_splits, void* stream) {
    DLP_REQUIRE(q && k_cache && v_cache && block_table && kv_len && out, "dlp_attn_paged_static: null argument");
    AttnParams p;
    p.q = (const __half*)q; p.kc = (const __half*)k_cache; p.vc = (const __half*)v_cache;
    p.block_table = block_table; p.bt_row = bt_row; p.kv_len = kv_len; p.out = (__half*)out;
    p.part_o = nullptr; p.part_ml = nullptr; p.counters = nullptr;
    p.nq = nq; p.Hq = Hq; p.Hkv = Hkv; p.max_pages = max_pages; p.page_size = page_size; p.n_splits = n_splits;
    return attn_paged_static(p, D, (cudaStream_t)stream);
}

int dlp_attn_prefill_tc(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table_row, void* out, int T, int pos0,
                        int Hq, int Hkv, int D, int page_size, int n_pages, int max_pages, void* stream) {
    DLP_REQUIRE(q && k_cache && v_cache && block_table_row && out, "dlp_attn_prefill_tc: null argument");
    return attn_prefill_tc((const __half*)q, (const __half*)k_cache, (const __half*)v_cache, block_table_row, (__half*)out, T, pos0, Hq, Hkv, D,
                           page_size, n_pages, max_pages, (cudaStream_t)stream);
}
int dlp_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table_row,
                     void* out, int T, int pos0, int Hq, int Hkv, int D, int page_size, void* stream) {
    DLP_REQUIRE(q && k_cache && v_cache && block_table_row && out, "dlp_attn_prefill: null argument");
    return attn_prefill((const __half*)q, (const __half*)k_cache, (const __half*)v_cache, block_table_row, (__half*)out,
                        T, pos0, Hq, Hkv, D, page_size, (cudaStream_t)stream);
}

int dlp_argmax(const float* logits, int R, int V, int ld, int32_t* out, void* stream) {
    DLP_REQUIRE(logits && out, "dlp_argmax: null argument");
    return argmax(logits, R, V, ld, out, (cudaStream_t)stream);
}
int dlp_sample(const float* logits, int R, int V, int ld, int top_k, float top_p, float min_p, float temp,
               uint32_t seed, const int32_t* step_ctr, int32_t* out, void* stream) {
    DLP_REQUIRE(logits && out, "dlp_sample: null argument");
    return sample(logits, R, V, ld, top_k, top_p, min_p, temp, seed, step_ctr, out, (cudaStream_t)stream);
}
int dlp_sample_dev(const float* logits, int R, int V, int ld, dlp_sample_params* params, const int32_t* row_slot, int32_t* out,
                   void* stream) {
    DLP_REQUIRE(logits && out && params, "dlp_sample_dev: null argument");
    static_assert(sizeof(dlp_sample_params) == sizeof(SampleParams), "sampler record layout");
    return sample_dev(logits, R, V, ld, reinterpret_cast<SampleParams*>(params), row_slot, out, (cudaStream_t)stream);
}
int dlp_f32_to_f16(const float* in, void* out, size_t n, void* stream) {
    DLP_REQUIRE(in && out, "dlp_f32_to_f16: null argument");
    return f32_to_f16(in, (__half*)out, n, (cudaStream_t)stream);
}

}  // extern "C"
