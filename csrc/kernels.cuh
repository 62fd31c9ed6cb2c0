// Internal kernel-launcher interface shared by the .cu translation units of libdlp_b200.so.
#pragma once
#include "quant.cuh"

namespace dlp {

enum GemvOut : int { GV_STORE_F16 = 0, GV_STORE_F32 = 1, GV_ADD_F32 = 2, GV_SWIGLU_F16 = 3, GV_PUSH_F32 = 4 };

// Fused tensor-parallel all-reduce around the decode mat-vecs (see gemv.cu, comm.cu, tpcomm.cuh).
// Every rank owns, per all-reduce site, a symmetric buffer  [2 parities][tp source ranks][AR_MAX_T token rows][K] fp32  that
// all peers can store into (peer pointers, or one NVLS multicast alias that the switch replicates to every rank).  A value is
// its own arrival signal: buffers are pre-filled with AR_EMPTY (the bit pattern of -0.0f, which a producer never stores), the
// consumer spins on the data words themselves, and it re-arms the *other* parity of its site for the step after next.  No
// flag, no system-scope fence, no counter: the latency of an all-reduce is one NVLink store.
constexpr uint32_t AR_EMPTY = 0x80000000u;
constexpr int AR_MAX_T = 4;
struct ArIn {                 // consumer: h_out = h_in + sum over ranks of buf[parity][r]
    float* buf;               // local buffer of this site (nullptr: no all-reduce to consume)
    const int* epoch;         // device-resident step counter (same value on every rank); parity = epoch & 1
    float* h_out;             // where CTA 0 stores the updated residual stream
    int* error;               // set to 1 on wait timeout
    int tp;
};
struct ArOut {                // producer: store this rank's partial rows into every rank's buffer
    float* peer_buf[8];       // peer_buf[p] = base of rank p's buffer for this site (own rank included)
    float* mc_buf;            // multicast alias of the same buffer (one multimem.st reaches all ranks), or nullptr
    const int* epoch;
    int rank, tp;
};
// Token exchange for sharded greedy sampling: slots [2 parities][tp][TOKX_MAX_R] of uint64, 0 = empty.
constexpr int TOKX_MAX_R = 64;
struct TokX {
    unsigned long long* peer_slots[8];
    unsigned long long* mc_slots;
    unsigned long long* slots;      // local
    const int* epoch;
    int* error;
    int rank, tp;
};

struct GemvParams {
    QWeight w0, w1;       // w1 only for the SwiGLU (gate, up) pair
    const void* x;        // fp16 [T, K]  or, with norm fusion, fp32 h [T, K]
    const float* norm_w;  // non-null => x is fp32 and RMSNorm(x) * norm_w is applied in the prologue
    float eps;
    void* y;
    int ldy;
    int out_mode;
    int T;
    ArIn ar_in;
    ArOut ar_out;
    // expert-indexed weights (MoE decode): weight row = row + (*row_off_dev) * row_off_mul, outputs scaled by *out_scale_dev.
    // Both values are produced on the device by the router, so the step stays graph-capturable.
    const int* row_off_dev;
    int row_off_mul;
    const float* out_scale_dev;
    unsigned long long* trace = nullptr;   // device timeline buffer (common.cuh trace_event), normally null
    int pdl_early = 0;                     // trigger the dependent launch right after the dependency wait (default on, DLP_PDL_EARLY=0)
};
int moe_route(const float* logits, int T, int n_expert, int top_k, int* ids, float* w, cudaStream_t s);

// Collectives fused into the tcgen05 GEMM (prefill / batched decode under tensor parallelism, sequence-parallel layout):
//   X side ("all-gather -> GEMM"): token rows [src*x_rows_per_src, +x_rows_per_src) of X were pushed by rank src into this
//     rank's gather buffer; the TMA producer waits for x_flags[src] >= x_epoch of exactly the owners its tile touches.
//   Y side ("GEMM -> reduce-scatter"): epilogue EPI_RS_PUSH_F32 stores the partial tile into the owner rank's reduce buffer
//     [src rank][rs_rows][N] and the last CTA publishes rs_epoch to every peer's flag array.
struct GemmCollective {
    const uint32_t* x_flags;
    int x_rows_per_src;
    uint32_t x_epoch;
    int* error;
    float* rs_peer_buf[8];
    uint32_t* rs_peer_flags[8];
    unsigned* rs_done;
    int rs_rows, rs_rank, rs_tp;
    uint32_t rs_epoch;
    // Epochs are `*step * SP_EPOCH_STRIDE + <the value above>` when `step` is given: the step counter lives on the device (bumped once
    // per forward pass), the values above are then positions inside the step - launch-time constants, so a captured CUDA graph
    // stays valid from step to step.  step == nullptr: the values above are absolute epochs.
    const uint32_t* step;
};
constexpr uint32_t SP_EPOCH_STRIDE = 4096;
__device__ __forceinline__ uint32_t sp_epoch(const uint32_t* step, uint32_t v) { return step ? *step * SP_EPOCH_STRIDE + v : v; }

// Reduce-scatter consumer + RMSNorm + all-gather producer (one CTA per owned token row), see comm.cu.
struct RsNormAg {
    float* h_shard;              // [rows, E] residual stream shard (updated in place)
    const float* red_buf;        // local reduce buffer [tp][rows][E] (nullptr: nothing to reduce)
    const uint32_t* red_flags;   // [tp]
    uint32_t red_epoch;
    const float* norm_w;         // nullptr: no norm / no gather (just fold the reduction into h_shard)
    float eps;
    __half* ag_peer_buf[8];      // peers' gather buffers [T, E] fp16
    uint32_t* ag_peer_flags[8];
    uint32_t ag_epoch;
    unsigned* done;
    int* error;
    int rank, tp, rows, E;
    const uint32_t* step;        // see GemmCollective::step
};
int rs_norm_ag(const RsNormAg& a, cudaStream_t s);

struct AttnParams {
    const __half* q;        // [nq, Hq, D]  (already scaled by 1/sqrt(D) and rotated)
    const __half* kc;       // [n_pages][Hkv][page][D]
    const __half* vc;
    const int* block_table; // [n_rows, max_pages]
    const int* bt_row;      // [nq] block-table row of each query (nullptr => row = query index)
    const int* kv_len;      // [nq] number of keys visible to each query
    __half* out;            // [nq, Hq, D]
    float* part_o;          // [nq, Hq, n_splits, D]   (n_splits > 1)
    float* part_ml;         // [nq, Hq, n_splits, 2]
    int* counters;          // [nq, Hkv * groups] zero-initialised; non-null => the last split CTA combines in-kernel
    int nq, Hq, Hkv, max_pages, page_size, n_splits;
    unsigned long long* trace = nullptr;   // device timeline buffer, normally null
    int pdl_early = 0;
    int keys_per_split = 64;               // one-launch decode kernel: target keys per cluster member
};

// Grouped (mixture-of-experts) form of the tcgen05 GEMM: the X rows are the expert-sorted token copies, the weight is the stack of
// all experts' matrices, and token tile j of the launch (blockIdx.y) belongs to expert tile_expert[j] and covers sorted rows
// [tile_row0[j], tile_row0[j] + tile_rows[j]).  The table is built on the device (moe_plan), so the step never visits the host;
// the launch uses an upper bound of tiles and surplus CTAs leave at once.
struct GemmGroup {
    const int* tile_expert = nullptr;   // nullptr: plain GEMM
    const int* tile_row0 = nullptr;
    const int* tile_rows = nullptr;
    const int* n_tiles = nullptr;
    int rows_per_expert = 0;            // output features of one expert (rows of the stacked weight per expert)
    int n_expert = 0;
    int max_tiles = 0;
    int bt = 0;                         // token-tile size the table was built for
};

int gemv(const GemvParams& p, int wtype, cudaStream_t stream);
int gemm_f16(const void* w0, const void* w1, int ldw, const void* x, int ldx, void* y, int ldy, int T, int N, int K,
             int epilogue, int is_bf16, int split_k, cudaStream_t stream, const GemmCollective* coll = nullptr,
             const GemmGroup* grp = nullptr, void* ws = nullptr, size_t ws_bytes = 0);
int gemm_quant(const QWeight& w0, const QWeight* w1, int wtype, const void* x, int ldx, void* y, int ldy, int T, int epilogue,
               int split_k, cudaStream_t stream, const GemmCollective* coll = nullptr, const GemmGroup* grp = nullptr, void* ws = nullptr,
               size_t ws_bytes = 0);
// MoE routing plan + data movement around the grouped GEMMs (moe.cu)
int moe_route_ex(const float* logits, int ld, int T, int n_expert, int top_k, int renorm, int* ids, float* w, cudaStream_t s);
int moe_plan(const int* ids, int n_assign, int n_expert, int bt, int* pos, int* src, int* bounds, int* tile_expert, int* tile_row0,
             int* tile_rows, int* n_tiles, cudaStream_t s);
int moe_gather(const __half* x, int ldx, const int* src, int top_k, int n_assign, int K, __half* xs, cudaStream_t s);
int moe_combine(const float* y, const int* pos, const float* w, int T, int top_k, int E, float* h, cudaStream_t s);
int dequant_rows_f32(const QWeight& w, int wtype, const int* row_ids, int n_rows, float* out, int ld_out, cudaStream_t s);
int dequant_rows_f16(const QWeight& w, int wtype, const int* row_ids, int n_rows, __half* out, int ld_out, cudaStream_t s);
int rmsnorm(float* h, const float* delta, const float* w, __half* out, int T, int E, float eps, cudaStream_t s);
int layernorm(float* h, const float* delta, const float* w, const float* b, __half* out, int T, int E, float eps, cudaStream_t s);
int bias_act(const float* x, const float* bias, __half* out, int T, int F, int act, cudaStream_t s);
int rope_kv(const void* qkv, int qkv_is_f16, const int* pos, const int* slot, const float* inv_freq, __half* q_out,
            __half* kc, __half* vc, int T, int Hq, int Hkv, int D, int n_rot, int page_size, int neox, float q_scale,
            cudaStream_t s, const float* bias = nullptr, const float* qn = nullptr, const float* kn = nullptr, float eps = 0.f,
            const void* pf0 = nullptr, size_t pf0_bytes = 0, const void* pf1 = nullptr, size_t pf1_bytes = 0);
int swiglu(const void* gu, int is_f16, __half* out, int T, int F, int ld_in, cudaStream_t s);
int advance(int* seq_len, const int* active, int* pos, int* slot, int* kv_len, const int* block_table, int max_pages,
            int page_size, int n, cudaStream_t s);
int argmax(const float* logits, int R, int V, int ld, int* out, cudaStream_t s);
int sample(const float* logits, int R, int V, int ld, int top_k, float top_p, float min_p, float temp, uint32_t seed,
           const int* step_ctr, int* out, cudaStream_t s);
// per-sequence sampler record in device memory (32 bytes; mirrors dlp_sample_params in include/dlp_b200.h)
struct SampleParams {
    int top_k;
    float top_p, min_p, temp;
    uint32_t seed;
    uint32_t ctr;       // draws made with this record so far (advanced by the kernel)
    uint32_t pad[2];
};
int sample_dev(const float* logits, int R, int V, int ld, SampleParams* params, const int* row_slot, int* out, cudaStream_t s);
int f32_to_f16(const float* in, __half* out, size_t n, cudaStream_t s);
int attn_paged(const AttnParams& p, int D, cudaStream_t s);
int attn_paged_static(AttnParams p, int D, cudaStream_t s);
// decode step in one launch: RoPE(Q, new K) + paged K/V write + split-KV attention + cluster/DSMEM combine (attention.cu)
int attn_decode_rope(const AttnParams& p, const float* qkv, const float* inv_freq, const float2* rope_cs, __half* kc_w,
                     __half* vc_w, int D, int neox, float q_scale, cudaStream_t s);
int advance_rope(int* seq_len, const int* active, int* pos, int* slot, int* kv_len, const int* block_table, int max_pages,
                 int page_size, int n, const float* inv_freq, int n_pairs, float2* rope_cs, cudaStream_t s);
int ar_consume(const float* h_in, float* h_out, const ArIn& a, int K, int T, cudaStream_t s);
int argmax_tp(const float* logits, int R, int V, int ld, int v0, const TokX& x, int* out, cudaStream_t s);
int fill_u32(uint32_t* p, size_t n, uint32_t v, cudaStream_t s);
int bump(int* ctr, cudaStream_t s);
int attn_prefill(const __half* q, const __half* kc, const __half* vc, const int* bt_row, __half* out, int T, int pos0,
                 int Hq, int Hkv, int D, int page_size, cudaStream_t s);

int attn_prefill_tc(const __half* q, const __half* kc, const __half* vc, const int* bt_row, __half* out, int T, int pos0, int Hq, int Hkv,
                    int D, int page_size, int n_pages, int max_pages, cudaStream_t s);

}  // namespace dlp
