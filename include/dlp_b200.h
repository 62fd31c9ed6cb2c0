/* dlp_b200 - C ABI of the Blackwell (sm_100a) kernel library behind distributed_llm_pipeline_b200.
 *
 * Every entry point returns 0 on success, a positive cudaError_t, or -1 for an argument error;
 * dlp_last_error() returns a thread-local description.  All work is enqueued on `stream`
 * (a cudaStream_t passed as void*); nothing here synchronises the device or allocates device memory.
 * Pointers are device pointers unless stated otherwise.
 *
 * Parity notes cite the reference implementation (llama.cpp vendored in un1c4on/Distributed-LLM-Pipeline).
 */
#ifndef DLP_B200_H
#define DLP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* weight storage types (values mirror GGML type ids) */
enum { DLP_W_F32 = 0, DLP_W_F16 = 1, DLP_W_Q8_0 = 8, DLP_W_Q4_K = 12, DLP_W_Q5_K = 13, DLP_W_Q6_K = 14, DLP_W_BF16 = 30 };

/* One [N, K] weight matrix in kernel layout (see csrc/quant.cuh):
 *   F16/BF16: base = dense [N][K];            Q4_K/Q5_K: base = native GGUF super-blocks;
 *   Q8_0:     base = int8 qs[N][K], aux0 = fp16 d[N][K/32]         (repacked at load);
 *   Q6_K:     base = ql[N][K/2], aux0 = qh[N][K/4], aux1 = int8 sc[N][K/16], aux2 = fp16 d[N][K/256]. */
typedef struct dlp_weight {
    const void* base;
    const void* aux0;
    const void* aux1;
    const void* aux2;
    int32_t wtype;
    int32_t N;
    int32_t K;
    int32_t reserved;
} dlp_weight;

/* output modes of dlp_gemv / dlp_linear */
enum { DLP_OUT_F16 = 0, DLP_OUT_F32 = 1, DLP_OUT_ADD_F32 = 2, DLP_OUT_SWIGLU_F16 = 3 };

const char* dlp_last_error(void);
int dlp_version(void);
int dlp_sm_count(void);
/* Programmatic dependent launch between consecutive kernels of a stream (default on; env DLP_PDL=0 or dlp_set_pdl(0)
 * turns it off for A/B runs).  Kernels call griddepcontrol.wait before reading their predecessor's output. */
int dlp_set_pdl(int on);
/* Install (or with NULL remove) a device timeline buffer of uint64: buf[0] = record count (zero it before a run), buf[1] =
 * capacity, then records (event id << 48 | %globaltimer ns) appended by CTA (0,0,0) of the decode-path kernels
 * (mat-vec, RoPE, attention: begin / dependency wait returned / main loop done / end).  Kernel launches issued - or graphs
 * captured - while a buffer is installed carry it; profiling aid only (tools/timeline.py). */
int dlp_set_trace(void* buf);

/* Decode mat-vec, T in {1,2,4}: y[T,N] (op)= x[T,K] . W^T.  If norm_w != NULL, x is the fp32 residual stream and
 * RMSNorm(x)*norm_w is fused into the prologue; otherwise x is fp16.  out_mode SWIGLU takes w0=gate, w1=up.
 * ref: mul_mat_vec_f / mul_mat_vec_q + GLU fusion (ggml-cuda/mmvf.cu, mmvq.cu:143,181-189). */
int dlp_gemv(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
             int ldy, int out_mode, int T, void* stream);

/* tcgen05/TMEM/TMA GEMM on dense fp16 (is_bf16=0) or bf16 operands: y[T,N] (op)= x[T,K] . w0[N,K]^T.
 * epilogue: 0 store f16, 1 store f32, 2 atomic-add f32 (allows split_k>1; 0 = auto), 3 SwiGLU f16 (w1 = up), 4 store bf16.
 * ref: cuBLAS GemmEx fallback + mma.sync MMQ (ggml-cuda/ggml-cuda.cu:1227-1371, mmq.cuh:3459). */
int dlp_gemm_f16(const void* w0, const void* w1, int ldw, const void* x, int ldx, void* y, int ldy, int T, int N,
                 int K, int epilogue, int is_bf16, int split_k, void* stream);

/* Any-T linear layer on any supported weight type.  T<=4 (T in {1,2,4}) routes to dlp_gemv.  Larger T runs the tcgen05
 * GEMM with the block-quant decode fused into its operand producer; x must be fp16 when T>4.  `scratch` is only used with
 * DLP_GEMM_DEQUANT_SCRATCH=1 (A/B aid: weights are first decoded to fp16 there, 2*N*K bytes per matrix, two for SwiGLU).
 * ref: ggml_cuda_mul_mat dispatch (ggml-cuda/ggml-cuda.cu:2182-2265) incl. dequantise-then-cuBLAS (:1287-1294). */
int dlp_linear(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
               int ldy, int out_mode, int T, void* scratch, size_t scratch_bytes, void* stream);
/* Same with a split-K workspace `ws` (caller-owned; its first 16 KB are tile counters that must be zero before the FIRST use and
 * return to zero after every call; 64 MB + 16 KB covers every shape here).  When a problem has fewer 128-row tiles than the GPU
 * has SMs (decode batches, narrow tensor-parallel shards) K is split, each split stores its fp32 partial tile in `ws`, and the
 * split that arrives last adds the partials in split order (deterministic) and runs the epilogue - any epilogue, including
 * the residual add.  Calls that share one `ws` must not overlap on the GPU.
 * ref: mul_mat_q stream-k + mul_mat_q_stream_k_fixup (ggml-cuda/mmq.cuh:3459,3702). */
int dlp_linear_ws(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
                  int ldy, int out_mode, int T, void* scratch, size_t scratch_bytes, void* ws, size_t ws_bytes, void* stream);

/* Decode rows of a weight matrix to fp32 / fp16 (row_ids == NULL => rows 0..n_rows-1).  Embedding gather + dequant.
 * ref: k_get_rows (ggml-cuda/getrows.cu), dequantize_block_* (convert.cu). */
int dlp_dequant_rows_f32(const dlp_weight* w, const int32_t* row_ids, int n_rows, float* out, int ld_out, void* stream);
int dlp_dequant_rows_f16(const dlp_weight* w, const int32_t* row_ids, int n_rows, void* out, int ld_out, void* stream);

/* h[T,E] (fp32) += delta (optional, fp32); out (fp16, optional) = RMSNorm(h) * w.   ref: rms_norm_f32 fused mul+add (norm.cu:107). */
int dlp_rmsnorm(float* h, const float* delta, const float* w, void* out_f16, int T, int E, float eps, void* stream);

/* LayerNorm families: h[T,E] (fp32) += delta (optional); out (fp16, optional) = (h - mean) * rsqrt(var + eps) * w + b, w and b optional.
 * ref: norm_f32 (norm.cu:5-45) + the mul / add nodes of build_norm (llama-graph.cpp:785). */
int dlp_layernorm(float* h, const float* delta, const float* w, const float* b, void* out_f16, int T, int E, float eps, void* stream);
/* out fp16 [T,F] = act(x fp32 [T,F] + bias[F] (optional)); act: 0 none, 1 GELU (tanh), 2 ReLU^2, 3 SiLU, 4 GELU (erf).  ref: unary.cu. */
int dlp_bias_act(const float* x, const float* bias, void* out_f16, int T, int F, int act, void* stream);

/* RoPE on Q,K + paged KV write of K,V.  qkv [T,(Hq+2Hkv)*D] fp32/fp16; q_out fp16 [T,Hq*D] scaled by q_scale;
 * caches [n_pages][Hkv][page_size][D] fp16; slot[t] = page*page_size+offset (<0: no write).
 * ref: rope_norm/rope_neox with fused set_rows (rope.cu:44-111,58-78), k_set_rows (set-rows.cu). */
int dlp_rope_kv(const void* qkv, int qkv_is_f16, const int32_t* pos, const int32_t* slot, const float* inv_freq,
                void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D, int n_rot, int page_size,
                int neox, float q_scale, void* stream);
/* Same, for the other architecture families: `bias` [(Hq+2Hkv)*D] fp32 is added to q|k|v first (qwen2,
 * llama.cpp/src/models/qwen2.cpp), `q_norm`/`k_norm` [D] fp32 are per-head RMSNorm weights applied before the rotation
 * (qwen3, llama.cpp/src/models/qwen3.cpp).  Any of them may be NULL (q_norm/k_norm only as a pair).
 * `pf0`/`pf1` (+ byte counts, 16 B aligned, < 4 GiB; NULL/0 = none) name up to two memory ranges - the planes of the weight
 * the next mat-vec will stream - that the kernel asks L2 to fetch (cp.async.bulk.prefetch.L2) before it waits on its
 * own inputs, so HBM keeps working through this kernel and the attention kernel that follows. */
int dlp_rope_kv_ex(const void* qkv, int qkv_is_f16, const int32_t* pos, const int32_t* slot, const float* inv_freq,
                   void* q_out, void* k_cache, void* v_cache, int T, int Hq, int Hkv, int D, int n_rot, int page_size,
                   int neox, float q_scale, const float* bias, const float* q_norm, const float* k_norm, float eps,
                   const void* pf0, size_t pf0_bytes, const void* pf1, size_t pf1_bytes, void* stream);

int dlp_swiglu(const void* gate_up, int is_f16, void* out_f16, int T, int F, int ld_in, void* stream);

/* Device-side decode bookkeeping for batch entries s < n, sequence slot sid = active[s] (active NULL: sid = s):
 * pos[s]=seq_len[sid]; slot[s]=block-table lookup; kv_len[s]=pos+1 (optional); seq_len[sid]++ .
 * ref: llama_kv_cache::find_slot + set_input_k_idxs on the host (llama-kv-cache.cpp:696,1240). */
int dlp_advance(int32_t* seq_len, const int32_t* active, int32_t* pos, int32_t* slot, int32_t* kv_len,
                const int32_t* block_table, int max_pages, int page_size, int n, void* stream);

/* Paged attention for nq queries (decode, batched decode, or one query per prefill token).
 * q/out fp16 [nq,Hq,D]; kv_len[q] keys visible; bt_row[q] selects the block-table row (NULL: q).
 * n_splits>1 uses split-KV with workspace part_o [nq*Hq*n_splits*D] / part_ml [nq*Hq*n_splits*2] floats.
 * ref: flash_attn_ext_vec + flash_attn_combine_results (ggml-cuda/fattn-vec.cuh, fattn-common.cuh:721-772). */
int dlp_attn_paged(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                   const int32_t* bt_row, const int32_t* kv_len, void* out, float* part_o, float* part_ml, int nq,
                   int Hq, int Hkv, int D, int max_pages, int page_size, int n_splits, void* stream);

/* Same, with the split-KV combine fused into the attention kernel: `counters` is an int32 array of nq * Hkv * (Hq/Hkv/G)
 * entries (G = 4, 2 or 1 query heads per CTA; nq * Hq entries always suffice), zero-initialised once; the last split CTA of
 * each (query, kv-head group) merges the partials and resets its counter.  One launch instead of two. */
int dlp_attn_paged_fused(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                         const int32_t* bt_row, const int32_t* kv_len, void* out, float* part_o, float* part_ml,
                         int32_t* counters, int nq, int Hq, int Hkv, int D, int max_pages, int page_size, int n_splits,
                         void* stream);

/* Whole decode-attention step in one launch (no reference counterpart as a single op: the reference runs ggml_rope_ext,
 * ggml_set_rows for K and V, flash_attn_ext_vec and flash_attn_combine_results as separate graph nodes,
 * llama.cpp/src/models/llama.cpp:60-110).  `qkv` is the fp32 [nq, (Hq+2Hkv)*D] output of the QKV mat-vec, not rotated; the
 * kernel rotates Q and the new token's K (full-width rotation, n_rot == D; `neox` = half-split pairs), writes that token's
 * K/V at position kv_len-1 of its sequence (page from the block table), attends over kv_len keys and writes `out` fp16
 * [nq, Hq*D].  The n_splits (1..16) KV splits of a (query, KV-head group) run as one thread-block cluster and combine
 * through distributed shared memory - no workspace.  kv_len counts the new token.  `rope_cs` (optional) is the step's
 * [nq][D/2] table of (cos, sin) pairs written by dlp_advance_rope; when NULL the kernel evaluates sincosf from inv_freq.
 * Ordering: with programmatic dependent launch on, the kernel reads kv_len, the block table and the cached K/V rows of earlier
 * positions BEFORE it waits for its predecessor; only `qkv` is read after the wait.  Those inputs must therefore not be produced
 * by the launch immediately before this one when that launch is one of this library's PDL-aware kernels (in the decode step
 * they come from dlp_advance_rope at the start of the step and from earlier steps). */
int dlp_advance_rope(int32_t* seq_len, const int32_t* active, int32_t* pos, int32_t* slot, int32_t* kv_len,
                     const int32_t* block_table, int max_pages, int page_size, int n, const float* inv_freq, int n_pairs,
                     float* rope_cs, void* stream);
int dlp_attn_decode_rope(const float* qkv, const float* inv_freq, const float* rope_cs, void* k_cache, void* v_cache, const int32_t* block_table,
                         const int32_t* bt_row, const int32_t* kv_len, void* out, int nq, int Hq, int Hkv, int D, int max_pages,
                         int page_size, int n_splits, int neox, float q_scale, void* stream);

/* dlp_attn_paged for callers that cannot supply the split-KV workspace.  (No counterpart in the reference: it always
 * allocates its own split buffers from the backend pool.)
 * Who needs it: hosts that are not allowed to allocate device memory when they call us - a server that split all of HBM
 * into weights and KV pages at start-up and froze its allocator, a plug-in under a "no allocation after init" rule, or
 * code running inside a stream capture, where cudaMalloc is illegal.  Inputs and outputs are still the caller's.
 * How: the partial results live in a fixed array compiled into the library's device image,
 * DLP_ATTN_STATIC_SCRATCH_FLOATS floats (4 MiB) that every GPU loading the library holds until the process exits.
 * Limit: nq*Hq*n_splits*(D+2) must not exceed DLP_ATTN_STATIC_SCRATCH_FLOATS, otherwise the call returns -1 and
 * dlp_last_error() says by how much (n_splits == 1 needs no scratch at all).
 * Rules: (1) the first call on each device looks the array's device address up (cudaGetSymbolAddress) - make that first
 * call outside any stream capture; (2) all calls on a device share the one array, so two calls whose kernels can overlap
 * on the GPU (different streams, concurrently replayed graphs) race: order them yourself; (3) nothing needs clearing
 * between calls - every slot the combine step reads is rewritten by the same call first, so leftovers are never seen.
 * The call itself performs no cudaMalloc/cudaFree and uses no allocating library routine. */
#define DLP_ATTN_STATIC_SCRATCH_FLOATS (1u << 20)
int dlp_attn_paged_static(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                          const int32_t* bt_row, const int32_t* kv_len, void* out, int nq, int Hq, int Hkv, int D,
                          int max_pages, int page_size, int n_splits, void* stream);

/* Causal flash-prefill over the paged cache for one sequence chunk: T query tokens at positions
 * [pos0, pos0+T) attending to keys [0, pos0+T).  q/out fp16 [T,Hq,D].  ref: flash_attn_ext_f16 (fattn-mma-f16.cuh). */
int dlp_attn_prefill(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table_row,
                     void* out, int T, int pos0, int Hq, int Hkv, int D, int page_size, void* stream);

/* The same operation on the 5th-generation tensor cores: tcgen05 MMAs for Q.K^T and P.V with the score tile and the output
 * accumulator in tensor memory, K / V pages fetched by TMA (V as the MN-major operand, straight from its cache layout).  q must be
 * [T, Hq*D] contiguous; k_cache / v_cache are the whole [n_pages][Hkv][page_size][D] arrays (page_size a multiple of 8 dividing
 * 128); block_table_row has max_pages entries.  ref: flash_attn_ext_f16 (fattn-mma-f16.cuh:38-104) is the mma.sync form. */
int dlp_attn_prefill_tc(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_table_row, void* out, int T, int pos0,
                        int Hq, int Hkv, int D, int page_size, int n_pages, int max_pages, void* stream);

/* Sampling on device, one row of fp32 logits per output token.  temp<=0 => greedy.
 * ref: llama_sampler chain top_k/top_p/min_p/temp/dist (llama.cpp/src/llama-sampling.cpp:1348-1918,1255), argmax.cu. */
int dlp_argmax(const float* logits, int R, int V, int ld, int32_t* out, void* stream);
int dlp_sample(const float* logits, int R, int V, int ld, int top_k, float top_p, float min_p, float temp,
               uint32_t seed, const int32_t* step_ctr, int32_t* out, void* stream);

/* Same chain with the parameters in device memory: params[slot] (32-byte records), slot = row_slot[row] (row_slot NULL:
 * slot = row).  temp <= 0 in a record makes that row greedy.  The kernel advances the record's own draw counter, so the
 * random stream belongs to the request (the host zeroes `ctr` when it writes the record) and nothing request-specific is
 * baked into the launch: one captured CUDA graph serves every request and rows with different parameters share a batch.
 * ref: one `common_sampler` + RNG per server slot (llama.cpp/common/sampling.cpp:179-311, tools/server/server-context.cpp:2742). */
typedef struct dlp_sample_params {
    int32_t top_k;
    float top_p, min_p, temp;
    uint32_t seed;
    uint32_t ctr;
    uint32_t reserved[2];
} dlp_sample_params;
int dlp_sample_dev(const float* logits, int R, int V, int ld, dlp_sample_params* params, const int32_t* row_slot, int32_t* out,
                   void* stream);

int dlp_f32_to_f16(const float* in, void* out, size_t n, void* stream);

/* ---- mixture of experts, decode path (graph-capturable: expert ids and weights never leave the device) ----
 * dlp_moe_route: softmax over all experts, top-k, renormalise (n_expert <= 32) -> ids[T*k], weights[T*k].
 * dlp_gemv_moe: one token's mat-vec against expert *expert_id_dev of a stacked [n_expert * rows_per_expert, K] weight;
 *   outputs are multiplied by *out_scale_dev when given (DLP_OUT_ADD_F32 accumulates the weighted expert output).
 * ref: build_moe_ffn (llama.cpp/src/llama-graph.cpp:999-1266), topk_moe_cuda (ggml-cuda/topk-moe.cu), mul_mat_vec_* _id forms. */
int dlp_moe_route(const float* logits, int T, int n_expert, int top_k, int32_t* ids, float* weights, void* stream);
int dlp_gemv_moe(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y, int ldy,
                 int out_mode, const int32_t* expert_id_dev, int rows_per_expert, const float* out_scale_dev, void* stream);

/* ---- mixture of experts, any token count: device-built routing plan + grouped tcgen05 GEMM (one launch per projection) ----
 * dlp_moe_route_ex: like dlp_moe_route for up to 256 experts and rows of `ld` floats; renorm = 0 keeps the raw softmax weights (qwen2moe).
 * dlp_moe_plan: from ids[n_assign] (n_assign = T * top_k, token-major) builds the expert-sorted order - pos[i] = sorted row of
 *   assignment i, src[row] = assignment at a sorted row, bounds[n_expert + 1] - and the token-tile table of the grouped GEMMs
 *   (expert, first sorted row, rows; *n_tiles entries; capacity n_assign / token_tile + n_expert).  Stable in token order.
 * dlp_moe_gather: xs[row] = x[src[row] / top_k] (fp16 rows of K elements): the expert-sorted B operand.
 * dlp_linear_grouped: y[row] (op)= xs[row] . W_e^T for every sorted row, e = the row's expert; w0 (w1 for SwiGLU) is the stack of
 *   all experts [n_expert * rows_per_expert, K] in any supported weight type (block-quant decode fused as in dlp_linear).
 *   Launches an upper bound of token tiles; the kernel reads the real count from the plan - nothing comes back to the host.
 * dlp_moe_combine: h[t] += sum_j weights[t*k+j] * y[pos[t*k+j]] in slot order (deterministic).
 * ref: mm_ids_helper + expert_bounds (ggml-cuda/mmid.cu), mul_mat_q ids_dst path (mmq.cuh:3459), topk_moe_cuda (topk-moe.cu),
 * build_moe_ffn (llama.cpp/src/llama-graph.cpp:999-1266). */
typedef struct dlp_moe_plan_bufs {
    int32_t* pos;          /* [n_assign] */
    int32_t* src;          /* [n_assign] */
    int32_t* bounds;       /* [n_expert + 1] */
    int32_t* tile_expert;  /* [n_assign / token_tile + n_expert] */
    int32_t* tile_row0;
    int32_t* tile_rows;
    int32_t* n_tiles;      /* [1] */
} dlp_moe_plan_bufs;
int dlp_moe_route_ex(const float* logits, int ld, int T, int n_expert, int top_k, int renorm, int32_t* ids, float* weights, void* stream);
int dlp_moe_plan(const int32_t* ids, int n_assign, int n_expert, int token_tile, dlp_moe_plan_bufs* bufs, void* stream);
int dlp_moe_gather(const void* x, int ldx, const int32_t* src, int top_k, int n_assign, int K, void* xs, void* stream);
int dlp_linear_grouped(const dlp_weight* w0, const dlp_weight* w1, const void* xs, void* y, int ldy, int out_mode, int n_assign, int n_expert,
                       int rows_per_expert, int token_tile, const dlp_moe_plan_bufs* bufs, void* ws, size_t ws_bytes, void* stream);
int dlp_moe_combine(const float* y, const int32_t* pos, const float* weights, int T, int top_k, int E, float* h, void* stream);

/* ---- tensor-parallel decode with the all-reduce fused into the mat-vecs (NVLink stores from inside the kernels, no NCCL) ----
 * Each rank owns, per all-reduce site, a symmetric buffer [2 parities][tp][DLP_AR_MAX_T][E] fp32 that every rank can store into:
 * through peer mappings (dlp_symm_alloc + dlp_ipc_export / dlp_ipc_import, or any symmetric-memory allocator) and, when the
 * fabric offers NVLS, through one multicast alias of all ranks' buffers.  The buffers start filled with DLP_AR_EMPTY
 * (dlp_fill_u32).  A row-parallel mat-vec launched with out_mode DLP_OUT_PUSH_F32 stores its partial rows into slot [rank] of
 * every rank - `multimem.st` on the alias (one store, replicated by the switch) or one store per peer pointer; a value is its
 * own arrival signal (a producer never stores the bit pattern of -0.0f), so there is no flag, fence or counter.  The next
 * mat-vec (or dlp_ar_consume) launched with a dlp_ar_in spins on the values of the current parity, folds
 *   h_out = h_in + sum_r buf[r]     into its RMSNorm prologue, and re-arms the other parity for the step after next.
 * `epoch` is a device-resident counter with the same value on every rank (dlp_bump once per forward pass); parity = epoch & 1.
 * Every rank must run the same sequence of sites each step; a peer that never stores shows up in `error` after ~9 s.
 * ref: the hub-and-spoke broadcast/gather around split matmuls (ggml-cuda/ggml-cuda.cu:1680-1744) - no reduce exists there. */
enum { DLP_OUT_PUSH_F32 = 4 };
#define DLP_AR_MAX_T 4
#define DLP_AR_EMPTY 0x80000000u
typedef struct dlp_ar_in {
    float* buf;              /* local buffer of this site (NULL: nothing to consume) */
    const int32_t* epoch;
    float* h_out;            /* updated residual stream (written by CTA 0) */
    int32_t* error;          /* optional: set to 1 if a peer never delivered */
    int32_t tp;
    int32_t reserved;
} dlp_ar_in;
typedef struct dlp_ar_out {
    float* peer_buf[8];      /* this site's buffer on every rank (own rank included) */
    float* mc_buf;           /* multicast alias of this site's buffer, or NULL (then peer_buf is used) */
    const int32_t* epoch;
    int32_t rank;
    int32_t tp;
} dlp_ar_out;
/* Sharded greedy sampling: `logits` [R, V] are this rank's vocabulary shard (global ids v0 .. v0+V); every rank ends up with the
 * same global arg-max in out[R] (R <= DLP_TOKX_MAX_R).  Slots: uint64 [2 parities][tp][DLP_TOKX_MAX_R], zero-initialised, symmetric
 * like the buffers above. */
#define DLP_TOKX_MAX_R 64
typedef struct dlp_tokx {
    void* peer_slots[8];
    void* mc_slots;          /* multicast alias or NULL */
    void* slots;             /* local */
    const int32_t* epoch;
    int32_t* error;
    int32_t rank;
    int32_t tp;
} dlp_tokx;
int dlp_argmax_tp(const float* logits, int R, int V, int ld, int v0, const dlp_tokx* x, int32_t* out, void* stream);
int dlp_fill_u32(void* p, size_t n_words, uint32_t value, void* stream);
int dlp_gemv_tp(const dlp_weight* w0, const dlp_weight* w1, const void* x, const float* norm_w, float eps, void* y,
                int ldy, int out_mode, int T, const dlp_ar_in* ar_in, const dlp_ar_out* ar_out, void* stream);
int dlp_ar_consume(const float* h_in, float* h_out, const dlp_ar_in* ar_in, int E, int T, void* stream);
int dlp_bump(int32_t* counter, void* stream);
/* ---- tensor-parallel prefill / batched decode: collectives fused into the tcgen05 GEMM (sequence-parallel layout) ----
 * The residual stream is sharded by token rows (rank r owns rows [r*rows, (r+1)*rows)).
 *   GEMM -> reduce-scatter: a row-parallel GEMM launched with epilogue DLP_EPI_RS_PUSH stores its partial tile directly
 *     into the owner rank's reduce buffer [src rank][rows][N] over NVLink; the last CTA publishes rs_epoch to all peers.
 *   dlp_rs_norm_ag: the owner folds the tp partials into its residual shard, applies RMSNorm and stores the fp16 rows into
 *     every peer's gather buffer [T][E]; the last CTA publishes ag_epoch.
 *   all-gather -> GEMM: the next column-parallel GEMM reads X from the gather buffer; its TMA producer waits, tile by tile,
 *     only for the owners whose rows the tile covers (x_flags[src] >= x_epoch).
 * ref: no counterpart kernel - the reference broadcasts and gathers around every split matmul through one hub GPU
 * (ggml-cuda/ggml-cuda.cu:1451-1769). */
enum { DLP_EPI_RS_PUSH = 5 };
typedef struct dlp_gemm_coll {
    const uint32_t* x_flags;     /* NULL: X is local */
    int32_t x_rows_per_src;
    uint32_t x_epoch;
    int32_t* error;
    float* rs_peer_buf[8];
    uint32_t* rs_peer_flags[8];
    uint32_t* rs_done;           /* local scratch counter, zero-initialised */
    int32_t rs_rows, rs_rank, rs_tp;
    uint32_t rs_epoch;
    const uint32_t* step;        /* optional device-resident forward-pass counter: epochs become *step * 4096 + the values above (their
                                    position inside the pass), which keeps a captured CUDA graph valid from step to step */
    void* ws;                    /* optional split-K workspace (see dlp_linear `scratch`): lets a narrow shard use all SMs */
    size_t ws_bytes;
} dlp_gemm_coll;
typedef struct dlp_rs_norm_ag_desc {
    float* h_shard;
    const float* red_buf;        /* NULL: nothing to reduce (first layer) */
    const uint32_t* red_flags;
    uint32_t red_epoch;
    uint32_t pad0;
    const float* norm_w;         /* NULL: only fold the reduction into h_shard */
    float eps;
    uint32_t pad1;
    void* ag_peer_buf[8];
    uint32_t* ag_peer_flags[8];
    uint32_t ag_epoch;
    uint32_t pad2;
    uint32_t* done;
    int32_t* error;
    int32_t rank, tp, rows, E;
    const uint32_t* step;        /* as in dlp_gemm_coll */
} dlp_rs_norm_ag_desc;
/* x fp16 [T, ldx]; epilogue: 0 f16, 1 f32, 2 atomic add f32, 3 SwiGLU f16 (w1 = up), 5 reduce-scatter push (y unused). */
int dlp_gemm_sp(const dlp_weight* w0, const dlp_weight* w1, const void* x, int ldx, void* y, int ldy, int epilogue, int T,
                const dlp_gemm_coll* coll, void* stream);
int dlp_rs_norm_ag(const dlp_rs_norm_ag_desc* d, void* stream);
/* symmetric memory plumbing (set-up time only; these do allocate / synchronise) */
int dlp_symm_alloc(size_t bytes, void** ptr);
int dlp_symm_free(void* ptr);
int dlp_ipc_export(void* ptr, void* handle64);
int dlp_ipc_import(const void* handle64, void** ptr);
int dlp_ipc_close(void* ptr);
int dlp_memcpy_d2h(void* dst, const void* src, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* DLP_B200_H */
