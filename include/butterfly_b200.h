/*
 * butterfly_b200 -- C ABI of the B200-native partitioned transformer block forward.
 *
 * Drop-in boundary.  The reference repository TensorHusker/Butterfly defines
 * no plugin / operator / FFI interface for this path: it contains no source at all
 * (ref:CLAUDE.md:13-15 "newly initialized repository ... Expected Architecture (To Be
 * Implemented)"; the intended engine / partitioning / communication components are one-line
 * bullets at ref:CLAUDE.md:19-21).  Each entry point below therefore cites the intended component
 * it realises instead of a reference function.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success
 * or a BFLY_E* code, and never throws; bfly_last_error() gives the message (per calling thread,
 * and sticky on the ctx).  One process per GPU rank, one host thread per ctx, one CUDA stream per
 * ctx; functions are not re-entrant on the same ctx.  The ctx owns all device memory (weights, KV
 * pages, workspaces); host buffers and device pointers passed to bfly_k_* are borrowed for the call.
 * There is no CPU fallback anywhere behind this header.
 */
#ifndef BUTTERFLY_B200_H
#define BUTTERFLY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFLY_OK 0
#define BFLY_EINVAL 1
#define BFLY_ECUDA 2
#define BFLY_ENCCL 3
#define BFLY_ESTATE 4
#define BFLY_ENOMEM 5

#define BFLY_FAMILY_GPT2 0
#define BFLY_FAMILY_LLAMA 1
#define BFLY_FAMILY_MIXTRAL 2

typedef struct bfly_ctx bfly_ctx;

/* Model shape + this rank's place in the tensor-parallel plan (butterfly_b200/plan.py). */
typedef struct bfly_model_desc {
  int32_t family;
  int32_t hidden, n_layers, n_heads, n_kv_heads, head_dim, ffn, vocab;
  int32_t n_experts, top_k;
  float norm_eps;
  float rope_theta;            /* 0 => learned absolute positions */
  int32_t max_pos;
  int32_t tied_head, bias;
  int32_t tp, rank;            /* tensor-parallel degree and this rank */
  int32_t page_size;           /* tokens per KV page (power of two >= 32; layout: butterfly_b200/kv_layout.py) */
  int32_t max_batch;           /* sequences resident in the KV cache */
  int32_t max_ctx;             /* KV capacity per sequence, tokens */
  int32_t max_tokens;          /* rows of one forward chunk (prefill is chunked to this) */
} bfly_model_desc;

typedef struct bfly_stats {
  uint64_t kernel_launches;    /* kernels of this library launched so far (graph replays included) */
  uint64_t weight_bytes;       /* device bytes held by weights */
  uint64_t kv_bytes;           /* device bytes held by the paged KV cache */
  uint64_t workspace_bytes;    /* device bytes held by activations / split-K scratch */
  int32_t sm_count;
  int32_t graph_nodes_decode;  /* kernel nodes in the last captured decode step */
} bfly_stats;

int bfly_version(void);
/* ctx may be NULL: returns the calling thread's last error. */
const char* bfly_last_error(const bfly_ctx* ctx);

/* ---- engine: intended "Distributed Inference Engine" (ref:CLAUDE.md:19) ------------------- */
int bfly_create(const bfly_model_desc* desc, int device, bfly_ctx** out);
int bfly_destroy(bfly_ctx* ctx);
int bfly_get_stats(const bfly_ctx* ctx, bfly_stats* out);

/* ---- weights: intended "Model Partitioning" (ref:CLAUDE.md:21); names = shard_spec() names -- */
int bfly_upload_weight(bfly_ctx* ctx, const char* name, const void* host_bf16, int64_t rows, int64_t cols);
int bfly_alloc_weight(bfly_ctx* ctx, const char* name, int64_t rows, int64_t cols);
/* Fill rows [dst_row, +rows) x cols [dst_col, +cols) of an allocated weight with the synthetic
 * block [r0.., c0..] of the logical tensor `key` (bit-identical to weights.synth_block). */
int bfly_fill_weight_block(bfly_ctx* ctx, const char* name, int64_t dst_row, int64_t dst_col, uint64_t key,
                           int64_t ncols_full, int64_t r0, int64_t c0, int64_t rows, int64_t cols, float scale,
                           float offset);
int bfly_read_weight(bfly_ctx* ctx, const char* name, void* host_bf16, int64_t max_bytes);
int bfly_weights_ready(bfly_ctx* ctx);   /* validates names/shapes; required before any forward */

/* ---- communication: intended "Communication Layer" (ref:CLAUDE.md:20), NCCL over NVLink ---- */
int bfly_comm_unique_id(void* out_id128);                       /* rank 0; 128 bytes */
int bfly_comm_init(bfly_ctx* ctx, const void* id128, int world, int rank);

/* Mode of the decode-time all-reduce over peer memory (tp > 1): -1 = chosen by size (two-shot at tp >= 8, and at
 * tp >= 4 when the rows of one call total >= 1 MiB), 0 = one-shot (every rank pushes its rows to every rank),
 * 1 = two-shot (row r is reduced by rank r % tp and pushed back).  Both modes add in rank order, so results
 * are bit-identical.  Call it at a quiescent point with the SAME value on every rank. */
int bfly_set_ar_mode(bfly_ctx* ctx, int mode);

/* Latency of the decode all-reduce in isolation (tp > 1, every rank calls it together): average microseconds per
 * call of the fused all-reduce + residual add + norm kernel over `iters` back-to-back calls on T rows of the
 * model's hidden size, and of the plain add + norm kernel on the same rows (the difference is the exchange). */
int bfly_bench_allreduce(bfly_ctx* ctx, int T, int iters, float* us_allreduce, float* us_plain);

/* ---- forward ------------------------------------------------------------------------------- */
/* Optional: install a page table int32[n_seqs][pages_per_seq] (default: sequence b owns pages
 * b*pages_per_seq ...).  Physical page ids must be unique and < max_batch*pages_per_seq. */
int bfly_set_page_table(bfly_ctx* ctx, const int32_t* table_host, int n_seqs, int pages_per_seq);
/* Prefill: tokens_host int32[B][S]; restarts the KV cache of sequences 0..B-1.  logits_host
 * (fp32 [B][vocab], last position, gathered over ranks) and next_tokens_host (int32[B]) may be NULL. */
int bfly_prefill(bfly_ctx* ctx, const int32_t* tokens_host, int B, int S, float* logits_host,
                 int32_t* next_tokens_host);
/* Ragged batches: restart only sequences seq0 .. seq0+n_seqs-1 with S tokens each (tokens_host
 * int32[n_seqs][S]); all other sequences keep their cache and length.  Load prompts of different
 * lengths group by group, then bfly_decode the whole batch: every kernel reads per-sequence lengths.
 * Outputs are for the n_seqs sequences of this call.  (No reference counterpart: the reference's
 * intended "Scheduling System", ref:CLAUDE.md:22, is unimplemented.) */
int bfly_prefill_at(bfly_ctx* ctx, int seq0, const int32_t* tokens_host, int n_seqs, int S, float* logits_host,
                    int32_t* next_tokens_host);
/* One decode step: tokens_host int32[B] (NULL = continue from the engine's own greedy tokens). */
int bfly_decode(bfly_ctx* ctx, const int32_t* tokens_host, int B, float* logits_host, int32_t* next_tokens_host);
/* n free-running greedy steps with everything resident on the device (one CUDA graph replay per
 * step, no host synchronisation until the end).  tokens_out_host int32[B][n_steps] = the token fed
 * at each step; may be NULL. */
int bfly_decode_greedy(bfly_ctx* ctx, int B, int n_steps, int32_t* tokens_out_host);
/* Same, but only enqueues the steps on the ctx stream (for device-side timing); no copy, no sync. */
int bfly_decode_enqueue(bfly_ctx* ctx, int B, int n_steps);
/* The tokens fed at the n_steps steps of the LAST bfly_decode_enqueue / bfly_decode_greedy of B sequences
 * (int32[B][n_steps]); synchronises the ctx stream. */
int bfly_read_history(bfly_ctx* ctx, int B, int n_steps, int32_t* tokens_out_host);
/* Profiling of the fused decode step (contexts created with BFLY_STEP_TRACE=1 in the environment): the
 * nanosecond timestamps at which CTA 0 saw each phase of the last step complete (entry 0 = kernel start, phase
 * order as in csrc/decode_step.cu); *n_entries = 0 when tracing is off. */
int bfly_read_step_trace(bfly_ctx* ctx, uint64_t* out_ns, int max_entries, int* n_entries, int* attn_splits);
int bfly_sync(bfly_ctx* ctx);
void* bfly_stream(bfly_ctx* ctx);        /* the ctx's cudaStream_t, for event timing by the caller */

/* ---- per-kernel entry points on raw device pointers (parity tests, profiling) -------------- */
/* All run on `stream` (a cudaStream_t, may be NULL = legacy default stream) and do not allocate.
 * kv_pages_bf16 is one layer's cache [n_pages][2][kv_local][page_size*head_dim] in the engine's layout:
 * K rows with the parity chunk swizzle, V transposed per 32-token block (csrc/kernels.cuh kv_k_offset /
 * kv_v_offset, mirrored by butterfly_b200/kv_layout.py). */
int bfly_k_fill_synth(void* dst_bf16, int64_t dst_pitch, uint64_t key, int64_t ncols_full, int64_t r0, int64_t c0,
                      int64_t rows, int64_t cols, float scale, float offset, void* stream);
int bfly_k_embed(const int32_t* tokens, const void* table_bf16, const void* pos_table_bf16, const int32_t* seq_lens,
                 int S, float* residual, int T, int h, int vocab, void* stream);
int bfly_k_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                    const void* bias_bf16, const void* gamma_bf16, const void* beta_bf16, float eps, int layer_norm,
                    void* out_bf16, int T, int h, void* stream);
int bfly_k_rope_kv(const float* qkv, int n_partials, int64_t partial_stride, const void* bias_bf16,
                   const float* rope_table, const int32_t* seq_lens, int S, int T, int nq, int nk, int head_dim,
                   void* q_out_bf16, void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                   int page_size, void* stream);
int bfly_k_attn_decode(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                       int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int S, int T, int nq,
                       float* part_o, float* part_ml, int n_splits, void* out_bf16, void* stream);
/* Decode step (one new token per sequence) as the engine runs it: RoPE + KV append + attention in ONE kernel.  qkv
 * holds the QKV projection's fp32 split-K slices [n_partials][T][(nq + 2 kv_local) * head_dim] (q | k | v columns,
 * partial_stride floats apart); the kernel folds them (+bias), rotates q and k (rope_table float2 [pos][head_dim/2],
 * NULL = no rotation), appends the new K/V row of sequence t at position seq_lens[t] to the cache, and attends
 * over seq_lens[t] + 1 positions.  Same result as bfly_k_rope_kv followed by bfly_k_attn_decode with S = 1. */
int bfly_k_attn_decode_rope(const float* qkv, int n_partials, int64_t partial_stride, const void* bias_bf16,
                            const float* rope_table, void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                            int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int T, int nq, float* part_o,
                            float* part_ml, int n_splits, void* out_bf16, void* stream);
/* Host-only (no GPU needed): the stream-K partition the fused decode step kernel uses for a weight matrix [M][K] on
 * `grid` CTAs -- first_out[t] = first CTA holding K blocks of 128-row tile t, nsl_out[t] = number of CTAs (= fp32 slices)
 * tile t is spread over; both arrays have ceil(M/128) entries.  CTA c owns units [c*U/g, (c+1)*U/g) of the U =
 * ceil(M/128)*ceil(K/64) (tile, K block) units in tile-major order, g = min(grid, U). */
int bfly_k_step_gemm_table(int M, int K, int grid, uint16_t* first_out, uint8_t* nsl_out, int* max_slices);
/* The context split the engine uses for a decode forward of T rows (0 on bad arguments): callers of
 * bfly_k_attn_decode size part_o = T*nq*n_splits*head_dim and part_ml = T*nq*n_splits*2 floats from it.
 * S > 1 is allowed (token s of sequence b attends to seq_lens[b] + s + 1 cached positions); seq_lens and
 * the page table must not be written by the kernel that precedes the call on the stream. */
int bfly_k_attn_decode_splits(int T, int kv_local, int max_ctx);
/* Causal attention of S new tokens per sequence (q is [B*S][nq*head_dim]); token s of sequence b sits at
 * position seq_lens[b] + s and sees keys 0 .. seq_lens[b] + s, which must already be in the cache. */
int bfly_k_attn_prefill(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                        int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int S, int B, int nq,
                        void* out_bf16, void* stream);
/* Same computation as bfly_k_attn_decode for callers that cannot supply or allocate the split-context
 * scratch (the host application's pool owns all of the GPU's memory, cudaMalloc is banned from the
 * hot path, or a binding forwards only inputs and outputs).  The scratch is a fixed 4 MiB array in the
 * library's device code, one per GPU.  With n_splits > 1 a call needs T*nq*n_splits*(head_dim+2) floats
 * <= 1,048,576, otherwise BFLY_EINVAL and nothing runs; n_splits == 1 uses no scratch and has no limit.
 * The first call on a device resolves the array's address, which (CUDA loads device code lazily) is also
 * when the 4 MiB are claimed on that GPU: make that call at start-up, before your pool takes the rest of
 * the memory and before any stream capture on that device; it returns BFLY_ECUDA if the memory is not
 * there.  Later calls obtain and release no device memory.  Calls on one device share the array
 * (contents of the previous call remain; nothing needs zeroing); calls that overlap on the GPU race --
 * serialise them.
 * No counterpart in the reference: this entry point is an addition of this library. */
int bfly_k_attn_decode_static(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table,
                              int pages_per_seq, int page_size, int kv_local, int head_dim, const int32_t* seq_lens,
                              int S, int T, int nq, int n_splits, void* out_bf16, void* stream);
int bfly_k_act(const float* in, int n_partials, int64_t partial_stride, const void* bias_bf16, int mode,
               int n_blocks, int f_block, const float* route_w, void* out_bf16, int T, void* stream);
int bfly_k_lm_finalize(const float* partials, int n_partials, int64_t partial_stride, int T, int v_local, int v0,
                       float* logits, int64_t logits_ld, uint64_t* keys, int32_t* tokens_out, void* stream);
int bfly_k_moe_router(const void* xn_bf16, const void* router_bf16, int T, int h, int n_experts, int top_k,
                      int first_local, int n_local, float* route_w, void* stream);
/* C[n][m] = sum_k W[m][k] X[n][k].  splits/bn/stages <= 0 pick the planner's choice.  K-slices of token
 * tiles <= 64 wide and splits <= 8 are reduced inside a thread-block cluster (one output slice);
 * otherwise out_f32 receives one fp32 slice per split, `split_stride` floats apart (row pitch ldc).
 * *slices_written tells the caller how many slices to fold.  out_bf16 needs a single complete sum. */
int bfly_k_gemm(const void* w_bf16, const void* x_bf16, int M, int N, int K, int64_t ldx, int splits, int bn,
                int stages, float* out_f32, int64_t split_stride, void* out_bf16, int64_t ldc, int* slices_written,
                void* stream);
/* act[n][b*f_block + j] = route_w[n][b] * silu(x . gate_row) * (x . up_row) for W = n_blocks blocks of
 * [gate(f_block) | up(f_block)] rows: the gate/up projection with the SwiGLU fused into the epilogue. */
int bfly_k_gemm_swiglu(const void* w_gate_up_bf16, const void* x_bf16, int N, int K, int n_blocks, int f_block,
                       const float* route_w, int splits, void* act_out_bf16, void* stream);
int bfly_k_gemm_plan(int M, int N, int K, int* splits, int* bn, int* stages);

#ifdef __cplusplus
}
#endif
#endif /* BUTTERFLY_B200_H */
