// Host-side launchers of every kernel on the partitioned block forward (SURVEY.md section 2.5,
// K1..K9).  All pointers are device pointers; all launches go to `stream`; nothing allocates.
#pragma once

#include "common.cuh"

namespace bfly {

typedef __nv_bfloat16 bf16;

// Paged KV cache of ONE layer on one rank: [n_pages][2 (K,V)][kv_local][page_size][head_dim] bf16,
// addressed through page_table[seq][logical page].
struct KvLayout {
  bf16* base;
  const int* page_table;
  int pages_per_seq;
  int page_size;
  int kv_local;
  int head_dim;
};

// Element offsets inside one (page, K|V, head) block of page_size*head_dim elements.
//   K: row-major [slot][d]; for d >= 64 the 16-byte chunks of a row are XOR-swizzled by token parity
//      (chunk ^ 4*(slot&1)) so the tensor-core fragment reads in attn_decode.cu are bank-conflict free
//      after a plain linear bulk copy into shared memory.
//   V: transposed per 32-token sub-block, [slot/32][d][slot%32], so a tile of V^T is contiguous and
//      token-major fragments are 16-byte reads.
__host__ __device__ inline int64_t kv_k_offset(int slot, int dd, int d) {
  int chunk = dd >> 3;
  if (d >= 64) chunk ^= (slot & 1) << 2;
  return (int64_t)slot * d + (chunk << 3) + (dd & 7);
}
__host__ __device__ inline int64_t kv_v_offset(int slot, int dd, int d) {
  return (int64_t)(slot >> 5) * d * 32 + (int64_t)dd * 32 + (slot & 31);
}

enum ActMode : int { ACT_SWIGLU = 0, ACT_GELU_TANH = 1 };

// Token t = b*S + s of a forward over S new tokens per sequence sits at position seq_lens[b] + s.

// synth.cu -- weights as a pure function of (key, row, col); bit-identical to weights.synth_block
int launch_fill_synth(bf16* dst, int64_t dst_pitch, uint64_t key, int64_t ncols_full, int64_t r0, int64_t c0,
                      int64_t rows, int64_t cols, float scale, float offset, cudaStream_t stream);

// elementwise.cu
// Clears argmax_keys[0..n_keys) and, when `history` is given, records the tokens fed at this step at
// history[*step_counter * T + t] (free-running greedy decode).
int launch_embed(const int* tokens, const bf16* table, const bf16* pos_table, const int* seq_lens, int S,
                 float* residual, int T, int h, int vocab, unsigned long long* argmax_keys, int n_keys,
                 int* history, const int* step_counter, cudaStream_t stream);
// partial_bf16 (optional): one more complete partial [T][h] in bf16, added like a slice (prefill at tp > 1)
int launch_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                    const bf16* bias, const bf16* gamma, const bf16* beta, float eps, int layer_norm, bf16* out,
                    int T, int h, cudaStream_t stream, const bf16* partial_bf16 = nullptr);
int launch_rope_kv(const float* qkv, int n_partials, int64_t partial_stride, const bf16* bias,
                   const float2* rope_table, const int* seq_lens, int S, int T, int nq, int nk, bf16* q_out,
                   KvLayout kv, cudaStream_t stream);
int launch_act(const float* in, int n_partials, int64_t partial_stride, const bf16* bias, int mode,
               int n_blocks, int f_block, const float* route_w, bf16* out, int T, cudaStream_t stream);
int launch_lm_finalize(const float* partials, int n_partials, int64_t partial_stride, int T, int v_local,
                       int v0, float* logits, int64_t logits_ld, unsigned long long* keys, cudaStream_t stream);
int launch_pick_tokens_pitched(const unsigned long long* keys, int n_ranks, int pitch, int T, int* tokens_out,
                               cudaStream_t stream);
int launch_advance(int* seq_lens, int B, int S, int* step_counter, unsigned* ar_step, cudaStream_t stream);
// routed-token grouping for MoE prefill: per local expert e, rows [off[e], off[e]+cnt[e]) of idx (token id), rw
// (routing weight) and xg (gathered activations); off has n_local + 1 entries
int launch_moe_group(const float* route_w, const bf16* xn, int T, int h, int n_local, int rows_cap, int* cnt, int* off,
                     int* idx, float* rw, bf16* xg, cudaStream_t stream);
int launch_moe_router(const bf16* xn, const bf16* router, int T, int h, int n_experts, int top_k,
                      int first_local, int n_local, float* route_w, cudaStream_t stream);
int launch_reduce_partials(const float* partials, int n_partials, int64_t partial_stride, float* out,
                           int64_t count, cudaStream_t stream);
int launch_reduce_partials_bf16(const float* partials, int n_partials, int64_t partial_stride, bf16* out, int64_t count,
                                cudaStream_t stream);

// allreduce_norm.cu -- decode-time all-reduce over NVLink peer memory fused with residual add + norm
constexpr int kMaxPeers = 8;
struct ArPeers {
  float* recv[kMaxPeers];        // rank r's receive buffer [2 parities][tp][max_rows][h] fp32 (peer-mapped)
  unsigned* flags[kMaxPeers];    // rank r's epoch flags    [2 parities][tp][max_rows]
  unsigned* cflags[kMaxPeers];   // rank r's per-chunk flags [2 parities][tp][max_rows][8] (fused decode step kernel)
  int tp, rank, max_rows;
};
int launch_ar_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                       const bf16* bias, const bf16* gamma, const bf16* beta, float eps, int layer_norm, bf16* out,
                       int T, int h, const ArPeers& peers, int call_idx, const unsigned* step_ptr, int calls_per_step,
                       int two_shot, int chunks, int bf16_payload, int lamport, cudaStream_t stream);
// lamport = 1: flag-free protocol (self-validating words, sentinel -0.0f); one CTA per row, fp32 payload.
int ar_fill_sentinel(float* recv, size_t n_floats, cudaStream_t stream);
// chunks: 0 = one CTA per row; C >= 1 = each row split over a cluster of C CTAs (ar_chunks_for picks C from h).
// bf16_payload (cluster variant only): partial sums cross NVLink as bf16.
int ar_chunks_for(int h);

// attn_decode.cu -- paged-KV attention, one query token per (virtual) sequence, split over the context
int attn_decode_splits(int T, int kv_local, int max_ctx, int sm_count);
int launch_attn_decode(const bf16* q, KvLayout kv, const int* seq_lens, int S, int T, int nq, float* part_o,
                       float* part_ml, int n_splits, bf16* out, cudaStream_t stream);

// decode only (S == 1): the same attention with the RoPE / KV-append step folded in -- reads the QKV projection's
// fp32 slices [n_partials][T][(nq + 2 kv_local) * head_dim] (+bias), appends the new K/V row to the cache itself
int launch_attn_decode_fused(const float* qkv, int n_partials, int64_t partial_stride, const bf16* bias,
                             const float2* rope_table, KvLayout kv, const int* seq_lens, int T, int nq, float* part_o,
                             float* part_ml, int n_splits, bf16* out, cudaStream_t stream);

// attn_prefill.cu -- causal attention of S new tokens per sequence over the paged cache (tensor cores)
int launch_attn_prefill(const bf16* q, KvLayout kv, const int* seq_lens, int S, int B, int nq, bf16* out,
                        cudaStream_t stream);

// gemm_tcgen05.cu -- C[n][m] (+= over split-K slices) = sum_k W[m][k] * X[n][k]; W, X bf16 K-major.
struct GemmPlan {
  int M, N, K;          // M = output features (dual: gated features = half the weight rows), N = tokens
  int bn;               // token tile (UMMA N), multiple of 16, <= 256
  int splits;           // split-K factor
  int cluster;          // 1: the K-slices form a thread-block cluster and are reduced through DSMEM into ONE
                        //    output (deterministic, rank order); 0: each slice is written to out + s * split_stride
  int pair;             // 1: cta_group::2 CTA pairs (tensor-bound regime)
  int persistent;       // 1: persistent kernel with a double-buffered TMEM accumulator (tensor-bound regime)
  int dual;             // 1: W = n_blocks x [gate(f_block) | up(f_block)] rows, fused SwiGLU epilogue (bf16 out)
  int stages;
  size_t smem_bytes;
};
struct GemmOut {
  float* out_f32;       // fp32 [slices][N][ldc]   (slices = 1 when the plan reduces in-cluster or splits == 1)
  int64_t split_stride;
  bf16* out_bf16;       // bf16 [N][ldc]           (needs one complete K sum)
  int64_t ldc;
  int n_blocks, f_block;    // dual only
  const float* route_w;     // dual + MoE: [N][n_blocks], nullptr otherwise
  // Grouped MoE prefill (persistent plans only): this launch handles the token rows
  // [*row_off_dev, *row_off_dev + min(N, *n_rows_dev)) of X -- counts that only exist on the device.
  const int* n_rows_dev;    // nullptr: all N rows, starting at row 0
  const int* row_off_dev;
  const int* row_map;       // fp32 output: row r of X accumulates (+=) into output row row_map[r]
  int64_t x_rows;           // rows X really has (TMA bound); 0 = N
  int64_t ldw;              // row pitch of W in elements; 0 = K
};
int gemm_init();        // resolves cuTensorMapEncodeTiled, sets kernel attributes (once per process)
GemmPlan gemm_plan(int M, int N, int K, int sm_count, int max_splits, int dual = 0, int force_persistent = 0);
// number of fp32 slices a consumer must fold for this plan
inline int gemm_slices(const GemmPlan& p) { return (p.cluster || p.splits == 1) ? 1 : p.splits; }
// ldx = row pitch of X in elements (>= K, multiple of 8).
int launch_gemm_ex(const bf16* W, const bf16* X, int64_t ldx, const GemmPlan& plan, const GemmOut& out,
                   cudaStream_t stream);
int launch_gemm(const bf16* W, const bf16* X, int64_t ldx, const GemmPlan& plan, float* out_f32,
                int64_t split_stride, bf16* out_bf16, int64_t ldc, cudaStream_t stream);

}  // namespace bfly
