// Engine state behind the opaque bfly_ctx (see include/butterfly_b200.h for the public contract).
#pragma once

#include <map>
#include <string>
#include <vector>

#include "../../include/butterfly_b200.h"
#include "kernels.cuh"
#include "decode_step.h"

namespace bfly {

constexpr int kMaxSplits = 16;           // split-K cap used by the planner inside the engine
constexpr int kMaxGreedySteps = 4096;    // history rows kept for bfly_decode_greedy
constexpr size_t kPinnedBytes = 1 << 20;
constexpr int kPrefillBf16MinRows = 256;     // forwards of at least this many rows all-reduce their partial sums as bf16
constexpr int kMoeGroupedMinTokens = 256;   // MoE forwards of at least this many rows gather tokens per expert

struct NcclId { char bytes[128]; };      // layout of ncclUniqueId
int nccl_unique_id(void* out128);

struct DevTensor {
  bf16* ptr;
  int64_t rows, cols;
};

struct LayerW {
  bf16 *attn_norm_w, *attn_norm_b, *ffn_norm_w, *ffn_norm_b;
  bf16 *wqkv, *bqkv, *wo, *bo;
  bf16 *w_in, *b_up, *w_down, *b_down;   // w_in = w_gate_up (gated) or w_up
  bf16* router;
};

struct GraphEntry {
  cudaGraphExec_t exec;
  int kernel_nodes;
};

// Device-side plan of the fused decode step kernel for one (batch, want_logits) pair.
struct StepPlan {
  StepParams p{};
  int grid = 0, n_counters = 0;
};

struct Engine {
  bfly_model_desc d{};
  int device = 0, sm_count = 0;
  cudaStream_t stream = nullptr;
  std::string last_error;

  // derived shard geometry (mirrors butterfly_b200/plan.py)
  int q_local = 0, kv_local = 0, qkv_rows = 0, experts_local = 0;
  int n_blocks = 1, f_block = 0, ffn_in_rows = 0, ffn_act_cols = 0;
  bool gated = true;
  bool fused_swiglu = true;                // gate/up GEMM applies the SwiGLU in its epilogue
  bool fused_rope = false;                 // decode attention folds the QKV slices, RoPE and the KV append itself
  bool moe_grouped = true;                 // MoE prefill runs each expert over its routed tokens only
  int moe_rows_cap = 0;                    // max_tokens * min(top_k, experts_local)
  int *moe_cnt = nullptr, *moe_off = nullptr, *moe_idx = nullptr;
  float *moe_rw = nullptr, *moe_out = nullptr;
  bf16* moe_xg = nullptr;
  int v_shard = 0, v0 = 0, v_local = 0;
  int pages_per_seq = 0;
  int64_t n_pages = 0, layer_kv_elems = 0;
  int host_len = 0;                       // tokens cached per sequence (uniform), tracked on the host

  // weights
  std::map<std::string, DevTensor> weights;
  std::vector<LayerW> layers;
  bf16 *embed = nullptr, *pos_embed = nullptr, *final_norm_w = nullptr, *final_norm_b = nullptr, *lm_head = nullptr;
  bool ready = false;

  // device state
  bf16* kv_pages = nullptr;
  int *page_table = nullptr, *seq_lens = nullptr, *step_counter = nullptr;
  float2* rope_table = nullptr;
  float *residual = nullptr, *ws = nullptr, *red = nullptr, *part_o = nullptr, *part_ml = nullptr, *route_w = nullptr;
  bf16 *xn = nullptr, *q = nullptr, *attn_out = nullptr, *act = nullptr, *red_bf = nullptr;
  bool prefill_ar_bf16 = false;
  float *logits = nullptr, *logits_all = nullptr;
  unsigned long long *keys = nullptr, *keys_all = nullptr;
  int *tokens_in = nullptr, *cur_tokens = nullptr, *history = nullptr;
  size_t ws_elems = 0, part_o_elems = 0, part_ml_elems = 0;   // scratch capacities, in elements
  void* pinned = nullptr;
  std::vector<void*> allocations;
  uint64_t weight_bytes = 0, kv_bytes = 0, workspace_bytes = 0;

  void* comm = nullptr;                   // ncclComm_t
  // decode-time all-reduce over peer memory (allreduce_norm.cu)
  ArPeers ar{};
  bool p2p_ready = false;
  unsigned* ar_step = nullptr;
  int ar_mode = -1;                       // -1 by size, 0 one-shot, 1 two-shot (set_ar_mode)
  int ar_lamport = 1;                     // flag-free decode all-reduce protocol (allreduce_norm.cu)
  int ar_chunks = 0, ar_bf16 = 0;         // cluster CTAs per row of the fused all-reduce (0 = one CTA per row); bf16 payload
  std::vector<void*> ipc_opened;
  int p2p_init();
  std::map<int, GraphEntry> graphs;       // key = B*2 + want_logits
  bool fused_step = true;                 // decode runs as ONE persistent kernel per step where the model allows it
  std::map<int, StepPlan> step_plans;     // same key
  bool step_supported(int B) const;
  int build_step_plan(int B, bool want_logits, StepPlan** out);
  int forward_step(int B, bool want_logits, bool record_history);
  int read_step_trace(unsigned long long* out, int max_entries, int* n_out, int* splits_out);
  int last_graph_nodes = 0;

  static int create(const bfly_model_desc& desc, int device, Engine** out);
  ~Engine();

  int alloc_weight(const std::string& name, int64_t rows, int64_t cols);
  int upload_weight(const std::string& name, const void* host, int64_t rows, int64_t cols);
  int fill_weight_block(const std::string& name, int64_t dst_row, int64_t dst_col, uint64_t key, int64_t ncols_full,
                        int64_t r0, int64_t c0, int64_t rows, int64_t cols, float scale, float offset);
  int read_weight(const std::string& name, void* host, int64_t max_bytes);
  int weights_ready();
  int comm_init(const void* id128, int world, int rank);
  int set_page_table(const int32_t* table_host, int n_seqs, int pps);
  int set_ar_mode(int mode);
  int ar_two_shot_by_size(int T) const;
  int bench_allreduce(int T, int iters, float* us_allreduce, float* us_plain);

  int prefill(const int32_t* tokens_host, int B, int S, float* logits_host, int32_t* next_host);
  int prefill_at(int seq0, const int32_t* tokens_host, int B, int S, float* logits_host, int32_t* next_host);
  int decode(const int32_t* tokens_host, int B, float* logits_host, int32_t* next_host);
  int decode_enqueue(int B, int n_steps);
  int decode_greedy(int B, int n_steps, int32_t* tokens_out_host);
  int read_history(int B, int n_steps, int32_t* tokens_out_host);

 private:
  int find(const std::string& name, int64_t rows, int64_t cols, bool required, bf16** out);
  int check_ready(int B);
  int gemm(const bf16* W, const bf16* X, int64_t ldx, int M, int N, int K, int* slices_out);
  int gemm_swiglu(const bf16* W, const bf16* X, int N);
  int moe_ffn_grouped(const bf16* w_in, const bf16* w_down, int T);
  int row_parallel_finish(int T, int splits, bool allow_fused, float* partials, const float** src, int* n_partials,
                          bool* fused, const bf16** src_bf16);
  int norm_after(int T, const float* src, int n, bool fused, int* ar_call, const bf16* bias, const bf16* gamma,
                 const bf16* beta, const bf16* src_bf16 = nullptr);
  int all_reduce_f32(float* buf, size_t count);
  int all_reduce_bf16(bf16* buf, size_t count);
  int all_gather_bytes(const void* src, void* dst, size_t bytes_per_rank);
  int forward(const int* tokens_dev, int seq0, int nb, int S, bool want_logits, bool record_history);
  int fetch_outputs(int seq0, int B, float* logits_host, int32_t* next_host);
  int decode_graph(int B, bool want_logits, GraphEntry** out);
};

}  // namespace bfly

struct bfly_ctx {
  bfly::Engine* engine;
};
