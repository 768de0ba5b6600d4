// extern "C" surface of libbutterfly_b200.so (declared in include/butterfly_b200.h).
// Nothing here throws across the ABI: every entry catches and converts to a BFLY_E* code.
#include <algorithm>
#include <exception>

#include "engine.h"

namespace bfly {
static thread_local std::string t_last_error;
void set_error(const std::string& msg) { t_last_error = msg; }
}  // namespace bfly

using namespace bfly;

namespace {
template <typename F>
int guarded(bfly_ctx* ctx, F&& f) {
  int rc;
  try {
    rc = f();
  } catch (const std::exception& ex) {
    set_error(std::string("internal exception: ") + ex.what());
    rc = BFLY_EINVAL;
  } catch (...) {
    set_error("internal exception");
    rc = BFLY_EINVAL;
  }
  if (rc != BFLY_OK && ctx && ctx->engine) ctx->engine->last_error = t_last_error;
  return rc;
}
inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline const bf16* cb(const void* p) { return reinterpret_cast<const bf16*>(p); }
inline bf16* mb(void* p) { return reinterpret_cast<bf16*>(p); }
#define NEED_CTX(ctx)                                        \
  if (!(ctx) || !(ctx)->engine) {                            \
    set_error("null bfly_ctx");                              \
    return BFLY_EINVAL;                                      \
  }
}  // namespace

extern "C" {

int bfly_version(void) { return 100; }

const char* bfly_last_error(const bfly_ctx* ctx) {
  if (ctx && ctx->engine && !ctx->engine->last_error.empty()) return ctx->engine->last_error.c_str();
  return t_last_error.c_str();
}

int bfly_create(const bfly_model_desc* desc, int device, bfly_ctx** out) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(desc != nullptr && out != nullptr, "bfly_create: null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
      set_error(std::string("no CUDA device available (butterfly_b200 has no CPU fallback): ") + cudaGetErrorString(e));
      return BFLY_ECUDA;
    }
    BFLY_CHECK(device >= 0 && device < n, "bfly_create: device index out of range");
    cudaDeviceProp prop;
    BFLY_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
      set_error(std::string("device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor) +
                "; this library contains sm_100a code only");
      return BFLY_ECUDA;
    }
    Engine* eng = nullptr;
    BFLY_TRY(Engine::create(*desc, device, &eng));
    *out = new bfly_ctx{eng};
    return BFLY_OK;
  });
}

int bfly_destroy(bfly_ctx* ctx) {
  if (!ctx) return BFLY_OK;
  return guarded(nullptr, [&]() -> int {
    if (ctx->engine) {
      cudaSetDevice(ctx->engine->device);
      cudaStreamSynchronize(ctx->engine->stream);
      delete ctx->engine;
    }
    delete ctx;
    return BFLY_OK;
  });
}

int bfly_get_stats(const bfly_ctx* ctx, bfly_stats* out) {
  NEED_CTX(ctx);
  if (!out) { set_error("null stats"); return BFLY_EINVAL; }
  const Engine* e = ctx->engine;
  out->kernel_launches = g_launch_count;
  out->weight_bytes = e->weight_bytes;
  out->kv_bytes = e->kv_bytes;
  out->workspace_bytes = e->workspace_bytes;
  out->sm_count = e->sm_count;
  out->graph_nodes_decode = e->last_graph_nodes;
  return BFLY_OK;
}

int bfly_upload_weight(bfly_ctx* ctx, const char* name, const void* host_bf16, int64_t rows, int64_t cols) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(name && host_bf16, "bfly_upload_weight: null argument");
    return ctx->engine->upload_weight(name, host_bf16, rows, cols);
  });
}

int bfly_alloc_weight(bfly_ctx* ctx, const char* name, int64_t rows, int64_t cols) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(name != nullptr, "bfly_alloc_weight: null name");
    return ctx->engine->alloc_weight(name, rows, cols);
  });
}

int bfly_fill_weight_block(bfly_ctx* ctx, const char* name, int64_t dst_row, int64_t dst_col, uint64_t key,
                           int64_t ncols_full, int64_t r0, int64_t c0, int64_t rows, int64_t cols, float scale,
                           float offset) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(name != nullptr, "bfly_fill_weight_block: null name");
    return ctx->engine->fill_weight_block(name, dst_row, dst_col, key, ncols_full, r0, c0, rows, cols, scale, offset);
  });
}

int bfly_read_weight(bfly_ctx* ctx, const char* name, void* host_bf16, int64_t max_bytes) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(name && host_bf16, "bfly_read_weight: null argument");
    return ctx->engine->read_weight(name, host_bf16, max_bytes);
  });
}

int bfly_weights_ready(bfly_ctx* ctx) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->weights_ready(); });
}

int bfly_comm_unique_id(void* out_id128) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(out_id128 != nullptr, "bfly_comm_unique_id: null buffer");
    return nccl_unique_id(out_id128);
  });
}

int bfly_comm_init(bfly_ctx* ctx, const void* id128, int world, int rank) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(id128 != nullptr, "bfly_comm_init: null id");
    return ctx->engine->comm_init(id128, world, rank);
  });
}

int bfly_set_page_table(bfly_ctx* ctx, const int32_t* table_host, int n_seqs, int pages_per_seq) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(table_host != nullptr, "bfly_set_page_table: null table");
    return ctx->engine->set_page_table(table_host, n_seqs, pages_per_seq);
  });
}

int bfly_set_ar_mode(bfly_ctx* ctx, int mode) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->set_ar_mode(mode); });
}

int bfly_read_step_trace(bfly_ctx* ctx, uint64_t* out_ns, int max_entries, int* n_entries, int* attn_splits) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CHECK(out_ns && n_entries && attn_splits && max_entries > 0, "read_step_trace: arguments");
    return ctx->engine->read_step_trace(reinterpret_cast<unsigned long long*>(out_ns), max_entries, n_entries, attn_splits);
  });
}

int bfly_bench_allreduce(bfly_ctx* ctx, int T, int iters, float* us_allreduce, float* us_plain) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->bench_allreduce(T, iters, us_allreduce, us_plain); });
}

int bfly_prefill(bfly_ctx* ctx, const int32_t* tokens_host, int B, int S, float* logits_host,
                 int32_t* next_tokens_host) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->prefill(tokens_host, B, S, logits_host, next_tokens_host); });
}

int bfly_prefill_at(bfly_ctx* ctx, int seq0, const int32_t* tokens_host, int n_seqs, int S, float* logits_host,
                    int32_t* next_tokens_host) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    return ctx->engine->prefill_at(seq0, tokens_host, n_seqs, S, logits_host, next_tokens_host);
  });
}

int bfly_decode(bfly_ctx* ctx, const int32_t* tokens_host, int B, float* logits_host, int32_t* next_tokens_host) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->decode(tokens_host, B, logits_host, next_tokens_host); });
}

int bfly_decode_greedy(bfly_ctx* ctx, int B, int n_steps, int32_t* tokens_out_host) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->decode_greedy(B, n_steps, tokens_out_host); });
}

int bfly_decode_enqueue(bfly_ctx* ctx, int B, int n_steps) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->decode_enqueue(B, n_steps); });
}

int bfly_read_history(bfly_ctx* ctx, int B, int n_steps, int32_t* tokens_out_host) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int { return ctx->engine->read_history(B, n_steps, tokens_out_host); });
}

int bfly_sync(bfly_ctx* ctx) {
  NEED_CTX(ctx);
  return guarded(ctx, [&]() -> int {
    BFLY_CUDA(cudaSetDevice(ctx->engine->device));
    BFLY_CUDA(cudaStreamSynchronize(ctx->engine->stream));
    return BFLY_OK;
  });
}

void* bfly_stream(bfly_ctx* ctx) { return (ctx && ctx->engine) ? (void*)ctx->engine->stream : nullptr; }

// ---- per-kernel entry points -----------------------------------------------------------------
int bfly_k_fill_synth(void* dst_bf16, int64_t dst_pitch, uint64_t key, int64_t ncols_full, int64_t r0, int64_t c0,
                      int64_t rows, int64_t cols, float scale, float offset, void* stream) {
  return guarded(nullptr, [&]() -> int {
    return launch_fill_synth(mb(dst_bf16), dst_pitch, key, ncols_full, r0, c0, rows, cols, scale, offset,
                             as_stream(stream));
  });
}

int bfly_k_embed(const int32_t* tokens, const void* table_bf16, const void* pos_table_bf16, const int32_t* seq_lens,
                 int S, float* residual, int T, int h, int vocab, void* stream) {
  return guarded(nullptr, [&]() -> int {
    return launch_embed(tokens, cb(table_bf16), cb(pos_table_bf16), seq_lens, S, residual, T, h, vocab, nullptr, 0,
                        nullptr, nullptr, as_stream(stream));
  });
}

int bfly_k_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                    const void* bias_bf16, const void* gamma_bf16, const void* beta_bf16, float eps, int layer_norm,
                    void* out_bf16, int T, int h, void* stream) {
  return guarded(nullptr, [&]() -> int {
    return launch_add_norm(residual, partials, n_partials, partial_stride, cb(bias_bf16), cb(gamma_bf16),
                           cb(beta_bf16), eps, layer_norm, mb(out_bf16), T, h, as_stream(stream));
  });
}

int bfly_k_rope_kv(const float* qkv, int n_partials, int64_t partial_stride, const void* bias_bf16,
                   const float* rope_table, const int32_t* seq_lens, int S, int T, int nq, int nk, int head_dim,
                   void* q_out_bf16, void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                   int page_size, void* stream) {
  return guarded(nullptr, [&]() -> int {
    KvLayout kv{mb(kv_pages_bf16), page_table, pages_per_seq, page_size, nk, head_dim};
    return launch_rope_kv(qkv, n_partials, partial_stride, cb(bias_bf16), reinterpret_cast<const float2*>(rope_table),
                          seq_lens, S, T, nq, nk, mb(q_out_bf16), kv, as_stream(stream));
  });
}

int bfly_k_attn_decode(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                       int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int S, int T, int nq,
                       float* part_o, float* part_ml, int n_splits, void* out_bf16, void* stream) {
  return guarded(nullptr, [&]() -> int {
    KvLayout kv{const_cast<bf16*>(cb(kv_pages_bf16)), page_table, pages_per_seq, page_size, kv_local, head_dim};
    return launch_attn_decode(cb(q_bf16), kv, seq_lens, S, T, nq, part_o, part_ml, n_splits, mb(out_bf16),
                              as_stream(stream));
  });
}

int bfly_k_attn_decode_rope(const float* qkv, int n_partials, int64_t partial_stride, const void* bias_bf16,
                            const float* rope_table, void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                            int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int T, int nq, float* part_o,
                            float* part_ml, int n_splits, void* out_bf16, void* stream) {
  return guarded(nullptr, [&]() -> int {
    KvLayout kv{mb(kv_pages_bf16), page_table, pages_per_seq, page_size, kv_local, head_dim};
    return launch_attn_decode_fused(qkv, n_partials, partial_stride, cb(bias_bf16), reinterpret_cast<const float2*>(rope_table),
                                    kv, seq_lens, T, nq, part_o, part_ml, n_splits, mb(out_bf16), as_stream(stream));
  });
}

int bfly_k_step_gemm_table(int M, int K, int grid, uint16_t* first_out, uint8_t* nsl_out, int* max_slices) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(M > 0 && K > 0 && grid > 0 && grid <= 65535 && first_out && nsl_out && max_slices, "step_gemm_table: arguments");
    StepGemm g{};
    g.M = M; g.K = K;
    g.nkb = ceil_div(K, 64);
    g.m_tiles = ceil_div(M, 128);
    g.units = g.m_tiles * g.nkb;
    g.ge = std::min(grid, g.units);
    int mx = 1;
    for (int t = 0; t < g.m_tiles; ++t) {
      const int a = sk_owner(g, t * g.nkb), b = sk_owner(g, t * g.nkb + g.nkb - 1);
      BFLY_CHECK(b - a + 1 <= 255, "step_gemm_table: more than 255 slices per tile");
      first_out[t] = (uint16_t)a;
      nsl_out[t] = (uint8_t)(b - a + 1);
      mx = std::max(mx, b - a + 1);
    }
    *max_slices = mx;
    return BFLY_OK;
  });
}

int bfly_k_attn_decode_splits(int T, int kv_local, int max_ctx) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) sms = v;
  } else {
    cudaGetLastError();
  }
  if (T < 1 || kv_local < 1 || max_ctx < 1) return 0;
  return attn_decode_splits(T, kv_local, max_ctx, sms);
}

int bfly_k_attn_prefill(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table, int pages_per_seq,
                        int page_size, int kv_local, int head_dim, const int32_t* seq_lens, int S, int B, int nq,
                        void* out_bf16, void* stream) {
  return guarded(nullptr, [&]() -> int {
    KvLayout kv{const_cast<bf16*>(cb(kv_pages_bf16)), page_table, pages_per_seq, page_size, kv_local, head_dim};
    return launch_attn_prefill(cb(q_bf16), kv, seq_lens, S, B, nq, mb(out_bf16), as_stream(stream));
  });
}

int bfly_k_act(const float* in, int n_partials, int64_t partial_stride, const void* bias_bf16, int mode,
               int n_blocks, int f_block, const float* route_w, void* out_bf16, int T, void* stream) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(mode == ACT_SWIGLU || mode == ACT_GELU_TANH, "act: mode");
    return launch_act(in, n_partials, partial_stride, cb(bias_bf16), mode, n_blocks, f_block, route_w, mb(out_bf16),
                      T, as_stream(stream));
  });
}

int bfly_k_lm_finalize(const float* partials, int n_partials, int64_t partial_stride, int T, int v_local, int v0,
                       float* logits, int64_t logits_ld, uint64_t* keys, int32_t* tokens_out, void* stream) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(keys != nullptr, "lm_finalize: keys scratch required");
    BFLY_CUDA(cudaMemsetAsync(keys, 0, (size_t)T * sizeof(uint64_t), as_stream(stream)));
    BFLY_TRY(launch_lm_finalize(partials, n_partials, partial_stride, T, v_local, v0, logits, logits_ld,
                                reinterpret_cast<unsigned long long*>(keys), as_stream(stream)));
    if (tokens_out)
      BFLY_TRY(launch_pick_tokens_pitched(reinterpret_cast<unsigned long long*>(keys), 1, T, T, tokens_out,
                                          as_stream(stream)));
    return BFLY_OK;
  });
}

int bfly_k_moe_router(const void* xn_bf16, const void* router_bf16, int T, int h, int n_experts, int top_k,
                      int first_local, int n_local, float* route_w, void* stream) {
  return guarded(nullptr, [&]() -> int {
    return launch_moe_router(cb(xn_bf16), cb(router_bf16), T, h, n_experts, top_k, first_local, n_local, route_w,
                             as_stream(stream));
  });
}

int bfly_k_gemm_plan(int M, int N, int K, int* splits, int* bn, int* stages) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(M > 0 && N > 0 && K > 0, "gemm_plan: shape");
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) {
      int v = 0;
      if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) sms = v;
    } else {
      cudaGetLastError();
    }
    const GemmPlan p = gemm_plan(M, N, K, sms, kMaxSplits);
    if (splits) *splits = p.splits;
    if (bn) *bn = p.bn;
    if (stages) *stages = p.stages;
    return BFLY_OK;
  });
}

int bfly_k_gemm(const void* w_bf16, const void* x_bf16, int M, int N, int K, int64_t ldx, int splits, int bn,
                int stages, float* out_f32, int64_t split_stride, void* out_bf16, int64_t ldc, int* slices_written,
                void* stream) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(w_bf16 && x_bf16 && M > 0 && N > 0 && K > 0, "gemm: null operand or empty shape");
    int dev = 0, sms = 148;
    BFLY_CUDA(cudaGetDevice(&dev));
    BFLY_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    GemmPlan p = gemm_plan(M, N, K, sms, kMaxSplits);
    if (bn > 0) p.bn = bn;
    if (splits > 0) {
      p.splits = splits;
      p.cluster = (splits > 1 && splits <= 8 && p.bn <= 64) ? 1 : 0;
    }
    if (stages > 0) p.stages = stages;
    if (p.splits != 1 || p.cluster || p.bn < 128 || p.bn % 32 != 0) { p.pair = 0; p.persistent = 0; }
    if (bn > 0 || stages > 0) p.persistent = 0;          // explicit tile overrides address the one-tile-per-CTA kernel
    if (stages <= 0) {                                   // the override may have changed the stage footprint
      const size_t sb = 16384 + (size_t)(p.bn >> p.pair) * 128;
      p.stages = std::max(2, std::min(p.stages, (int)((200 * 1024) / sb)));
    }
    if (slices_written) *slices_written = gemm_slices(p);
    return launch_gemm(cb(w_bf16), cb(x_bf16), ldx, p, out_f32, split_stride, mb(out_bf16), ldc, as_stream(stream));
  });
}

int bfly_k_gemm_swiglu(const void* w_gate_up_bf16, const void* x_bf16, int N, int K, int n_blocks, int f_block,
                       const float* route_w, int splits, void* act_out_bf16, void* stream) {
  return guarded(nullptr, [&]() -> int {
    BFLY_CHECK(w_gate_up_bf16 && x_bf16 && act_out_bf16 && N > 0 && K > 0 && n_blocks > 0 && f_block > 0,
               "gemm_swiglu: null operand or empty shape");
    int dev = 0, sms = 148;
    BFLY_CUDA(cudaGetDevice(&dev));
    BFLY_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    GemmPlan p = gemm_plan(n_blocks * f_block, N, K, sms, kMaxSplits, 1);
    if (splits > 0) {
      BFLY_CHECK(splits == 1 || (splits <= 8 && p.bn <= 64), "gemm_swiglu: splits > 1 needs a cluster (<= 8, N <= 64)");
      p.splits = splits;
      p.cluster = splits > 1 ? 1 : 0;
    }
    GemmOut o{};
    o.out_bf16 = mb(act_out_bf16);
    o.ldc = (int64_t)n_blocks * f_block;
    o.n_blocks = n_blocks; o.f_block = f_block; o.route_w = route_w;
    return launch_gemm_ex(cb(w_gate_up_bf16), cb(x_bf16), K, p, o, as_stream(stream));
  });
}

}  // extern "C"
