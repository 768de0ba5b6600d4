// Engine: owns device memory (weights, paged KV cache, workspaces), runs the partitioned block
// forward for prefill and decode on one CUDA stream, captures the decode step in a CUDA graph and
// calls NCCL for the row-parallel all-reduce / vocabulary all-gather (tp > 1).
// Realises the reference's intended "Distributed Inference Engine" (ref:CLAUDE.md:19), which has
// no implementation to follow; the structure is SURVEY.md section 3 (call stacks 3 and 4).
#include "engine.h"

#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

namespace bfly {

std::atomic<unsigned long long> g_launch_count{0};
thread_local unsigned long long t_launch_count = 0;

// ---- NCCL, resolved at run time so the library loads (and exports its ABI) without NCCL/GPU ----
namespace {
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

std::mutex g_nccl_mutex;

int load_nccl() {
  std::lock_guard<std::mutex> lock(g_nccl_mutex);
  if (g_nccl.lib) return BFLY_OK;
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);   // the copy torch already loaded, if any
  if (!lib) {
    set_error(std::string("cannot load libnccl.so.2: ") + dlerror());
    return BFLY_ENCCL;
  }
  g_nccl.GetUniqueId = (int (*)(void*))dlsym(lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(lib, "ncclCommInitRank");
  g_nccl.CommDestroy = (int (*)(void*))dlsym(lib, "ncclCommDestroy");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(lib, "ncclAllReduce");
  g_nccl.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(lib, "ncclAllGather");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllReduce || !g_nccl.AllGather) {
    set_error("libnccl.so.2 lacks a required symbol");
    return BFLY_ENCCL;
  }
  g_nccl.lib = lib;
  return BFLY_OK;
}

// ncclDataType_t / ncclRedOp_t values (nccl.h): ncclInt8 = 0, ncclFloat32 = 7, ncclSum = 0
constexpr int kNcclInt8 = 0, kNcclFloat = 7, kNcclBfloat16 = 9, kNcclSum = 0;

#define BFLY_NCCL(expr)                                                                          \
  do {                                                                                           \
    int _r = (expr);                                                                             \
    if (_r != 0) {                                                                               \
      set_error(std::string(#expr) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "nccl error")); \
      return BFLY_ENCCL;                                                                         \
    }                                                                                            \
  } while (0)

template <typename T>
int dev_alloc(Engine* e, T** p, size_t count, uint64_t* account) {
  void* q = nullptr;
  const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
  cudaError_t err = cudaMalloc(&q, bytes);
  if (err != cudaSuccess) {
    set_error("cudaMalloc of " + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(err));
    return BFLY_ENOMEM;
  }
  e->allocations.push_back(q);
  if (account) *account += bytes;
  *p = (T*)q;
  return BFLY_OK;
}
}  // namespace

int nccl_unique_id(void* out128) {
  BFLY_TRY(load_nccl());
  BFLY_NCCL(g_nccl.GetUniqueId(out128));
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
int Engine::create(const bfly_model_desc& desc, int device, Engine** out) {
  const bfly_model_desc& c = desc;
  BFLY_CHECK(c.family >= 0 && c.family <= 2, "family");
  BFLY_CHECK(c.hidden > 0 && c.hidden % 8 == 0, "hidden % 8 == 0");
  BFLY_CHECK(c.head_dim == 32 || c.head_dim == 64 || c.head_dim == 128, "head_dim in {32,64,128}");
  BFLY_CHECK(c.tp >= 1 && c.rank >= 0 && c.rank < c.tp, "0 <= rank < tp");
  BFLY_CHECK(c.n_heads % c.tp == 0 && c.n_kv_heads % c.tp == 0 && c.n_heads % c.n_kv_heads == 0, "heads vs tp");
  BFLY_CHECK(c.n_experts == 0 ? c.ffn % c.tp == 0 : (c.n_experts % c.tp == 0 && c.n_experts <= 32), "ffn/experts vs tp");
  BFLY_CHECK(c.n_experts == 0 || (c.top_k >= 1 && c.top_k <= c.n_experts), "top_k");
  BFLY_CHECK(c.page_size >= 32 && (c.page_size & (c.page_size - 1)) == 0, "page_size power of two >= 32");
  BFLY_CHECK(c.max_batch >= 1 && c.max_ctx >= 1 && c.max_tokens >= c.max_batch, "max_batch/max_ctx/max_tokens");
  BFLY_CHECK(c.rope_theta > 0.f || c.max_pos >= c.max_ctx, "learned positions need max_pos >= max_ctx");
  BFLY_CUDA(cudaSetDevice(device));
  BFLY_TRY(gemm_init());

  std::unique_ptr<Engine> owner(new Engine());        // freed (with everything it allocated) on any early return
  Engine* e = owner.get();
  e->d = c;
  e->device = device;
  cudaDeviceProp prop;
  BFLY_CUDA(cudaGetDeviceProperties(&prop, device));
  e->sm_count = prop.multiProcessorCount;
  BFLY_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));

  e->q_local = c.n_heads / c.tp;
  e->kv_local = c.n_kv_heads / c.tp;
  e->qkv_rows = (e->q_local + 2 * e->kv_local) * c.head_dim;
  e->experts_local = c.n_experts ? c.n_experts / c.tp : 0;
  e->n_blocks = c.n_experts ? e->experts_local : 1;
  e->f_block = c.n_experts ? c.ffn : c.ffn / c.tp;
  e->gated = c.family != BFLY_FAMILY_GPT2;
  e->ffn_in_rows = e->n_blocks * e->f_block * (e->gated ? 2 : 1);
  e->ffn_act_cols = e->n_blocks * e->f_block;
  e->v_shard = (c.vocab + c.tp - 1) / c.tp;
  e->v0 = std::min(c.vocab, c.rank * e->v_shard);
  e->v_local = std::min(c.vocab, (c.rank + 1) * e->v_shard) - e->v0;
  e->pages_per_seq = (c.max_ctx + c.page_size - 1) / c.page_size;
  e->n_pages = (int64_t)c.max_batch * e->pages_per_seq;
  e->layer_kv_elems = e->n_pages * 2 * e->kv_local * c.page_size * c.head_dim;
  e->layers.resize(c.n_layers);
  {
    const char* env = getenv("BFLY_FUSED_SWIGLU");      // tuning knob; default on
    e->fused_swiglu = !(env && env[0] == '0');
    env = getenv("BFLY_MOE_GROUPED");                   // tuning knob; default on
    e->moe_grouped = !(env && env[0] == '0') && e->gated && e->fused_swiglu;
    // The fused decode step kernel (decode_step.cu) is parity-green but measured SLOWER than the kernel-per-op
    // path so far, so it is opt-in: BFLY_FUSED_STEP=1.
    env = getenv("BFLY_FUSED_STEP");
    e->fused_step = env && env[0] == '1';
    // RoPE + KV append inside the decode attention kernel (one kernel less per layer) measured SLOWER than the
    // separate rope_kv kernel under programmatic dependent launch (tp=1 4.85 vs 4.77 ms/step, tp=8 shard 1.74 vs
    // 1.70): the fold sits on the attention kernel's critical path instead of overlapping its KV prefetch.  Opt-in.
    env = getenv("BFLY_FUSED_ROPE");
    e->fused_rope = env && env[0] == '1';
    // bf16 payload for the prefill all-reduce: +7 % prefill at tp=2, +15 % at tp=8 (152.6 k -> 175.9 k tok/s), but NCCL
    // rounds the running sum to bf16 at every ring hop and the logit error at tp=8 rose to 6.75 % of sigma against the
    // 7 % gate (5.5 % with fp32) -- too little margin for a result that depends on NCCL's algorithm choice.  Opt-in.
    env = getenv("BFLY_PREFILL_AR_BF16");
    e->prefill_ar_bf16 = env && env[0] == '1';
    env = getenv("BFLY_AR_CLUSTER");                    // A/B knob: 0 = one CTA per row (round-1 kernel)
    // (rows split over a cluster of CTAs measured SLOWER than one CTA per row at tp=2 and tp=8 -- more flags in
    //  flight, cluster launch cost -- so it is off unless asked for: "a" = automatic chunk count, "1".."8" = that many)
    e->ar_chunks = 0;
    if (env && env[0] == 'a') e->ar_chunks = ar_chunks_for(c.hidden);
    if (env && env[0] >= '1' && env[0] <= '8' && c.hidden % ((env[0] - '0') * 4) == 0) e->ar_chunks = env[0] - '0';
    env = getenv("BFLY_AR_BF16");                       // A/B knob: partial sums cross NVLink as bf16
    e->ar_bf16 = (env && env[0] == '1' && e->ar_chunks > 0) ? 1 : 0;
    env = getenv("BFLY_AR_PROTO");                      // "flags" = round-1 epoch-flag protocol; default: flag-free (Lamport)
    e->ar_lamport = (env && env[0] == 'f') ? 0 : 1;
    if (e->fused_step || e->ar_chunks > 0) e->ar_lamport = 0;     // those kernels speak the flag protocol
    env = getenv("BFLY_AR_TWO_SHOT");                   // initial all-reduce mode; bfly_set_ar_mode() changes it
    e->ar_mode = env ? (env[0] == '1' ? 1 : 0) : -1;
  }

  const int h = c.hidden, maxT = c.max_tokens, maxB = c.max_batch;
  // split-K scratch: worst case over every GEMM shape and every token count a forward can see
  const int gemm_m[5] = {e->qkv_rows, h, e->ffn_in_rows, h, std::max(e->v_local, 1)};
  const int gemm_k[5] = {h, e->q_local * c.head_dim, h, e->ffn_act_cols, h};
  size_t ws = 0;
  for (int g = 0; g < 5; ++g) {
    const int n_cap = (g == 4) ? maxB : maxT;
    for (int n = 1; n <= n_cap; ++n) {          // every row count: the planner is not monotone in n
      const GemmPlan p = gemm_plan(gemm_m[g], n, gemm_k[g], e->sm_count, kMaxSplits);
      ws = std::max(ws, (size_t)p.splits * n * gemm_m[g]);
    }
  }
  e->ws_elems = ws + 64;
  // split-context attention scratch: attn_decode_splits() is not monotone in T, so take the maximum over
  // EVERY row count a forward can see (forward() checks the request against these capacities)
  size_t po = 0, pml = 0;
  for (int t = 1; t <= maxT; ++t) {
    const int s = attn_decode_splits(t, e->kv_local, c.max_ctx, e->sm_count);
    po = std::max(po, (size_t)t * e->q_local * s * c.head_dim);
    pml = std::max(pml, (size_t)t * e->q_local * s * 2);
  }
  e->part_o_elems = po;
  e->part_ml_elems = pml;

  int rc = BFLY_OK;
  uint64_t* wsb = &e->workspace_bytes;
  auto A = [&](int r) { if (rc == BFLY_OK) rc = r; };
  A(dev_alloc(e, &e->kv_pages, (size_t)c.n_layers * e->layer_kv_elems, &e->kv_bytes));
  A(dev_alloc(e, &e->page_table, (size_t)e->n_pages, wsb));
  A(dev_alloc(e, &e->seq_lens, (size_t)maxB, wsb));
  A(dev_alloc(e, &e->step_counter, 1, wsb));
  A(dev_alloc(e, &e->residual, (size_t)maxT * h, wsb));
  A(dev_alloc(e, &e->xn, (size_t)maxT * h, wsb));
  A(dev_alloc(e, &e->q, (size_t)maxT * e->q_local * c.head_dim, wsb));
  A(dev_alloc(e, &e->attn_out, (size_t)maxT * e->q_local * c.head_dim, wsb));
  A(dev_alloc(e, &e->act, (size_t)maxT * e->ffn_act_cols, wsb));
  A(dev_alloc(e, &e->ws, e->ws_elems, wsb));
  A(dev_alloc(e, &e->part_o, po, wsb));
  A(dev_alloc(e, &e->part_ml, pml, wsb));
  if (c.tp > 1) A(dev_alloc(e, &e->red, (size_t)maxT * h, wsb));
  if (c.tp > 1) A(dev_alloc(e, &e->red_bf, (size_t)maxT * h, wsb));
  if (c.n_experts) A(dev_alloc(e, &e->route_w, (size_t)maxT * e->experts_local, wsb));
  if (c.n_experts && e->moe_grouped && maxT >= kMoeGroupedMinTokens) {
    e->moe_rows_cap = maxT * std::min(c.top_k, e->experts_local);
    A(dev_alloc(e, &e->moe_cnt, (size_t)e->experts_local, wsb));
    A(dev_alloc(e, &e->moe_off, (size_t)e->experts_local + 1, wsb));
    A(dev_alloc(e, &e->moe_idx, (size_t)e->moe_rows_cap, wsb));
    A(dev_alloc(e, &e->moe_rw, (size_t)e->moe_rows_cap, wsb));
    A(dev_alloc(e, &e->moe_xg, (size_t)e->moe_rows_cap * h, wsb));
    A(dev_alloc(e, &e->moe_out, (size_t)maxT * h, wsb));
  }
  A(dev_alloc(e, &e->logits, (size_t)maxB * e->v_shard, wsb));
  if (c.tp > 1) A(dev_alloc(e, &e->logits_all, (size_t)c.tp * maxB * e->v_shard, wsb));
  A(dev_alloc(e, &e->keys, (size_t)maxB, wsb));
  A(dev_alloc(e, &e->keys_all, (size_t)c.tp * maxB, wsb));
  A(dev_alloc(e, &e->tokens_in, (size_t)maxT, wsb));
  A(dev_alloc(e, &e->cur_tokens, (size_t)maxB, wsb));
  A(dev_alloc(e, &e->history, (size_t)kMaxGreedySteps * maxB, wsb));
  if (c.rope_theta > 0.f) A(dev_alloc(e, &e->rope_table, (size_t)c.max_ctx * (c.head_dim / 2), wsb));
  if (rc != BFLY_OK) return rc;

  // identity page table, zero lengths
  std::vector<int> table((size_t)e->n_pages);
  for (int64_t i = 0; i < e->n_pages; ++i) table[(size_t)i] = (int)i;
  BFLY_CUDA(cudaMemcpy(e->page_table, table.data(), table.size() * sizeof(int), cudaMemcpyHostToDevice));
  BFLY_CUDA(cudaMemset(e->seq_lens, 0, (size_t)maxB * sizeof(int)));
  // stale cache slots are multiplied by an exact 0 probability: they must never hold NaN/Inf bits
  BFLY_CUDA(cudaMemset(e->kv_pages, 0, (size_t)c.n_layers * e->layer_kv_elems * sizeof(bf16)));
  BFLY_CUDA(cudaMemset(e->step_counter, 0, sizeof(int)));
  BFLY_CUDA(cudaMemset(e->cur_tokens, 0, (size_t)maxB * sizeof(int)));
  if (c.rope_theta > 0.f) {
    // cos/sin in double, as the oracle does (oracle/ops.py::rope), stored fp32
    const int half = c.head_dim / 2;
    std::vector<float2> tab((size_t)c.max_ctx * half);
    for (int i = 0; i < half; ++i) {
      const double inv = std::pow((double)c.rope_theta, -2.0 * i / (double)c.head_dim);
      for (int p = 0; p < c.max_ctx; ++p) {
        const double a = (double)p * inv;
        tab[(size_t)p * half + i] = make_float2((float)std::cos(a), (float)std::sin(a));
      }
    }
    BFLY_CUDA(cudaMemcpy(e->rope_table, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice));
  }
  BFLY_CUDA(cudaHostAlloc((void**)&e->pinned, kPinnedBytes, cudaHostAllocDefault));
  *out = owner.release();
  return BFLY_OK;
}

Engine::~Engine() {
  cudaSetDevice(device);
  if (stream) cudaStreamSynchronize(stream);
  for (auto& kv : graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  for (void* p : ipc_opened) cudaIpcCloseMemHandle(p);
  if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
  for (void* p : allocations) cudaFree(p);
  if (pinned) cudaFreeHost(pinned);
  if (stream) cudaStreamDestroy(stream);
}

// ---- weights -----------------------------------------------------------------------------------
int Engine::alloc_weight(const std::string& name, int64_t rows, int64_t cols) {
  BFLY_CHECK(rows > 0 && cols > 0, "weight shape");
  BFLY_CUDA(cudaSetDevice(device));
  auto it = weights.find(name);
  if (it != weights.end()) {
    BFLY_CHECK(it->second.rows == rows && it->second.cols == cols, "weight re-registered with another shape: " + name);
    return BFLY_OK;
  }
  DevTensor t{};
  t.rows = rows; t.cols = cols;
  BFLY_TRY(dev_alloc(this, &t.ptr, (size_t)(rows * cols), &weight_bytes));
  weights[name] = t;
  ready = false;
  return BFLY_OK;
}

int Engine::upload_weight(const std::string& name, const void* host, int64_t rows, int64_t cols) {
  BFLY_TRY(alloc_weight(name, rows, cols));
  BFLY_CUDA(cudaMemcpy(weights[name].ptr, host, (size_t)(rows * cols) * sizeof(bf16), cudaMemcpyHostToDevice));
  return BFLY_OK;
}

int Engine::fill_weight_block(const std::string& name, int64_t dst_row, int64_t dst_col, uint64_t key,
                              int64_t ncols_full, int64_t r0, int64_t c0, int64_t rows, int64_t cols, float scale,
                              float offset) {
  auto it = weights.find(name);
  if (it == weights.end()) { set_error("unknown weight: " + name); return BFLY_ESTATE; }
  const DevTensor& t = it->second;
  BFLY_CHECK(dst_row >= 0 && dst_col >= 0 && dst_row + rows <= t.rows && dst_col + cols <= t.cols,
             "fill_weight_block: block outside " + name);
  BFLY_CUDA(cudaSetDevice(device));
  return launch_fill_synth(t.ptr + dst_row * t.cols + dst_col, t.cols, key, ncols_full, r0, c0, rows, cols, scale,
                           offset, stream);
}

int Engine::read_weight(const std::string& name, void* host, int64_t max_bytes) {
  auto it = weights.find(name);
  if (it == weights.end()) { set_error("unknown weight: " + name); return BFLY_ESTATE; }
  const int64_t bytes = it->second.rows * it->second.cols * (int64_t)sizeof(bf16);
  BFLY_CHECK(max_bytes >= bytes, "read_weight: buffer too small");
  BFLY_CUDA(cudaStreamSynchronize(stream));
  BFLY_CUDA(cudaMemcpy(host, it->second.ptr, (size_t)bytes, cudaMemcpyDeviceToHost));
  return BFLY_OK;
}

int Engine::find(const std::string& name, int64_t rows, int64_t cols, bool required, bf16** out) {
  auto it = weights.find(name);
  if (it == weights.end()) {
    *out = nullptr;
    if (!required) return BFLY_OK;
    set_error("missing weight: " + name);
    return BFLY_ESTATE;
  }
  if (it->second.rows != rows || it->second.cols != cols) {
    set_error("weight " + name + " has shape [" + std::to_string(it->second.rows) + "," + std::to_string(it->second.cols) +
              "], expected [" + std::to_string(rows) + "," + std::to_string(cols) + "]");
    return BFLY_EINVAL;
  }
  *out = it->second.ptr;
  return BFLY_OK;
}

int Engine::weights_ready() {
  const bfly_model_desc& c = d;
  const int h = c.hidden;
  const bool ln = c.family == BFLY_FAMILY_GPT2;
  BFLY_TRY(find("embed", c.vocab, h, true, &embed));
  BFLY_TRY(find("pos_embed", c.max_pos, h, c.rope_theta == 0.f, &pos_embed));
  BFLY_TRY(find("final_norm.w", h, 1, true, &final_norm_w));
  BFLY_TRY(find("final_norm.b", h, 1, ln, &final_norm_b));
  BFLY_TRY(find("lm_head", v_local, h, v_local > 0, &lm_head));
  for (int l = 0; l < c.n_layers; ++l) {
    const std::string p = "layers." + std::to_string(l) + ".";
    LayerW& w = layers[l];
    BFLY_TRY(find(p + "attn_norm.w", h, 1, true, &w.attn_norm_w));
    BFLY_TRY(find(p + "attn_norm.b", h, 1, ln, &w.attn_norm_b));
    BFLY_TRY(find(p + "ffn_norm.w", h, 1, true, &w.ffn_norm_w));
    BFLY_TRY(find(p + "ffn_norm.b", h, 1, ln, &w.ffn_norm_b));
    BFLY_TRY(find(p + "wqkv", qkv_rows, h, true, &w.wqkv));
    BFLY_TRY(find(p + "bqkv", qkv_rows, 1, c.bias != 0, &w.bqkv));
    BFLY_TRY(find(p + "wo", h, q_local * c.head_dim, true, &w.wo));
    BFLY_TRY(find(p + "bo", h, 1, c.bias != 0, &w.bo));
    BFLY_TRY(find(p + (gated ? "w_gate_up" : "w_up"), ffn_in_rows, h, true, &w.w_in));
    BFLY_TRY(find(p + "b_up", ffn_act_cols, 1, c.bias != 0 && !gated, &w.b_up));
    BFLY_TRY(find(p + "w_down", h, ffn_act_cols, true, &w.w_down));
    BFLY_TRY(find(p + "b_down", h, 1, c.bias != 0, &w.b_down));
    BFLY_TRY(find(p + "router", c.n_experts, h, c.n_experts > 0, &w.router));
  }
  BFLY_CUDA(cudaStreamSynchronize(stream));
  ready = true;
  return BFLY_OK;
}

// ---- communication -----------------------------------------------------------------------------
int Engine::comm_init(const void* id128, int world, int rank) {
  BFLY_CHECK(world == d.tp && rank == d.rank, "comm_init: world/rank must equal the plan's tp/rank");
  if (world == 1) return BFLY_OK;
  BFLY_TRY(load_nccl());
  BFLY_CUDA(cudaSetDevice(device));
  NcclId id;
  std::memcpy(id.bytes, id128, 128);
  BFLY_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
  const char* env = getenv("BFLY_P2P_ALLREDUCE");            // tuning knob; default on
  if (!(env && env[0] == '0')) BFLY_TRY(p2p_init());
  return BFLY_OK;
}

// Peer-mapped receive buffers for the fused decode all-reduce: every rank allocates one buffer,
// the CUDA IPC handles travel through an NCCL all-gather, and each rank maps its peers' buffers.
int Engine::p2p_init() {
  const int tp = d.tp, max_rows = d.max_batch, h = d.hidden;
  const size_t row_flags_bytes = (((size_t)2 * tp * max_rows * sizeof(unsigned)) + 255) / 256 * 256;
  const size_t flags_bytes = row_flags_bytes * 9;        // per-row flags, then 8 per-chunk flags per row
  const size_t recv_bytes = (size_t)2 * tp * max_rows * h * sizeof(float);
  char* base = nullptr;
  BFLY_TRY(dev_alloc(this, &base, flags_bytes + recv_bytes, &workspace_bytes));
  BFLY_CUDA(cudaMemset(base, 0, flags_bytes + recv_bytes));
  if (ar_lamport) {
    BFLY_TRY(ar_fill_sentinel(reinterpret_cast<float*>(base + flags_bytes), recv_bytes / sizeof(float), stream));
    BFLY_CUDA(cudaStreamSynchronize(stream));
  }
  BFLY_TRY(dev_alloc(this, &ar_step, 1, &workspace_bytes));
  BFLY_CUDA(cudaMemset(ar_step, 0, sizeof(unsigned)));
  cudaIpcMemHandle_t mine;
  BFLY_CUDA(cudaIpcGetMemHandle(&mine, base));
  char *h_dev = nullptr, *all_dev = nullptr;
  BFLY_TRY(dev_alloc(this, &h_dev, sizeof(mine), &workspace_bytes));
  BFLY_TRY(dev_alloc(this, &all_dev, sizeof(mine) * tp, &workspace_bytes));
  BFLY_CUDA(cudaMemcpy(h_dev, &mine, sizeof(mine), cudaMemcpyHostToDevice));
  BFLY_CUDA(cudaDeviceSynchronize());                         // the zeroed buffer is in place before anyone can push
  BFLY_TRY(all_gather_bytes(h_dev, all_dev, sizeof(mine)));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  std::vector<cudaIpcMemHandle_t> all((size_t)tp);
  BFLY_CUDA(cudaMemcpy(all.data(), all_dev, sizeof(mine) * tp, cudaMemcpyDeviceToHost));
  ar.tp = tp; ar.rank = d.rank; ar.max_rows = max_rows;
  for (int r = 0; r < tp; ++r) {
    char* p = base;
    if (r != d.rank) {
      void* q = nullptr;
      BFLY_CUDA(cudaIpcOpenMemHandle(&q, all[(size_t)r], cudaIpcMemLazyEnablePeerAccess));
      ipc_opened.push_back(q);
      p = (char*)q;
    }
    ar.flags[r] = (unsigned*)p;
    ar.cflags[r] = (unsigned*)(p + row_flags_bytes);
    ar.recv[r] = (float*)(p + flags_bytes);
  }
  // nobody starts pushing before every rank has mapped its peers
  BFLY_TRY(all_gather_bytes(h_dev, all_dev, sizeof(mine)));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  p2p_ready = true;
  return BFLY_OK;
}

// Profiling aid: BFLY_SKIP_COLLECTIVES=1 lets ONE process run rank r's shard of a tp > 1 plan with the
// collectives left out (results are then wrong by construction; bench.py --shard-of labels them so).
static bool skip_collectives() {
  static const bool v = [] { const char* e = getenv("BFLY_SKIP_COLLECTIVES"); return e && e[0] == '1'; }();
  return v;
}

int Engine::all_reduce_f32(float* buf, size_t count) {
  if (d.tp == 1) return BFLY_OK;
  if (!comm && skip_collectives()) return BFLY_OK;
  if (!comm) { set_error("tp > 1 but bfly_comm_init was not called"); return BFLY_ESTATE; }
  BFLY_NCCL(g_nccl.AllReduce(buf, buf, count, kNcclFloat, kNcclSum, comm, stream));
  return BFLY_OK;
}

int Engine::all_reduce_bf16(bf16* buf, size_t count) {
  if (d.tp == 1) return BFLY_OK;
  if (!comm && skip_collectives()) return BFLY_OK;
  if (!comm) { set_error("tp > 1 but bfly_comm_init was not called"); return BFLY_ESTATE; }
  BFLY_NCCL(g_nccl.AllReduce(buf, buf, count, kNcclBfloat16, kNcclSum, comm, stream));
  return BFLY_OK;
}

int Engine::all_gather_bytes(const void* src, void* dst, size_t bytes_per_rank) {
  if (!comm && skip_collectives()) return BFLY_OK;
  if (!comm) { set_error("tp > 1 but bfly_comm_init was not called"); return BFLY_ESTATE; }
  BFLY_NCCL(g_nccl.AllGather(src, dst, bytes_per_rank, kNcclInt8, comm, stream));
  return BFLY_OK;
}

// ---- forward -----------------------------------------------------------------------------------
int Engine::gemm(const bf16* W, const bf16* X, int64_t ldx, int M, int N, int K, int* slices_out) {
  const GemmPlan p = gemm_plan(M, N, K, sm_count, kMaxSplits);
  const int slices = gemm_slices(p);
  if ((size_t)slices * N * M > ws_elems) { set_error("split-K scratch too small (internal sizing error)"); return BFLY_ESTATE; }
  *slices_out = slices;
  return launch_gemm(W, X, ldx, p, ws, (int64_t)N * M, nullptr, M, stream);
}

// gate/up projection with the SwiGLU (and the MoE routing weight) fused into the GEMM epilogue
int Engine::gemm_swiglu(const bf16* W, const bf16* X, int N) {
  const GemmPlan p = gemm_plan(ffn_act_cols, N, d.hidden, sm_count, kMaxSplits, 1);
  GemmOut o{};
  o.out_bf16 = act;
  o.ldc = ffn_act_cols;
  o.n_blocks = n_blocks;
  o.f_block = f_block;
  o.route_w = d.n_experts ? route_w : nullptr;
  return launch_gemm_ex(W, X, d.hidden, p, o, stream);
}

// Row-parallel output: fold slices and all-reduce when tp > 1; tells the consumer where to read.
int Engine::row_parallel_finish(int T, int slices, bool allow_fused, float* partials, const float** src,
                                int* n_partials, bool* fused, const bf16** src_bf16) {
  const int h = d.hidden;
  *fused = false;
  *src_bf16 = nullptr;
  if (d.tp == 1) { *src = partials; *n_partials = slices; return BFLY_OK; }
  if (allow_fused && p2p_ready && T <= ar.max_rows) {      // decode: the consumer kernel does the all-reduce
    *src = partials; *n_partials = slices; *fused = true;
    return BFLY_OK;
  }
  // Bandwidth regime (prefill): the rank's folded partial crosses NVLink as bf16 -- half the bytes of the fp32
  // all-reduce that dominated prefill at tp = 8 -- and is added to the fp32 residual by the norm kernel.
  if (prefill_ar_bf16 && T >= kPrefillBf16MinRows) {
    BFLY_TRY(launch_reduce_partials_bf16(partials, slices, (int64_t)T * h, red_bf, (int64_t)T * h, stream));
    BFLY_TRY(all_reduce_bf16(red_bf, (size_t)T * h));
    *src = nullptr; *n_partials = 0; *src_bf16 = red_bf;
    return BFLY_OK;
  }
  float* buf = partials;
  if (slices > 1) {
    BFLY_TRY(launch_reduce_partials(partials, slices, (int64_t)T * h, red, (int64_t)T * h, stream));
    buf = red;
  }
  BFLY_TRY(all_reduce_f32(buf, (size_t)T * h));
  *src = buf; *n_partials = 1;
  return BFLY_OK;
}

// MoE prefill: every local expert runs its gate/up (+SwiGLU, x routing weight) and down projections over
// the tokens routed to it only (top-k of E: 1/4 of the dense work for Mixtral), on rows gathered into
// expert order; the down projection's epilogue adds each row back into its token's row of moe_out.
// Row counts stay on the device: the persistent GEMM reads them and skips the tiles beyond.
int Engine::moe_ffn_grouped(const bf16* w_in, const bf16* w_down, int T) {
  const int h = d.hidden, f = d.ffn;
  BFLY_TRY(launch_moe_group(route_w, xn, T, h, experts_local, moe_rows_cap, moe_cnt, moe_off, moe_idx, moe_rw, moe_xg,
                            stream));
  BFLY_CUDA(cudaMemsetAsync(moe_out, 0, (size_t)T * h * sizeof(float), stream));
  for (int e = 0; e < experts_local; ++e) {
    const GemmPlan pu = gemm_plan(f, T, h, sm_count, kMaxSplits, 1, 1);
    const GemmPlan pd = gemm_plan(h, T, f, sm_count, kMaxSplits, 0, 1);
    if (!pu.persistent || !pd.persistent) { set_error("grouped MoE needs persistent GEMM plans (internal)"); return BFLY_ESTATE; }
    GemmOut ou{};
    ou.out_bf16 = act;                       // [rows][f] in expert order (the buffer holds max_tokens x E_l x f)
    ou.ldc = f;
    ou.n_blocks = 1;
    ou.f_block = f;
    ou.route_w = moe_rw;
    ou.n_rows_dev = moe_cnt + e;
    ou.row_off_dev = moe_off + e;
    ou.x_rows = moe_rows_cap;
    BFLY_TRY(launch_gemm_ex(w_in + (int64_t)e * 2 * f * h, moe_xg, h, pu, ou, stream));
    GemmOut od{};
    od.out_f32 = moe_out;
    od.ldc = h;
    od.n_rows_dev = moe_cnt + e;
    od.row_off_dev = moe_off + e;
    od.row_map = moe_idx;
    od.x_rows = moe_rows_cap;
    od.ldw = (int64_t)experts_local * f;
    BFLY_TRY(launch_gemm_ex(w_down + (int64_t)e * f, act, f, pd, od, stream));
  }
  return BFLY_OK;
}

// Size rule of the decode all-reduce (mode -1).  Measured with the flag-free protocol: at tp=8 two-shot wins even for
// the 32 x 4096 rows of the 8B shape (isolated exchange 11.8 vs 13.3 us, decode step 2.13 vs 2.20 ms) because one-shot
// moves tp x the bytes; at tp=2 one-shot wins (6.8 vs 8.4 us); tp=4 keeps the round-1 threshold of 1 MiB per call.
int Engine::ar_two_shot_by_size(int T) const {
  if (d.tp >= 8) return 1;
  return d.tp >= 4 && (size_t)T * d.hidden * sizeof(float) >= (1u << 20);
}

// residual += (all-reduced) row-parallel output + bias; xn = norm(residual)
int Engine::norm_after(int T, const float* src, int n, bool fused, int* ar_call, const bf16* bias, const bf16* gamma,
                       const bf16* beta, const bf16* src_bf16) {
  const int h = d.hidden;
  const int ln = d.family == BFLY_FAMILY_GPT2;
  if (fused) {
    const int idx = (*ar_call)++;
    // every rank takes the same branch: T, h, tp and the mode are identical across the group
    const int two_shot = ar_mode >= 0 ? ar_mode : ar_two_shot_by_size(T);
    return launch_ar_add_norm(residual, src, n, (int64_t)T * h, bias, gamma, beta, d.norm_eps, ln, xn, T, h, ar, idx,
                              ar_step, 2 * d.n_layers, two_shot, ar_chunks, ar_bf16, ar_lamport, stream);
  }
  return launch_add_norm(residual, src, n, (int64_t)T * h, bias, gamma, beta, d.norm_eps, ln, xn, T, h, stream, src_bf16);
}

int Engine::forward(const int* tokens_dev, int seq0, int nb, int S, bool want_logits, bool record_history) {
  const bfly_model_desc& c = d;
  const int T = nb * S, h = c.hidden, dh = c.head_dim;
  const bool ln = c.family == BFLY_FAMILY_GPT2;
  int* lens = seq_lens + seq0;
  const int* table = page_table + (int64_t)seq0 * pages_per_seq;
  const int attn_splits = attn_decode_splits(T, kv_local, c.max_ctx, sm_count);

  BFLY_TRY(launch_embed(tokens_dev, embed, pos_embed, lens, S, residual, T, h, c.vocab, keys, nb,
                        record_history ? history : nullptr, step_counter, stream));
  const float* pend_src = nullptr;
  int pend_n = 0, ar_call = 0;
  bool pend_fused = false, any_fused = false;
  const bool allow_fused = (S == 1);
  const bf16* pend_bias = nullptr;
  const bf16* pend_bf16 = nullptr;
  for (int l = 0; l < c.n_layers; ++l) {
    const LayerW& w = layers[l];
    KvLayout kv{kv_pages + (int64_t)l * layer_kv_elems, table, pages_per_seq, c.page_size, kv_local, dh};
    int s = 1;
    BFLY_TRY(norm_after(T, pend_src, pend_n, pend_fused, &ar_call, pend_bias, w.attn_norm_w, w.attn_norm_b, pend_bf16));
    BFLY_TRY(gemm(w.wqkv, xn, h, qkv_rows, T, h, &s));
    const bool fuse_rope = S == 1 && fused_rope;
    if (!fuse_rope)
      BFLY_TRY(launch_rope_kv(ws, s, (int64_t)T * qkv_rows, w.bqkv, rope_table, lens, S, T, q_local, kv_local, q, kv,
                              stream));
    if (S > 1)
      BFLY_TRY(launch_attn_prefill(q, kv, lens, S, nb, q_local, attn_out, stream));
    else if (fuse_rope) {
      if ((size_t)T * q_local * attn_splits * dh > part_o_elems || (size_t)T * q_local * attn_splits * 2 > part_ml_elems) {
        set_error("split-attention scratch too small (internal sizing error)");
        return BFLY_ESTATE;
      }
      BFLY_TRY(launch_attn_decode_fused(ws, s, (int64_t)T * qkv_rows, w.bqkv, rope_table, kv, lens, T, q_local, part_o,
                                        part_ml, attn_splits, attn_out, stream));
    } else {
      if ((size_t)T * q_local * attn_splits * dh > part_o_elems || (size_t)T * q_local * attn_splits * 2 > part_ml_elems) {
        set_error("split-attention scratch too small (internal sizing error)");
        return BFLY_ESTATE;
      }
      BFLY_TRY(launch_attn_decode(q, kv, lens, S, T, q_local, part_o, part_ml, attn_splits, attn_out, stream));
    }
    BFLY_TRY(gemm(w.wo, attn_out, q_local * dh, h, T, q_local * dh, &s));
    BFLY_TRY(row_parallel_finish(T, s, allow_fused, ws, &pend_src, &pend_n, &pend_fused, &pend_bf16));
    any_fused |= pend_fused;
    BFLY_TRY(norm_after(T, pend_src, pend_n, pend_fused, &ar_call, w.bo, w.ffn_norm_w, w.ffn_norm_b, pend_bf16));
    if (c.n_experts)
      BFLY_TRY(launch_moe_router(xn, w.router, T, h, c.n_experts, c.top_k, c.rank * experts_local, experts_local,
                                 route_w, stream));
    // Fused SwiGLU epilogue: +5 % at prefill (no fp32 gate/up round trip); at decode the cluster it needs
    // to complete the K sum costs more than the separate activation kernel (measured), so T <= 64 stays unfused.
    float* ffn_out = ws;
    if (c.n_experts && moe_out != nullptr && T >= kMoeGroupedMinTokens) {
      BFLY_TRY(moe_ffn_grouped(w.w_in, w.w_down, T));
      ffn_out = moe_out;
      s = 1;
    } else {
      if (gated && fused_swiglu && T > 64) {
        BFLY_TRY(gemm_swiglu(w.w_in, xn, T));
      } else {
        BFLY_TRY(gemm(w.w_in, xn, h, ffn_in_rows, T, h, &s));
        BFLY_TRY(launch_act(ws, s, (int64_t)T * ffn_in_rows, w.b_up, gated ? ACT_SWIGLU : ACT_GELU_TANH, n_blocks,
                            f_block, c.n_experts ? route_w : nullptr, act, T, stream));
      }
      BFLY_TRY(gemm(w.w_down, act, ffn_act_cols, h, T, ffn_act_cols, &s));
    }
    BFLY_TRY(row_parallel_finish(T, s, allow_fused, ffn_out, &pend_src, &pend_n, &pend_fused, &pend_bf16));
    any_fused |= pend_fused;
    pend_bias = w.b_down;
  }
  BFLY_TRY(norm_after(T, pend_src, pend_n, pend_fused, &ar_call, pend_bias, final_norm_w, final_norm_b, pend_bf16));
  // LM head on the last position of each sequence (row pitch S*h picks rows b*S + S-1)
  int s = 1;
  if (v_local > 0) {
    BFLY_TRY(gemm(lm_head, xn + (int64_t)(S - 1) * h, (int64_t)S * h, v_local, nb, h, &s));
    BFLY_TRY(launch_lm_finalize(ws, s, (int64_t)nb * v_local, nb, v_local, v0,
                                want_logits ? logits + (int64_t)seq0 * v_shard : nullptr, v_shard, keys, stream));
  }
  if (c.tp > 1) {
    BFLY_TRY(all_gather_bytes(keys, keys_all, (size_t)d.max_batch * sizeof(unsigned long long)));
    // keys_all is [tp][max_batch]; pick over ranks with that pitch
    BFLY_TRY(launch_pick_tokens_pitched(keys_all, c.tp, d.max_batch, nb, cur_tokens + seq0, stream));
  } else {
    BFLY_TRY(launch_pick_tokens_pitched(keys, 1, d.max_batch, nb, cur_tokens + seq0, stream));
  }
  BFLY_TRY(launch_advance(lens, nb, S, record_history ? step_counter : nullptr, any_fused ? ar_step : nullptr, stream));
  return BFLY_OK;
}

// ---- fused decode step (decode_step.cu) --------------------------------------------------------
bool Engine::step_supported(int B) const {
  const bfly_model_desc& c = d;
  if (!fused_step || c.family != BFLY_FAMILY_LLAMA || c.n_experts != 0 || c.bias != 0 || c.rope_theta <= 0.f) return false;
  if (c.head_dim != 128 || c.hidden % 128 != 0 || B > kStepMaxTokens || f_block % 4 != 0 || v_local <= 0) return false;
  if (c.tp > 1 && !p2p_ready && !(comm == nullptr && getenv("BFLY_SKIP_COLLECTIVES"))) return false;
  if (c.tp > 1 && p2p_ready && B > ar.max_rows) return false;
  return true;
}

namespace {
// stream-K tables of one GEMM: first CTA and number of CTAs (= slices) per weight tile
int make_step_gemm(Engine* e, int M, int K, int T, int grid, StepGemm* out) {
  StepGemm g{};
  g.M = M; g.K = K;
  g.nkb = ceil_div(K, 64);
  g.m_tiles = ceil_div(M, 128);
  g.units = g.m_tiles * g.nkb;
  g.ge = std::min(grid, g.units);
  std::vector<unsigned short> first((size_t)g.m_tiles);
  std::vector<unsigned char> nsl((size_t)g.m_tiles);
  int max_sl = 1;
  for (int t = 0; t < g.m_tiles; ++t) {
    const int a = sk_owner(g, t * g.nkb), b = sk_owner(g, t * g.nkb + g.nkb - 1);
    first[(size_t)t] = (unsigned short)a;
    nsl[(size_t)t] = (unsigned char)(b - a + 1);
    max_sl = std::max(max_sl, b - a + 1);
  }
  BFLY_CHECK(max_sl <= 255 && grid <= 65535, "decode_step: too many slices per tile");
  g.max_slices = max_sl;
  unsigned short* d_first = nullptr;
  unsigned char* d_nsl = nullptr;
  BFLY_TRY(dev_alloc(e, &d_first, first.size(), &e->workspace_bytes));
  BFLY_TRY(dev_alloc(e, &d_nsl, nsl.size(), &e->workspace_bytes));
  BFLY_CUDA(cudaMemcpy(d_first, first.data(), first.size() * sizeof(unsigned short), cudaMemcpyHostToDevice));
  BFLY_CUDA(cudaMemcpy(d_nsl, nsl.data(), nsl.size(), cudaMemcpyHostToDevice));
  g.first = d_first; g.nsl = d_nsl;
  BFLY_TRY(dev_alloc(e, &g.out, (size_t)max_sl * T * M, &e->workspace_bytes));
  *out = g;
  return BFLY_OK;
}
}  // namespace

int Engine::build_step_plan(int B, bool want_logits, StepPlan** out) {
  const int key = B * 2 + (want_logits ? 1 : 0);
  auto it = step_plans.find(key);
  if (it != step_plans.end()) { *out = &it->second; return BFLY_OK; }
  const bfly_model_desc& c = d;
  BFLY_CUDA(cudaSetDevice(device));
  BFLY_TRY(decode_step_init());
  StepPlan sp{};
  StepParams& p = sp.p;
  sp.grid = sm_count;
  const int T = B, h = c.hidden, dh = c.head_dim;
  p.T = T;
  p.bn = std::max(16, ceil_div(T, 16) * 16);
  p.stages = std::min(10, (int)(kStepRingBytes / (16384 + (size_t)p.bn * 128)));
  if (const char* env = getenv("BFLY_STEP_STAGES")) p.stages = std::max(2, std::min(p.stages, atoi(env)));   // tuning knob
  p.n_layers = c.n_layers; p.h = h; p.nq = q_local; p.nk = kv_local; p.f_local = f_block;
  p.qkv_rows = qkv_rows; p.v_local = v_local; p.v0 = v0; p.vocab = c.vocab;
  p.eps = c.norm_eps;
  p.attn_scale_log2 = 1.4426950408889634f / std::sqrt((float)dh);
  if (const char* env = getenv("BFLY_STEP_TUNE")) p.tune = atoi(env);
  p.n_producers = 3;
  if (const char* env = getenv("BFLY_STEP_PRODUCERS")) p.n_producers = std::max(1, std::min(3, atoi(env)));   // A/B knob
  // norm phases: as many (row, chunk) items as fit the grid, chunks tile-aligned
  p.chunks = 1;
  while (p.chunks < 8 && T * p.chunks * 2 <= sp.grid && h % (p.chunks * 2 * 128) == 0) p.chunks *= 2;
  // context splits of the attention items: fewest rounds x tiles per item (+3 tiles of ramp per item)
  {
    const int tiles = std::max(1, ceil_div(c.max_ctx, 32));
    int best = 1;
    long best_cost = -1;
    for (int s = 1; s <= 16 && (s == 1 || tiles / s >= 4); ++s) {
      const long cost = (long)ceil_div(T * kv_local * s, sp.grid) * (ceil_div(tiles, s) + 3);
      if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = s; }
    }
    p.attn_splits = best;
  }
  BFLY_TRY(make_step_gemm(this, qkv_rows, h, T, sp.grid, &p.gemm[SG_QKV]));
  BFLY_TRY(make_step_gemm(this, h, q_local * dh, T, sp.grid, &p.gemm[SG_WO]));
  BFLY_TRY(make_step_gemm(this, 2 * f_block, h, T, sp.grid, &p.gemm[SG_GU]));
  BFLY_TRY(make_step_gemm(this, h, f_block, T, sp.grid, &p.gemm[SG_DOWN]));
  BFLY_TRY(make_step_gemm(this, v_local, h, T, sp.grid, &p.gemm[SG_HEAD]));
  // tensor maps
  const int n_maps = c.n_layers * 4 + 4;
  std::vector<char> maps((size_t)n_maps * 128);
  auto M = [&](int i) { return (void*)(maps.data() + (size_t)i * 128); };
  for (int l = 0; l < c.n_layers; ++l) {
    const LayerW& w = layers[l];
    BFLY_TRY(make_tensor_map_2d(M(l * 4 + 0), w.wqkv, qkv_rows, h, h, 128));
    BFLY_TRY(make_tensor_map_2d(M(l * 4 + 1), w.wo, h, q_local * dh, q_local * dh, 128));
    BFLY_TRY(make_tensor_map_2d(M(l * 4 + 2), w.w_in, 2 * f_block, h, h, 128));
    BFLY_TRY(make_tensor_map_2d(M(l * 4 + 3), w.w_down, h, f_block, f_block, 128));
  }
  BFLY_TRY(make_tensor_map_2d(M(c.n_layers * 4 + 0), lm_head, v_local, h, h, 128));
  BFLY_TRY(make_tensor_map_2d(M(c.n_layers * 4 + 1), xn, T, h, h, p.bn));
  BFLY_TRY(make_tensor_map_2d(M(c.n_layers * 4 + 2), attn_out, T, q_local * dh, q_local * dh, p.bn));
  BFLY_TRY(make_tensor_map_2d(M(c.n_layers * 4 + 3), act, T, f_block, f_block, p.bn));
  std::memcpy(p.map_head_param, M(c.n_layers * 4 + 0), 128);
  std::memcpy(p.map_xg_param, M(c.n_layers * 4 + 1), 128);
  char* d_maps = nullptr;
  BFLY_TRY(dev_alloc(this, &d_maps, maps.size(), &workspace_bytes));
  BFLY_CUDA(cudaMemcpy(d_maps, maps.data(), maps.size(), cudaMemcpyHostToDevice));
  p.maps = d_maps;
  std::vector<StepLayer> sl((size_t)c.n_layers);
  for (int l = 0; l < c.n_layers; ++l)
    sl[(size_t)l] = StepLayer{layers[l].attn_norm_w, layers[l].ffn_norm_w, kv_pages + (int64_t)l * layer_kv_elems};
  StepLayer* d_layers = nullptr;
  BFLY_TRY(dev_alloc(this, &d_layers, sl.size(), &workspace_bytes));
  BFLY_CUDA(cudaMemcpy(d_layers, sl.data(), sl.size() * sizeof(StepLayer), cudaMemcpyHostToDevice));
  p.layers = d_layers;
  p.final_norm_w = final_norm_w;
  p.residual = residual; p.xg = xn; p.q = q; p.attn_out = attn_out; p.act = act;
  BFLY_TRY(dev_alloc(this, &p.ssq, (size_t)3 * T * p.chunks, &workspace_bytes));
  BFLY_TRY(dev_alloc(this, &p.part_o, (size_t)T * q_local * p.attn_splits * dh, &workspace_bytes));
  BFLY_TRY(dev_alloc(this, &p.part_ml, (size_t)T * q_local * p.attn_splits * 2, &workspace_bytes));
  p.tokens = cur_tokens; p.embed = embed; p.history = history; p.step_counter = step_counter; p.keys = keys;
  p.logits = want_logits ? logits : nullptr;
  p.logits_ld = v_shard;
  p.seq_lens = seq_lens; p.page_table = page_table; p.pages_per_seq = pages_per_seq; p.page_size = c.page_size;
  p.rope_table = rope_table;
  sp.n_counters = step_phase_count(c.n_layers, p.attn_splits);
  BFLY_TRY(dev_alloc(this, &p.counters, (size_t)sp.n_counters, &workspace_bytes));
  if (getenv("BFLY_STEP_TRACE")) {               // profiling: per-phase completion times of the last step
    const size_t n_trace = (size_t)sp.n_counters + 2 + (size_t)sp.n_counters * sp.grid;
    BFLY_TRY(dev_alloc(this, &p.trace, n_trace, &workspace_bytes));
    BFLY_CUDA(cudaMemset(p.trace, 0, n_trace * sizeof(unsigned long long)));
  }
  p.peers = ar;
  p.ar_step = ar_step;
  p.ar_on = (c.tp > 1 && p2p_ready) ? 1 : 0;
  BFLY_CUDA(cudaDeviceSynchronize());
  step_plans[key] = sp;
  *out = &step_plans[key];
  return BFLY_OK;
}

// One decode step of B sequences as the fused kernel + the greedy-token exchange.
int Engine::forward_step(int B, bool want_logits, bool record_history) {
  StepPlan* sp = nullptr;
  BFLY_TRY(build_step_plan(B, want_logits, &sp));
  StepParams p = sp->p;
  if (!record_history) p.history = nullptr;
  BFLY_CUDA(cudaMemsetAsync(p.counters, 0, (size_t)sp->n_counters * sizeof(unsigned), stream));
  BFLY_TRY(launch_decode_step(p, sp->grid, stream));
  if (d.tp > 1) {
    BFLY_TRY(all_gather_bytes(keys, keys_all, (size_t)d.max_batch * sizeof(unsigned long long)));
    BFLY_TRY(launch_pick_tokens_pitched(keys_all, d.tp, d.max_batch, B, cur_tokens, stream));
  } else {
    BFLY_TRY(launch_pick_tokens_pitched(keys, 1, d.max_batch, B, cur_tokens, stream));
  }
  BFLY_TRY(launch_advance(seq_lens, B, 1, record_history ? step_counter : nullptr, p.ar_on ? ar_step : nullptr, stream));
  return BFLY_OK;
}

int Engine::check_ready(int B) {
  if (!ready) { set_error("bfly_weights_ready() has not succeeded yet"); return BFLY_ESTATE; }
  BFLY_CHECK(B >= 1 && B <= d.max_batch, "batch must be in [1, max_batch]");
  BFLY_CUDA(cudaSetDevice(device));
  return BFLY_OK;
}

int Engine::fetch_outputs(int seq0, int B, float* logits_host, int32_t* next_host) {
  const bfly_model_desc& c = d;
  if (logits_host) {
    if (c.tp > 1) {
      BFLY_TRY(all_gather_bytes(logits, logits_all, (size_t)c.max_batch * v_shard * sizeof(float)));
      for (int r = 0; r < c.tp; ++r) {
        const int r0 = std::min(c.vocab, r * v_shard), r1 = std::min(c.vocab, (r + 1) * v_shard);
        if (r1 > r0)
          BFLY_CUDA(cudaMemcpy2DAsync(logits_host + r0, (size_t)c.vocab * sizeof(float),
                                      logits_all + ((size_t)r * c.max_batch + seq0) * v_shard,
                                      (size_t)v_shard * sizeof(float), (size_t)(r1 - r0) * sizeof(float), B,
                                      cudaMemcpyDeviceToHost, stream));
      }
    } else {
      BFLY_CUDA(cudaMemcpy2DAsync(logits_host, (size_t)c.vocab * sizeof(float), logits + (size_t)seq0 * v_shard,
                                  (size_t)v_shard * sizeof(float), (size_t)c.vocab * sizeof(float), B,
                                  cudaMemcpyDeviceToHost, stream));
    }
  }
  if (next_host)
    BFLY_CUDA(cudaMemcpyAsync(next_host, cur_tokens + seq0, (size_t)B * sizeof(int), cudaMemcpyDeviceToHost, stream));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  return BFLY_OK;
}

int Engine::prefill(const int32_t* tokens_host, int B, int S, float* logits_host, int32_t* next_host) {
  host_len = 0;
  return prefill_at(0, tokens_host, B, S, logits_host, next_host);
}

// Restarts sequences [seq0, seq0 + B) with S tokens each; every other sequence keeps its cache and its
// length, so prompts of different lengths are loaded group by group and then decoded as one batch.
int Engine::prefill_at(int seq0, const int32_t* tokens_host, int B, int S, float* logits_host, int32_t* next_host) {
  BFLY_TRY(check_ready(B));
  BFLY_CHECK(seq0 >= 0 && seq0 + B <= d.max_batch, "prefill: sequences must lie in [0, max_batch)");
  BFLY_CHECK(tokens_host != nullptr && S >= 1 && S <= d.max_ctx && S <= d.max_tokens, "prefill: 1 <= S <= min(max_ctx, max_tokens)");
  BFLY_CUDA(cudaMemsetAsync(seq_lens + seq0, 0, (size_t)B * sizeof(int), stream));
  host_len = std::max(host_len, S);               // longest sequence: what the KV-overflow checks go by
  const int per_chunk = std::max(1, d.max_tokens / S);
  for (int b0 = 0; b0 < B; b0 += per_chunk) {
    const int nb = std::min(per_chunk, B - b0);
    BFLY_CUDA(cudaMemcpyAsync(tokens_in, tokens_host + (size_t)b0 * S, (size_t)nb * S * sizeof(int),
                              cudaMemcpyHostToDevice, stream));
    BFLY_TRY(forward(tokens_in, seq0 + b0, nb, S, logits_host != nullptr, false));
    // tokens_in is reused by the next chunk: the H2D copy is stream-ordered after this forward
  }
  return fetch_outputs(seq0, B, logits_host, next_host);
}

int Engine::decode_graph(int B, bool want_logits, GraphEntry** out) {
  const int key = B * 2 + (want_logits ? 1 : 0);
  auto it = graphs.find(key);
  if (it != graphs.end()) { *out = &it->second; return BFLY_OK; }
  if (step_supported(B)) {                  // allocations and tensor maps happen before the capture begins
    StepPlan* sp = nullptr;
    BFLY_TRY(build_step_plan(B, want_logits, &sp));
  }
  GraphEntry ge{};
  const unsigned long long before = t_launch_count;
  cudaGraph_t graph = nullptr;
  BFLY_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
  const int rc = step_supported(B) ? forward_step(B, want_logits, true) : forward(cur_tokens, 0, B, 1, want_logits, true);
  const cudaError_t ce = cudaStreamEndCapture(stream, &graph);
  ge.kernel_nodes = (int)(t_launch_count - before);
  g_launch_count.fetch_sub((unsigned long long)ge.kernel_nodes, std::memory_order_relaxed);   // captured, not executed
  if (rc != BFLY_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
  BFLY_CUDA(ce);
  const cudaError_t ie = cudaGraphInstantiate(&ge.exec, graph, 0);
  cudaGraphDestroy(graph);
  BFLY_CUDA(ie);
  graphs[key] = ge;
  last_graph_nodes = ge.kernel_nodes;
  *out = &graphs[key];
  return BFLY_OK;
}

int Engine::decode(const int32_t* tokens_host, int B, float* logits_host, int32_t* next_host) {
  BFLY_TRY(check_ready(B));
  BFLY_CHECK(host_len + 1 <= d.max_ctx, "decode: KV cache is full (max_ctx reached)");
  GraphEntry* ge = nullptr;
  BFLY_TRY(decode_graph(B, logits_host != nullptr, &ge));
  host_len += 1;
  if (tokens_host)
    BFLY_CUDA(cudaMemcpyAsync(cur_tokens, tokens_host, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, stream));
  BFLY_CUDA(cudaMemsetAsync(step_counter, 0, sizeof(int), stream));
  BFLY_CUDA(cudaGraphLaunch(ge->exec, stream));
  g_launch_count += (unsigned long long)ge->kernel_nodes;
  return fetch_outputs(0, B, logits_host, next_host);
}

int Engine::decode_enqueue(int B, int n_steps) {
  BFLY_TRY(check_ready(B));
  BFLY_CHECK(n_steps >= 1 && n_steps <= kMaxGreedySteps, "n_steps in [1, 4096]");
  BFLY_CHECK(host_len + n_steps <= d.max_ctx, "decode: KV cache would overflow max_ctx");
  GraphEntry* ge = nullptr;
  BFLY_TRY(decode_graph(B, false, &ge));
  host_len += n_steps;
  BFLY_CUDA(cudaMemsetAsync(step_counter, 0, sizeof(int), stream));
  for (int i = 0; i < n_steps; ++i) BFLY_CUDA(cudaGraphLaunch(ge->exec, stream));
  g_launch_count += (unsigned long long)ge->kernel_nodes * n_steps;
  return BFLY_OK;
}

// The tokens fed at each step of the last bfly_decode_enqueue / bfly_decode_greedy of B sequences.
int Engine::read_history(int B, int n_steps, int32_t* tokens_out_host) {
  BFLY_CHECK(tokens_out_host != nullptr && B >= 1 && B <= d.max_batch && n_steps >= 1 && n_steps <= kMaxGreedySteps,
             "read_history: shape");
  BFLY_CUDA(cudaSetDevice(device));
  // history is [step][B]; the caller wants [B][n_steps]
  std::vector<int> tmp((size_t)n_steps * B);
  BFLY_CUDA(cudaMemcpyAsync(tmp.data(), history, tmp.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  for (int s = 0; s < n_steps; ++s)
    for (int b = 0; b < B; ++b) tokens_out_host[(size_t)b * n_steps + s] = tmp[(size_t)s * B + b];
  return BFLY_OK;
}

int Engine::decode_greedy(int B, int n_steps, int32_t* tokens_out_host) {
  BFLY_TRY(decode_enqueue(B, n_steps));
  if (tokens_out_host) return read_history(B, n_steps, tokens_out_host);
  BFLY_CUDA(cudaStreamSynchronize(stream));
  return BFLY_OK;
}

// Mode of the fused decode all-reduce: -1 = by size (two-shot for rows >= 1 MiB at tp >= 4), 0 = one-shot,
// 1 = two-shot.  Must be set identically on every rank.  Captured decode graphs embed the mode: drop them.
int Engine::set_ar_mode(int mode) {
  BFLY_CHECK(mode >= -1 && mode <= 1, "set_ar_mode: mode in {-1, 0, 1}");
  BFLY_CUDA(cudaSetDevice(device));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  if (mode != ar_mode) {
    for (auto& kv : graphs)
      if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    graphs.clear();
  }
  ar_mode = mode;
  return BFLY_OK;
}

// Microbenchmark of C1 (profiles/r02_c1_latency.txt): `iters` back-to-back calls of the fused all-reduce +
// residual add + norm kernel on T rows (all ranks call this together), and of the plain add+norm kernel on
// the same rows, timed with CUDA events on the ctx stream.  Average microseconds per call.
int Engine::bench_allreduce(int T, int iters, float* us_allreduce, float* us_plain) {
  BFLY_CHECK(p2p_ready, "bench_allreduce: needs tp > 1 with the peer-memory all-reduce initialised");
  BFLY_CHECK(T >= 1 && T <= ar.max_rows && T <= d.max_tokens && iters >= 1 && us_allreduce && us_plain, "bench_allreduce: arguments");
  BFLY_CUDA(cudaSetDevice(device));
  const int h = d.hidden, calls = 2 * d.n_layers, ln = d.family == BFLY_FAMILY_GPT2;
  const int two_shot = ar_mode >= 0 ? ar_mode : ar_two_shot_by_size(T);
  BFLY_CUDA(cudaMemsetAsync(ws, 0, (size_t)T * h * sizeof(float), stream));
  BFLY_CUDA(cudaMemsetAsync(residual, 0, (size_t)T * h * sizeof(float), stream));
  struct Events {                               // destroyed on every return path
    cudaEvent_t a = nullptr, b = nullptr;
    ~Events() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
  } ev;
  BFLY_CUDA(cudaEventCreate(&ev.a));
  BFLY_CUDA(cudaEventCreate(&ev.b));
  cudaEvent_t e0 = ev.a, e1 = ev.b;
  auto run = [&](int n, bool with_ar) -> int {
    for (int i = 0; i < n; ++i) {
      if (with_ar) {
        BFLY_TRY(launch_ar_add_norm(residual, ws, 1, (int64_t)T * h, nullptr, final_norm_w, final_norm_b, d.norm_eps, ln, xn, T,
                                    h, ar, i % calls, ar_step, calls, two_shot, ar_chunks, ar_bf16, ar_lamport, stream));
        if (i % calls == calls - 1) BFLY_TRY(launch_advance(seq_lens, 0, 0, nullptr, ar_step, stream));
      } else {
        BFLY_TRY(launch_add_norm(residual, ws, 1, (int64_t)T * h, nullptr, final_norm_w, final_norm_b, d.norm_eps, ln, xn, T, h,
                                 stream));
      }
    }
    return BFLY_OK;
  };
  const int n = ceil_div(iters, calls) * calls;          // whole epochs, so the step counter ends consistent
  float ms = 0.f;
  BFLY_TRY(run(calls, true));
  BFLY_CUDA(cudaEventRecord(e0, stream));
  BFLY_TRY(run(n, true));
  BFLY_CUDA(cudaEventRecord(e1, stream));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  BFLY_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *us_allreduce = ms * 1e3f / (float)n;
  BFLY_TRY(run(calls, false));
  BFLY_CUDA(cudaEventRecord(e0, stream));
  BFLY_TRY(run(n, false));
  BFLY_CUDA(cudaEventRecord(e1, stream));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  BFLY_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  *us_plain = ms * 1e3f / (float)n;
  return BFLY_OK;
}

// Profiling (BFLY_STEP_TRACE=1): nanosecond timestamps at which CTA 0 saw each phase of the LAST fused decode
// step complete (entry 0 = kernel start); returns the number of entries, 0 when tracing is off.
int Engine::read_step_trace(unsigned long long* out, int max_entries, int* n_out, int* splits_out) {
  *n_out = 0;
  *splits_out = 0;
  for (auto& kv : step_plans) {
    const StepPlan& sp = kv.second;
    if (sp.p.trace == nullptr) continue;
    const int n = std::min(max_entries, sp.n_counters + 2 + sp.n_counters * sp.grid);
    BFLY_CUDA(cudaStreamSynchronize(stream));
    BFLY_CUDA(cudaMemcpy(out, sp.p.trace, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    *n_out = n;
    *splits_out = sp.p.attn_splits;
    if (kv.first % 2 == 0) break;            // prefer the no-logits plan (the timed one)
  }
  return BFLY_OK;
}

int Engine::set_page_table(const int32_t* table_host, int n_seqs, int pps) {
  BFLY_CHECK(table_host != nullptr && n_seqs >= 1 && n_seqs <= d.max_batch && pps == pages_per_seq, "set_page_table: shape");
  std::vector<char> seen((size_t)n_pages, 0);
  for (int64_t i = 0; i < (int64_t)n_seqs * pps; ++i) {
    const int p = table_host[i];
    BFLY_CHECK(p >= 0 && p < n_pages && !seen[(size_t)p], "set_page_table: page ids must be unique and in range");
    seen[(size_t)p] = 1;
  }
  BFLY_CUDA(cudaSetDevice(device));
  BFLY_CUDA(cudaStreamSynchronize(stream));
  BFLY_CUDA(cudaMemcpy(page_table, table_host, (size_t)n_seqs * pps * sizeof(int), cudaMemcpyHostToDevice));
  return BFLY_OK;
}

}  // namespace bfly
