// bfly_k_attn_decode_static -- paged decode attention for callers that cannot allocate device memory.
//
// The ordinary entry point (bfly_k_attn_decode) needs two fp32 scratch buffers whenever the context
// is split over several CTAs (n_splits > 1): the per-split partial outputs and their (max, sum)
// pairs, which a second kernel merges.  Somebody has to own that memory.  Some users of this library
// have nowhere to get it from: a host application whose memory pool has already claimed the whole
// GPU (a further cudaMalloc simply fails), a serving loop that must not call cudaMalloc/cudaFree on
// its hot path because those calls synchronise the device, or a thin language binding that only
// forwards the caller's input and output buffers and has no allocator of its own.  For them the
// scratch lives here, in a fixed array that is part of the loaded library image.
//
// What to know before using it:
//   * The array is kStaticScratchBytes (4 MiB) per GPU and belongs to the library's device code, not to
//     any allocator.  CUDA loads device code lazily by default, so the 4 MiB are claimed on a GPU when
//     this entry point first runs there (the address lookup below loads the module that holds the
//     array); from then on they stay until the process exits.  Setting CUDA_MODULE_LOADING=EAGER moves
//     that moment to library load.  Either way the FIRST call on a device needs 4 MiB of free device
//     memory: make it during start-up, before your pool takes the rest of the GPU.  If the memory is not
//     there, that call returns BFLY_ECUDA with the runtime's message and launches nothing.
//   * A request with n_splits > 1 needs T * nq * n_splits * (head_dim + 2) floats of the array.  Larger
//     requests are refused with BFLY_EINVAL and a message that states both numbers; nothing is launched.
//     n_splits == 1 writes the output directly and uses no scratch, so it has no such limit.
//   * The first call on a given device looks up the array's device address (cudaGetSymbolAddress) with
//     that device current and remembers it.  Make that first call before any stream capture is begun
//     on that device.
//   * All calls on a device share the one array and it keeps whatever the previous call left in it.
//     Nothing in it needs to start at zero: every slot a call reads has been written by the same call
//     first (each split's CTA writes its slice before the merge kernel reads it).  Two calls whose
//     kernels overlap on the GPU (different streams, or different host threads) race on the array and
//     produce garbage -- serialise them.
//   * After that first call, no path of this function obtains or releases device memory: it makes no
//     cudaMalloc / cudaFree call and uses no library routine that does.
//
// The reference has no counterpart (it has no code at all); this is an addition of this library.
#include <mutex>

#include "kernels.cuh"

namespace {

constexpr size_t kStaticScratchBytes = 4u << 20;
constexpr size_t kStaticScratchFloats = kStaticScratchBytes / sizeof(float);
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime's symbol table, never a device address.
static __device__ __align__(256) float g_attn_static_scratch[kStaticScratchFloats];

float* g_scratch_addr[kMaxDevices] = {};      // device address of the array, per device ordinal
std::mutex g_scratch_mutex;

// Resolve (once per device, with that device current) where the array lives on the current device.
int scratch_for_current_device(float** out) {
  int dev = -1;
  BFLY_CUDA(cudaGetDevice(&dev));
  BFLY_CHECK(dev >= 0 && dev < kMaxDevices, "attn_decode_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_attn_static_scratch);
    if (err != cudaSuccess) {
      bfly::set_error(std::string("attn_decode_static: cannot place the built-in scratch array on device ") +
                      std::to_string(dev) + " (it needs " + std::to_string(kStaticScratchBytes >> 20) +
                      " MiB of free device memory on first use): " + cudaGetErrorString(err));
      return BFLY_ECUDA;
    }
    BFLY_CHECK(p != nullptr, "attn_decode_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return BFLY_OK;
}

}  // namespace

extern "C" int bfly_k_attn_decode_static(const void* q_bf16, const void* kv_pages_bf16, const int32_t* page_table,
                                         int pages_per_seq, int page_size, int kv_local, int head_dim,
                                         const int32_t* seq_lens, int S, int T, int nq, int n_splits, void* out_bf16,
                                         void* stream) {
  using namespace bfly;
  try {
    BFLY_CHECK(q_bf16 && kv_pages_bf16 && page_table && seq_lens && out_bf16, "attn_decode_static: null pointer");
    BFLY_CHECK(T > 0 && nq > 0 && n_splits >= 1 && head_dim > 0 && S > 0, "attn_decode_static: empty shape");
    const size_t slots = (size_t)T * (size_t)nq * (size_t)n_splits;
    const size_t need = slots * ((size_t)head_dim + 2);
    if (n_splits > 1 && need > kStaticScratchFloats) {       // a single split writes bf16 directly: no scratch
      set_error("attn_decode_static: T*nq*n_splits*(head_dim+2) = " + std::to_string(need) +
                " floats of scratch needed, the built-in array holds " + std::to_string(kStaticScratchFloats) + " (" +
                std::to_string(kStaticScratchBytes >> 20) +
                " MiB); lower n_splits or the batch, or use bfly_k_attn_decode with your own scratch");
      return BFLY_EINVAL;
    }
    float* scratch = nullptr;
    BFLY_TRY(scratch_for_current_device(&scratch));
    float* part_o = scratch;                                  // [T][nq][n_splits][head_dim]
    float* part_ml = scratch + slots * (size_t)head_dim;      // [T][nq][n_splits][2]
    KvLayout kv{const_cast<bf16*>(reinterpret_cast<const bf16*>(kv_pages_bf16)), page_table, pages_per_seq, page_size,
                kv_local, head_dim};
    return launch_attn_decode(reinterpret_cast<const bf16*>(q_bf16), kv, seq_lens, S, T, nq, part_o, part_ml, n_splits,
                              reinterpret_cast<bf16*>(out_bf16), reinterpret_cast<cudaStream_t>(stream));
  } catch (...) {
    set_error("attn_decode_static: internal exception");
    return BFLY_EINVAL;
  }
}
