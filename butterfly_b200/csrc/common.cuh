// Shared device/host helpers for libbutterfly_b200 (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <mutex>
#include <string>
#include <utility>

#include "../../include/butterfly_b200.h"   // BFLY_OK / BFLY_E* codes

namespace bfly {

// ---- error plumbing -----------------------------------------------------------------------
// Every C-ABI entry returns an int code; the message is kept per thread and per ctx.
void set_error(const std::string& msg);   // thread-local last error (api.cu)

#define BFLY_CUDA(expr)                                                                        \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      ::bfly::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                        ":" + std::to_string(__LINE__) + ")");                                 \
      return BFLY_ECUDA;                                                               \
    }                                                                                          \
  } while (0)

#define BFLY_CHECK(cond, msg)                                  \
  do {                                                         \
    if (!(cond)) {                                             \
      ::bfly::set_error(std::string("invalid argument: ") + (msg)); \
      return BFLY_EINVAL;                              \
    }                                                          \
  } while (0)

#define BFLY_TRY(expr)            \
  do {                            \
    int _rc = (expr);             \
    if (_rc != 0) return _rc;     \
  } while (0)

// Function attributes (cudaFuncSetAttribute) belong to a device, not to the process: a host that drives
// several GPUs from one process must set them once on each.  run(fn) calls fn() the first time it is
// reached with a given device current (and again if fn failed).
struct PerDeviceOnce {
  static constexpr int kMaxDevices = 64;
  std::mutex mu;
  bool done[kMaxDevices] = {};
  template <class Fn>
  int run(Fn&& fn) {
    int dev = -1;
    BFLY_CUDA(cudaGetDevice(&dev));
    BFLY_CHECK(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!done[dev]) {
      BFLY_TRY(fn());
      done[dev] = true;
    }
    return BFLY_OK;
  }
};

// Counts kernel launches made by this library (bench.py reports it as gpu_launches).
extern std::atomic<unsigned long long> g_launch_count;
extern thread_local unsigned long long t_launch_count;      // this thread's share (graph-capture accounting)
inline void count_launch(int n = 1) {
  g_launch_count.fetch_add((unsigned long long)n, std::memory_order_relaxed);
  t_launch_count += (unsigned long long)n;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

#ifdef __CUDACC__
// ---- programmatic dependent launch (PDL) ------------------------------------------------------
// Every kernel of the forward is launched with programmatic stream serialization: it may start while
// its predecessor is still running, calls pdl_trigger() at once (so its own successor can do the
// same), does whatever does not depend on earlier kernels (barrier init, TMEM allocation, WEIGHT
// prefetch), and calls pdl_wait() before the first access to memory another kernel produces or still
// reads.  Every kernel calls pdl_wait(), so completion stays transitive along the stream.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(std::forward<Args>(args))...);
}

// ---- device helpers -------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// block-wide sum for up to 1024 threads; `red` is a 32-float shared scratch.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();           // protect `red` from the previous use
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  return warp_sum(r);
}

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);   // .x = a (low half), .y = b
  return *reinterpret_cast<uint32_t*>(&v);
}

// streaming 128-bit load that does not pollute L1 (weights / KV are read once)
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint16_t ld_stream_u16(const void* p) {
  uint16_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}
#endif  // __CUDACC__

}  // namespace bfly
