// Microbenchmark: cost of one grid-wide phase boundary (arrive + wait) for 148 CTAs x 192 threads, in the
// variants considered for csrc/decode_step.cu.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gb grid_barrier.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ unsigned ld_relaxed_gpu(const unsigned* p) {
  unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void red_release_gpu(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

template <int V>
__global__ void __launch_bounds__(192, 1) k(unsigned* counters, int n, float* data, float* out) {
  extern __shared__ char smem[];
  float acc = 0.f;
  for (int i = 0; i < n; ++i) {
    unsigned* c = counters + i;
    // a little "work": every thread writes one float and (after the barrier) reads another CTA's
    data[(size_t)(i & 1) * gridDim.x * 192 + blockIdx.x * 192 + threadIdx.x] = (float)i;
    // ---- arrive ----
    if (V == 0) fence_proxy_async();
    if (V == 5) fence_proxy_async_global();
    __syncthreads();
    if (threadIdx.x == 0) {
      if (V == 0 || V == 1 || V == 5) { __threadfence(); atomicAdd(c, 1u); }
      else red_release_gpu(c, 1u);
    }
    // ---- wait ----
    if (threadIdx.x == 0) {
      if (V == 4) { while (ld_relaxed_gpu(c) < gridDim.x) {} __threadfence(); }
      else { while (ld_acquire_gpu(c) < gridDim.x) {} }
      if (V == 0 || V == 1 || V == 5) __threadfence();
      if (V == 0) fence_proxy_async();
      if (V == 5) fence_proxy_async_global();
    }
    __syncthreads();
    acc += __ldcg(data + (size_t)(i & 1) * gridDim.x * 192 + ((blockIdx.x + 1) % gridDim.x) * 192 + threadIdx.x);
  }
  if (acc == -1.f) out[0] = acc;
}

template <int V>
void run(const char* name, int grid, int n) {
  unsigned* counters; float *data, *out;
  cudaMalloc(&counters, n * sizeof(unsigned)); cudaMalloc(&data, 2 * grid * 192 * sizeof(float)); cudaMalloc(&out, 4);
  cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaMemset(counters, 0, n * sizeof(unsigned));
    void* args[] = {&counters, &n, &data, &out};
    cudaEventRecord(e0);
    cudaLaunchCooperativeKernel((void*)k<V>, dim3(grid), dim3(192), args, 200 * 1024, 0);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(err)); return; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  printf("%-62s %.3f us per phase boundary\n", name, best * 1e3f / n);
  cudaFree(counters); cudaFree(data); cudaFree(out);
}

int main() {
  int grid = 148, n = 2000;
  run<0>("V0 fence.proxy.async x2, threadfence+atomicAdd, acquire poll+fence", grid, n);
  run<1>("V1 no proxy fences, threadfence+atomicAdd, acquire poll+fence", grid, n);
  run<5>("V5 fence.proxy.async.global x2, threadfence+atomicAdd, poll+fence", grid, n);
  run<2>("V2 red.release.gpu, acquire poll", grid, n);
  run<4>("V4 red.release.gpu, relaxed poll + one fence", grid, n);
  return 0;
}
