// Synthetic weight fill: every element is a pure function of (key, row, col), bit-identical to
// butterfly_b200/weights.py::synth_block (integer hash + one IEEE multiply/add + RNE to bf16), so a
// rank creates its shard directly in HBM.  Setup path, not the hot path.
#include "kernels.cuh"

namespace bfly {

__device__ __forceinline__ float synth_value(uint64_t key, uint64_t idx, float scale, float offset) {
  uint64_t z = key + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  const int s = (int)(z & 0xFFFF) + (int)((z >> 16) & 0xFFFF) + (int)((z >> 32) & 0xFFFF) + (int)(z >> 48);
  float v = __fmul_rn(__int2float_rn(s - 131070), scale);     // explicit rn ops: never contracted to FMA
  if (offset != 0.f) v = __fadd_rn(v, offset);
  return v;
}

__global__ void fill_synth_kernel(bf16* __restrict__ dst, int64_t pitch, uint64_t key, int64_t ncols_full,
                                  int64_t r0, int64_t c0, int64_t rows, int64_t cols, float scale, float offset) {
  const int64_t total = rows * cols;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols, c = i - r * cols;
    const uint64_t idx = (uint64_t)(r0 + r) * (uint64_t)ncols_full + (uint64_t)(c0 + c);
    dst[r * pitch + c] = __float2bfloat16_rn(synth_value(key, idx, scale, offset));
  }
}

int launch_fill_synth(bf16* dst, int64_t dst_pitch, uint64_t key, int64_t ncols_full, int64_t r0, int64_t c0,
                      int64_t rows, int64_t cols, float scale, float offset, cudaStream_t stream) {
  BFLY_CHECK(dst != nullptr && rows >= 0 && cols >= 0 && dst_pitch >= cols, "fill_synth: bad block");
  if (rows * cols == 0) return BFLY_OK;
  const int64_t total = rows * cols;
  const int block = 256;
  const int grid = (int)std::min<int64_t>((total + block - 1) / block, 148 * 16);
  fill_synth_kernel<<<grid, block, 0, stream>>>(dst, dst_pitch, key, ncols_full, r0, c0, rows, cols, scale, offset);
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

}  // namespace bfly
