// Bandwidth-bound kernels of the block forward: 128-bit
// coalesced accesses, warp-shuffle reductions, fp32 math, bf16 only where the numerics contract
// stores a GEMM operand or the KV cache.  Each consumer also folds the split-K slices of the GEMM
// that produced its input (and the bias / residual add), so no separate reduction pass exists.
#include <algorithm>

#include "kernels.cuh"

namespace bfly {

// ---------------------------------------------------------------------------------------------
// K8: embedding gather (+ learned positions) -> fp32 residual stream; also clears the argmax keys
// ---------------------------------------------------------------------------------------------
__global__ void embed_kernel(const int* __restrict__ tokens, const bf16* __restrict__ table,
                             const bf16* __restrict__ pos_table, const int* __restrict__ seq_lens, int S,
                             float* __restrict__ residual, int h, int vocab, unsigned long long* argmax_keys,
                             int n_keys, int* history, const int* step_counter) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x;
  int tok = tokens[t];
  if (history != nullptr && threadIdx.x == 0) history[(int64_t)(*step_counter) * gridDim.x + t] = tok;
  tok = min(max(tok, 0), vocab - 1);
  const bf16* row = table + (int64_t)tok * h;
  const bf16* prow = nullptr;
  if (pos_table != nullptr) prow = pos_table + (int64_t)(seq_lens[t / S] + t % S) * h;
  float* dst = residual + (int64_t)t * h;
  for (int i = threadIdx.x * 8; i < h; i += blockDim.x * 8) {
    const uint4 v = *reinterpret_cast<const uint4*>(row + i);
    float f[8] = {bf16_lo(v.x), bf16_hi(v.x), bf16_lo(v.y), bf16_hi(v.y),
                  bf16_lo(v.z), bf16_hi(v.z), bf16_lo(v.w), bf16_hi(v.w)};
    if (prow != nullptr) {
      const uint4 p = *reinterpret_cast<const uint4*>(prow + i);
      f[0] += bf16_lo(p.x); f[1] += bf16_hi(p.x); f[2] += bf16_lo(p.y); f[3] += bf16_hi(p.y);
      f[4] += bf16_lo(p.z); f[5] += bf16_hi(p.z); f[6] += bf16_lo(p.w); f[7] += bf16_hi(p.w);
    }
    *reinterpret_cast<float4*>(dst + i) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(dst + i + 4) = make_float4(f[4], f[5], f[6], f[7]);
  }
  if (argmax_keys != nullptr && threadIdx.x == 0 && t < n_keys) argmax_keys[t] = 0ull;
}

int launch_embed(const int* tokens, const bf16* table, const bf16* pos_table, const int* seq_lens, int S,
                 float* residual, int T, int h, int vocab, unsigned long long* argmax_keys, int n_keys,
                 int* history, const int* step_counter, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && h % 8 == 0 && S > 0, "embed: T>0, S>0, h%8==0");
  BFLY_CUDA(launch_pdl(embed_kernel, dim3(T), dim3(128), 0, stream, tokens, table, pos_table, seq_lens, S, residual, h, vocab, argmax_keys,
                                      n_keys, history, step_counter));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K3: residual += sum(split-K slices) + bias;  out = norm(residual) as the next GEMM's bf16 operand
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) add_norm_kernel(
    float* __restrict__ residual, const float* __restrict__ partials, int n_partials, int64_t partial_stride,
    const bf16* __restrict__ bias, const bf16* __restrict__ gamma, const bf16* __restrict__ beta, float eps,
    int layer_norm, bf16* __restrict__ out, int h, const bf16* __restrict__ partial_bf16) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) float srow[];
  __shared__ float red[32];
  const int row = blockIdx.x;
  float* res = residual + (int64_t)row * h;
  float sum = 0.f, sumsq = 0.f;
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    float4 v = *reinterpret_cast<const float4*>(res + i);
    if (partial_bf16 != nullptr) {                // one complete bf16 partial (prefill: all-reduced over the ranks)
      const uint2 pb = *reinterpret_cast<const uint2*>(partial_bf16 + (int64_t)row * h + i);
      v.x += bf16_lo(pb.x); v.y += bf16_hi(pb.x); v.z += bf16_lo(pb.y); v.w += bf16_hi(pb.y);
    }
    const float* pp = partials + (int64_t)row * h + i;
    // slices are folded in index order, four loads in flight at a time (latency, not bandwidth, bound)
    int p = 0;
    for (; p + 4 <= n_partials; p += 4) {
      const float4 a0 = *reinterpret_cast<const float4*>(pp + (p + 0) * partial_stride);
      const float4 a1 = *reinterpret_cast<const float4*>(pp + (p + 1) * partial_stride);
      const float4 a2 = *reinterpret_cast<const float4*>(pp + (p + 2) * partial_stride);
      const float4 a3 = *reinterpret_cast<const float4*>(pp + (p + 3) * partial_stride);
      v.x = (((v.x + a0.x) + a1.x) + a2.x) + a3.x; v.y = (((v.y + a0.y) + a1.y) + a2.y) + a3.y;
      v.z = (((v.z + a0.z) + a1.z) + a2.z) + a3.z; v.w = (((v.w + a0.w) + a1.w) + a2.w) + a3.w;
    }
    for (; p < n_partials; ++p) {
      const float4 a = *reinterpret_cast<const float4*>(pp + p * partial_stride);
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    }
    if (bias != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(bias + i);
      v.x += bf16_lo(b.x); v.y += bf16_hi(b.x); v.z += bf16_lo(b.y); v.w += bf16_hi(b.y);
    }
    if (n_partials > 0 || bias != nullptr || partial_bf16 != nullptr) *reinterpret_cast<float4*>(res + i) = v;
    *reinterpret_cast<float4*>(srow + i) = v;
    sum += (v.x + v.y) + (v.z + v.w);
    sumsq += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  float mean = 0.f, rstd;
  if (layer_norm) {
    mean = block_sum(sum, red) / (float)h;
    float var = 0.f;
    for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
      const float4 v = *reinterpret_cast<const float4*>(srow + i);
      const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
      var += (a * a + b * b) + (c * c + d * d);
    }
    var = block_sum(var, red) / (float)h;
    rstd = 1.0f / sqrtf(var + eps);
  } else {
    rstd = 1.0f / sqrtf(block_sum(sumsq, red) / (float)h + eps);
  }
  bf16* o = out + (int64_t)row * h;
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    const float4 v = *reinterpret_cast<const float4*>(srow + i);
    const uint2 g = *reinterpret_cast<const uint2*>(gamma + i);
    float y0 = (v.x - mean) * rstd * bf16_lo(g.x), y1 = (v.y - mean) * rstd * bf16_hi(g.x);
    float y2 = (v.z - mean) * rstd * bf16_lo(g.y), y3 = (v.w - mean) * rstd * bf16_hi(g.y);
    if (beta != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(beta + i);
      y0 += bf16_lo(b.x); y1 += bf16_hi(b.x); y2 += bf16_lo(b.y); y3 += bf16_hi(b.y);
    }
    *reinterpret_cast<uint2*>(o + i) = make_uint2(pack_bf16(y0, y1), pack_bf16(y2, y3));
  }
}

int launch_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                    const bf16* bias, const bf16* gamma, const bf16* beta, float eps, int layer_norm, bf16* out,
                    int T, int h, cudaStream_t stream, const bf16* partial_bf16) {
  BFLY_CHECK(T > 0 && h % 4 == 0 && h * 4 <= 160 * 1024, "add_norm: T>0, h%4==0, h<=40960");
  BFLY_CHECK(n_partials == 0 || partials != nullptr, "add_norm: partials missing");
  const size_t smem = (size_t)h * sizeof(float);
  static PerDeviceOnce attr;
  BFLY_TRY(attr.run([] {
    BFLY_CUDA(cudaFuncSetAttribute(add_norm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    return (int)BFLY_OK;
  }));
  // few rows (decode): one wide CTA per row minimises latency; many rows (prefill): narrow CTAs keep many more
  // rows in flight per SM
  const int threads = T >= 512 ? 256 : std::min(1024, std::max(128, ((h / 4 + 31) / 32) * 32));
  BFLY_CUDA(launch_pdl(add_norm_kernel, dim3(T), dim3(threads), smem, stream, residual, partials, n_partials, partial_stride, bias, gamma, beta,
                                            eps, layer_norm, out, h, partial_bf16));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K4: fold QKV split-K slices (+bias), rotate q/k (half-split pairs), write q (bf16) and append
//     k, v to the paged cache
// ---------------------------------------------------------------------------------------------
__global__ void rope_kv_kernel(const float* __restrict__ qkv, int n_partials, int64_t partial_stride,
                               const bf16* __restrict__ bias, const float2* __restrict__ rope_table,
                               const int* __restrict__ seq_lens, int S, int nq, int nk, bf16* __restrict__ q_out,
                               KvLayout kv) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x, head = blockIdx.y, i = threadIdx.x;
  const int d = kv.head_dim, half = d >> 1;
  const int ld = (nq + 2 * nk) * d;
  const int c0 = head * d + i, c1 = c0 + half;
  float x0 = 0.f, x1 = 0.f;
  {
    // slices are folded in index order, four pairs of loads in flight at a time (latency bound)
    const float* src = qkv + (int64_t)t * ld;
    int p = 0;
    for (; p + 4 <= n_partials; p += 4) {
      float a[4], b2[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k] = src[(p + k) * partial_stride + c0];
        b2[k] = src[(p + k) * partial_stride + c1];
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) { x0 += a[k]; x1 += b2[k]; }
    }
    for (; p < n_partials; ++p) {
      x0 += src[p * partial_stride + c0];
      x1 += src[p * partial_stride + c1];
    }
  }
  if (bias != nullptr) {
    x0 += __bfloat162float(bias[c0]);
    x1 += __bfloat162float(bias[c1]);
  }
  const int b = t / S;
  const int pos = seq_lens[b] + (t - b * S);
  if (rope_table != nullptr && head < nq + nk) {
    const float2 cs = rope_table[(int64_t)pos * half + i];
    const float y0 = x0 * cs.x - x1 * cs.y;
    const float y1 = x1 * cs.x + x0 * cs.y;
    x0 = y0; x1 = y1;
  }
  if (head < nq) {
    bf16* dst = q_out + (int64_t)t * nq * d + head * d;
    dst[i] = __float2bfloat16_rn(x0);
    dst[i + half] = __float2bfloat16_rn(x1);
  } else {
    const int which = (head < nq + nk) ? 0 : 1;
    const int kh = (head - nq) - which * nk;
    const int page = kv.page_table[(int64_t)b * kv.pages_per_seq + pos / kv.page_size];
    const int slot = pos % kv.page_size;
    bf16* blk = kv.base + (((int64_t)page * 2 + which) * kv.kv_local + kh) * kv.page_size * d;
    if (which == 0) {
      // K block: [page_size][d] rows; 16-byte chunks XOR-swizzled by token parity (see attn_decode.cu)
      blk[kv_k_offset(slot, i, d)] = __float2bfloat16_rn(x0);
      blk[kv_k_offset(slot, i + half, d)] = __float2bfloat16_rn(x1);
    } else {
      // V block: transposed per 32-token sub-block, [page_size/32][d][32]
      blk[kv_v_offset(slot, i, d)] = __float2bfloat16_rn(x0);
      blk[kv_v_offset(slot, i + half, d)] = __float2bfloat16_rn(x1);
    }
  }
}

// Prefill variant: a CTA owns 32 consecutive tokens of one sequence and one head, stages the rotated
// values in shared memory and writes q rows, K rows and the TRANSPOSED V block with consecutive threads
// on consecutive addresses (the per-token kernel above would scatter 2-byte V stores 64 bytes apart).
constexpr int kRopeTile = 32;
__global__ void __launch_bounds__(256) rope_kv_tile_kernel(
    const float* __restrict__ qkv, int n_partials, int64_t partial_stride, const bf16* __restrict__ bias,
    const float2* __restrict__ rope_table, const int* __restrict__ seq_lens, int S, int nq, int nk,
    bf16* __restrict__ q_out, KvLayout kv) {
  extern __shared__ float tile[];                     // [32][d + 1]
  pdl_trigger();
  pdl_wait();
  const int s0 = blockIdx.x * kRopeTile, head = blockIdx.y, b = blockIdx.z;
  const int d = kv.head_dim, half = d >> 1, ldt = d + 1;
  const int ld = (nq + 2 * nk) * d;
  const int n_tok = min(kRopeTile, S - s0);
  const int pos0 = seq_lens[b] + s0;
  const bool rotate = rope_table != nullptr && head < nq + nk;
  const int q4 = half >> 2;                            // float4 groups per half row (half % 4 == 0)
#pragma unroll 2
  for (int idx = threadIdx.x; idx < n_tok * q4; idx += blockDim.x) {
    const int tok = idx / q4, i = (idx - tok * q4) * 4;
    const int c0 = head * d + i, c1 = c0 + half;
    const float* src = qkv + (int64_t)(b * S + s0 + tok) * ld;
    float4 a = *reinterpret_cast<const float4*>(src + c0), bb = *reinterpret_cast<const float4*>(src + c1);
    for (int p = 1; p < n_partials; ++p) {
      const float4 a2 = *reinterpret_cast<const float4*>(src + p * partial_stride + c0);
      const float4 b2 = *reinterpret_cast<const float4*>(src + p * partial_stride + c1);
      a.x += a2.x; a.y += a2.y; a.z += a2.z; a.w += a2.w;
      bb.x += b2.x; bb.y += b2.y; bb.z += b2.z; bb.w += b2.w;
    }
    float x0[4] = {a.x, a.y, a.z, a.w}, x1[4] = {bb.x, bb.y, bb.z, bb.w};
    if (bias != nullptr) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        x0[k] += __bfloat162float(bias[c0 + k]);
        x1[k] += __bfloat162float(bias[c1 + k]);
      }
    }
    if (rotate) {
      const float4* cs = reinterpret_cast<const float4*>(rope_table + (int64_t)(pos0 + tok) * half + i);
      const float4 cs01 = cs[0], cs23 = cs[1];           // (cos, sin) of i, i+1 | i+2, i+3
      const float cc[4] = {cs01.x, cs01.z, cs23.x, cs23.z}, sn[4] = {cs01.y, cs01.w, cs23.y, cs23.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float y0 = x0[k] * cc[k] - x1[k] * sn[k], y1 = x1[k] * cc[k] + x0[k] * sn[k];
        x0[k] = y0; x1[k] = y1;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tile[tok * ldt + i + k] = x0[k];
      tile[tok * ldt + i + half + k] = x1[k];
    }
  }
  __syncthreads();
  if (head < nq) {
    for (int idx = threadIdx.x; idx < n_tok * half; idx += blockDim.x) {
      const int tok = idx / half, dd = (idx - tok * half) * 2;
      bf16* dst = q_out + ((int64_t)(b * S + s0 + tok) * nq + head) * d + dd;
      *reinterpret_cast<uint32_t*>(dst) = pack_bf16(tile[tok * ldt + dd], tile[tok * ldt + dd + 1]);
    }
    return;
  }
  const int which = (head < nq + nk) ? 0 : 1;
  const int kh = (head - nq) - which * nk;
  const int ps = kv.page_size;
  if (which == 0) {
    for (int idx = threadIdx.x; idx < n_tok * half; idx += blockDim.x) {
      const int tok = idx / half, dd = (idx - tok * half) * 2;
      const int pos = pos0 + tok;
      const int page = kv.page_table[(int64_t)b * kv.pages_per_seq + pos / ps];
      bf16* blk = kv.base + (((int64_t)page * 2) * kv.kv_local + kh) * ps * d;
      *reinterpret_cast<uint32_t*>(blk + kv_k_offset(pos % ps, dd, d)) =
          pack_bf16(tile[tok * ldt + dd], tile[tok * ldt + dd + 1]);
    }
  } else {
    for (int idx = threadIdx.x; idx < d * kRopeTile; idx += blockDim.x) {
      const int dd = idx / kRopeTile, tok = idx - dd * kRopeTile;
      if (tok >= n_tok) continue;
      const int pos = pos0 + tok;
      const int page = kv.page_table[(int64_t)b * kv.pages_per_seq + pos / ps];
      bf16* blk = kv.base + (((int64_t)page * 2 + 1) * kv.kv_local + kh) * ps * d;
      blk[kv_v_offset(pos % ps, dd, d)] = __float2bfloat16_rn(tile[tok * ldt + dd]);
    }
  }
}

int launch_rope_kv(const float* qkv, int n_partials, int64_t partial_stride, const bf16* bias,
                   const float2* rope_table, const int* seq_lens, int S, int T, int nq, int nk, bf16* q_out,
                   KvLayout kv, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && n_partials >= 1 && kv.head_dim % 2 == 0 && kv.head_dim <= 2048, "rope_kv: bad shape");
  BFLY_CHECK(nk == kv.kv_local, "rope_kv: nk must equal the cache's kv heads");
  if (S >= kRopeTile && T % S == 0) {
    dim3 tgrid(ceil_div(S, kRopeTile), nq + 2 * nk, T / S);
    BFLY_CHECK(T / S <= 65535, "rope_kv: too many sequences in one call");
    const size_t smem = (size_t)kRopeTile * (kv.head_dim + 1) * sizeof(float);
    BFLY_CUDA(launch_pdl(rope_kv_tile_kernel, tgrid, dim3(256), smem, stream, qkv, n_partials, partial_stride, bias,
                         rope_table, seq_lens, S, nq, nk, q_out, kv));
    count_launch();
    return BFLY_OK;
  }
  dim3 grid(T, nq + 2 * nk);
  BFLY_CUDA(launch_pdl(rope_kv_kernel, dim3(grid), dim3(kv.head_dim / 2), 0, stream, qkv, n_partials, partial_stride, bias, rope_table, seq_lens,
                                                      S, nq, nk, q_out, kv));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K5: fold gate/up split-K slices, SwiGLU or GELU-tanh (+bias), optional MoE routing weight
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float silu_mul(float g, float u) { return g / (1.0f + expf(-g)) * u; }
__device__ __forceinline__ float gelu_tanh_f(float x) {
  return 0.5f * x * (1.0f + tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x)));
}

__global__ void act_kernel(const float* __restrict__ in, int n_partials, int64_t partial_stride,
                           const bf16* __restrict__ bias, int mode, int n_blocks, int f_block,
                           const float* __restrict__ route_w, bf16* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.y;
  const int f_out = n_blocks * f_block;
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (j >= f_out) return;
  const int blk = j / f_block, jj = j - blk * f_block;
  float4 y;
  if (mode == ACT_SWIGLU) {
    const int ld = 2 * f_out;
    const int64_t off_g = (int64_t)t * ld + (int64_t)blk * 2 * f_block + jj;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f), u = g;
    int p = 0;
    for (; p + 4 <= n_partials; p += 4) {          // index order, four slice pairs in flight
      float4 a[4], b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k] = *reinterpret_cast<const float4*>(in + (p + k) * partial_stride + off_g);
        b[k] = *reinterpret_cast<const float4*>(in + (p + k) * partial_stride + off_g + f_block);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        g.x += a[k].x; g.y += a[k].y; g.z += a[k].z; g.w += a[k].w;
        u.x += b[k].x; u.y += b[k].y; u.z += b[k].z; u.w += b[k].w;
      }
    }
    for (; p < n_partials; ++p) {
      const float4 a = *reinterpret_cast<const float4*>(in + p * partial_stride + off_g);
      const float4 b = *reinterpret_cast<const float4*>(in + p * partial_stride + off_g + f_block);
      g.x += a.x; g.y += a.y; g.z += a.z; g.w += a.w;
      u.x += b.x; u.y += b.y; u.z += b.z; u.w += b.w;
    }
    y = make_float4(silu_mul(g.x, u.x), silu_mul(g.y, u.y), silu_mul(g.z, u.z), silu_mul(g.w, u.w));
  } else {
    const int64_t off = (int64_t)t * f_out + j;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < n_partials; ++p) {
      const float4 a = *reinterpret_cast<const float4*>(in + p * partial_stride + off);
      x.x += a.x; x.y += a.y; x.z += a.z; x.w += a.w;
    }
    if (bias != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(bias + j);
      x.x += bf16_lo(b.x); x.y += bf16_hi(b.x); x.z += bf16_lo(b.y); x.w += bf16_hi(b.y);
    }
    y = make_float4(gelu_tanh_f(x.x), gelu_tanh_f(x.y), gelu_tanh_f(x.z), gelu_tanh_f(x.w));
  }
  if (route_w != nullptr) {
    const float w = route_w[(int64_t)t * n_blocks + blk];
    y.x *= w; y.y *= w; y.z *= w; y.w *= w;
  }
  *reinterpret_cast<uint2*>(out + (int64_t)t * f_out + j) = make_uint2(pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
}

int launch_act(const float* in, int n_partials, int64_t partial_stride, const bf16* bias, int mode,
               int n_blocks, int f_block, const float* route_w, bf16* out, int T, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && n_partials >= 1 && n_blocks >= 1 && f_block % 4 == 0, "act: bad shape (f_block%4==0)");
  const int f_out = n_blocks * f_block;
  dim3 grid(ceil_div(f_out / 4, 256), T);
  BFLY_CUDA(launch_pdl(act_kernel, dim3(grid), dim3(256), 0, stream, in, n_partials, partial_stride, bias, mode, n_blocks, f_block, route_w, out));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K8 (head side): fold LM-head split-K slices into fp32 logits and a per-row greedy key.
// key = (order-preserving bits of the logit) << 32 | (0xFFFFFFFF - global vocab index), so an
// atomicMax is exact and deterministic: highest logit wins, ties go to the lowest index.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
  uint32_t u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)idx);
}

constexpr int LM_CHUNK = 2048;

__global__ void __launch_bounds__(256) lm_finalize_kernel(
    const float* __restrict__ partials, int n_partials, int64_t partial_stride, int v_local, int v0,
    float* __restrict__ logits, int64_t logits_ld, unsigned long long* __restrict__ keys) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.y;
  const int begin = blockIdx.x * LM_CHUNK, end = min(v_local, begin + LM_CHUNK);
  unsigned long long best = 0ull;
  for (int j = begin + threadIdx.x; j < end; j += blockDim.x) {
    float v = 0.f;
    for (int p = 0; p < n_partials; ++p) v += partials[p * partial_stride + (int64_t)t * v_local + j];
    if (logits != nullptr) logits[(int64_t)t * logits_ld + j] = v;
    const unsigned long long k = argmax_key(v, v0 + j);
    best = k > best ? k : best;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
    best = other > best ? other : best;
  }
  __shared__ unsigned long long sbest[8];
  if ((threadIdx.x & 31) == 0) sbest[threadIdx.x >> 5] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) best = sbest[w] > best ? sbest[w] : best;
    atomicMax(keys + t, best);
  }
}

int launch_lm_finalize(const float* partials, int n_partials, int64_t partial_stride, int T, int v_local,
                       int v0, float* logits, int64_t logits_ld, unsigned long long* keys, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && v_local > 0 && n_partials >= 1 && keys != nullptr, "lm_finalize: bad shape");
  dim3 grid(ceil_div(v_local, LM_CHUNK), T);
  BFLY_CUDA(launch_pdl(lm_finalize_kernel, dim3(grid), dim3(256), 0, stream, partials, n_partials, partial_stride, v_local, v0, logits,
                                              logits_ld, keys));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

__global__ void pick_tokens_kernel(const unsigned long long* __restrict__ keys, int n_ranks, int pitch, int T,
                                   int* __restrict__ tokens_out) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  unsigned long long best = 0ull;
  for (int r = 0; r < n_ranks; ++r) {
    const unsigned long long k = keys[(int64_t)r * pitch + t];
    best = k > best ? k : best;
  }
  tokens_out[t] = (int)(0xFFFFFFFFu - (uint32_t)(best & 0xFFFFFFFFull));
}

// keys is [n_ranks][pitch]; row r holds rank r's per-token keys
int launch_pick_tokens_pitched(const unsigned long long* keys, int n_ranks, int pitch, int T, int* tokens_out,
                               cudaStream_t stream) {
  BFLY_CHECK(T > 0 && n_ranks >= 1 && pitch >= T, "pick_tokens: bad shape");
  BFLY_CUDA(launch_pdl(pick_tokens_kernel, dim3(ceil_div(T, 128)), dim3(128), 0, stream, keys, n_ranks, pitch, T, tokens_out));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

__global__ void advance_kernel(int* seq_lens, int B, int S, int* step_counter, unsigned* ar_step) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) seq_lens[b] += S;
  if (b == 0 && step_counter != nullptr) *step_counter += 1;
  if (b == 0 && ar_step != nullptr) *ar_step += 1u;
}

int launch_advance(int* seq_lens, int B, int S, int* step_counter, unsigned* ar_step, cudaStream_t stream) {
  BFLY_CUDA(launch_pdl(advance_kernel, dim3(std::max(1, ceil_div(B, 128))), dim3(128), 0, stream, seq_lens, B, S, step_counter, ar_step));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K9 router: softmax(x . Wr^T) -> top-k -> renormalise; emits the weight of each LOCAL expert per
// token (0 when the token is not routed there), consumed by act_kernel.
// ---------------------------------------------------------------------------------------------
__global__ void moe_router_kernel(const bf16* __restrict__ xn, const bf16* __restrict__ router, int h,
                                  int n_experts, int top_k, int first_local, int n_local,
                                  float* __restrict__ route_w) {
  pdl_trigger();
  pdl_wait();
  __shared__ float logit[32];
  const int t = blockIdx.x, e = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bf16* x = xn + (int64_t)t * h;
  const bf16* w = router + (int64_t)e * h;
  float acc = 0.f;
  for (int i = lane * 8; i < h; i += 32 * 8) {
    const uint4 a = *reinterpret_cast<const uint4*>(x + i);
    const uint4 b = *reinterpret_cast<const uint4*>(w + i);
    acc += bf16_lo(a.x) * bf16_lo(b.x) + bf16_hi(a.x) * bf16_hi(b.x) + bf16_lo(a.y) * bf16_lo(b.y) +
           bf16_hi(a.y) * bf16_hi(b.y) + bf16_lo(a.z) * bf16_lo(b.z) + bf16_hi(a.z) * bf16_hi(b.z) +
           bf16_lo(a.w) * bf16_lo(b.w) + bf16_hi(a.w) * bf16_hi(b.w);
  }
  acc = warp_sum(acc);
  if (lane == 0) logit[e] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = -INFINITY;
    for (int i = 0; i < n_experts; ++i) m = fmaxf(m, logit[i]);
    float p[32];
    for (int i = 0; i < n_experts; ++i) p[i] = expf(logit[i] - m);
    unsigned chosen = 0u;
    float denom = 0.f;
    for (int k = 0; k < top_k; ++k) {
      int best = -1;
      for (int i = 0; i < n_experts; ++i)
        if (!((chosen >> i) & 1u) && (best < 0 || p[i] > p[best])) best = i;
      chosen |= 1u << best;
      denom += p[best];
    }
    for (int j = 0; j < n_local; ++j) {
      const int i = first_local + j;
      route_w[(int64_t)t * n_local + j] = ((chosen >> i) & 1u) ? p[i] / denom : 0.f;
    }
  }
}

int launch_moe_router(const bf16* xn, const bf16* router, int T, int h, int n_experts, int top_k,
                      int first_local, int n_local, float* route_w, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && h % 8 == 0 && n_experts >= 1 && n_experts <= 32 && top_k >= 1 && top_k <= n_experts,
             "moe_router: bad shape");
  BFLY_CUDA(launch_pdl(moe_router_kernel, dim3(T), dim3(32 * n_experts), 0, stream, xn, router, h, n_experts, top_k, first_local, n_local, route_w));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// K9 grouping (prefill): the tokens routed to each LOCAL expert, compacted in token order.  Rows
// [off[e], off[e] + cnt[e]) of idx / rw belong to local expert e; off[n_local] = rows in total
// (<= T * min(top_k, n_local)).  One CTA: T * n_local is at most a few 10^4 flags.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) moe_compact_kernel(const float* __restrict__ route_w, int T, int n_local,
                                                           int* __restrict__ cnt, int* __restrict__ off,
                                                           int* __restrict__ idx, float* __restrict__ rw) {
  pdl_trigger();
  pdl_wait();
  __shared__ int warp_tot[32];
  __shared__ int base_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
  if (threadIdx.x == 0) base_s = 0;
  __syncthreads();
  for (int e = 0; e < n_local; ++e) {
    const int start = base_s;
    if (threadIdx.x == 0) off[e] = start;
    for (int t0 = 0; t0 < T; t0 += blockDim.x) {
      const int t = t0 + (int)threadIdx.x;
      const float w = t < T ? route_w[(int64_t)t * n_local + e] : 0.f;
      const bool on = w > 0.f;
      const unsigned m = __ballot_sync(0xffffffffu, on);
      if (lane == 0) warp_tot[warp] = __popc(m);
      __syncthreads();
      int before = 0, total = 0;
      for (int i = 0; i < n_warps; ++i) {
        const int c = warp_tot[i];
        if (i < warp) before += c;
        total += c;
      }
      const int base = base_s;
      if (on) {
        const int pos = base + before + __popc(m & ((1u << lane) - 1u));
        idx[pos] = t;
        rw[pos] = w;
      }
      __syncthreads();
      if (threadIdx.x == 0) base_s = base + total;
      __syncthreads();
    }
    if (threadIdx.x == 0) cnt[e] = base_s - start;
  }
  if (threadIdx.x == 0) off[n_local] = base_s;
}

// gathers the routed rows of the normed activations into expert order: xg[r] = xn[idx[r]]
__global__ void moe_gather_kernel(const bf16* __restrict__ xn, const int* __restrict__ idx, const int* __restrict__ off,
                                  int n_local, bf16* __restrict__ xg, int h) {
  pdl_trigger();
  pdl_wait();
  const int total = off[n_local];
  for (int r = blockIdx.x; r < total; r += gridDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(xn + (int64_t)idx[r] * h);
    uint4* dst = reinterpret_cast<uint4*>(xg + (int64_t)r * h);
    for (int i = threadIdx.x; i < h / 8; i += blockDim.x) dst[i] = src[i];
  }
}

int launch_moe_group(const float* route_w, const bf16* xn, int T, int h, int n_local, int rows_cap, int* cnt, int* off,
                     int* idx, float* rw, bf16* xg, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && h % 8 == 0 && n_local >= 1 && rows_cap >= T, "moe_group: bad shape");
  BFLY_CUDA(launch_pdl(moe_compact_kernel, dim3(1), dim3(1024), 0, stream, route_w, T, n_local, cnt, off, idx, rw));
  const int grid = std::min(rows_cap, 148 * 8);
  BFLY_CUDA(launch_pdl(moe_gather_kernel, dim3(grid), dim3(128), 0, stream, xn, (const int*)idx, (const int*)off, n_local, xg, h));
  count_launch(2);
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

// ---------------------------------------------------------------------------------------------
// fold split-K slices into one fp32 buffer (only used in front of an NCCL all-reduce, TP > 1)
// ---------------------------------------------------------------------------------------------
__global__ void reduce_partials_kernel(const float* __restrict__ partials, int n_partials, int64_t partial_stride,
                                       float* __restrict__ out, int64_t count4) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= count4) return;
  float4 v = reinterpret_cast<const float4*>(partials)[i];
  for (int p = 1; p < n_partials; ++p) {
    const float4 a = reinterpret_cast<const float4*>(partials + p * partial_stride)[i];
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
  }
  reinterpret_cast<float4*>(out)[i] = v;
}

// same fold, rounded once to bf16 (the payload of the prefill all-reduce at tp > 1)
__global__ void reduce_partials_bf16_kernel(const float* __restrict__ partials, int n_partials, int64_t partial_stride,
                                            bf16* __restrict__ out, int64_t count4) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= count4) return;
  float4 v = reinterpret_cast<const float4*>(partials)[i];
  for (int p = 1; p < n_partials; ++p) {
    const float4 a = reinterpret_cast<const float4*>(partials + p * partial_stride)[i];
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
  }
  reinterpret_cast<uint2*>(out)[i] = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
}

int launch_reduce_partials_bf16(const float* partials, int n_partials, int64_t partial_stride, bf16* out, int64_t count,
                                cudaStream_t stream) {
  BFLY_CHECK(count > 0 && count % 4 == 0 && n_partials >= 1 && partial_stride % 4 == 0, "reduce_partials_bf16: bad shape");
  const int64_t c4 = count / 4;
  BFLY_CUDA(launch_pdl(reduce_partials_bf16_kernel, dim3((unsigned)ceil_div64(c4, 256)), dim3(256), 0, stream, partials,
                       n_partials, partial_stride, out, c4));
  count_launch();
  return BFLY_OK;
}

int launch_reduce_partials(const float* partials, int n_partials, int64_t partial_stride, float* out,
                           int64_t count, cudaStream_t stream) {
  BFLY_CHECK(count > 0 && count % 4 == 0 && n_partials >= 1 && partial_stride % 4 == 0, "reduce_partials: bad shape");
  const int64_t c4 = count / 4;
  BFLY_CUDA(launch_pdl(reduce_partials_kernel, dim3((unsigned)ceil_div64(c4, 256)), dim3(256), 0, stream, partials, n_partials, partial_stride, out, c4));
  count_launch();
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

}  // namespace bfly
