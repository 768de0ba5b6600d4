// K6: paged-KV attention for one query token per (virtual) sequence, GQA-aware, split over the
// context.  HBM-bound: K and V rows are streamed exactly once with
// coalesced row-contiguous loads (a warp reads one 2*D-byte row per instruction, TB rows in flight),
// all math in fp32 with an online softmax in the exp2 domain.
//
// Work split: grid (split, kv head, token).  A CTA's 4 warps interleave over batches of TB = 32/G
// cached tokens; the TB*G = 32 partial dot products a lane accumulates are folded with a 31-shuffle
// transpose-reduction so lane L ends up owning the score of (token L/G, head L%G).  Warps are merged
// through shared memory, splits through a small fp32 scratch (part_o / part_ml) and a combine kernel.
// Also used for prefill (token t of sequence b sees seq_lens[b] + s + 1 keys) until a tiled
// tensor-core prefill kernel replaces it.
#include "kernels.cuh"

namespace bfly {

namespace {

template <int N>
struct TransposeReduce {
  // v[0..N) per lane -> after log2(N) exchange steps v[0] of lane L holds sum over lanes of v[L's index]
  static __device__ __forceinline__ void run(float* v, int lane) {
    constexpr int H = N / 2;
    const bool upper = (lane & H) != 0;
#pragma unroll
    for (int i = 0; i < H; ++i) {
      const float send = upper ? v[i] : v[i + H];
      const float keep = upper ? v[i + H] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, H);
    }
    TransposeReduce<H>::run(v, lane);
  }
};
template <>
struct TransposeReduce<1> {
  static __device__ __forceinline__ void run(float*, int) {}
};

template <int VEC> struct RawVec;
template <> struct RawVec<1> { uint16_t r; };
template <> struct RawVec<2> { uint32_t r; };
template <> struct RawVec<4> { uint2 r; };

template <int VEC>
__device__ __forceinline__ void load_row(const bf16* p, RawVec<VEC>& raw);
template <> __device__ __forceinline__ void load_row<1>(const bf16* p, RawVec<1>& raw) { raw.r = ld_stream_u16(p); }
template <> __device__ __forceinline__ void load_row<2>(const bf16* p, RawVec<2>& raw) { raw.r = ld_stream_u32(p); }
template <> __device__ __forceinline__ void load_row<4>(const bf16* p, RawVec<4>& raw) { raw.r = ld_stream_v2(p); }

template <int VEC>
__device__ __forceinline__ void unpack(const RawVec<VEC>& raw, float* f);
template <> __device__ __forceinline__ void unpack<1>(const RawVec<1>& raw, float* f) {
  f[0] = __uint_as_float((uint32_t)raw.r << 16);
}
template <> __device__ __forceinline__ void unpack<2>(const RawVec<2>& raw, float* f) {
  f[0] = bf16_lo(raw.r); f[1] = bf16_hi(raw.r);
}
template <> __device__ __forceinline__ void unpack<4>(const RawVec<4>& raw, float* f) {
  f[0] = bf16_lo(raw.r.x); f[1] = bf16_hi(raw.r.x); f[2] = bf16_lo(raw.r.y); f[3] = bf16_hi(raw.r.y);
}

constexpr int kWarps = 4;

template <int D, int G>
__global__ void __launch_bounds__(kWarps * 32) attn_decode_kernel(
    const bf16* __restrict__ q, KvLayout kv, const int* __restrict__ seq_lens, int S, int nq,
    float* __restrict__ part_o, float* __restrict__ part_ml, int n_splits, float scale_log2) {
  constexpr int VEC = D / 32;
  constexpr int TB = 32 / G;
  const int split = blockIdx.x, kvh = blockIdx.y, t = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = t / S;
  const int ctx = seq_lens[b] + (t - b * S) + 1;
  const int ps = kv.page_size;
  const int n_pages = (ctx + ps - 1) / ps;
  const int pages_per_split = (n_pages + n_splits - 1) / n_splits;
  const int tok_begin = split * pages_per_split * ps;
  const int tok_end = min(ctx, tok_begin + pages_per_split * ps);
  const int* table = kv.page_table + (int64_t)b * kv.pages_per_seq;
  const int64_t head_stride = (int64_t)ps * D;                       // one (page, K|V, head) block
  const int64_t page_stride = 2 * (int64_t)kv.kv_local * head_stride;

  float qr[G][VEC];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    RawVec<VEC> raw;
    load_row<VEC>(q + ((int64_t)t * nq + kvh * G + g) * D + lane * VEC, raw);
    unpack<VEC>(raw, qr[g]);
#pragma unroll
    for (int v = 0; v < VEC; ++v) qr[g][v] *= scale_log2;
  }
  float o[G][VEC];
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int v = 0; v < VEC; ++v) o[g][v] = 0.f;
  float m = -INFINITY, l = 0.f;

  for (int base = tok_begin + warp * TB; base < tok_end; base += kWarps * TB) {
    RawVec<VEC> kraw[TB], vraw[TB];
#pragma unroll
    for (int j = 0; j < TB; ++j) {
      const int tok = base + j;
      if (tok < tok_end) {
        const int page = table[tok / ps];
        const bf16* kp = kv.base + (int64_t)page * page_stride + (int64_t)kvh * head_stride +
                         (int64_t)(tok % ps) * D + lane * VEC;
        load_row<VEC>(kp, kraw[j]);
        load_row<VEC>(kp + (int64_t)kv.kv_local * head_stride, vraw[j]);
      } else {
        kraw[j] = RawVec<VEC>{};
        vraw[j] = RawVec<VEC>{};
      }
    }
    float acc[32];
#pragma unroll
    for (int j = 0; j < TB; ++j) {
      float kf[VEC];
      unpack<VEC>(kraw[j], kf);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float a = 0.f;
#pragma unroll
        for (int v = 0; v < VEC; ++v) a += qr[g][v] * kf[v];
        acc[j * G + g] = a;
      }
    }
    TransposeReduce<32>::run(acc, lane);
    const bool mine_valid = (base + lane / G) < tok_end;
    const float s = mine_valid ? acc[0] : -INFINITY;
    float mx = s;
#pragma unroll
    for (int off = G; off < 32; off <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    const float m_new = fmaxf(m, mx);                     // finite: token `base` is valid for every head
    const float corr = exp2f(m - m_new);                  // m = -inf on the first batch -> 0
    const float p = mine_valid ? exp2f(s - m_new) : 0.f;
    l = l * corr + p;
    m = m_new;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float cg = __shfl_sync(0xffffffffu, corr, g);   // lane g owns head g
#pragma unroll
      for (int v = 0; v < VEC; ++v) o[g][v] *= cg;
    }
#pragma unroll
    for (int j = 0; j < TB; ++j) {
      float vf[VEC];
      unpack<VEC>(vraw[j], vf);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float pj = __shfl_sync(0xffffffffu, p, j * G + g);
#pragma unroll
        for (int v = 0; v < VEC; ++v) o[g][v] += pj * vf[v];
      }
    }
  }
  // per-head l over the lanes that share the head
#pragma unroll
  for (int off = G; off < 32; off <<= 1) l += __shfl_xor_sync(0xffffffffu, l, off);

  __shared__ float sm_m[kWarps][G], sm_l[kWarps][G];
  __shared__ float sm_o[kWarps][G][D];
  if (lane < G) { sm_m[warp][lane] = m; sm_l[warp][lane] = l; }
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int v = 0; v < VEC; ++v) sm_o[warp][g][lane * VEC + v] = o[g][v];
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * D; idx += kWarps * 32) {
    const int g = idx / D, dd = idx - g * D;
    float M = sm_m[0][g];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) M = fmaxf(M, sm_m[w][g]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < kWarps; ++w) {
        const float f = exp2f(sm_m[w][g] - M);
        L += sm_l[w][g] * f;
        O += sm_o[w][g][dd] * f;
      }
    }
    const int64_t slot = ((int64_t)t * nq + kvh * G + g) * n_splits + split;
    part_o[slot * D + dd] = O;
    if (dd == 0) { part_ml[slot * 2] = M; part_ml[slot * 2 + 1] = L; }
  }
}

template <int D>
__global__ void attn_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                    int n_splits, bf16* __restrict__ out) {
  const int64_t th = blockIdx.x;                 // t * nq + head
  const int dd = threadIdx.x;
  float M = -INFINITY;
  for (int s = 0; s < n_splits; ++s) M = fmaxf(M, part_ml[(th * n_splits + s) * 2]);
  float L = 0.f, O = 0.f;
  for (int s = 0; s < n_splits; ++s) {
    const float ms = part_ml[(th * n_splits + s) * 2];
    if (ms > -INFINITY) {
      const float f = exp2f(ms - M);
      L += part_ml[(th * n_splits + s) * 2 + 1] * f;
      O += part_o[(th * n_splits + s) * D + dd] * f;
    }
  }
  out[th * D + dd] = __float2bfloat16_rn(O / L);
}

template <int D, int G>
int launch_impl(const bf16* q, const KvLayout& kv, const int* seq_lens, int S, int T, int nq, float* part_o,
                float* part_ml, int n_splits, bf16* out, cudaStream_t stream) {
  dim3 grid(n_splits, kv.kv_local, T);
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)D);
  attn_decode_kernel<D, G><<<grid, kWarps * 32, 0, stream>>>(q, kv, seq_lens, S, nq, part_o, part_ml, n_splits,
                                                            scale_log2);
  BFLY_CUDA(cudaGetLastError());
  attn_combine_kernel<D><<<T * nq, D, 0, stream>>>(part_o, part_ml, n_splits, out);
  BFLY_CUDA(cudaGetLastError());
  count_launch(2);
  return BFLY_OK;
}

}  // namespace

int attn_decode_splits(int T, int kv_local, int max_ctx, int sm_count) {
  const int want = ceil_div(sm_count * 8, std::max(1, T * kv_local));
  const int cap = std::max(1, max_ctx / 64);
  return std::max(1, std::min(std::min(want, cap), 64));
}

int launch_attn_decode(const bf16* q, KvLayout kv, const int* seq_lens, int S, int T, int nq, float* part_o,
                       float* part_ml, int n_splits, bf16* out, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && T <= 65535 && n_splits >= 1 && kv.kv_local >= 1 && nq % kv.kv_local == 0,
             "attn_decode: bad shape (T<=65535, nq % kv_local == 0)");
  const int G = nq / kv.kv_local, D = kv.head_dim;
#define BFLY_ATTN_CASE(DD, GG) \
  if (D == DD && G == GG) return launch_impl<DD, GG>(q, kv, seq_lens, S, T, nq, part_o, part_ml, n_splits, out, stream)
  BFLY_ATTN_CASE(128, 1); BFLY_ATTN_CASE(128, 2); BFLY_ATTN_CASE(128, 4); BFLY_ATTN_CASE(128, 8);
  BFLY_ATTN_CASE(64, 1);  BFLY_ATTN_CASE(64, 2);  BFLY_ATTN_CASE(64, 4);  BFLY_ATTN_CASE(64, 8);
  BFLY_ATTN_CASE(32, 1);  BFLY_ATTN_CASE(32, 2);  BFLY_ATTN_CASE(32, 4);  BFLY_ATTN_CASE(32, 8);
#undef BFLY_ATTN_CASE
  set_error("attn_decode: unsupported (head_dim, q-per-kv) = (" + std::to_string(D) + ", " + std::to_string(G) +
            "); supported head_dim {32,64,128}, group {1,2,4,8}");
  return BFLY_EINVAL;
}

}  // namespace bfly
