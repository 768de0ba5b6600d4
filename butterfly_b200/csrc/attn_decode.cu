// K6: paged-KV attention for one query token per (virtual) sequence, GQA-aware, split over the
// context.  HBM-bound: every K / V byte is read exactly once.
//
// Data movement: a 32-token tile of one KV head is two contiguous 32*D*2-byte blocks in the paged
// cache (K rows, and V stored TRANSPOSED [D][32]); each warp streams its tiles with two bulk async
// copies (cp.async.bulk -> UBLKCP, mbarrier complete_tx) into a private 3-stage shared-memory ring,
// so 4 warps keep up to 192 KB in flight per SM with no registers tied up.
//
// Math: tensor cores through mma.sync m16n8k16 (bf16 in, fp32 accumulate) in the transposed
// formulation  S^T[token, head] = K[token, :] . Q[head, :]^T  and  O^T[d, head] = V^T[d, token] . P^T,
// i.e. the cache tiles are always the row-major A operand and the <= 8 query heads of the GQA group
// are the N = 8 columns, so no accumulator rows are padding.  The contraction indices (d for QK,
// token for PV) are permuted so every thread reads 16 contiguous bytes per fragment; K rows carry
// a 16-byte-chunk XOR swizzle (chunk ^ 4*(token&1)), applied when the cache is written, that makes
// those reads bank-conflict free.  P goes from the S^T accumulator layout to the B-operand layout
// with movmatrix.  Online softmax in the exp2 domain, fp32.
//
// Work split: grid (split, kv head, token); the 4 warps of a CTA interleave over the CTA's tiles,
// are merged through shared memory, and splits are merged by attn_combine_kernel through the fp32
// scratch part_o / part_ml.
#include "attn_common.cuh"

namespace bfly {

namespace {

constexpr int kWarps = 4;
constexpr int kTile = 32;          // tokens per tile
constexpr int kStages = 3;

template <int D>
constexpr size_t attn_smem_bytes() {
  return (size_t)kWarps * kStages * 2 * kTile * D * 2      // K + V^T rings
         + (size_t)kWarps * 8 * D * 4                      // per-warp O^T for the merge
         + (size_t)kWarps * 8 * 2 * 4                      // per-warp (m, l)
         + (size_t)kWarps * kStages * 8 + 128              // mbarriers + alignment slack
         + (size_t)8 * D * 2;                              // fused variant: rotated q of the GQA group (bf16)
}

// Fused variant (kFused, decode only: S == 1): the kernel also does what rope_kv_kernel did -- it folds the QKV
// GEMM's split-K slices (+bias) for its own query heads, rotates them (RoPE) and keeps them in shared memory, and
// the CTA that owns the tile of the NEW token folds/rotates that token's K and V row, appends it to the paged
// cache and only then fetches that tile.  One kernel and one dependency less per layer.
struct FusedQkv {
  const float* qkv;          // fp32 [n_partials][T][(nq + 2 nk) * D] slices of the QKV projection
  int n_partials;
  int64_t partial_stride;
  const bf16* bias;          // [(nq + 2 nk) * D] or nullptr
  const float2* rope_table;  // [pos][D/2] (cos, sin) or nullptr (learned positions)
};

template <int D, int G, bool kFused>
__global__ void __launch_bounds__(kWarps * 32, 1) attn_decode_kernel(
    const bf16* __restrict__ q, KvLayout kv, const int* __restrict__ seq_lens, int S, int nq,
    float* __restrict__ part_o, float* __restrict__ part_ml, int n_splits, float scale_log2,
    bf16* __restrict__ out_direct, FusedQkv fq) {
  static_assert(G >= 1 && G <= 8 && (D == 32 || D == 64 || D == 128), "unsupported shape");
  constexpr int KSTEP_PAIRS = D / 32;         // 16-byte chunks per thread per K row = pairs of k-steps
  constexpr int MT_O = D / 16;                // m-tiles of O^T
  constexpr uint32_t kBlockBytes = kTile * D * 2;
  constexpr bool kSwizzle = D >= 64;

  extern __shared__ __align__(128) uint8_t smem[];
  pdl_trigger();
  // Before the dependency on the predecessor (the RoPE / KV-append kernel) resolves, only state that
  // older kernels finished long ago is touched: seq_lens (last written by the previous step) and cache
  // tiles that do not contain the token being appended right now.
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane >> 2, c = lane & 3;
  const int split = blockIdx.x, kvh = blockIdx.y, t = blockIdx.z;
  const int b = t / S;
  const int ctx = seq_lens[b] + (t - b * S) + 1;
  const int n_tiles = (ctx + kTile - 1) / kTile;
  const int tiles_per_split = (n_tiles + n_splits - 1) / n_splits;
  const int tile_begin = split * tiles_per_split;
  const int tile_end = min(n_tiles, tile_begin + tiles_per_split);
  const int my_tiles = (tile_end - tile_begin - warp + kWarps - 1) / kWarps;    // may be <= 0

  uint8_t* ring = smem;                                                          // [warp][stage][K | V^T]
  float* sm_o = reinterpret_cast<float*>(smem + (size_t)kWarps * kStages * 2 * kBlockBytes);   // [warp][8][D]
  float* sm_ml = sm_o + kWarps * 8 * D;                                          // [warp][8][2]
  const uint32_t bars = smem_u32(sm_ml + kWarps * 8 * 2);                        // [warp][stage]
  const uint32_t my_ring = smem_u32(ring) + (uint32_t)warp * kStages * 2 * kBlockBytes;
  const uint32_t my_bars = bars + (uint32_t)warp * kStages * 8;

  if (lane == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(my_bars + 8u * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  const int* table = kv.page_table + (int64_t)b * kv.pages_per_seq;
  const int ps = kv.page_size;
  const int64_t head_stride = (int64_t)ps * D;                          // elements per (page, K|V, head) block
  const int64_t page_stride = 2 * (int64_t)kv.kv_local * head_stride;
  auto issue = [&](int i) {                                             // lane 0 only
    const int tile = tile_begin + warp + i * kWarps;
    const int tok0 = tile * kTile;
    const int page = table[tok0 / ps];
    const int sub = (tok0 % ps) / kTile;                                // 32-token sub-block inside the page
    const bf16* kp = kv.base + (int64_t)page * page_stride + (int64_t)kvh * head_stride + (int64_t)sub * kTile * D;
    const bf16* vp = kp + (int64_t)kv.kv_local * head_stride;
    const int s = i % kStages;
    const uint32_t bar = my_bars + 8u * s;
    const uint32_t dst = my_ring + (uint32_t)s * 2 * kBlockBytes;
    mbar_expect_tx(bar, 2 * kBlockBytes);
    bulk_copy_g2s(dst, kp, kBlockBytes, bar);
    bulk_copy_g2s(dst + kBlockBytes, vp, kBlockBytes, bar);
  };
  {
    // S == 1: only the last tile holds K/V appended by the predecessor kernel.  S > 1: the S new tokens of a
    // sequence may sit in earlier tiles too, so nothing is fetched ahead of the dependency.
    const int last_tile = n_tiles - 1;
    const int n_pre = min(kStages, my_tiles);
    if (lane == 0 && S == 1)
      for (int i = 0; i < n_pre; ++i)
        if (tile_begin + warp + i * kWarps != last_tile) issue(i);
    pdl_wait();                                        // q and the new K/V rows are ready from here on
    if constexpr (kFused) {
      // ---- what rope_kv_kernel did, for this CTA's heads only (S == 1: token t sits at position ctx - 1) ----
      bf16* sm_q = reinterpret_cast<bf16*>(smem + attn_smem_bytes<D>() - (size_t)8 * D * 2);
      constexpr int half = D / 2;
      const int nk = kv.kv_local, pos = ctx - 1;
      const int ld = (nq + 2 * nk) * D;
      const float* row = fq.qkv + (int64_t)t * ld;
      auto fold_pair = [&](int c0, float& x0, float& x1) {        // columns c0 and c0 + half, slices in index order
        x0 = 0.f; x1 = 0.f;
        int p = 0;
        for (; p + 4 <= fq.n_partials; p += 4) {
          float a[4], b[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            a[k] = row[(p + k) * fq.partial_stride + c0];
            b[k] = row[(p + k) * fq.partial_stride + c0 + half];
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) { x0 += a[k]; x1 += b[k]; }
        }
        for (; p < fq.n_partials; ++p) {
          x0 += row[p * fq.partial_stride + c0];
          x1 += row[p * fq.partial_stride + c0 + half];
        }
        if (fq.bias != nullptr) {
          x0 += __bfloat162float(fq.bias[c0]);
          x1 += __bfloat162float(fq.bias[c0 + half]);
        }
      };
      auto rotate = [&](int i, float& x0, float& x1) {
        if (fq.rope_table != nullptr) {
          const float2 cs = fq.rope_table[(int64_t)pos * half + i];
          const float y0 = x0 * cs.x - x1 * cs.y, y1 = x1 * cs.x + x0 * cs.y;
          x0 = y0; x1 = y1;
        }
      };
      for (int idx = threadIdx.x; idx < G * half; idx += kWarps * 32) {
        const int g = idx / half, i = idx - g * half;
        float x0, x1;
        fold_pair((kvh * G + g) * D + i, x0, x1);
        rotate(i, x0, x1);
        sm_q[g * D + i] = __float2bfloat16_rn(x0);
        sm_q[g * D + i + half] = __float2bfloat16_rn(x1);
      }
      const bool owns_new = last_tile >= tile_begin && last_tile < tile_end;      // CTA-uniform
      if (owns_new) {
        const int page = table[pos / ps];
        const int slot = pos % ps;
        bf16* kblk = kv.base + (int64_t)page * page_stride + (int64_t)kvh * head_stride;
        bf16* vblk = kblk + (int64_t)kv.kv_local * head_stride;
        for (int idx = threadIdx.x; idx < 2 * half; idx += kWarps * 32) {
          const int which = idx / half, i = idx - which * half;              // 0: K pair (rotated), 1: V pair
          float x0, x1;
          fold_pair((nq + which * nk + kvh) * D + i, x0, x1);
          if (which == 0) {
            rotate(i, x0, x1);
            kblk[kv_k_offset(slot, i, D)] = __float2bfloat16_rn(x0);
            kblk[kv_k_offset(slot, i + half, D)] = __float2bfloat16_rn(x1);
          } else {
            vblk[kv_v_offset(slot, i, D)] = __float2bfloat16_rn(x0);
            vblk[kv_v_offset(slot, i + half, D)] = __float2bfloat16_rn(x1);
          }
        }
        asm volatile("fence.proxy.async;" ::: "memory");     // the bulk copy below (async proxy) must see these rows
      }
      __syncthreads();
    }
    if (lane == 0)
      for (int i = 0; i < n_pre; ++i)
        if (S != 1 || tile_begin + warp + i * kWarps == last_tile) issue(i);
  }

  // Q as B fragments: head r (< G), the same d-permutation as the K chunks (chunk c of every 64 bytes)
  uint32_t qb[KSTEP_PAIRS][4];
#pragma unroll
  for (int j = 0; j < KSTEP_PAIRS; ++j) {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if constexpr (kFused) {
      const bf16* sm_q = reinterpret_cast<const bf16*>(smem + attn_smem_bytes<D>() - (size_t)8 * D * 2);
      if (r < G) v = *reinterpret_cast<const uint4*>(sm_q + r * D + c * 8 + 32 * j);
    } else {
      if (r < G) v = *reinterpret_cast<const uint4*>(q + ((int64_t)t * nq + kvh * G + r) * D + c * 8 + 32 * j);
    }
    qb[j][0] = v.x; qb[j][1] = v.y; qb[j][2] = v.z; qb[j][3] = v.w;
  }

  float oacc[MT_O][4];
#pragma unroll
  for (int mo = 0; mo < MT_O; ++mo)
#pragma unroll
    for (int e = 0; e < 4; ++e) oacc[mo][e] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};      // heads 2c, 2c+1

  for (int i = 0; i < my_tiles; ++i) {
    const int s = i % kStages;
    mbar_wait(my_bars + 8u * s, (uint32_t)(i / kStages) & 1u);
    const uint32_t k_s = my_ring + (uint32_t)s * 2 * kBlockBytes;
    const uint32_t v_s = k_s + kBlockBytes;
    const int tok0 = (tile_begin + warp + i * kWarps) * kTile;

    // ---- S^T = K . Q^T : 2 m-tiles of 16 token rows; row r of m-tile mt is token 8*(r>>1)+4*mt+(r&1) (+2 for r+8)
    float sacc[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) sacc[mt][e] = 0.f;
      const int ta = 8 * (r >> 1) + 4 * mt + (r & 1), tb = ta + 2;
      const uint32_t swz = kSwizzle ? (uint32_t)((ta & 1) << 2) : 0u;   // ta and tb have the same parity
#pragma unroll
      for (int j = 0; j < KSTEP_PAIRS; ++j) {
        const uint32_t chunk = ((uint32_t)(c + 4 * j)) ^ swz;
        const uint4 x = lds128(k_s + (uint32_t)ta * (D * 2) + chunk * 16);
        const uint4 y = lds128(k_s + (uint32_t)tb * (D * 2) + chunk * 16);
        mma_bf16(sacc[mt], x.x, y.x, x.y, y.y, qb[j][0], qb[j][1]);
        mma_bf16(sacc[mt], x.z, y.z, x.w, y.w, qb[j][2], qb[j][3]);
      }
    }
    // ---- online softmax over this tile (columns = heads 2c, 2c+1)
    float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int ta = tok0 + 8 * (r >> 1) + 4 * mt + (r & 1);
      const bool va = ta < ctx, vb = (ta + 2) < ctx;
      sacc[mt][0] = va ? sacc[mt][0] * scale_log2 : -INFINITY;
      sacc[mt][1] = va ? sacc[mt][1] * scale_log2 : -INFINITY;
      sacc[mt][2] = vb ? sacc[mt][2] * scale_log2 : -INFINITY;
      sacc[mt][3] = vb ? sacc[mt][3] * scale_log2 : -INFINITY;
      tmax[0] = fmaxf(tmax[0], fmaxf(sacc[mt][0], sacc[mt][2]));
      tmax[1] = fmaxf(tmax[1], fmaxf(sacc[mt][1], sacc[mt][3]));
    }
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      tmax[0] = fmaxf(tmax[0], __shfl_xor_sync(0xffffffffu, tmax[0], off));
      tmax[1] = fmaxf(tmax[1], __shfl_xor_sync(0xffffffffu, tmax[1], off));
    }
    float corr[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float m_new = fmaxf(m_run[e], tmax[e]);     // finite: every tile in range has a valid token
      corr[e] = exp2f(m_run[e] - m_new);
      m_run[e] = m_new;
    }
    uint32_t pb[2][2];
    float psum[2] = {0.f, 0.f};
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const float p0 = exp2f(sacc[mt][0] - m_run[0]), p1 = exp2f(sacc[mt][1] - m_run[1]);
      const float p2 = exp2f(sacc[mt][2] - m_run[0]), p3 = exp2f(sacc[mt][3] - m_run[1]);
      psum[0] += p0 + p2;
      psum[1] += p1 + p3;
      pb[mt][0] = movmatrix_trans(pack_bf16(p0, p1));   // rows r      -> B k-slots 2c,2c+1   of k-step mt
      pb[mt][1] = movmatrix_trans(pack_bf16(p2, p3));   // rows r + 8  -> B k-slots 2c+8,2c+9 of k-step mt
    }
    l_run[0] = l_run[0] * corr[0] + psum[0];
    l_run[1] = l_run[1] * corr[1] + psum[1];
    // ---- O^T += V^T . P^T : m-tiles over d, 2 k-steps over the 32 tokens
#pragma unroll
    for (int mo = 0; mo < MT_O; ++mo) {
      oacc[mo][0] *= corr[0]; oacc[mo][1] *= corr[1]; oacc[mo][2] *= corr[0]; oacc[mo][3] *= corr[1];
      const uint4 x = lds128(v_s + (uint32_t)(16 * mo + r) * (kTile * 2) + c * 16);
      const uint4 y = lds128(v_s + (uint32_t)(16 * mo + r + 8) * (kTile * 2) + c * 16);
      mma_bf16(oacc[mo], x.x, y.x, x.y, y.y, pb[0][0], pb[0][1]);
      mma_bf16(oacc[mo], x.z, y.z, x.w, y.w, pb[1][0], pb[1][1]);
    }
    __syncwarp();                                        // every lane is done reading stage s
    if (lane == 0 && i + kStages < my_tiles) issue(i + kStages);
  }

  // ---- merge the CTA's warps ----
#pragma unroll
  for (int off = 4; off < 32; off <<= 1) {
    l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
    l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
  }
  if (r == 0) {
    sm_ml[(warp * 8 + 2 * c) * 2] = m_run[0];     sm_ml[(warp * 8 + 2 * c) * 2 + 1] = l_run[0];
    sm_ml[(warp * 8 + 2 * c + 1) * 2] = m_run[1]; sm_ml[(warp * 8 + 2 * c + 1) * 2 + 1] = l_run[1];
  }
#pragma unroll
  for (int mo = 0; mo < MT_O; ++mo) {
    float* o0 = sm_o + ((size_t)warp * 8 + 2 * c) * D + 16 * mo + r;
    o0[0] = oacc[mo][0]; o0[D] = oacc[mo][1]; o0[8] = oacc[mo][2]; o0[D + 8] = oacc[mo][3];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * D; idx += kWarps * 32) {
    const int g = idx / D, dd = idx - g * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_ml[(w * 8 + g) * 2]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < kWarps; ++w) {
        const float mw = sm_ml[(w * 8 + g) * 2];
        if (mw > -INFINITY) {
          const float f = exp2f(mw - M);
          L += sm_ml[(w * 8 + g) * 2 + 1] * f;
          O += sm_o[((size_t)w * 8 + g) * D + dd] * f;
        }
      }
    }
    if (out_direct != nullptr) {          // single split: this CTA saw the whole context
      out_direct[((int64_t)t * nq + kvh * G + g) * D + dd] = __float2bfloat16_rn(O / L);
    } else {
      const int64_t slot = ((int64_t)t * nq + kvh * G + g) * n_splits + split;
      part_o[slot * D + dd] = O;
      if (dd == 0) { part_ml[slot * 2] = M; part_ml[slot * 2 + 1] = L; }
    }
  }
}

template <int D>
__global__ void attn_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                    int n_splits, bf16* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
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

template <int D, int G, bool kFused>
int launch_impl(const bf16* q, const KvLayout& kv, const int* seq_lens, int S, int T, int nq, float* part_o,
                float* part_ml, int n_splits, bf16* out, const FusedQkv& fq, cudaStream_t stream) {
  static PerDeviceOnce attr;
  constexpr size_t smem = attn_smem_bytes<D>();
  BFLY_TRY(attr.run([] {
    BFLY_CUDA(cudaFuncSetAttribute(attn_decode_kernel<D, G, kFused>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return (int)BFLY_OK;
  }));
  dim3 grid(n_splits, kv.kv_local, T);
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)D);
  BFLY_CUDA(launch_pdl(attn_decode_kernel<D, G, kFused>, grid, dim3(kWarps * 32), smem, stream, q, kv, seq_lens, S, nq, part_o,
                       part_ml, n_splits, scale_log2, n_splits == 1 ? out : nullptr, fq));
  count_launch();
  if (n_splits > 1) {
    BFLY_CUDA(launch_pdl(attn_combine_kernel<D>, dim3(T * nq), dim3(D), 0, stream, part_o, part_ml, n_splits, out));
    count_launch();
  }
  return BFLY_OK;
}

template <bool kFused>
int dispatch(const bf16* q, const KvLayout& kv, const int* seq_lens, int S, int T, int nq, float* part_o, float* part_ml,
             int n_splits, bf16* out, const FusedQkv& fq, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && T <= 65535 && n_splits >= 1 && kv.kv_local >= 1 && nq % kv.kv_local == 0,
             "attn_decode: bad shape (T<=65535, nq % kv_local == 0)");
  BFLY_CHECK(kv.page_size % kTile == 0, "attn_decode: page_size must be a multiple of 32");
  const int G = nq / kv.kv_local, D = kv.head_dim;
#define BFLY_ATTN_CASE(DD, GG) \
  if (D == DD && G == GG) return launch_impl<DD, GG, kFused>(q, kv, seq_lens, S, T, nq, part_o, part_ml, n_splits, out, fq, stream)
  BFLY_ATTN_CASE(128, 1); BFLY_ATTN_CASE(128, 2); BFLY_ATTN_CASE(128, 4); BFLY_ATTN_CASE(128, 8);
  BFLY_ATTN_CASE(64, 1);  BFLY_ATTN_CASE(64, 2);  BFLY_ATTN_CASE(64, 4);  BFLY_ATTN_CASE(64, 8);
  BFLY_ATTN_CASE(32, 1);  BFLY_ATTN_CASE(32, 2);  BFLY_ATTN_CASE(32, 4);  BFLY_ATTN_CASE(32, 8);
#undef BFLY_ATTN_CASE
  set_error("attn_decode: unsupported (head_dim, q-per-kv) = (" + std::to_string(D) + ", " + std::to_string(G) +
            "); supported head_dim {32,64,128}, group {1,2,4,8}");
  return BFLY_EINVAL;
}

}  // namespace

// Choose the context split so that (token x kv-head x split) CTAs fill whole waves of SMs (one CTA per
// SM: the rings take ~200 KB of shared memory) while every CTA keeps at least a few tiles per warp.
int attn_decode_splits(int T, int kv_local, int max_ctx, int sm_count) {
  (void)sm_count;
  const int base = std::max(1, T * kv_local);
  const int tiles = std::max(1, (max_ctx + kTile - 1) / kTile);
  // Measured on B200 (tools/tune_kernels.py attn, B=32 ctx=2048): 256 or 128 (sequence, kv head) CTAs run
  // fastest unsplit (79-81 % of HBM peak), 64 want 2 splits, 32 want 3 -- every split adds a ring
  // ramp-up and a slice for the combine kernel, so split only until ~96 CTAs exist.
  int n = 1;
  while (base * n < 96 && n < 32 && tiles / (n + 1) >= 2 * kWarps) ++n;
  return n;
}

int launch_attn_decode(const bf16* q, KvLayout kv, const int* seq_lens, int S, int T, int nq, float* part_o,
                       float* part_ml, int n_splits, bf16* out, cudaStream_t stream) {
  return dispatch<false>(q, kv, seq_lens, S, T, nq, part_o, part_ml, n_splits, out, FusedQkv{}, stream);
}

// Decode step (one new token per sequence): RoPE + KV append + attention in one kernel, straight from the QKV
// projection's fp32 slices.
int launch_attn_decode_fused(const float* qkv, int n_partials, int64_t partial_stride, const bf16* bias,
                             const float2* rope_table, KvLayout kv, const int* seq_lens, int T, int nq, float* part_o,
                             float* part_ml, int n_splits, bf16* out, cudaStream_t stream) {
  BFLY_CHECK(qkv != nullptr && n_partials >= 1, "attn_decode_fused: qkv slices missing");
  FusedQkv fq{qkv, n_partials, partial_stride, bias, rope_table};
  return dispatch<true>(nullptr, kv, seq_lens, 1, T, nq, part_o, part_ml, n_splits, out, fq, stream);
}

}  // namespace bfly
