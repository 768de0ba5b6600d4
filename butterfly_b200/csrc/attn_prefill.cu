// K7: causal prefill attention over the paged KV cache, tensor cores.
//
// Same transposed formulation and cache layouts as the decode kernel (attn_decode.cu):
//     S^T[key, col] = K[key, :] . Q[col, :]^T        O^T[d, col] = V^T[d, key] . P^T[key, col]
// but the N columns are now (query token, head-of-the-GQA-group) pairs: a CTA owns 64 columns =
// 64/G consecutive query tokens of one sequence and one KV head, its 4 warps take 16 columns each, and
// all warps share ONE shared-memory ring of K / V^T tiles filled by bulk async copies (one elected
// producer thread, full/empty mbarriers).  Every KV tile read from L2/HBM therefore feeds 64 columns
// (16x the arithmetic intensity of decode), and the key loop stops at the causal diagonal.
// K and V for the new tokens were already appended to the cache by rope_kv_kernel, so prefill with a
// non-empty cache (chunked prefill, prefix reuse) needs nothing special: query i of the chunk sits at
// position seq_lens[b] + i and sees keys 0 .. seq_lens[b] + i.
#include "attn_common.cuh"

namespace bfly {

namespace {

constexpr int kWarps = 4;
constexpr int kTile = 32;            // keys per tile
constexpr int kCols = 64;            // (query, head) columns per CTA
constexpr int kStages = 4;

template <int D>
constexpr size_t prefill_smem_bytes() {
  return (size_t)kStages * 2 * kTile * D * 2 + 2 * kStages * 8 + 128;
}

template <int D, int G>
__global__ void __launch_bounds__(kWarps * 32) attn_prefill_kernel(const bf16* __restrict__ q, KvLayout kv,
                                                                   const int* __restrict__ seq_lens, int S, int nq,
                                                                   bf16* __restrict__ out, float scale_log2) {
  static_assert(G >= 1 && G <= 8 && (D == 32 || D == 64 || D == 128), "unsupported shape");
  constexpr int QB = kCols / G;                 // query tokens per CTA
  constexpr int KSTEP_PAIRS = D / 32;
  constexpr int MT_O = D / 16;
  constexpr uint32_t kBlockBytes = kTile * D * 2;
  constexpr bool kSwizzle = D >= 64;

  extern __shared__ __align__(128) uint8_t smem[];
  pdl_trigger();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane >> 2, c = lane & 3;
  const int qblk = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int s0 = qblk * QB;                      // first query (index inside the chunk) of this CTA
  const int nq_here = min(QB, S - s0);
  const int p0 = seq_lens[b] + s0;               // absolute position of that query
  const int n_tiles = (p0 + nq_here + kTile - 1) / kTile;     // keys 0 .. p0 + nq_here - 1

  const uint32_t ring = smem_u32(smem);
  const uint32_t full_bars = ring + (uint32_t)kStages * 2 * kBlockBytes;
  const uint32_t empty_bars = full_bars + 8u * kStages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bars + 8u * s, 1);
      mbar_init(empty_bars + 8u * s, kWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const int* table = kv.page_table + (int64_t)b * kv.pages_per_seq;
  const int ps = kv.page_size;
  const int64_t head_stride = (int64_t)ps * D;
  const int64_t page_stride = 2 * (int64_t)kv.kv_local * head_stride;
  auto issue = [&](int tile) {                   // thread 0 only
    const int tok0 = tile * kTile;
    const int page = table[tok0 / ps];
    const int sub = (tok0 % ps) / kTile;
    const bf16* kp = kv.base + (int64_t)page * page_stride + (int64_t)kvh * head_stride + (int64_t)sub * kTile * D;
    const bf16* vp = kp + (int64_t)kv.kv_local * head_stride;
    const int s = tile % kStages;
    const uint32_t bar = full_bars + 8u * s;
    const uint32_t dst = ring + (uint32_t)s * 2 * kBlockBytes;
    mbar_expect_tx(bar, 2 * kBlockBytes);
    bulk_copy_g2s(dst, kp, kBlockBytes, bar);
    bulk_copy_g2s(dst + kBlockBytes, vp, kBlockBytes, bar);
  };
  if (threadIdx.x == 0)
    for (int i = 0; i < min(kStages, n_tiles); ++i) issue(i);

  // this warp's 16 columns = 2 n-tiles; B fragments of Q for column n = 16*warp + 8*nt + r
  uint32_t qb[2][KSTEP_PAIRS][4];
  int qi_mine[2][2];                             // query index (inside the CTA) of columns 2c, 2c+1 per n-tile
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    const int col = 16 * warp + 8 * nt + r;
    const int qi = col / G, g = col - qi * G;
#pragma unroll
    for (int j = 0; j < KSTEP_PAIRS; ++j) {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (qi < nq_here)
        v = *reinterpret_cast<const uint4*>(q + ((int64_t)(b * S + s0 + qi) * nq + kvh * G + g) * D + c * 8 + 32 * j);
      qb[nt][j][0] = v.x; qb[nt][j][1] = v.y; qb[nt][j][2] = v.z; qb[nt][j][3] = v.w;
    }
    qi_mine[nt][0] = (16 * warp + 8 * nt + 2 * c) / G;
    qi_mine[nt][1] = (16 * warp + 8 * nt + 2 * c + 1) / G;
  }

  float oacc[2][MT_O][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int mo = 0; mo < MT_O; ++mo)
#pragma unroll
      for (int e = 0; e < 4; ++e) oacc[nt][mo][e] = 0.f;
  float m_run[2][2], l_run[2][2];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) { m_run[nt][0] = m_run[nt][1] = -INFINITY; l_run[nt][0] = l_run[nt][1] = 0.f; }

  for (int tile = 0; tile < n_tiles; ++tile) {
    const int s = tile % kStages;
    const uint32_t ph = (uint32_t)(tile / kStages) & 1u;
    mbar_wait(full_bars + 8u * s, ph);
    const uint32_t k_s = ring + (uint32_t)s * 2 * kBlockBytes;
    const uint32_t v_s = k_s + kBlockBytes;
    const int tok0 = tile * kTile;
    const bool diag = tok0 + kTile > p0;          // only tiles touching the causal diagonal need masks

    float sacc[2][2][4];                          // [n-tile][m-tile][..]
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int e = 0; e < 4; ++e) sacc[nt][mt][e] = 0.f;
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int ta = 8 * (r >> 1) + 4 * mt + (r & 1), tb = ta + 2;
      const uint32_t swz = kSwizzle ? (uint32_t)((ta & 1) << 2) : 0u;
#pragma unroll
      for (int j = 0; j < KSTEP_PAIRS; ++j) {
        const uint32_t chunk = ((uint32_t)(c + 4 * j)) ^ swz;
        const uint4 x = lds128(k_s + (uint32_t)ta * (D * 2) + chunk * 16);
        const uint4 y = lds128(k_s + (uint32_t)tb * (D * 2) + chunk * 16);
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          mma_bf16(sacc[nt][mt], x.x, y.x, x.y, y.y, qb[nt][j][0], qb[nt][j][1]);
          mma_bf16(sacc[nt][mt], x.z, y.z, x.w, y.w, qb[nt][j][2], qb[nt][j][3]);
        }
      }
    }
    uint32_t pb[2][2][2];                          // [n-tile][k-step][lo/hi]
    float corr[2][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const int ta = tok0 + 8 * (r >> 1) + 4 * mt + (r & 1);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int key = ta + ((e >> 1) << 1);
          const bool ok = !diag || key <= p0 + qi_mine[nt][e & 1];
          sacc[nt][mt][e] = ok ? sacc[nt][mt][e] * scale_log2 : -INFINITY;
        }
        tmax[0] = fmaxf(tmax[0], fmaxf(sacc[nt][mt][0], sacc[nt][mt][2]));
        tmax[1] = fmaxf(tmax[1], fmaxf(sacc[nt][mt][1], sacc[nt][mt][3]));
      }
#pragma unroll
      for (int off = 4; off < 32; off <<= 1) {
        tmax[0] = fmaxf(tmax[0], __shfl_xor_sync(0xffffffffu, tmax[0], off));
        tmax[1] = fmaxf(tmax[1], __shfl_xor_sync(0xffffffffu, tmax[1], off));
      }
      float msafe[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float m_new = fmaxf(m_run[nt][e], tmax[e]);
        // a column can be fully masked in a tile beyond its own diagonal (or be a padding column)
        msafe[e] = (m_new == -INFINITY) ? 0.f : m_new;
        corr[nt][e] = exp2f(m_run[nt][e] - msafe[e]);
        m_run[nt][e] = m_new;
      }
      float psum[2] = {0.f, 0.f};
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const float p0v = exp2f(sacc[nt][mt][0] - msafe[0]), p1v = exp2f(sacc[nt][mt][1] - msafe[1]);
        const float p2v = exp2f(sacc[nt][mt][2] - msafe[0]), p3v = exp2f(sacc[nt][mt][3] - msafe[1]);
        psum[0] += p0v + p2v;
        psum[1] += p1v + p3v;
        pb[nt][mt][0] = movmatrix_trans(pack_bf16(p0v, p1v));
        pb[nt][mt][1] = movmatrix_trans(pack_bf16(p2v, p3v));
      }
      l_run[nt][0] = l_run[nt][0] * corr[nt][0] + psum[0];
      l_run[nt][1] = l_run[nt][1] * corr[nt][1] + psum[1];
    }
#pragma unroll
    for (int mo = 0; mo < MT_O; ++mo) {
      const uint4 x = lds128(v_s + (uint32_t)(16 * mo + r) * (kTile * 2) + c * 16);
      const uint4 y = lds128(v_s + (uint32_t)(16 * mo + r + 8) * (kTile * 2) + c * 16);
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        oacc[nt][mo][0] *= corr[nt][0]; oacc[nt][mo][1] *= corr[nt][1];
        oacc[nt][mo][2] *= corr[nt][0]; oacc[nt][mo][3] *= corr[nt][1];
        mma_bf16(oacc[nt][mo], x.x, y.x, x.y, y.y, pb[nt][0][0], pb[nt][0][1]);
        mma_bf16(oacc[nt][mo], x.z, y.z, x.w, y.w, pb[nt][1][0], pb[nt][1][1]);
      }
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty_bars + 8u * s) : "memory");
    if (threadIdx.x == 0 && tile + kStages < n_tiles) {
      mbar_wait(empty_bars + 8u * s, ph);           // all 4 warps are done with this stage
      issue(tile + kStages);
    }
  }

  // ---- finalize: O^T / l, bf16, out[token][head][d]
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      l_run[nt][0] += __shfl_xor_sync(0xffffffffu, l_run[nt][0], off);
      l_run[nt][1] += __shfl_xor_sync(0xffffffffu, l_run[nt][1], off);
    }
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int col = 16 * warp + 8 * nt + 2 * c + e;
      const int qi = col / G, g = col - qi * G;
      if (qi >= nq_here) continue;
      const float inv = 1.0f / l_run[nt][e];
      bf16* o = out + ((int64_t)(b * S + s0 + qi) * nq + kvh * G + g) * D;
#pragma unroll
      for (int mo = 0; mo < MT_O; ++mo) {
        o[16 * mo + r] = __float2bfloat16_rn(oacc[nt][mo][e] * inv);
        o[16 * mo + r + 8] = __float2bfloat16_rn(oacc[nt][mo][2 + e] * inv);
      }
    }
  }
}

template <int D, int G>
int launch_impl(const bf16* q, const KvLayout& kv, const int* seq_lens, int S, int B, int nq, bf16* out,
                cudaStream_t stream) {
  static PerDeviceOnce attr;
  constexpr size_t smem = prefill_smem_bytes<D>();
  BFLY_TRY(attr.run([] {
    BFLY_CUDA(cudaFuncSetAttribute(attn_prefill_kernel<D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return (int)BFLY_OK;
  }));
  constexpr int QB = kCols / G;
  dim3 grid(ceil_div(S, QB), kv.kv_local, B);
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)D);
  BFLY_CUDA(launch_pdl(attn_prefill_kernel<D, G>, grid, dim3(kWarps * 32), smem, stream, q, kv, seq_lens, S, nq, out,
                       scale_log2));
  count_launch();
  return BFLY_OK;
}

}  // namespace

int launch_attn_prefill(const bf16* q, KvLayout kv, const int* seq_lens, int S, int B, int nq, bf16* out,
                        cudaStream_t stream) {
  BFLY_CHECK(S > 0 && B > 0 && B <= 65535 && kv.kv_local >= 1 && kv.kv_local <= 65535 && nq % kv.kv_local == 0,
             "attn_prefill: bad shape");
  BFLY_CHECK(kv.page_size % kTile == 0, "attn_prefill: page_size must be a multiple of 32");
  const int G = nq / kv.kv_local, D = kv.head_dim;
#define BFLY_ATTN_CASE(DD, GG) \
  if (D == DD && G == GG) return launch_impl<DD, GG>(q, kv, seq_lens, S, B, nq, out, stream)
  BFLY_ATTN_CASE(128, 1); BFLY_ATTN_CASE(128, 2); BFLY_ATTN_CASE(128, 4); BFLY_ATTN_CASE(128, 8);
  BFLY_ATTN_CASE(64, 1);  BFLY_ATTN_CASE(64, 2);  BFLY_ATTN_CASE(64, 4);  BFLY_ATTN_CASE(64, 8);
  BFLY_ATTN_CASE(32, 1);  BFLY_ATTN_CASE(32, 2);  BFLY_ATTN_CASE(32, 4);  BFLY_ATTN_CASE(32, 8);
#undef BFLY_ATTN_CASE
  set_error("attn_prefill: unsupported (head_dim, q-per-kv) = (" + std::to_string(D) + ", " + std::to_string(G) + ")");
  return BFLY_EINVAL;
}

}  // namespace bfly
