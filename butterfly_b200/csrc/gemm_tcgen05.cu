// K1/K2: column- and row-parallel projections as one hand-written sm_100a GEMM
//     C[n][m] = sum_k W[m][k] * X[n][k]      W = weight shard [M, K], X = activations [N, K],
// both bf16 K-major (row-major [rows, K]), fp32 accumulation in tensor memory.
//
// The weight matrix is always the UMMA "A" operand (128 weight rows = 128 TMEM lanes) and the
// tokens are the UMMA N dimension (16..256, a runtime value), i.e. the "swap-AB" orientation: at
// decode (N = batch <= 64) the tensor core still runs full 128-row tiles while the kernel streams
// the weights exactly once (HBM-bound); at prefill the same kernel runs 128x256 tiles
// (tensor-bound).  Split-K slices (blockIdx.z) are written as fp32 slices that the consumer kernel
// folds in a fixed order, so results are deterministic (no atomics).
//
// Pipeline per CTA (192 threads): warp 0 = TMA producer (cp.async.bulk.tensor, 128B swizzle,
// mbarrier complete_tx), warp 1 = TMEM allocator + single-thread tcgen05.mma issuer
// (tcgen05.commit releases smem stages / publishes the accumulator), warps 2..5 = epilogue
// (tcgen05.ld 32x32b -> registers -> coalesced global stores, lane = weight row so each warp store
// writes 32 consecutive features of one token).
#include <cuda.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <mutex>

#include "kernels.cuh"

namespace bfly {

namespace {

constexpr int BM = 128;                 // weight rows per tile (UMMA M)
constexpr int BK = 64;                  // K per stage: 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int kThreads = 192;
constexpr uint32_t kABytes = BM * BK * 2;   // 16 KiB per stage
constexpr int kMaxStages = 8;
constexpr int kMaxCluster = 8;           // portable thread-block cluster size
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;   // L2 policy: stream (weights)
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;    // L2 policy: keep (activations)

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must surface as a trap (launch error), never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile in shared memory, 128-byte swizzle: rows of 128 B, 8-row atoms 1024 B apart.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFFu)        // start address, 16-byte units
         | ((uint64_t)(1024u >> 4) << 32)         // stride byte offset between 8-row atoms
         | (1ull << 46)                           // descriptor version (sm_100)
         | (2ull << 61);                          // layout: SWIZZLE_128B
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

struct GemmArgs {
  int M, N, K;                 // M = output features (dual: gated features), N = tokens
  int bn, stages, splits;
  int cluster;                 // 1: the `splits` CTAs along z form a cluster and reduce through DSMEM
  int dual;                    // 1: two accumulators (gate, up) and a fused SwiGLU epilogue
  int f_block, n_blocks;       // dual: W rows are n_blocks x [gate(f_block) | up(f_block)]
  uint32_t tmem_cols;
  float* out_f32;
  bf16* out_bf16;
  const float* route_w;        // dual + MoE: [N][n_blocks] routing weights (0 = not routed), else nullptr
  int64_t split_stride;
  int64_t ldc;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
  return v;
}
__device__ __forceinline__ float silu_mul(float g, float u) { return g / (1.0f + expf(-g)) * u; }

__global__ void __launch_bounds__(kThreads) gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                           const __grid_constant__ CUtensorMap tmap_x,
                                                           const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // SWIZZLE_128B tiles: 1 KiB aligned
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = args.stages, bn = args.bn, dual = args.dual;
  const uint32_t a_bytes = kABytes * (uint32_t)(1 + dual);
  const uint32_t b_bytes = (uint32_t)bn * BK * 2;
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + (uint32_t)stages * a_bytes;
  const uint32_t bar_base = b_base + (uint32_t)stages * b_bytes;           // 8-byte aligned by construction
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(kMaxStages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * kMaxStages);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 1);

  // tile -> weight rows.  dual: features [feat0, feat0+128) of block `blk`; gate rows then up rows
  int m0 = blockIdx.x * BM, m_up = 0, feat0 = m0, blk = 0, feat_limit = args.M;
  if (dual) {
    const int tiles_per_blk = (args.f_block + BM - 1) / BM;
    blk = blockIdx.x / tiles_per_blk;
    const int in_blk = (blockIdx.x - blk * tiles_per_blk) * BM;
    m0 = blk * 2 * args.f_block + in_blk;
    m_up = m0 + args.f_block;
    feat0 = blk * args.f_block + in_blk;
    feat_limit = (blk + 1) * args.f_block;
  }
  const int n0 = blockIdx.y * bn;
  const int nkb_total = (args.K + BK - 1) / BK;
  const int kb_begin = (int)((int64_t)nkb_total * blockIdx.z / args.splits);
  const int kb_end = (int)((int64_t)nkb_total * (blockIdx.z + 1) / args.splits);
  const int nkb = kb_end - kb_begin;

  pdl_trigger();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "r"(args.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  // Epilogue store of one element (token n, tile-local weight row ml).
  auto store = [&](int n, int ml, float g, float u, int slice) {
    const int feat = feat0 + ml;
    if (n >= args.N || feat >= feat_limit) return;
    if (dual) {
      float y = silu_mul(g, u);
      if (args.route_w != nullptr) y *= args.route_w[(int64_t)n * args.n_blocks + blk];
      args.out_bf16[(int64_t)n * args.ldc + feat] = __float2bfloat16_rn(y);
    } else if (args.out_f32 != nullptr) {
      args.out_f32[(int64_t)slice * args.split_stride + (int64_t)n * args.ldc + feat] = g;
    } else {
      args.out_bf16[(int64_t)n * args.ldc + feat] = __float2bfloat16_rn(g);
    }
  };

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = a_bytes + b_bytes;
      // Weights never depend on earlier kernels: fill the whole ring with weight tiles BEFORE waiting
      // for the predecessor (programmatic dependent launch), then add the activation tiles.
      auto load_a = [&](int i, int s) {
        const int k_elem = (kb_begin + i) * BK;
        tma_load_2d(a_base + (uint32_t)s * a_bytes, &tmap_w, full_bar(s), k_elem, m0, kEvictFirst);
        if (dual) tma_load_2d(a_base + (uint32_t)s * a_bytes + kABytes, &tmap_w, full_bar(s), k_elem, m_up, kEvictFirst);
      };
      const int pre = min(stages, nkb);
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), tx);
        load_a(i, i);
      }
      pdl_wait();
      for (int i = 0; i < pre; ++i)
        tma_load_2d(b_base + (uint32_t)i * b_bytes, &tmap_x, full_bar(i), (kb_begin + i) * BK, n0, kEvictLast);
      for (int i = pre; i < nkb; ++i) {
        const int s = i % stages;
        const uint32_t ph = (uint32_t)(i / stages) & 1u;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), tx);
        load_a(i, s);
        tma_load_2d(b_base + (uint32_t)s * b_bytes, &tmap_x, full_bar(s), (kb_begin + i) * BK, n0, kEvictLast);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N = bn, M = 128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % stages;
        const uint32_t ph = (uint32_t)(i / stages) & 1u;
        mbar_wait(full_bar(s), ph);
        tcgen05_fence_after();
        const uint64_t da = umma_smem_desc(a_base + (uint32_t)s * a_bytes);
        const uint64_t db = umma_smem_desc(b_base + (uint32_t)s * b_bytes);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advance 32 bytes (16 bf16) inside the 128-byte swizzle row: +2 in 16-byte units
          umma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((i | k) != 0));
        }
        if (dual) {
          const uint64_t du = umma_smem_desc(a_base + (uint32_t)s * a_bytes + kABytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma_bf16(tmem_base + (uint32_t)bn, du + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                      (uint32_t)((i | k) != 0));
        }
        umma_commit(empty_bar(s));         // frees the smem stage once these MMAs have read it
      }
      umma_commit(tmem_full_bar);          // accumulators complete
    }
  } else {
    pdl_wait();      // outputs alias scratch that an earlier consumer kernel may still be reading
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int ml = quarter * 32 + lane;                 // tile-local weight row
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    if (!args.cluster) {
      for (int c0 = 0; c0 < bn; c0 += 16) {
        if (n0 + c0 >= args.N) break;                   // warp-uniform
        uint32_t r[16], ru[16];
        tmem_ld16(taddr + (uint32_t)c0, r);
        if (dual) tmem_ld16(taddr + (uint32_t)(bn + c0), ru);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j)
          store(n0 + c0 + j, ml, __uint_as_float(r[j]), dual ? __uint_as_float(ru[j]) : 0.f, (int)blockIdx.z);
      }
    } else {
      // dump this CTA's partial accumulators into its own shared memory: red[acc][col][row] (the operand
      // ring is free: every MMA that read it has completed)
      for (int acc = 0; acc <= dual; ++acc) {
        for (int c0 = 0; c0 < bn; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + (uint32_t)(acc * bn + c0), r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const uint32_t off = ((uint32_t)((acc * bn + c0 + j) * BM + ml)) * 4u;
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(a_base + off), "r"(r[j]) : "memory");
          }
        }
      }
    }
  }
  if (args.cluster) {
    // split-K reduction across the cluster through distributed shared memory, in rank order
    // (deterministic); CTA `rank` finalises token columns rank, rank + splits, ...
    __syncwarp();
    cluster_sync_all();
    if (warp >= 2) {
      const uint32_t rank = cluster_ctarank();
      const int ml = (warp & 3) * 32 + lane;
      // peers' reduction buffers, resolved once (mapa), then all loads of a column are issued together
      uint32_t peer[kMaxCluster];
#pragma unroll
      for (int k = 0; k < kMaxCluster; ++k)
        asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(peer[k]) : "r"(a_base), "r"((uint32_t)(k < args.splits ? k : 0)));
      for (int c = (int)rank; c < bn; c += args.splits) {
        if (n0 + c >= args.N) break;
        float vg[kMaxCluster], vu[kMaxCluster];
        const uint32_t off_g = ((uint32_t)(c * BM + ml)) * 4u, off_u = ((uint32_t)((bn + c) * BM + ml)) * 4u;
#pragma unroll
        for (int k = 0; k < kMaxCluster; ++k) {
          vg[k] = 0.f; vu[k] = 0.f;
          if (k < args.splits) {
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(vg[k]) : "r"(peer[k] + off_g));
            if (dual) asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(vu[k]) : "r"(peer[k] + off_u));
          }
        }
        float g = 0.f, u = 0.f;
#pragma unroll
        for (int k = 0; k < kMaxCluster; ++k) { g += vg[k]; u += vu[k]; }     // rank order: deterministic
        store(n0 + c, ml, g, u, 0);
      }
    }
    __syncwarp();
    cluster_sync_all();            // nobody may exit while a peer still reads its shared memory
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(args.tmem_cols)
                 : "memory");
  }
}

// ---- host side --------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
std::once_flag g_init_once;
int g_init_rc = BFLY_OK;

size_t smem_needed(int bn, int stages, int dual) {
  return 1024 /*align slack*/ + (size_t)stages * (kABytes * (size_t)(1 + dual) + (size_t)bn * BK * 2) +
         8 * (2 * kMaxStages + 2);
}

int make_map(CUtensorMap* map, const bf16* ptr, int64_t rows, int cols, int64_t pitch, int box_rows) {
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch * sizeof(bf16)};
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(ptr), dims, strides, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r) + " for [" +
              std::to_string(rows) + "," + std::to_string(cols) + "] box_rows=" + std::to_string(box_rows));
    return BFLY_ECUDA;
  }
  return BFLY_OK;
}

}  // namespace

int gemm_init() {
  std::call_once(g_init_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
      set_error(std::string("cannot resolve cuTensorMapEncodeTiled: ") + cudaGetErrorString(e));
      g_init_rc = BFLY_ECUDA;
      return;
    }
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    e = cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      set_error(std::string("gemm: cudaFuncSetAttribute: ") + cudaGetErrorString(e));
      g_init_rc = BFLY_ECUDA;
    }
  });
  return g_init_rc;
}

// M = output features (dual: gated features, i.e. half the weight rows).
GemmPlan gemm_plan(int M, int N, int K, int sm_count, int max_splits, int dual) {
  GemmPlan p{};
  p.M = M; p.N = N; p.K = K;
  p.dual = dual;
  p.bn = std::min(256, ((N + 15) / 16) * 16);
  const int m_tiles = ceil_div(M, BM), n_tiles = ceil_div(N, p.bn), nkb = ceil_div(K, BK);
  const int base = m_tiles * n_tiles;
  // Weight-streaming regime (few token tiles): aim for ~2 CTAs per SM so that HBM requests from
  // co-resident CTAs hide wave quantisation; keep >= 4 K-blocks per slice.  (Measured on B200,
  // tools/tune_kernels.py: one wave of ~2 CTAs/SM beats any multi-wave split for these shapes.)
  int splits = 1;
  if (base < 2 * sm_count) {
    splits = std::max(1, (2 * sm_count) / base);          // never more CTAs than ~2 per SM: a second wave costs more
    splits = std::min(splits, std::max(1, nkb / 4));
    splits = std::min(splits, std::max(1, max_splits));
  }
  // K-slices of token tiles <= 64 wide CAN be reduced inside a thread-block cluster through DSMEM
  // (one final output, fused epilogue).  Measured: cluster scheduling (all CTAs of a cluster co-resident
  // in one GPC) costs 1.5-10 us per GEMM against fp32 slices folded by the consumer, so plain
  // projections keep the slices; only the fused SwiGLU projection, which needs the complete K sum,
  // uses the cluster at decode.  BFLY_GEMM_CLUSTER=1 forces it everywhere (experiments).
  static const bool force_cluster = [] { const char* e = getenv("BFLY_GEMM_CLUSTER"); return e && e[0] == '1'; }();
  p.cluster = 0;
  if (dual) {
    if (p.bn <= 64 && base < sm_count) {
      splits = std::min(kMaxCluster, std::max(1, (3 * sm_count / 2 + base - 1) / base));
      splits = std::min(splits, std::max(1, nkb / 4));
      p.cluster = splits > 1 ? 1 : 0;
    } else {
      splits = 1;
    }
  } else if (force_cluster && p.bn <= 64 && splits > 1) {
    splits = std::min(splits, kMaxCluster);
    p.cluster = 1;
  }
  p.splits = splits;
  const size_t stage_bytes = kABytes * (size_t)(1 + dual) + (size_t)p.bn * BK * 2;
  int stages = (p.bn <= 64) ? ((dual || p.cluster) ? 3 : 4) : (int)std::min<size_t>(kMaxStages, (200 * 1024) / stage_bytes);
  stages = std::max(2, std::min(stages, kMaxStages));
  p.stages = stages;
  p.smem_bytes = smem_needed(p.bn, stages, dual);
  return p;
}

int launch_gemm_ex(const bf16* W, const bf16* X, int64_t ldx, const GemmPlan& plan, const GemmOut& out,
                   cudaStream_t stream) {
  BFLY_TRY(gemm_init());
  const int dual = plan.dual ? 1 : 0;
  BFLY_CHECK(plan.M > 0 && plan.N > 0 && plan.K > 0 && plan.K % 8 == 0, "gemm: M,N,K > 0 and K % 8 == 0");
  BFLY_CHECK(plan.bn >= 16 && plan.bn <= 256 && plan.bn % 16 == 0, "gemm: bn must be a multiple of 16 in [16,256]");
  BFLY_CHECK(plan.stages >= 2 && plan.stages <= kMaxStages, "gemm: stages in [2,8]");
  BFLY_CHECK(plan.splits >= 1 && plan.splits <= ceil_div(plan.K, BK), "gemm: splits in [1, K/64]");
  BFLY_CHECK(!plan.cluster || (plan.splits <= kMaxCluster && plan.bn <= 64), "gemm: cluster reduce needs splits <= 8, bn <= 64");
  BFLY_CHECK(((uintptr_t)W & 15) == 0 && ((uintptr_t)X & 15) == 0, "gemm: operands must be 16-byte aligned");
  BFLY_CHECK(ldx >= plan.K && ldx % 8 == 0, "gemm: ldx >= K and ldx % 8 == 0");
  const bool single_output = plan.splits == 1 || plan.cluster;
  if (dual) {
    BFLY_CHECK(out.out_bf16 != nullptr && out.out_f32 == nullptr, "gemm: the fused SwiGLU epilogue writes bf16 only");
    BFLY_CHECK(single_output, "gemm: the fused SwiGLU epilogue needs splits == 1 or a cluster reduction");
    BFLY_CHECK(out.n_blocks >= 1 && out.f_block >= 1 && out.n_blocks * out.f_block == plan.M, "gemm: dual blocks");
    BFLY_CHECK(2 * plan.bn <= 512, "gemm: dual accumulators exceed tensor memory");
  } else {
    BFLY_CHECK((out.out_f32 != nullptr) != (out.out_bf16 != nullptr), "gemm: exactly one of out_f32 / out_bf16");
    BFLY_CHECK(out.out_bf16 == nullptr || single_output, "gemm: bf16 output needs one complete K sum");
  }
  const size_t smem = smem_needed(plan.bn, plan.stages, dual);
  BFLY_CHECK(smem <= 227 * 1024, "gemm: tile does not fit shared memory");
  BFLY_CHECK(!plan.cluster || (size_t)(1 + dual) * plan.bn * BM * 4 <= (size_t)plan.stages * kABytes * (1 + dual),
             "gemm: cluster reduction buffer does not fit the operand ring");
  CUtensorMap mw, mx;
  const int64_t w_rows = dual ? 2 * (int64_t)plan.M : plan.M;
  BFLY_TRY(make_map(&mw, W, w_rows, plan.K, plan.K, BM));
  BFLY_TRY(make_map(&mx, X, plan.N, plan.K, ldx, plan.bn));
  GemmArgs a{};
  a.M = plan.M; a.N = plan.N; a.K = plan.K;
  a.bn = plan.bn; a.stages = plan.stages; a.splits = plan.splits;
  a.cluster = plan.cluster && plan.splits > 1 ? 1 : 0;
  a.dual = dual;
  a.f_block = dual ? out.f_block : plan.M;
  a.n_blocks = dual ? out.n_blocks : 1;
  uint32_t cols = 32;
  while ((int)cols < plan.bn * (1 + dual)) cols <<= 1;
  a.tmem_cols = cols;
  a.out_f32 = out.out_f32; a.out_bf16 = out.out_bf16; a.route_w = out.route_w;
  a.split_stride = out.split_stride; a.ldc = out.ldc;
  int m_tiles = ceil_div(plan.M, BM);
  if (dual) m_tiles = out.n_blocks * ceil_div(out.f_block, BM);
  dim3 grid(m_tiles, ceil_div(plan.N, plan.bn), plan.splits);

  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.numAttrs = 1;
  if (a.cluster) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 1;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = plan.splits;
    cfg.numAttrs = 2;
  }
  cfg.attrs = attr;
  BFLY_CUDA(cudaLaunchKernelEx(&cfg, gemm_tn_kernel, mw, mx, a));
  count_launch();
  return BFLY_OK;
}

int launch_gemm(const bf16* W, const bf16* X, int64_t ldx, const GemmPlan& plan, float* out_f32,
                int64_t split_stride, bf16* out_bf16, int64_t ldc, cudaStream_t stream) {
  GemmOut o{};
  o.out_f32 = out_f32; o.split_stride = split_stride; o.out_bf16 = out_bf16; o.ldc = ldc;
  return launch_gemm_ex(W, X, ldx, plan, o, stream);
}

}  // namespace bfly
