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
  int M, N, K;
  int bn, stages, splits;
  uint32_t tmem_cols;
  float* out_f32;
  bf16* out_bf16;
  int64_t split_stride;
  int64_t ldc;
};

__global__ void __launch_bounds__(kThreads) gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                           const __grid_constant__ CUtensorMap tmap_x,
                                                           const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // SWIZZLE_128B tiles: 1 KiB aligned
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = args.stages, bn = args.bn;
  const uint32_t b_bytes = (uint32_t)bn * BK * 2;
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + (uint32_t)stages * kABytes;
  const uint32_t bar_base = b_base + (uint32_t)stages * b_bytes;           // 8-byte aligned by construction
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(kMaxStages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * kMaxStages);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 1);

  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * bn;
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

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = kABytes + b_bytes;
      // Weights never depend on earlier kernels: fill the whole ring with weight tiles BEFORE waiting
      // for the predecessor (programmatic dependent launch), then add the activation tiles.
      const int pre = min(stages, nkb);
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), tx);
        tma_load_2d(a_base + (uint32_t)i * kABytes, &tmap_w, full_bar(i), (kb_begin + i) * BK, m0, kEvictFirst);
      }
      pdl_wait();
      for (int i = 0; i < pre; ++i)
        tma_load_2d(b_base + (uint32_t)i * b_bytes, &tmap_x, full_bar(i), (kb_begin + i) * BK, n0, kEvictLast);
      for (int i = pre; i < nkb; ++i) {
        const int s = i % stages;
        const uint32_t ph = (uint32_t)(i / stages) & 1u;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), tx);
        const int k_elem = (kb_begin + i) * BK;
        tma_load_2d(a_base + (uint32_t)s * kABytes, &tmap_w, full_bar(s), k_elem, m0, kEvictFirst);
        tma_load_2d(b_base + (uint32_t)s * b_bytes, &tmap_x, full_bar(s), k_elem, n0, kEvictLast);
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
        const uint64_t da = umma_smem_desc(a_base + (uint32_t)s * kABytes);
        const uint64_t db = umma_smem_desc(b_base + (uint32_t)s * b_bytes);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advance 32 bytes (16 bf16) inside the 128-byte swizzle row: +2 in 16-byte units
          umma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((i | k) != 0));
        }
        umma_commit(empty_bar(s));         // frees the smem stage once these MMAs have read it
      }
      umma_commit(tmem_full_bar);          // accumulator complete
    }
  } else {
    pdl_wait();      // outputs alias scratch that an earlier consumer kernel may still be reading
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int m = m0 + quarter * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const bool m_ok = m < args.M;
    float* of = args.out_f32 ? args.out_f32 + (int64_t)blockIdx.z * args.split_stride + m : nullptr;
    bf16* ob = args.out_bf16 ? args.out_bf16 + m : nullptr;
    for (int c0 = 0; c0 < bn; c0 += 16) {
      if (n0 + c0 >= args.N) break;                     // warp-uniform
      uint32_t r[16];
      tmem_ld16(taddr + (uint32_t)c0, r);
      tmem_ld_wait();
      if (m_ok) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int n = n0 + c0 + j;
          if (n < args.N) {
            if (of) of[(int64_t)n * args.ldc] = __uint_as_float(r[j]);
            else ob[(int64_t)n * args.ldc] = __float2bfloat16_rn(__uint_as_float(r[j]));
          }
        }
      }
    }
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

size_t smem_needed(int bn, int stages) {
  return 1024 /*align slack*/ + (size_t)stages * (kABytes + (size_t)bn * BK * 2) + 8 * (2 * kMaxStages + 2);
}

int make_map(CUtensorMap* map, const bf16* ptr, int rows, int cols, int64_t pitch, int box_rows) {
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

GemmPlan gemm_plan(int M, int N, int K, int sm_count, int max_splits) {
  GemmPlan p{};
  p.M = M; p.N = N; p.K = K;
  p.bn = std::min(256, ((N + 15) / 16) * 16);
  const int m_tiles = ceil_div(M, BM), n_tiles = ceil_div(N, p.bn), nkb = ceil_div(K, BK);
  const int base = m_tiles * n_tiles;
  // Weight-streaming regime (few token tiles): aim for ~2 CTAs per SM so that HBM requests from
  // co-resident CTAs hide wave quantisation; keep >= 4 K-blocks per slice.
  int splits = 1;
  if (base < 2 * sm_count) {
    splits = std::max(1, (2 * sm_count + base / 2) / base);
    splits = std::min(splits, std::max(1, nkb / 4));
    splits = std::min(splits, std::max(1, max_splits));
  }
  p.splits = splits;
  const size_t stage_bytes = kABytes + (size_t)p.bn * BK * 2;
  int stages = (p.bn <= 64) ? 4 : (int)std::min<size_t>(kMaxStages, (200 * 1024) / stage_bytes);
  stages = std::max(2, std::min(stages, kMaxStages));
  p.stages = stages;
  p.smem_bytes = smem_needed(p.bn, stages);
  return p;
}

int launch_gemm(const bf16* W, const bf16* X, int64_t ldx, const GemmPlan& plan, float* out_f32,
                int64_t split_stride, bf16* out_bf16, int64_t ldc, cudaStream_t stream) {
  BFLY_TRY(gemm_init());
  BFLY_CHECK(plan.M > 0 && plan.N > 0 && plan.K > 0 && plan.K % 8 == 0, "gemm: M,N,K > 0 and K % 8 == 0");
  BFLY_CHECK(plan.bn >= 16 && plan.bn <= 256 && plan.bn % 16 == 0, "gemm: bn must be a multiple of 16 in [16,256]");
  BFLY_CHECK(plan.stages >= 2 && plan.stages <= kMaxStages, "gemm: stages in [2,8]");
  BFLY_CHECK(plan.splits >= 1 && plan.splits <= ceil_div(plan.K, BK), "gemm: splits in [1, K/64]");
  BFLY_CHECK((out_f32 != nullptr) != (out_bf16 != nullptr), "gemm: exactly one of out_f32 / out_bf16");
  BFLY_CHECK(out_bf16 == nullptr || plan.splits == 1, "gemm: bf16 output needs splits == 1");
  BFLY_CHECK(((uintptr_t)W & 15) == 0 && ((uintptr_t)X & 15) == 0, "gemm: operands must be 16-byte aligned");
  const size_t smem = smem_needed(plan.bn, plan.stages);
  BFLY_CHECK(smem <= 227 * 1024, "gemm: tile does not fit shared memory");
  CUtensorMap mw, mx;
  BFLY_CHECK(ldx >= plan.K && ldx % 8 == 0, "gemm: ldx >= K and ldx % 8 == 0");
  BFLY_TRY(make_map(&mw, W, plan.M, plan.K, plan.K, BM));
  BFLY_TRY(make_map(&mx, X, plan.N, plan.K, ldx, plan.bn));
  GemmArgs a{};
  a.M = plan.M; a.N = plan.N; a.K = plan.K;
  a.bn = plan.bn; a.stages = plan.stages; a.splits = plan.splits;
  uint32_t cols = 32;
  while ((int)cols < plan.bn) cols <<= 1;
  a.tmem_cols = cols;
  a.out_f32 = out_f32; a.out_bf16 = out_bf16; a.split_stride = split_stride; a.ldc = ldc;
  dim3 grid(ceil_div(plan.M, BM), ceil_div(plan.N, plan.bn), plan.splits);
  BFLY_CUDA(launch_pdl(gemm_tn_kernel, grid, dim3(kThreads), smem, stream, mw, mx, a));
  count_launch();
  return BFLY_OK;
}

}  // namespace bfly
