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

#include "decode_step.h"
#include "tcgen05_common.cuh"

namespace bfly {

namespace {

// kPair is a compile-time switch: a kernel that contains cta_group::2 instructions can only be launched as
// a cluster, so the single-CTA instantiation must not contain them.
template <bool kPair>
__global__ void __launch_bounds__(kThreads) gemm_tn_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                           const __grid_constant__ CUtensorMap tmap_x,
                                                           const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // SWIZZLE_128B tiles: 1 KiB aligned
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = args.stages, bn = args.bn, dual = args.dual;
  constexpr int pair = kPair ? 1 : 0;
  const uint32_t cta_rank = pair ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  const uint32_t a_bytes = kABytes * (uint32_t)(1 + dual);
  const uint32_t b_bytes = (uint32_t)(bn >> pair) * BK * 2;       // pair: each CTA stages half of the tokens
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
    feat_limit = blk < args.n_blocks ? (blk + 1) * args.f_block : 0;     // padding tile of an odd pair: stores nothing
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
      mbar_init(full_bar(s), 1 + pair);           // pair: both CTAs' producers arrive on the leader's barrier
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if constexpr (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                   "r"(args.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                   "r"(args.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tcgen05_fence_before();
  if (pair) cluster_sync_all(); else __syncthreads();   // pair: the peer's barriers must exist before its TMA signals them
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
      auto load = [&](uint32_t dst, const CUtensorMap* map, int s, int k_elem, int row, uint64_t policy) {
        if constexpr (kPair) tma_load_2d_pair(dst, map, full_bar(s), k_elem, row, policy);
        else tma_load_2d(dst, map, full_bar(s), k_elem, row, policy);
      };
      auto arm = [&](int s) {      // the leader's barrier counts the bytes of both CTAs
        if (leader) mbar_expect_tx(full_bar(s), tx << pair);
        else mbar_arrive_remote(full_bar(s), 0u);
      };
      auto load_a = [&](int i, int s) {
        const int k_elem = (kb_begin + i) * BK;
        load(a_base + (uint32_t)s * a_bytes, &tmap_w, s, k_elem, m0, kEvictFirst);
        if (dual) load(a_base + (uint32_t)s * a_bytes + kABytes, &tmap_w, s, k_elem, m_up, kEvictFirst);
      };
      const int n_load = n0 + (int)cta_rank * (bn >> 1);          // pair: this CTA's half of the token tile
      const int pre = min(stages, nkb);
      for (int i = 0; i < pre; ++i) {
        arm(i);
        load_a(i, i);
      }
      // ... and pull the next weight tiles into L2 while the predecessor (a small latency-bound kernel:
      // norm, RoPE, activation) leaves HBM idle
      const int pf_end = min(nkb, pre + args.l2_prefetch);
      for (int i = pre; i < pf_end; ++i) {
        tma_prefetch_l2_2d(&tmap_w, (kb_begin + i) * BK, m0);
        if (dual) tma_prefetch_l2_2d(&tmap_w, (kb_begin + i) * BK, m_up);
      }
      pdl_wait();
      for (int i = 0; i < pre; ++i)
        load(b_base + (uint32_t)i * b_bytes, &tmap_x, i, (kb_begin + i) * BK, n_load, kEvictLast);
      for (int i = pre; i < nkb; ++i) {
        const int s = i % stages;
        const uint32_t ph = (uint32_t)(i / stages) & 1u;
        mbar_wait(empty_bar(s), ph ^ 1u);
        arm(s);
        load_a(i, s);
        load(b_base + (uint32_t)s * b_bytes, &tmap_x, s, (kb_begin + i) * BK, n_load, kEvictLast);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N = bn, M = 128 (256 for a CTA pair)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)((BM << pair) >> 4) << 24);
      auto mma = [&](uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t acc) {
        if constexpr (kPair) umma_bf16_pair(d_tmem, a_desc, b_desc, idesc, acc);
        else umma_bf16(d_tmem, a_desc, b_desc, idesc, acc);
      };
      auto commit = [&](uint32_t bar) {
        if constexpr (kPair) umma_commit_pair(bar);
        else umma_commit(bar);
      };
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
          mma(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), (uint32_t)((i | k) != 0));
        }
        if (dual) {
          const uint64_t du = umma_smem_desc(a_base + (uint32_t)s * a_bytes + kABytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            mma(tmem_base + (uint32_t)bn, du + (uint64_t)(2 * k), db + (uint64_t)(2 * k), (uint32_t)((i | k) != 0));
        }
        commit(empty_bar(s));              // frees the smem stage (in both CTAs of a pair) once these MMAs have read it
      }
      commit(tmem_full_bar);               // accumulators complete
    }
  } else {
    pdl_wait();      // outputs alias scratch that an earlier consumer kernel may still be reading
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int ml = quarter * 32 + lane;                 // tile-local weight row
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    if (!args.cluster) {
      // 32 accumulator columns (and, in dual mode, the matching 32 of the second accumulator) per TMEM
      // round trip: the tcgen05.ld latency is exposed once per chunk, so wide chunks keep the epilogue short
      int c0 = 0;
      for (; c0 + 32 <= bn; c0 += 32) {
        if (n0 + c0 >= args.N) break;                   // warp-uniform
        uint32_t r[32], ru[32];
        tmem_ld32(taddr + (uint32_t)c0, r);
        if (dual) tmem_ld32(taddr + (uint32_t)(bn + c0), ru);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j)
          store(n0 + c0 + j, ml, __uint_as_float(r[j]), dual ? __uint_as_float(ru[j]) : 0.f, (int)blockIdx.z);
      }
      if (c0 < bn && n0 + c0 < args.N) {                // bn % 32 == 16 tail
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
  if (pair) { __syncwarp(); cluster_sync_all(); } else __syncthreads();
  if (warp == 1) {
    if constexpr (kPair)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(args.tmem_cols)
                   : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(args.tmem_cols)
                   : "memory");
  }
}


// -----------------------------------------------------------------------------------------------
// Persistent variant for the tensor-bound regime (>= 128 tokens per tile): one CTA (pair) per SM (pair)
// loops over (weight tile, token tile) units; the accumulator is DOUBLE-BUFFERED in tensor memory so the
// epilogue of unit i (tcgen05.ld -> SwiGLU/convert -> global stores) overlaps the MMAs of unit i+1, and
// the TMA ring runs ahead across unit boundaries.  Units are ordered token-tile-fastest, so the CTAs of
// a wave share one weight tile through L2 and every weight byte leaves HBM once (see tile_of for the
// 2D-blocked refinement that also keeps the token tiles L2-resident).
// -----------------------------------------------------------------------------------------------
constexpr int kPersistEpiWarps = 8;
constexpr int kBandUnits = 8;             // weight units (CTA-pair tiles) per L2 band of the persistent schedule
constexpr int kPersistThreads = 64 + 32 * kPersistEpiWarps;

template <bool kPair>
__global__ void __launch_bounds__(kPersistThreads) gemm_persistent_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                                   const __grid_constant__ CUtensorMap tmap_x,
                                                                   const GemmArgs args) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = args.stages, bn = args.bn, dual = args.dual;
  constexpr int pair = kPair ? 1 : 0;
  const uint32_t cta_rank = pair ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  const uint32_t a_bytes = kABytes * (uint32_t)(1 + dual);
  const uint32_t b_bytes = (uint32_t)(bn >> pair) * BK * 2;
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + (uint32_t)stages * a_bytes;
  const uint32_t bar_base = b_base + (uint32_t)stages * b_bytes;
  auto full_bar = [&](int s) { return bar_base + 8u * (uint32_t)s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (uint32_t)(kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (uint32_t)(2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (uint32_t)(2 * kMaxStages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 4);

  const int n_groups = (int)gridDim.x >> pair;            // CTAs (or CTA pairs) in the grid
  const int my_group = (int)blockIdx.x >> pair;
  // token rows of this launch: static, or (grouped MoE) read from the device once the producer kernel is done
  int n_rows = args.N, row_off = 0, n_tiles = args.n_tiles;
  int total_units = args.m_units * n_tiles;
  const int nkb = (args.K + BK - 1) / BK;
  const uint32_t acc_cols = (uint32_t)bn * (uint32_t)(1 + dual);

  // unit -> tile geometry
  struct Tile { int m0, m_up, feat0, feat_limit, blk, n0; };
  auto tile_of = [&](int u) {
    Tile t;
    // 2D-blocked unit order: bands of kBandUnits weight units; inside a band the weight unit varies fastest,
    // so a wave of ~74 units touches ~8 weight units x ~9 token tiles (both L2-resident) instead of
    // re-streaming every token tile for each weight unit.
    const int per_band = kBandUnits * n_tiles;
    const int band = u / per_band, rem = u - band * per_band;
    const int band_m = min(kBandUnits, args.m_units - band * kBandUnits);
    const int n_idx = rem / band_m;
    const int m_tile = (band * kBandUnits + (rem - n_idx * band_m)) * (1 + pair) + (int)cta_rank;
    t.n0 = n_idx * bn;
    t.m0 = m_tile * BM; t.m_up = 0; t.feat0 = t.m0; t.blk = 0; t.feat_limit = args.M;
    if (dual) {
      const int tiles_per_blk = (args.f_block + BM - 1) / BM;
      t.blk = m_tile / tiles_per_blk;
      const int in_blk = (m_tile - t.blk * tiles_per_blk) * BM;
      t.m0 = t.blk * 2 * args.f_block + in_blk;
      t.m_up = t.m0 + args.f_block;
      t.feat0 = t.blk * args.f_block + in_blk;
      t.feat_limit = t.blk < args.n_blocks ? (t.blk + 1) * args.f_block : 0;
    }
    return t;
  };

  pdl_trigger();
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_x) : "memory");
    for (int s = 0; s < stages; ++s) {
      mbar_init(full_bar(s), 1 + pair);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), kPersistEpiWarps * (1 + pair));   // one arrival per epilogue warp of every CTA of the pair
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if constexpr (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                   "r"(args.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                   "r"(args.tmem_cols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  tcgen05_fence_before();
  if (pair) cluster_sync_all(); else __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  if (args.n_rows_dev != nullptr) {
    pdl_wait();
    n_rows = min(args.N, *args.n_rows_dev);
    row_off = args.row_off_dev != nullptr ? *args.row_off_dev : 0;
    n_tiles = (n_rows + bn - 1) / bn;
    total_units = args.m_units * n_tiles;
  }

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = a_bytes + b_bytes;
      auto load = [&](uint32_t dst, const CUtensorMap* map, int s, int k_elem, int row, uint64_t policy) {
        if constexpr (kPair) tma_load_2d_pair(dst, map, full_bar(s), k_elem, row, policy);
        else tma_load_2d(dst, map, full_bar(s), k_elem, row, policy);
      };
      auto arm = [&](int s) {
        if (leader) mbar_expect_tx(full_bar(s), tx << pair);
        else mbar_arrive_remote(full_bar(s), 0u);
      };
      int it = 0;
      bool waited = false;
      for (int u = my_group; u < total_units; u += n_groups) {
        const Tile t = tile_of(u);
        const int n_load = row_off + t.n0 + (int)cta_rank * (bn >> 1) * pair;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % stages;
          const uint32_t ph = (uint32_t)(it / stages) & 1u;
          mbar_wait(empty_bar(s), ph ^ 1u);
          arm(s);
          const int k_elem = kb * BK;
          load(a_base + (uint32_t)s * a_bytes, &tmap_w, s, k_elem, t.m0, kEvictLast);
          if (dual) load(a_base + (uint32_t)s * a_bytes + kABytes, &tmap_w, s, k_elem, t.m_up, kEvictLast);
          if (!waited) { pdl_wait(); waited = true; }     // weights never depend on the previous kernel; tokens do
          load(b_base + (uint32_t)s * b_bytes, &tmap_x, s, k_elem, n_load, kEvictLast);
        }
      }
      if (!waited) pdl_wait();
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)((BM << pair) >> 4) << 24);
      auto mma = [&](uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t acc) {
        if constexpr (kPair) umma_bf16_pair(d_tmem, a_desc, b_desc, idesc, acc);
        else umma_bf16(d_tmem, a_desc, b_desc, idesc, acc);
      };
      auto commit = [&](uint32_t bar) {
        if constexpr (kPair) umma_commit_pair(bar);
        else umma_commit(bar);
      };
      int it = 0, ui = 0;
      for (int u = my_group; u < total_units; u += n_groups, ++ui) {
        const int acc = ui & 1;
        mbar_wait(tempty_bar(acc), (((uint32_t)ui >> 1) & 1u) ^ 1u);     // epilogue has drained this accumulator
        tcgen05_fence_after();
        const uint32_t d0 = tmem_base + (uint32_t)acc * acc_cols;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % stages;
          mbar_wait(full_bar(s), (uint32_t)(it / stages) & 1u);
          tcgen05_fence_after();
          const uint64_t da = umma_smem_desc(a_base + (uint32_t)s * a_bytes);
          const uint64_t db = umma_smem_desc(b_base + (uint32_t)s * b_bytes);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            mma(d0, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), (uint32_t)((kb | k) != 0));
          if (dual) {
            const uint64_t du = umma_smem_desc(a_base + (uint32_t)s * a_bytes + kABytes);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              mma(d0 + (uint32_t)bn, du + (uint64_t)(2 * k), db + (uint64_t)(2 * k), (uint32_t)((kb | k) != 0));
          }
          commit(empty_bar(s));
        }
        commit(tfull_bar(acc));
      }
    }
  } else {
    pdl_wait();
    // 8 epilogue warps: two per TMEM lane quarter, each draining half of the token columns
    const int quarter = warp & 3;
    const int col_half = (warp - 2) >> 2;
    const int ml = quarter * 32 + lane;
    const int c_begin = col_half * (bn >> 1), c_end = c_begin + (bn >> 1);
    int ui = 0;
    for (int u = my_group; u < total_units; u += n_groups, ++ui) {
      const Tile t = tile_of(u);
      const int acc = ui & 1;
      mbar_wait(tfull_bar(acc), ((uint32_t)ui >> 1) & 1u);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)acc * acc_cols;
      const int feat = t.feat0 + ml;
      const bool row_ok = feat < t.feat_limit;
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        if (t.n0 + c0 >= n_rows) break;
        uint32_t r[32], ru[32];
        tmem_ld32(taddr + (uint32_t)c0, r);
        if (dual) tmem_ld32(taddr + (uint32_t)(bn + c0), ru);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (t.n0 + c0 + j >= n_rows) break;
            const int n = row_off + t.n0 + c0 + j;
            const float g = __uint_as_float(r[j]);
            if (dual) {
              float y = silu_mul(g, __uint_as_float(ru[j]));
              if (args.route_w != nullptr) y *= args.route_w[(int64_t)n * args.n_blocks + t.blk];
              args.out_bf16[(int64_t)n * args.ldc + feat] = __float2bfloat16_rn(y);
            } else if (args.out_f32 != nullptr) {
              if (args.row_map != nullptr) args.out_f32[(int64_t)args.row_map[n] * args.ldc + feat] += g;
              else args.out_f32[(int64_t)n * args.ldc + feat] = g;
            } else {
              args.out_bf16[(int64_t)n * args.ldc + feat] = __float2bfloat16_rn(g);
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {                                   // hand the accumulator back to the MMA issuer
        if constexpr (kPair) mbar_arrive_remote(tempty_bar(acc), 0u);
        else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty_bar(acc)) : "memory");
      }
    }
  }
  tcgen05_fence_before();
  if (pair) { __syncwarp(); cluster_sync_all(); } else __syncthreads();
  if (warp == 1) {
    if constexpr (kPair)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(args.tmem_cols)
                   : "memory");
    else
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

size_t smem_needed(int bn, int stages, int dual, int pair) {
  return 1024 /*align slack*/ + (size_t)stages * (kABytes * (size_t)(1 + dual) + (size_t)(bn >> pair) * BK * 2) +
         8 * (2 * kMaxStages + 6);
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

int make_tensor_map_2d(void* out128, const bf16* ptr, int64_t rows, int cols, int64_t pitch, int box_rows) {
  BFLY_TRY(gemm_init());
  return make_map(reinterpret_cast<CUtensorMap*>(out128), ptr, rows, cols, pitch, box_rows);
}

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
  });
  BFLY_TRY(g_init_rc);
  static PerDeviceOnce attr;                       // the shared-memory opt-in is per device
  return attr.run([] {
    constexpr int kSmem = 227 * 1024;
    BFLY_CUDA(cudaFuncSetAttribute(gemm_tn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    BFLY_CUDA(cudaFuncSetAttribute(gemm_tn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    BFLY_CUDA(cudaFuncSetAttribute(gemm_persistent_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    BFLY_CUDA(cudaFuncSetAttribute(gemm_persistent_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    return (int)BFLY_OK;
  });
}

// M = output features (dual: gated features, i.e. half the weight rows).
GemmPlan gemm_plan(int M, int N, int K, int sm_count, int max_splits, int dual, int force_persistent) {
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
  if (force_persistent && N >= 128) {        // grouped MoE: the row count lives on the device, only the persistent
    splits = 1;                              // kernel can follow it
    p.cluster = 0;
  }
  p.splits = splits;
  // Tensor-bound regime (>= 128 tokens per tile, no split): pair adjacent weight tiles with cta_group::2 so
  // each SM reads only half of the token tile from shared memory (a single CTA's 128x256 UMMA needs
  // 192 B/clk of operand reads against 128 B/clk of shared-memory bandwidth).  BFLY_GEMM_PAIR=0 disables.
  static const bool allow_pair = [] { const char* e = getenv("BFLY_GEMM_PAIR"); return !(e && e[0] == '0'); }();
  p.pair = (allow_pair && splits == 1 && !p.cluster && p.bn >= 128 && p.bn % 32 == 0 && m_tiles >= 2) ? 1 : 0;
  // ... and run it persistently with a double-buffered accumulator (2 x bn x (1+dual) <= 512 TMEM columns):
  // token tiles of 256 (plain) / 128 (fused SwiGLU).  BFLY_GEMM_PERSISTENT=0 disables.
  static const bool allow_persist = [] { const char* e = getenv("BFLY_GEMM_PERSISTENT"); return !(e && e[0] == '0'); }();
  p.persistent = 0;
  if ((allow_persist || force_persistent) && splits == 1 && !p.cluster && N >= 128) {
    p.persistent = 1;
    p.bn = dual ? 128 : std::min(256, ((N + 63) / 64) * 64);
    p.pair = (allow_pair && m_tiles >= 2) ? 1 : 0;
  }
  const size_t stage_bytes = kABytes * (size_t)(1 + dual) + (size_t)(p.bn >> p.pair) * BK * 2;
  int stages = (p.bn <= 64) ? ((dual || p.cluster) ? 3 : 4) : (int)std::min<size_t>(kMaxStages, (200 * 1024) / stage_bytes);
  // weight-streaming regime: a 5th stage where two CTAs still share an SM (token tiles <= 32: 2 x 104 KB) -- measured
  // on the isolated kernel (tools/microbench/gemm_probe.py): gate/up 5822 -> 6035 GB/s, LM head 6274 -> 6395 GB/s
  static const int small_stages = [] { const char* e = getenv("BFLY_GEMM_STAGES_SMALL"); return e ? atoi(e) : 0; }();
  if (p.bn <= 64 && !dual && !p.cluster) {
    int want = small_stages > 0 ? small_stages : 5;        // decode step: tp=1 4.744 -> 4.727 ms, tp=8 shard 1.762 -> 1.735 ms
    if (2 * smem_needed(p.bn, want, dual, 0) > 227 * 1024) want = 4;
    stages = want;
  }
  static const int big_stages = [] { const char* e = getenv("BFLY_GEMM_STAGES_BIG"); return e ? atoi(e) : 0; }();
  if (p.bn > 64 && big_stages > 0) stages = big_stages;         // experiment knob: 3 lets two pair-CTAs share an SM
  stages = std::max(2, std::min(stages, kMaxStages));
  p.stages = stages;
  p.smem_bytes = smem_needed(p.bn, stages, dual, p.pair);
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
  const int pair = plan.pair ? 1 : 0;
  BFLY_CHECK(!pair || (plan.splits == 1 && !plan.cluster && plan.bn % 32 == 0), "gemm: CTA pairs need splits == 1, bn % 32 == 0");
  BFLY_CHECK(!plan.persistent || (plan.splits == 1 && !plan.cluster && plan.bn % 64 == 0 && 2 * plan.bn * (1 + dual) <= 512),
             "gemm: the persistent kernel needs splits == 1, bn % 32 == 0 and two accumulators in 512 TMEM columns");
  const size_t smem = smem_needed(plan.bn, plan.stages, dual, pair);
  BFLY_CHECK(smem <= 227 * 1024, "gemm: tile does not fit shared memory");
  BFLY_CHECK(!plan.cluster || (size_t)(1 + dual) * plan.bn * BM * 4 <= (size_t)plan.stages * kABytes * (1 + dual),
             "gemm: cluster reduction buffer does not fit the operand ring");
  CUtensorMap mw, mx;
  const int64_t w_rows = dual ? 2 * (int64_t)plan.M : plan.M;
  const bool grouped = out.n_rows_dev != nullptr;
  BFLY_CHECK(!grouped || plan.persistent, "gemm: device-side row counts need the persistent kernel");
  BFLY_CHECK(grouped || (out.row_off_dev == nullptr && out.row_map == nullptr), "gemm: row_off / row_map need n_rows_dev");
  BFLY_CHECK(out.row_map == nullptr || (out.out_f32 != nullptr && !dual), "gemm: row_map needs the fp32 output");
  BFLY_CHECK(out.ldw == 0 || (out.ldw >= plan.K && out.ldw % 8 == 0), "gemm: ldw >= K and ldw % 8 == 0");
  BFLY_CHECK(out.x_rows == 0 || out.x_rows >= plan.N, "gemm: x_rows >= N");
  BFLY_TRY(make_map(&mw, W, w_rows, plan.K, out.ldw ? out.ldw : plan.K, BM));
  BFLY_TRY(make_map(&mx, X, out.x_rows ? out.x_rows : plan.N, plan.K, ldx, plan.bn >> pair));
  GemmArgs a{};
  a.M = plan.M; a.N = plan.N; a.K = plan.K;
  a.bn = plan.bn; a.stages = plan.stages; a.splits = plan.splits;
  a.cluster = plan.cluster && plan.splits > 1 ? 1 : 0;
  a.dual = dual;
  a.pair = pair;
  // Measured (8B, tp=1): 0 / 6 / 12 / 24 / 64 K-blocks of L2 prefetch give 4.72 / 4.72 / 4.79 / 4.87 / 5.01 ms
  // per step -- the prefetch competes with the HBM-bound kernel it overlaps -- so it is off by default; behind
  // the cross-GPU all-reduce (tp=2 and tp=8, 0..96 MB) it made no measurable difference either.
  static const int l2pf = [] { const char* e = getenv("BFLY_GEMM_L2PF"); return e ? atoi(e) : 0; }();
  a.l2_prefetch = plan.bn <= 64 ? l2pf : 0;          // weight-streaming regime only
  a.f_block = dual ? out.f_block : plan.M;
  a.n_blocks = dual ? out.n_blocks : 1;
  uint32_t cols = 32;
  while ((int)cols < plan.bn * (1 + dual) * (plan.persistent ? 2 : 1)) cols <<= 1;
  a.tmem_cols = cols;
  a.out_f32 = out.out_f32; a.out_bf16 = out.out_bf16; a.route_w = out.route_w;
  a.n_rows_dev = out.n_rows_dev; a.row_off_dev = out.row_off_dev; a.row_map = out.row_map;
  a.split_stride = out.split_stride; a.ldc = out.ldc;
  int m_tiles = ceil_div(plan.M, BM);
  if (dual) m_tiles = out.n_blocks * ceil_div(out.f_block, BM);
  if (pair) m_tiles = (m_tiles + 1) & ~1;        // an odd last tile is paired with an all-out-of-range one
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
  if (a.cluster || pair) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = pair ? 2 : 1;
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = pair ? 1 : plan.splits;
    cfg.numAttrs = 2;
  }
  cfg.attrs = attr;
  if (plan.persistent) {
    int dev = 0, sms = 148;
    BFLY_CUDA(cudaGetDevice(&dev));
    BFLY_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    a.m_units = m_tiles >> pair;
    a.n_tiles = ceil_div(plan.N, plan.bn);
    const int groups = std::min(a.m_units * a.n_tiles, sms >> pair);
    cfg.gridDim = dim3((unsigned)(groups << pair), 1, 1);
    cfg.blockDim = dim3(kPersistThreads);
    if (pair) BFLY_CUDA(cudaLaunchKernelEx(&cfg, gemm_persistent_kernel<true>, mw, mx, a));
    else BFLY_CUDA(cudaLaunchKernelEx(&cfg, gemm_persistent_kernel<false>, mw, mx, a));
  } else if (pair) {
    BFLY_CUDA(cudaLaunchKernelEx(&cfg, gemm_tn_kernel<true>, mw, mx, a));
  } else {
    BFLY_CUDA(cudaLaunchKernelEx(&cfg, gemm_tn_kernel<false>, mw, mx, a));
  }
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
