// Fused decode step (SURVEY.md rows a3..a14 at decode): ONE persistent cooperative kernel per step.
//
// Why: at tensor-parallel degree 4-8 a rank's layer is ~13 us of HBM traffic spread over 10-11 dependent
// kernels; every kernel boundary costs 3-5 us (grid drain, dependency resolution, pipeline ramp), so the
// step is bound by the length of the chain, not by bandwidth.  Here one CTA per SM stays resident for the
// whole step and the kernel boundaries become PHASE COUNTERS in global memory: every CTA adds 1 to the
// phase's counter when its share of the phase is done (release), and whoever needs the phase's results
// waits until the counter equals the grid size (acquire).  Arrival and waiting are separate, so a CTA with
// no work in a small phase (norm, RoPE) runs ahead into the next GEMM phase and streams that phase's WEIGHT
// tiles into its shared-memory ring while the small phase is still running elsewhere -- HBM does not idle
// across a phase boundary.
//
// Phases per layer: norm | QKV GEMM | RoPE + KV append | paged attention | [merge of context splits] |
// attention-out GEMM | norm (+ tensor-parallel all-reduce over NVLink peer memory) | gate/up GEMM | SwiGLU |
// down GEMM; then final norm, LM-head GEMM, logits/argmax keys.
//
// GEMM phases: tcgen05.mma (UMMA M = 128 weight rows, N = tokens), operands by TMA into a 128B-swizzled
// ring, accumulators double-buffered in tensor memory, partitioned STREAM-K (decode_step.h): each CTA gets an
// equal run of (weight tile, K block) units, so 148 SMs stream equal bytes whatever the tile count; partial
// sums go to fp32 slices that the consumer phase folds in K order (deterministic, no atomics).
// Norm phases are split over column chunks (T x chunks items) so ~all SMs take part; a chunk only produces a
// partial sum of squares, so the 1/rms factor is NOT applied to the GEMM operand (xg = x * gamma in bf16):
// it is a per-token scalar and the consumer of the following GEMM multiplies it back in (RoPE, SwiGLU,
// logits).  Attention is the tensor-core kernel of attn_decode.cu run over a static list of (token, kv head,
// context split) items with the KV ring kept full across item boundaries.
//
// Restrictions (the engine falls back to the kernel-per-op path otherwise): RMSNorm + SwiGLU + RoPE models
// without biases (Llama family), head_dim 128, T <= 64 rows, dense FFN.
#include <cuda.h>

#include <algorithm>

#include "decode_step.h"
#include "tcgen05_common.cuh"
#include "attn_common.cuh"

namespace bfly {

namespace {

constexpr int D = 128;                       // head dim
constexpr int kAW = 4;                       // attention warps (warps 4..7)
constexpr int kTile = 32;                    // tokens per KV tile
constexpr int kAStages = 3;
constexpr uint32_t kBlockBytes = kTile * D * 2;              // one K (or V^T) tile
constexpr uint32_t kAttnRingBytes = kAW * kAStages * 2 * kBlockBytes;   // 196608
constexpr uint32_t kSmO = kAttnRingBytes;                                // [warp][8][D] fp32
constexpr uint32_t kSmMl = kSmO + kAW * 8 * D * 4;                       // [warp][8][2] fp32
static_assert(kSmMl + kAW * 8 * 2 * 4 <= kStepRingBytes, "attention buffers exceed the ring");
constexpr int kMaxChunks = 8;
constexpr int kProducerWarps = 3;             // warps 0..2; warp 3 issues the MMAs; warps 4..7 drain the accumulators
constexpr int kRingStages = 10;               // most operand-ring stages (bn = 32: 10 x 20 KB)

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// generic-proxy writes to global memory must be ordered against later async-proxy reads (TMA, bulk copies)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// ---- phase counters ---------------------------------------------------------------------------
// Every CTA arrives exactly once per phase, after all of its threads finished the phase.  Writers order their
// generic-proxy global writes against later async-proxy reads (TMA / bulk copies issued by OTHER CTAs) with
// fence.proxy.async BEFORE the CTA barrier; the release at gpu scope is cumulative over the barrier.  The
// waiting side needs no proxy fence -- and must not have one: after the weight prefetch it would wait for the
// tiles in flight (measured: +3 us per GEMM phase).  Cost of one boundary: ~1.4 us (tools/microbench).
__device__ __forceinline__ void phase_arrive(unsigned* counter, unsigned long long* cta_trace = nullptr) {
  fence_proxy_async();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (cta_trace != nullptr) *cta_trace = globaltimer_ns();
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(counter), "r"(1u) : "memory");
  }
}
// One thread waits until all CTAs of the grid have arrived; bounded (a protocol bug must trap, not hang).
__device__ __forceinline__ void phase_wait(const unsigned* counter, unsigned long long* trace = nullptr) {
  const unsigned target = gridDim.x;
  if (ld_acquire_gpu_u32(counter) < target) {
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while (ld_acquire_gpu_u32(counter) < target) {
      if ((++spins & 255u) == 0u && globaltimer_ns() - t0 > 20000000000ull) __trap();
    }
  }
  if (trace != nullptr && blockIdx.x == 0) *trace = globaltimer_ns();
}
// whole CTA: thread 0 waits, everybody continues after it
__device__ __forceinline__ void phase_wait_cta(const unsigned* counter, unsigned long long* trace = nullptr) {
  if (threadIdx.x == 0) phase_wait(counter, trace);
  __syncthreads();
}

// per-token 1/rms from the chunk sums of squares of a norm phase
__device__ __forceinline__ float rstd_of(const float* ssq, int t, int chunks, int h, float eps) {
  float s = 0.f;
  for (int c = 0; c < chunks; ++c) s += __ldcg(ssq + t * chunks + c);
  return 1.0f / sqrtf(s / (float)h + eps);
}

// Sum of the first n_sl of a GEMM's max_sl fp32 slices at pp (slices `stride` floats apart), added in slice
// (= K) order.  The loads of ALL max_sl slices are issued before the first add and do not depend on n_sl (a
// table lookup that would otherwise sit in front of them): one L2 round trip per batch instead of one per slice.
constexpr int kFoldBatch = 8;
__device__ __forceinline__ float4 fold4(const float* pp, int64_t stride, int n_sl, int max_sl) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int p0 = 0; p0 < max_sl; p0 += kFoldBatch) {
    float4 v[kFoldBatch];
#pragma unroll
    for (int k = 0; k < kFoldBatch; ++k)
      v[k] = (p0 + k < max_sl) ? __ldcg(reinterpret_cast<const float4*>(pp + (int64_t)(p0 + k) * stride)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < kFoldBatch; ++k)
      if (p0 + k < n_sl) { acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w; }
  }
  return acc;
}
__device__ __forceinline__ float2 fold2(const float* pp, int64_t stride, int n_sl, int max_sl) {
  float2 acc = make_float2(0.f, 0.f);
  for (int p0 = 0; p0 < max_sl; p0 += kFoldBatch) {
    float2 v[kFoldBatch];
#pragma unroll
    for (int k = 0; k < kFoldBatch; ++k)
      v[k] = (p0 + k < max_sl) ? __ldcg(reinterpret_cast<const float2*>(pp + (int64_t)(p0 + k) * stride)) : make_float2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k < kFoldBatch; ++k)
      if (p0 + k < n_sl) { acc.x += v[k].x; acc.y += v[k].y; }
  }
  return acc;
}

struct Smem {
  uint32_t ring;        // operand ring / attention buffers
  uint32_t full, empty; // GEMM ring barriers [kRingStages]
  uint32_t tfull, tempty;   // accumulator barriers [2]
  uint32_t abar;        // attention ring barriers [kAW][kAStages]
  uint32_t tmem_slot;
  uint8_t* ring_ptr;
};

struct Pipe {           // running use counts: they fix the mbarrier parities across phases
  uint32_t ring_it;     // K blocks through the GEMM ring so far (this CTA)
  uint32_t acc_it;      // accumulator uses so far
  uint32_t attn_it;     // tiles through this warp's attention ring so far
};

// ---------------------------------------------------------------------------------------------
// GEMM phase: out[slice][n][m] = sum over this CTA's K blocks of W[m][k] * X[n][k]
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void gemm_phase(const StepParams& P, const StepGemm& g, const CUtensorMap* map_w,
                                           const CUtensorMap* map_x, const unsigned* dep, const Smem& sm, Pipe& pipe,
                                           uint32_t tmem_base) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x;
  int u0 = 0, u1 = 0;
  if (c < g.ge) { u0 = sk_lo(g, c); u1 = sk_lo(g, c + 1); }
  const int n_units = u1 - u0;
  const int n_seg = n_units > 0 ? ((u1 - 1) / g.nkb - u0 / g.nkb + 1) : 0;
  const int stages = P.stages, bn = P.bn;
  const uint32_t b_bytes = (uint32_t)bn * BK * 2;
  const uint32_t stage_bytes = kABytes + b_bytes;
  const int chains = (P.tune & 32) ? 4 : ((P.tune & 16) ? 2 : 1);
  const uint64_t pol_w = (P.tune & 1) ? 0x1000000000000000ull : kEvictFirst;     // evict_normal vs evict_first

  if (warp < kProducerWarps) {
    // One thread's TMA issues are serialised (measured: a single producer thread caps an SM at ~33 GB/s whatever the
    // ring depth, two independent pipelines per SM reach ~43 GB/s), so the K blocks are dealt round-robin to
    // n_producers warps, each with its own elected lane.
    const int np = P.n_producers;
    if (lane == 0 && warp < np) {
      auto load_w = [&](int i) {
        const uint32_t it = pipe.ring_it + (uint32_t)i;
        const int s = (int)(it % (uint32_t)stages);
        mbar_wait(sm.empty + 8u * s, ((it / (uint32_t)stages) & 1u) ^ 1u);
        mbar_expect_tx(sm.full + 8u * s, (P.tune & 4) ? kABytes : stage_bytes);
        const int u = u0 + i, tile = u / g.nkb, kb = u - tile * g.nkb;
        tma_load_2d(sm.ring + (uint32_t)s * stage_bytes, map_w, sm.full + 8u * s, kb * BK, tile * BM, pol_w);
      };
      auto load_x = [&](int i) {
        if (P.tune & 4) return;                    // timing experiment: weights only (results are garbage)
        const uint32_t it = pipe.ring_it + (uint32_t)i;
        const int s = (int)(it % (uint32_t)stages);
        const int u = u0 + i, kb = u % g.nkb;
        tma_load_2d(sm.ring + (uint32_t)s * stage_bytes + kABytes, map_x, sm.full + 8u * s, kb * BK, 0, kEvictLast);
      };
      // weights do not depend on the previous phase: fill the ring with weight tiles first, then wait
      const int pre = min(stages, n_units);
      for (int i = warp; i < pre; i += np) load_w(i);
      if (dep != nullptr) phase_wait(dep, (P.trace && warp == 0) ? P.trace + (dep - P.counters) + 1 : nullptr);
      for (int i = warp; i < pre; i += np) load_x(i);
      int i = warp;
      while (i < pre) i += np;
      for (; i < n_units; i += np) {
        load_w(i);
        load_x(i);
      }
    }
  } else if (warp == kProducerWarps) {
    if (lane == 0 && n_units > 0) {
      // instruction descriptor: D = f32, A = B = bf16, both K-major, N = bn, M = 128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      uint32_t acc_it = pipe.acc_it;
      for (int i = 0; i < n_units; ++i) {
        const int u = u0 + i, tile = u / g.nkb, kb = u - tile * g.nkb;
        const bool seg_first = (i == 0) || (kb == 0);
        const bool seg_last = (i == n_units - 1) || (kb == g.nkb - 1);
        const uint32_t acc = acc_it & 1u;
        if (seg_first) {
          mbar_wait(sm.tempty + 8u * acc, ((acc_it >> 1) & 1u) ^ 1u);       // epilogue drained this accumulator
          tcgen05_fence_after();
        }
        const uint32_t it = pipe.ring_it + (uint32_t)i;
        const int s = (int)(it % (uint32_t)stages);
        mbar_wait(sm.full + 8u * s, (it / (uint32_t)stages) & 1u);
        tcgen05_fence_after();
        const uint64_t da = umma_smem_desc(sm.ring + (uint32_t)s * stage_bytes);
        const uint64_t db = umma_smem_desc(sm.ring + (uint32_t)s * stage_bytes + kABytes);
        // independent accumulation chains: consecutive MMAs do not wait on each other's accumulator
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)(bn * chains);
        if (P.tune & 8) {                           // timing experiment: no MMAs, the stage is handed straight back
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(sm.empty + 8u * s) : "memory");
          if (seg_last) {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(sm.tfull + 8u * acc) : "memory");
            ++acc_it;
          }
          continue;
        }
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k)
          umma_bf16(d_tmem + (uint32_t)((k & (chains - 1)) * bn), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                    (uint32_t)(!seg_first || k >= chains));
        umma_commit(sm.empty + 8u * s);
        if (seg_last) {
          umma_commit(sm.tfull + 8u * acc);
          ++acc_it;
        }
      }
    }
  } else if (n_units > 0) {
    const int quarter = warp & 3;
    const int ml = quarter * 32 + lane;
    uint32_t acc_it = pipe.acc_it;
    int u = u0;
    while (u < u1) {
      const int tile = u / g.nkb, kb0 = u - tile * g.nkb;
      const int len = min(g.nkb - kb0, u1 - u);
      const int slice = c - (int)g.first[tile];
      const uint32_t acc = acc_it & 1u;
      mbar_wait(sm.tfull + 8u * acc, (acc_it >> 1) & 1u);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * (uint32_t)(bn * chains);
      const int m = tile * BM + ml;
      float* dst = g.out + ((int64_t)slice * P.T) * g.M + m;
      for (int c0 = 0; c0 < bn; c0 += 16) {
        if (c0 >= P.T) break;
        uint32_t r[16];
        tmem_ld16(taddr + (uint32_t)c0, r);
        tmem_ld_wait();
        for (int ch = 1; ch < chains; ++ch) {       // fold the accumulation chains in chain order
          uint32_t r2[16];
          tmem_ld16(taddr + (uint32_t)(ch * bn + c0), r2);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
        }
        if (m < g.M) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (c0 + j < P.T) dst[(int64_t)(c0 + j) * g.M] = __uint_as_float(r[j]);
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(sm.tempty + 8u * acc) : "memory");
      ++acc_it;
      u += len;
    }
  }
  pipe.ring_it += (uint32_t)n_units;
  pipe.acc_it += (uint32_t)n_seg;
}

// ---------------------------------------------------------------------------------------------
// Norm phase over (row, column chunk) items: residual += folded slices of the previous row-parallel GEMM
// (summed over the tensor-parallel ranks through peer memory when ar_call >= 0); xg = bf16(residual * gamma);
// ssq[row][chunk] = sum of squares of the chunk.  `prev` == nullptr: the row comes from the embedding table.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void norm_phase(const StepParams& P, const StepGemm* prev, const bf16* gamma, float* ssq,
                                           int ar_call, float* red) {
  const int C = P.chunks, h = P.h, cw = h / C, n_items = P.T * C;
  const int tid = threadIdx.x;
  const bool ar = ar_call >= 0 && P.ar_on;
  const int tp = P.peers.tp, me = P.peers.rank;
  unsigned epoch = 0;
  int parity = 0;
  if (ar) {
    epoch = __ldcg(P.ar_step) * (unsigned)(2 * P.n_layers) + (unsigned)ar_call + 1u;
    parity = ar_call & 1;
  }
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int row = item / C, ch = item - row * C;
    const int col0 = ch * cw;
    float* res = P.residual + (int64_t)row * h;
    if (prev == nullptr) {
      int tok = P.tokens[row];
      if (ch == 0 && tid == 0) {
        if (P.history != nullptr) P.history[(int64_t)(*P.step_counter) * P.T + row] = tok;
        P.keys[row] = 0ull;
      }
      tok = min(max(tok, 0), P.vocab - 1);
      const bf16* erow = P.embed + (int64_t)tok * h;
      float ss = 0.f;
      for (int i = col0 + tid * 4; i < col0 + cw; i += kStepThreads * 4) {
        const uint2 e = *reinterpret_cast<const uint2*>(erow + i);
        const float4 v = make_float4(bf16_lo(e.x), bf16_hi(e.x), bf16_lo(e.y), bf16_hi(e.y));
        *reinterpret_cast<float4*>(res + i) = v;
        const uint2 gm = *reinterpret_cast<const uint2*>(gamma + i);
        *reinterpret_cast<uint2*>(P.xg + (int64_t)row * h + i) =
            make_uint2(pack_bf16(v.x * bf16_lo(gm.x), v.y * bf16_hi(gm.x)), pack_bf16(v.z * bf16_lo(gm.y), v.w * bf16_hi(gm.y)));
        ss += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
      }
      ss = block_sum(ss, red);
      if (tid == 0) ssq[row * C + ch] = ss;
      continue;
    }
    const int64_t slice_stride = (int64_t)P.T * prev->M;
    const int64_t slot_me = (int64_t)(parity * tp + me) * P.peers.max_rows + row;
    if (ar) {
      // 1) fold this rank's slices of the chunk and push it into every rank's receive buffer
      for (int i = col0 + tid * 4; i < col0 + cw; i += kStepThreads * 4) {
        const float4 v = fold4(prev->out + (int64_t)row * prev->M + i, slice_stride, prev->nsl[i >> 7], prev->max_slices);
        for (int r = 0; r < tp; ++r) *reinterpret_cast<float4*>(P.peers.recv[r] + slot_me * h + i) = v;
      }
      __syncthreads();        // orders every thread's pushes before the release stores of the flags
      if (tid < tp) st_release_sys_u32(P.peers.cflags[tid] + slot_me * kMaxChunks + ch, epoch);
      if (tid < tp) {
        const unsigned* f = P.peers.cflags[me] + ((int64_t)(parity * tp + tid) * P.peers.max_rows + row) * kMaxChunks + ch;
        if (ld_acquire_sys_u32(f) != epoch) {
          const unsigned long long t0 = globaltimer_ns();
          unsigned spins = 0;
          while (ld_acquire_sys_u32(f) != epoch)
            if ((++spins & 1023u) == 0u && globaltimer_ns() - t0 > 60000000000ull) __trap();
        }
      }
      __syncthreads();
    }
    // 2) residual += sum; operand and partial sum of squares
    float ss = 0.f;
    for (int i = col0 + tid * 4; i < col0 + cw; i += kStepThreads * 4) {
      float4 v = *reinterpret_cast<const float4*>(res + i);
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ar) {
        for (int r = 0; r < tp; ++r) {        // rank order: every rank computes the same bits
          const float4 x = ld_volatile_f4(P.peers.recv[me] + ((int64_t)(parity * tp + r) * P.peers.max_rows + row) * h + i);
          a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
        }
      } else {
        a = fold4(prev->out + (int64_t)row * prev->M + i, slice_stride, prev->nsl[i >> 7], prev->max_slices);
      }
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
      *reinterpret_cast<float4*>(res + i) = v;
      const uint2 gm = *reinterpret_cast<const uint2*>(gamma + i);
      *reinterpret_cast<uint2*>(P.xg + (int64_t)row * h + i) =
          make_uint2(pack_bf16(v.x * bf16_lo(gm.x), v.y * bf16_hi(gm.x)), pack_bf16(v.z * bf16_lo(gm.y), v.w * bf16_hi(gm.y)));
      ss += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    }
    ss = block_sum(ss, red);
    if (tid == 0) ssq[row * C + ch] = ss;
  }
}

// ---------------------------------------------------------------------------------------------
// RoPE + KV append: one warp per (token, head of q|k|v); D = 128 so a head is exactly one weight tile
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void rope_phase(const StepParams& P, const StepLayer& L, const float* ssq) {
  const StepGemm& g = P.gemm[SG_QKV];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int heads = P.nq + 2 * P.nk, n_items = P.T * heads;
  const int64_t slice_stride = (int64_t)P.T * g.M;
  for (int item = blockIdx.x * (kStepThreads / 32) + warp; item < n_items; item += gridDim.x * (kStepThreads / 32)) {
    const int t = item / heads, head = item - t * heads;
    // the lane owns dims (2 lane, 2 lane + 1) of the first half and their rotation partners 64 further on
    const float* src = g.out + (int64_t)t * g.M + head * D + 2 * lane;
    const int n_sl = g.nsl[head];
    float2 lo = fold2(src, slice_stride, n_sl, g.max_slices), hi = fold2(src + D / 2, slice_stride, n_sl, g.max_slices);
    const float rstd = rstd_of(ssq, t, P.chunks, P.h, P.eps);
    const int pos = P.seq_lens[t];
    lo.x *= rstd; lo.y *= rstd; hi.x *= rstd; hi.y *= rstd;
    if (head < P.nq + P.nk) {
      const float4 cs = *reinterpret_cast<const float4*>(P.rope_table + (int64_t)pos * (D / 2) + 2 * lane);   // (cos, sin) x 2
      const float a0 = lo.x * cs.x - hi.x * cs.y, b0 = hi.x * cs.x + lo.x * cs.y;
      const float a1 = lo.y * cs.z - hi.y * cs.w, b1 = hi.y * cs.z + lo.y * cs.w;
      lo = make_float2(a0, a1);
      hi = make_float2(b0, b1);
    }
    if (head < P.nq) {
      bf16* dst = P.q + ((int64_t)t * P.nq + head) * D + 2 * lane;
      *reinterpret_cast<uint32_t*>(dst) = pack_bf16(lo.x, lo.y);
      *reinterpret_cast<uint32_t*>(dst + D / 2) = pack_bf16(hi.x, hi.y);
    } else {
      const int which = head < P.nq + P.nk ? 0 : 1;
      const int kh = (head - P.nq) - which * P.nk;
      const int page = P.page_table[(int64_t)t * P.pages_per_seq + pos / P.page_size];
      const int slot = pos % P.page_size;
      bf16* blk = L.kv + (((int64_t)page * 2 + which) * P.nk + kh) * P.page_size * D;
      if (which == 0) {           // K row: the chunk swizzle keeps the two dims of a pair adjacent
        *reinterpret_cast<uint32_t*>(blk + kv_k_offset(slot, 2 * lane, D)) = pack_bf16(lo.x, lo.y);
        *reinterpret_cast<uint32_t*>(blk + kv_k_offset(slot, 2 * lane + D / 2, D)) = pack_bf16(hi.x, hi.y);
      } else {                    // V^T: consecutive dims are 32 tokens apart
        blk[kv_v_offset(slot, 2 * lane, D)] = __float2bfloat16_rn(lo.x);
        blk[kv_v_offset(slot, 2 * lane + 1, D)] = __float2bfloat16_rn(lo.y);
        blk[kv_v_offset(slot, 2 * lane + D / 2, D)] = __float2bfloat16_rn(hi.x);
        blk[kv_v_offset(slot, 2 * lane + 1 + D / 2, D)] = __float2bfloat16_rn(hi.y);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Paged decode attention over (token, kv head, context split) items; see attn_decode.cu for the math layout.
// Warps 2..5 each stream their tiles of the CTA's item list through a private 3-stage ring that stays full
// across item boundaries.
// ---------------------------------------------------------------------------------------------
struct AttnCur {
  int item, j, my_tiles, tile_begin, t, kvh, split, ctx;
};

__device__ __forceinline__ void attn_phase(const StepParams& P, const StepLayer& L, const Smem& sm, Pipe& pipe) {
  const int warp = threadIdx.x >> 5;
  if (warp < 4) return;
  const int aw = warp - 4, lane = threadIdx.x & 31, atid = threadIdx.x - 128;
  const int r = lane >> 2, c = lane & 3;
  const int ns = P.attn_splits, nk = P.nk, nq = P.nq, G = nq / nk;
  const int n_items = P.T * nk * ns;
  const int ps = P.page_size;
  const int64_t head_stride = (int64_t)ps * D;
  const int64_t page_stride = 2 * (int64_t)nk * head_stride;
  const uint32_t my_ring = sm.ring + (uint32_t)aw * kAStages * 2 * kBlockBytes;
  const uint32_t my_bars = sm.abar + (uint32_t)aw * kAStages * 8;
  float* sm_o = reinterpret_cast<float*>(sm.ring_ptr + kSmO);
  float* sm_ml = reinterpret_cast<float*>(sm.ring_ptr + kSmMl);

  auto load_item = [&](AttnCur& q) {
    const int tk = q.item / ns;
    q.split = q.item - tk * ns;
    q.t = tk / nk;
    q.kvh = tk - q.t * nk;
    q.ctx = P.seq_lens[q.t] + 1;
    const int n_tiles = (q.ctx + kTile - 1) / kTile;
    const int tps = (n_tiles + ns - 1) / ns;
    q.tile_begin = q.split * tps;
    const int tile_end = min(n_tiles, q.tile_begin + tps);
    const int span = tile_end - q.tile_begin - aw;
    q.my_tiles = span > 0 ? (span + kAW - 1) / kAW : 0;
    q.j = 0;
  };
  AttnCur ic, cc;
  ic.item = cc.item = blockIdx.x;
  if (ic.item < n_items) { load_item(ic); cc = ic; }
  uint32_t issued = pipe.attn_it, consumed = pipe.attn_it;

  auto pump = [&]() {          // keep the ring full: issue tiles of this and the following items
    while (ic.item < n_items && issued < consumed + kAStages) {
      if (ic.j < ic.my_tiles) {
        if (lane == 0) {
          const int tile = ic.tile_begin + aw + ic.j * kAW;
          const int tok0 = tile * kTile;
          const int page = P.page_table[(int64_t)ic.t * P.pages_per_seq + tok0 / ps];
          const int sub = (tok0 % ps) / kTile;
          const bf16* kp = L.kv + (int64_t)page * page_stride + (int64_t)ic.kvh * head_stride + (int64_t)sub * kTile * D;
          const bf16* vp = kp + (int64_t)nk * head_stride;
          const uint32_t s = issued % kAStages;
          const uint32_t bar = my_bars + 8u * s;
          const uint32_t dst = my_ring + s * 2 * kBlockBytes;
          mbar_expect_tx(bar, 2 * kBlockBytes);
          bulk_copy_g2s(dst, kp, kBlockBytes, bar);
          bulk_copy_g2s(dst + kBlockBytes, vp, kBlockBytes, bar);
        }
        ++issued;
        ++ic.j;
      } else {
        ic.item += gridDim.x;
        if (ic.item < n_items) load_item(ic);
      }
    }
  };
  pump();

  for (; cc.item < n_items; cc.item += gridDim.x) {
    if (cc.item != (int)blockIdx.x) load_item(cc);
    const int t = cc.t, kvh = cc.kvh, ctx = cc.ctx;
    // Q as B fragments: head r (< G), the d-permutation of the K chunks (chunk c of every 64 bytes)
    uint32_t qb[D / 32][4];
#pragma unroll
    for (int j = 0; j < D / 32; ++j) {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (r < G) v = __ldcg(reinterpret_cast<const uint4*>(P.q + ((int64_t)t * nq + kvh * G + r) * D + c * 8 + 32 * j));
      qb[j][0] = v.x; qb[j][1] = v.y; qb[j][2] = v.z; qb[j][3] = v.w;
    }
    float oacc[D / 16][4];
#pragma unroll
    for (int mo = 0; mo < D / 16; ++mo)
#pragma unroll
      for (int e = 0; e < 4; ++e) oacc[mo][e] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    for (int i = 0; i < cc.my_tiles; ++i) {
      const uint32_t s = consumed % kAStages;
      mbar_wait(my_bars + 8u * s, (consumed / kAStages) & 1u);
      const uint32_t k_s = my_ring + s * 2 * kBlockBytes;
      const uint32_t v_s = k_s + kBlockBytes;
      const int tok0 = (cc.tile_begin + aw + i * kAW) * kTile;
      float sacc[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) sacc[mt][e] = 0.f;
        const int ta = 8 * (r >> 1) + 4 * mt + (r & 1), tb = ta + 2;
        const uint32_t swz = (uint32_t)((ta & 1) << 2);
#pragma unroll
        for (int j = 0; j < D / 32; ++j) {
          const uint32_t chunk = ((uint32_t)(c + 4 * j)) ^ swz;
          const uint4 x = lds128(k_s + (uint32_t)ta * (D * 2) + chunk * 16);
          const uint4 y = lds128(k_s + (uint32_t)tb * (D * 2) + chunk * 16);
          mma_bf16(sacc[mt], x.x, y.x, x.y, y.y, qb[j][0], qb[j][1]);
          mma_bf16(sacc[mt], x.z, y.z, x.w, y.w, qb[j][2], qb[j][3]);
        }
      }
      float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const int ta = tok0 + 8 * (r >> 1) + 4 * mt + (r & 1);
        const bool va = ta < ctx, vb = (ta + 2) < ctx;
        sacc[mt][0] = va ? sacc[mt][0] * P.attn_scale_log2 : -INFINITY;
        sacc[mt][1] = va ? sacc[mt][1] * P.attn_scale_log2 : -INFINITY;
        sacc[mt][2] = vb ? sacc[mt][2] * P.attn_scale_log2 : -INFINITY;
        sacc[mt][3] = vb ? sacc[mt][3] * P.attn_scale_log2 : -INFINITY;
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
        const float m_new = fmaxf(m_run[e], tmax[e]);
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
        pb[mt][0] = movmatrix_trans(pack_bf16(p0, p1));
        pb[mt][1] = movmatrix_trans(pack_bf16(p2, p3));
      }
      l_run[0] = l_run[0] * corr[0] + psum[0];
      l_run[1] = l_run[1] * corr[1] + psum[1];
#pragma unroll
      for (int mo = 0; mo < D / 16; ++mo) {
        oacc[mo][0] *= corr[0]; oacc[mo][1] *= corr[1]; oacc[mo][2] *= corr[0]; oacc[mo][3] *= corr[1];
        const uint4 x = lds128(v_s + (uint32_t)(16 * mo + r) * (kTile * 2) + c * 16);
        const uint4 y = lds128(v_s + (uint32_t)(16 * mo + r + 8) * (kTile * 2) + c * 16);
        mma_bf16(oacc[mo], x.x, y.x, x.y, y.y, pb[0][0], pb[0][1]);
        mma_bf16(oacc[mo], x.z, y.z, x.w, y.w, pb[1][0], pb[1][1]);
      }
      __syncwarp();                       // every lane is done reading stage s
      ++consumed;
      pump();
    }

    // ---- merge the four warps of the item ----
#pragma unroll
    for (int off = 4; off < 32; off <<= 1) {
      l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], off);
      l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], off);
    }
    if (r == 0) {
      sm_ml[(aw * 8 + 2 * c) * 2] = m_run[0];     sm_ml[(aw * 8 + 2 * c) * 2 + 1] = l_run[0];
      sm_ml[(aw * 8 + 2 * c + 1) * 2] = m_run[1]; sm_ml[(aw * 8 + 2 * c + 1) * 2 + 1] = l_run[1];
    }
#pragma unroll
    for (int mo = 0; mo < D / 16; ++mo) {
      float* o0 = sm_o + ((size_t)aw * 8 + 2 * c) * D + 16 * mo + r;
      o0[0] = oacc[mo][0]; o0[D] = oacc[mo][1]; o0[8] = oacc[mo][2]; o0[D + 8] = oacc[mo][3];
    }
    named_bar_sync(1, kAW * 32);
    for (int idx = atid; idx < G * D; idx += kAW * 32) {
      const int g = idx / D, dd = idx - g * D;
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kAW; ++w) M = fmaxf(M, sm_ml[(w * 8 + g) * 2]);
      float Lsum = 0.f, O = 0.f;
      if (M > -INFINITY) {
#pragma unroll
        for (int w = 0; w < kAW; ++w) {
          const float mw = sm_ml[(w * 8 + g) * 2];
          if (mw > -INFINITY) {
            const float f = exp2f(mw - M);
            Lsum += sm_ml[(w * 8 + g) * 2 + 1] * f;
            O += sm_o[((size_t)w * 8 + g) * D + dd] * f;
          }
        }
      }
      if (ns == 1) {
        P.attn_out[((int64_t)t * nq + kvh * G + g) * D + dd] = __float2bfloat16_rn(O / Lsum);
      } else {
        const int64_t slot = ((int64_t)t * nq + kvh * G + g) * ns + cc.split;
        P.part_o[slot * D + dd] = O;
        if (dd == 0) { P.part_ml[slot * 2] = M; P.part_ml[slot * 2 + 1] = Lsum; }
      }
    }
    named_bar_sync(1, kAW * 32);          // the merge buffers are reused by the next item
  }
  pipe.attn_it = consumed;
}

// merge of the context splits: one warp per (token, head)
__device__ __forceinline__ void combine_phase(const StepParams& P) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ns = P.attn_splits, n_items = P.T * P.nq;
  for (int item = blockIdx.x * (kStepThreads / 32) + warp; item < n_items; item += gridDim.x * (kStepThreads / 32)) {
    const int64_t th = item;
    // every split wrote its slot (an empty one as O = 0, l = 0, m = -inf): all loads go out at once
    float ms[16], ls[16];
    float4 o[16];
#pragma unroll
    for (int s2 = 0; s2 < 16; ++s2) {
      if (s2 < ns) {
        const float2 ml = __ldcg(reinterpret_cast<const float2*>(P.part_ml + (th * ns + s2) * 2));
        ms[s2] = ml.x; ls[s2] = ml.y;
        o[s2] = __ldcg(reinterpret_cast<const float4*>(P.part_o + (th * ns + s2) * D + 4 * lane));
      } else {
        ms[s2] = -INFINITY; ls[s2] = 0.f; o[s2] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    float M = -INFINITY;
#pragma unroll
    for (int s2 = 0; s2 < 16; ++s2) M = fmaxf(M, ms[s2]);
    float Lsum = 0.f;
    float4 O = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int s2 = 0; s2 < 16; ++s2) {
      const float f = ms[s2] > -INFINITY ? exp2f(ms[s2] - M) : 0.f;
      Lsum += ls[s2] * f;
      O.x += o[s2].x * f; O.y += o[s2].y * f; O.z += o[s2].z * f; O.w += o[s2].w * f;
    }
    const float inv = 1.0f / Lsum;
    *reinterpret_cast<uint2*>(P.attn_out + th * D + 4 * lane) = make_uint2(pack_bf16(O.x * inv, O.y * inv), pack_bf16(O.z * inv, O.w * inv));
  }
}

// SwiGLU over the folded gate/up slices: act[t][j] = silu(rstd * gate) * (rstd * up)
__device__ __forceinline__ void act_phase(const StepParams& P, const float* ssq) {
  const StepGemm& g = P.gemm[SG_GU];
  const int f = P.f_local, f4 = f >> 2;
  const int64_t n_items = (int64_t)P.T * f4;
  const int64_t slice_stride = (int64_t)P.T * g.M;
  for (int64_t idx = (int64_t)blockIdx.x * kStepThreads + threadIdx.x; idx < n_items; idx += (int64_t)gridDim.x * kStepThreads) {
    const int t = (int)(idx / f4), j = (int)(idx - (int64_t)t * f4) * 4;
    const float* pg = g.out + (int64_t)t * g.M + j;
    float4 a = fold4(pg, slice_stride, g.nsl[j >> 7], g.max_slices);
    float4 b = fold4(pg + f, slice_stride, g.nsl[(f + j) >> 7], g.max_slices);
    const float rstd = rstd_of(ssq, t, P.chunks, P.h, P.eps);
    a.x *= rstd; a.y *= rstd; a.z *= rstd; a.w *= rstd;
    b.x *= rstd; b.y *= rstd; b.z *= rstd; b.w *= rstd;
    const float y0 = a.x / (1.0f + expf(-a.x)) * b.x, y1 = a.y / (1.0f + expf(-a.y)) * b.y;
    const float y2 = a.z / (1.0f + expf(-a.z)) * b.z, y3 = a.w / (1.0f + expf(-a.w)) * b.w;
    *reinterpret_cast<uint2*>(P.act + (int64_t)t * f + j) = make_uint2(pack_bf16(y0, y1), pack_bf16(y2, y3));
  }
}

__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
  uint32_t u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)idx);
}

// logits = rstd * folded LM-head slices; exact greedy key per token (elementwise.cu: lm_finalize_kernel)
__device__ __forceinline__ void finalize_phase(const StepParams& P, const float* ssq) {
  const StepGemm& g = P.gemm[SG_HEAD];
  constexpr int kChunk = 512;                        // columns per warp item: 4 float4 per lane, all in flight at once
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunks = (P.v_local + kChunk - 1) / kChunk, n_items = P.T * chunks;
  const int64_t slice_stride = (int64_t)P.T * g.M;
  const bool vec = (g.M & 3) == 0;                   // rows of the slices are 16-byte aligned
  for (int item = blockIdx.x * (kStepThreads / 32) + warp; item < n_items; item += gridDim.x * (kStepThreads / 32)) {
    const int t = item / chunks, begin = (item - t * chunks) * kChunk;
    const float rstd = rstd_of(ssq, t, P.chunks, P.h, P.eps);
    unsigned long long best = 0ull;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int j = begin + (k * 32 + lane) * 4;
      if (j >= P.v_local) continue;
      const float* pp = g.out + (int64_t)t * g.M + j;
      float v[4];
      if (vec && j + 3 < P.v_local) {
        const float4 x = fold4(pp, slice_stride, g.nsl[j >> 7], g.max_slices);
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          v[e] = 0.f;
          if (j + e < P.v_local) {
            const int n_sl = g.nsl[(j + e) >> 7];
            for (int p = 0; p < n_sl; ++p) v[e] += __ldcg(pp + e + p * slice_stride);
          }
        }
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (j + e < P.v_local) {
          const float x = v[e] * rstd;
          if (P.logits != nullptr) P.logits[(int64_t)t * P.logits_ld + j + e] = x;
          const unsigned long long key = argmax_key(x, P.v0 + j + e);
          best = key > best ? key : best;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other > best ? other : best;
    }
    if (lane == 0) atomicMax(P.keys + t, best);
  }
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kStepThreads, 1) decode_step_kernel(const __grid_constant__ StepParams P) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ float red[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  Smem sm;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  sm.ring = base;
  sm.ring_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bars = base + kStepRingBytes;
  sm.full = bars;
  sm.empty = bars + 8u * kRingStages;
  sm.tfull = bars + 8u * (2 * kRingStages);
  sm.tempty = bars + 8u * (2 * kRingStages + 2);
  sm.abar = bars + 8u * (2 * kRingStages + 4);
  sm.tmem_slot = bars + 8u * (2 * kRingStages + 4 + kAW * kAStages);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kRingStages; ++s) {
      mbar_init(sm.full + 8u * s, 1);
      mbar_init(sm.empty + 8u * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(sm.tfull + 8u * a, 1);
      mbar_init(sm.tempty + 8u * a, 4);                    // one arrival per epilogue warp
    }
    for (int i = 0; i < kAW * kAStages; ++i) mbar_init(sm.abar + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kProducerWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sm.tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(P.maps);
    for (int i = 0; i < 4; ++i) asm volatile("prefetch.tensormap [%0];" ::"l"(maps + P.n_layers * 4 + i) : "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(sm.tmem_slot) : "memory");

  const CUtensorMap* maps = reinterpret_cast<const CUtensorMap*>(P.maps);
  const CUtensorMap* map_head = maps + P.n_layers * 4;
  const CUtensorMap* map_xg = map_head + 1;
  const CUtensorMap* map_ao = map_head + 2;
  const CUtensorMap* map_act = map_head + 3;
  const int TC = P.T * P.chunks;
  float* ssq_attn = P.ssq;
  float* ssq_ffn = P.ssq + TC;
  float* ssq_fin = P.ssq + 2 * TC;
  Pipe pipe{0u, 0u, 0u};
  unsigned* cnt = P.counters;      // cnt[-1] of a phase = the counter of the phase before it
  const bool split = P.attn_splits > 1;

  const int n_phases = P.n_layers * (9 + (split ? 1 : 0)) + 2;
  if (P.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) P.trace[0] = globaltimer_ns();
  for (int l = 0; l < P.n_layers; ++l) {
    const StepLayer L = P.layers[l];
    // ---- attention norm (layer 0: embedding) ----
    if (l > 0) phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    norm_phase(P, l == 0 ? nullptr : &P.gemm[SG_DOWN], L.attn_norm_w, ssq_attn, l == 0 ? -1 : 2 * l - 1, red);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- QKV ----
    gemm_phase(P, P.gemm[SG_QKV], maps + l * 4 + 0, map_xg, cnt - 1, sm, pipe, tmem_base);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- RoPE + KV append ----
    phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    rope_phase(P, L, ssq_attn);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- attention ----
    phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    attn_phase(P, L, sm, pipe);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    if (split) {
      phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
      combine_phase(P);
      phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    }
    // ---- attention out ----
    gemm_phase(P, P.gemm[SG_WO], maps + l * 4 + 1, map_ao, cnt - 1, sm, pipe, tmem_base);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- ffn norm (+ all-reduce) ----
    phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    norm_phase(P, &P.gemm[SG_WO], L.ffn_norm_w, ssq_ffn, 2 * l, red);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- gate / up ----
    gemm_phase(P, P.gemm[SG_GU], maps + l * 4 + 2, map_xg, cnt - 1, sm, pipe, tmem_base);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- SwiGLU ----
    phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    act_phase(P, ssq_ffn);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    // ---- down ----
    gemm_phase(P, P.gemm[SG_DOWN], maps + l * 4 + 3, map_act, cnt - 1, sm, pipe, tmem_base);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
  }
  // ---- final norm, LM head, logits / greedy keys ----
  phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
  norm_phase(P, &P.gemm[SG_DOWN], P.final_norm_w, ssq_fin, 2 * P.n_layers - 1, red);
  phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
  if (P.v_local > 0) {
    if (P.tune & 2)
      gemm_phase(P, P.gemm[SG_HEAD], reinterpret_cast<const CUtensorMap*>(P.map_head_param),
                 reinterpret_cast<const CUtensorMap*>(P.map_xg_param), cnt - 1, sm, pipe, tmem_base);
    else
      gemm_phase(P, P.gemm[SG_HEAD], map_head, map_xg, cnt - 1, sm, pipe, tmem_base);
    phase_arrive(cnt, P.trace ? P.trace + n_phases + 2 + (size_t)(cnt - P.counters) * gridDim.x + blockIdx.x : nullptr); ++cnt;
    phase_wait_cta(cnt - 1, P.trace ? P.trace + (cnt - 1 - P.counters) + 1 : nullptr);
    finalize_phase(P, ssq_fin);
    __syncthreads();
    if (P.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) P.trace[(cnt - P.counters) + 1] = globaltimer_ns();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == kProducerWarps)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

constexpr size_t kStepSmem = 1024 + kStepRingBytes + 8 * (2 * kRingStages + 4 + kAW * kAStages) + 64;

}  // namespace

int step_phase_count(int n_layers, int attn_splits) { return n_layers * (9 + (attn_splits > 1 ? 1 : 0)) + 2; }

int decode_step_init() {
  static PerDeviceOnce attr;
  return attr.run([] {
    BFLY_CUDA(cudaFuncSetAttribute(decode_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStepSmem));
    return (int)BFLY_OK;
  });
}

int launch_decode_step(const StepParams& p, int grid, cudaStream_t stream) {
  BFLY_TRY(decode_step_init());
  BFLY_CHECK(p.T >= 1 && p.T <= kStepMaxTokens && p.bn % 16 == 0 && p.bn >= p.T && p.bn <= 64, "decode_step: T in [1, 64]");
  BFLY_CHECK(p.stages >= 2 && p.stages <= kRingStages && (size_t)p.stages * (kABytes + (size_t)p.bn * BK * 2) <= kStepRingBytes,
             "decode_step: ring does not fit");
  BFLY_CHECK(p.chunks >= 1 && p.chunks <= kMaxChunks && p.h % (p.chunks * 128) == 0, "decode_step: chunks");
  BFLY_CHECK(p.nq % p.nk == 0 && p.nq / p.nk <= 8 && p.page_size % kTile == 0, "decode_step: attention geometry");
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kStepThreads);
  cfg.dynamicSmemBytes = kStepSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;           // all CTAs co-resident: the phase counters need it
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BFLY_CUDA(cudaLaunchKernelEx(&cfg, decode_step_kernel, p));
  count_launch();
  return BFLY_OK;
}

}  // namespace bfly
