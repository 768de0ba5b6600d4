// C1 fused into K3 for decode: the partial-sum all-reduce after a row-parallel GEMM, the residual
// add and the following norm in ONE kernel over NVLink peer memory (SURVEY.md rows a10 + a4).
//
// Every rank runs this kernel with one CTA per token row.  The CTA folds its rank's split-K slices of
// the row, PUSHES the fp32 row into a slot of every rank's receive buffer (peer-mapped stores through
// NVSwitch), publishes an epoch flag next to it with release semantics at system scope, waits until
// the same row has arrived from every rank, and then sums the tp contributions from its OWN memory in
// rank order -- so every rank computes a bit-identical residual -- adds the bias, normalises and writes
// the next GEMM's bf16 operand.  Synchronisation is per row (no grid barrier); buffers are double
// buffered by call parity and flags carry a monotonically increasing epoch, so nothing is ever reset.
// A rank can be at most one call ahead of a peer (its next call needs that peer's flag), which is what
// makes two parities sufficient.  Latency regime only (T <= max_batch rows); prefill keeps NCCL.
//
// Two modes.  ONE-SHOT (small rows, e.g. 32 x 4096): every rank pushes its row to all tp ranks -- one
// NVLink hop, tp x the bytes.  TWO-SHOT (T*h*4 >= 1 MiB and tp >= 4, e.g. the 64 x 8192 rows of the 70B
// shape, where one-shot spends ~38 us per call moving 14 MB per rank): row r is OWNED by rank r % tp;
// the others push their partial row to the owner only, the owner sums in rank order and pushes the sum
// back to everybody -- two hops, 2*(tp-1)/tp of a row per rank instead of tp-1 rows.  Both modes use the
// same buffers: slot [parity][src][row] of the owner holds src's partial; slot [parity][owner][row] of a
// non-owner (unused by the first hop) receives the reduced row.  Either way every rank adds the same
// fp32 sum to its replicated residual, so the ranks stay bit-identical.
#include <algorithm>
#include <cstdlib>

#include "tcgen05_common.cuh"

namespace bfly {

namespace {

__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// spin until a peer has published this call's epoch; a lost peer must not hang the GPU
// (60 s of wall clock: ranks may legitimately be seconds apart, e.g. while one of them instantiates a graph)
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void wait_epoch(const unsigned* f, unsigned epoch) {
  if (ld_acquire_sys_u32(f) == epoch) return;
  const unsigned long long t0 = globaltimer_ns();
  unsigned spins = 0;
  while (ld_acquire_sys_u32(f) != epoch) {
    if ((++spins & 1023u) == 0u && globaltimer_ns() - t0 > 60000000000ull) __trap();
  }
}

// the receive buffer is written by other GPUs: read it around L1
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

__global__ void __launch_bounds__(1024) ar_add_norm_kernel(
    float* __restrict__ residual, const float* __restrict__ partials, int n_partials, int64_t partial_stride,
    const bf16* __restrict__ bias, const bf16* __restrict__ gamma, const bf16* __restrict__ beta, float eps,
    int layer_norm, bf16* __restrict__ out, int h, ArPeers peers, int call_idx, const unsigned* __restrict__ step_ptr,
    int calls_per_step, int two_shot) {
  extern __shared__ __align__(16) float srow[];
  __shared__ float red[32];
  pdl_trigger();
  pdl_wait();
  const int row = blockIdx.x, tp = peers.tp, me = peers.rank;
  // The step counter is bumped by the last kernel of the previous step.  Read it around L1: under programmatic
  // dependent launch this CTA may share its SM with a CTA of an EARLIER all-reduce that cached the old value,
  // and a stale epoch deadlocks both ranks (seen with back-to-back calls in tools/bench_c1.py).
  const unsigned epoch = __ldcg(step_ptr) * (unsigned)calls_per_step + (unsigned)call_idx + 1u;
  const int parity = call_idx & 1;
  const int64_t slot_me = ((int64_t)(parity * tp + me) * peers.max_rows + row);

  const int owner = two_shot ? row % tp : -1;       // two-shot: the rank that reduces this row
  const bool own = owner == me;

  // 1) fold this rank's slices of the row; push it to every rank (one-shot) or to the row's owner
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    const float* pp = partials + (int64_t)row * h + i;
    float4 v = *reinterpret_cast<const float4*>(pp);
    int p = 1;
    for (; p + 4 <= n_partials; p += 4) {            // index order, four slices in flight
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
    if (!two_shot) {
      for (int r = 0; r < tp; ++r) *reinterpret_cast<float4*>(peers.recv[r] + slot_me * h + i) = v;
    } else if (own) {
      *reinterpret_cast<float4*>(srow + i) = v;      // the owner's own partial never leaves the SM
    } else {
      *reinterpret_cast<float4*>(peers.recv[owner] + slot_me * h + i) = v;
    }
  }
  // The CTA barrier orders every thread's pushes before the flag stores below, and those are
  // st.release at system scope (cumulative), so no per-thread system fence is needed.
  __syncthreads();
  if (!two_shot) {
    // 2) publish the epoch flag at every rank, 3) wait for every rank's row
    if ((int)threadIdx.x < tp) st_release_sys_u32(peers.flags[threadIdx.x] + slot_me, epoch);
    if ((int)threadIdx.x < tp) wait_epoch(peers.flags[me] + ((int64_t)(parity * tp + (int)threadIdx.x) * peers.max_rows + row), epoch);
    __syncthreads();
  } else if (own) {
    // 2) wait for the tp-1 partial rows, 3) sum in rank order, push the sum to everybody
    if ((int)threadIdx.x < tp && (int)threadIdx.x != me)
      wait_epoch(peers.flags[me] + ((int64_t)(parity * tp + (int)threadIdx.x) * peers.max_rows + row), epoch);
    __syncthreads();
    for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < tp; ++r) {
        const float4 x = r == me ? *reinterpret_cast<const float4*>(srow + i)
                                 : ld_volatile_f4(peers.recv[me] + ((int64_t)(parity * tp + r) * peers.max_rows + row) * h + i);
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
      *reinterpret_cast<float4*>(srow + i) = a;      // same thread read srow[i..i+3] above
      for (int r = 0; r < tp; ++r)
        if (r != me) *reinterpret_cast<float4*>(peers.recv[r] + slot_me * h + i) = a;
    }
    __syncthreads();
    if ((int)threadIdx.x < tp && (int)threadIdx.x != me) st_release_sys_u32(peers.flags[threadIdx.x] + slot_me, epoch);
  } else {
    // 2) tell the owner the partial row is there, 3) wait for the reduced row to come back
    if (threadIdx.x == 0) {
      st_release_sys_u32(peers.flags[owner] + slot_me, epoch);
      wait_epoch(peers.flags[me] + ((int64_t)(parity * tp + owner) * peers.max_rows + row), epoch);
    }
    __syncthreads();
  }

  // 4) sum in rank order from local memory, residual add, norm
  float* res = residual + (int64_t)row * h;
  float sum = 0.f, sumsq = 0.f;
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    float4 v = *reinterpret_cast<const float4*>(res + i);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!two_shot) {
      for (int r = 0; r < tp; ++r) {
        const float4 x = ld_volatile_f4(peers.recv[me] + ((int64_t)(parity * tp + r) * peers.max_rows + row) * h + i);
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
    } else if (own) {
      a = *reinterpret_cast<const float4*>(srow + i);
    } else {
      a = ld_volatile_f4(peers.recv[me] + ((int64_t)(parity * tp + owner) * peers.max_rows + row) * h + i);
    }
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    if (bias != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(bias + i);
      v.x += bf16_lo(b.x); v.y += bf16_hi(b.x); v.z += bf16_lo(b.y); v.w += bf16_hi(b.y);
    }
    *reinterpret_cast<float4*>(res + i) = v;
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


// -----------------------------------------------------------------------------------------------
// Column-chunked variant: a row is split over the C CTAs of a thread-block cluster (grid (C, T)), so C x T
// SMs issue NVLink stores instead of T, and each CTA's serial work shrinks by C.  The exchange is the same
// protocol per (row, chunk) -- flags in the per-chunk array, the same receive buffers -- and the norm's row
// statistics are combined through distributed shared memory in cluster-rank order (deterministic).
// kBf16: the partial sums travel as bf16 (half the NVLink bytes); every rank rounds its own folded partial once
// and all ranks add the same rounded values in rank order, so they stay bit-identical.
// -----------------------------------------------------------------------------------------------
constexpr int kArMaxChunks = 8;

// A slot is `h` fp32 wide in the receive buffer whatever the payload type (bf16 uses the first half of it), so
// the fp32 rows that two-shot sends back never overlap another source's bf16 fan-in.
template <bool kBf16>
__device__ __forceinline__ void push4(float* base, int64_t slot_elem0, int col, const float4& v) {
  if constexpr (kBf16) {
    *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(base + slot_elem0) + col) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  } else {
    *reinterpret_cast<float4*>(base + slot_elem0 + col) = v;
  }
}
template <bool kBf16>
__device__ __forceinline__ float4 pull4(const float* base, int64_t slot_elem0, int col) {
  if constexpr (kBf16) {
    uint2 u;
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(u.x), "=r"(u.y) : "l"(reinterpret_cast<const bf16*>(base + slot_elem0) + col));
    return make_float4(bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y));
  } else {
    return ld_volatile_f4(base + slot_elem0 + col);
  }
}
__device__ __forceinline__ float4 round4_bf16(const float4& v) {
  const uint32_t a = pack_bf16(v.x, v.y), b = pack_bf16(v.z, v.w);
  return make_float4(bf16_lo(a), bf16_hi(a), bf16_lo(b), bf16_hi(b));
}

// sum over the cluster of one float per CTA, in cluster-rank order; `slot` is a shared-memory float of this CTA
__device__ __forceinline__ float cluster_sum(float v, float* slot, int C) {
  if (threadIdx.x == 0) *slot = v;
  cluster_sync_all();
  float total = 0.f;
  const uint32_t addr = smem_u32(slot);
  for (int r = 0; r < C; ++r) total += ld_dsmem_f32(addr, (uint32_t)r);
  cluster_sync_all();                 // the slot may be rewritten (and the CTA may exit) only after every peer has read it
  return total;
}

template <bool kBf16>
__global__ void __launch_bounds__(1024) ar_add_norm_cluster_kernel(
    float* __restrict__ residual, const float* __restrict__ partials, int n_partials, int64_t partial_stride,
    const bf16* __restrict__ bias, const bf16* __restrict__ gamma, const bf16* __restrict__ beta, float eps,
    int layer_norm, bf16* __restrict__ out, int h, ArPeers peers, int call_idx, const unsigned* __restrict__ step_ptr,
    int calls_per_step, int two_shot) {
  extern __shared__ __align__(16) float srow[];        // this CTA's chunk of the row
  __shared__ float red[32];
  __shared__ float cl_slot[2];
  pdl_trigger();
  pdl_wait();
  const int C = (int)gridDim.x, ch = (int)blockIdx.x, row = (int)blockIdx.y;
  const int cw = h / C, col0 = ch * cw;
  const int tp = peers.tp, me = peers.rank;
  // The step counter is bumped by the last kernel of the previous step.  Read it around L1: under programmatic
  // dependent launch this CTA may share its SM with a CTA of an EARLIER all-reduce that cached the old value,
  // and a stale epoch deadlocks both ranks (seen with back-to-back calls in tools/bench_c1.py).
  const unsigned epoch = __ldcg(step_ptr) * (unsigned)calls_per_step + (unsigned)call_idx + 1u;
  const int parity = call_idx & 1;
  const int64_t slot_me = ((int64_t)(parity * tp + me) * peers.max_rows + row);
  const int owner = two_shot ? (row + ch) % tp : -1;
  const bool own = owner == me;
  auto flag_of = [&](int at_rank, int src) {
    return peers.cflags[at_rank] + ((int64_t)(parity * tp + src) * peers.max_rows + row) * kArMaxChunks + ch;
  };

  // 1) fold this rank's slices of the chunk; push it to every rank (one-shot) or to the chunk's owner
  for (int i = threadIdx.x * 4; i < cw; i += blockDim.x * 4) {
    const float* pp = partials + (int64_t)row * h + col0 + i;
    float4 v = *reinterpret_cast<const float4*>(pp);
    int p = 1;
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
    if (!two_shot) {
      for (int r = 0; r < tp; ++r) push4<kBf16>(peers.recv[r], slot_me * h, col0 + i, v);
    } else if (own) {
      *reinterpret_cast<float4*>(srow + i) = kBf16 ? round4_bf16(v) : v;     // the owner's partial stays on the SM
    } else {
      push4<kBf16>(peers.recv[owner], slot_me * h, col0 + i, v);
    }
  }
  __syncthreads();        // orders every thread's pushes before the release stores of the flags
  if (!two_shot) {
    if ((int)threadIdx.x < tp) st_release_sys_u32(flag_of((int)threadIdx.x, me), epoch);
    if ((int)threadIdx.x < tp) wait_epoch(flag_of(me, (int)threadIdx.x), epoch);
    __syncthreads();
  } else if (own) {
    if ((int)threadIdx.x < tp && (int)threadIdx.x != me) wait_epoch(flag_of(me, (int)threadIdx.x), epoch);
    __syncthreads();
    for (int i = threadIdx.x * 4; i < cw; i += blockDim.x * 4) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < tp; ++r) {
        const float4 x = r == me ? *reinterpret_cast<const float4*>(srow + i)
                                 : pull4<kBf16>(peers.recv[me], ((int64_t)(parity * tp + r) * peers.max_rows + row) * h, col0 + i);
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
      // bf16 payload: the reduced chunk travels back as bf16 too (fan-in and fan-out of different chunks share a
      // slot, so they must use the same element pitch); the owner keeps the same rounded value
      if (kBf16) a = round4_bf16(a);
      *reinterpret_cast<float4*>(srow + i) = a;
      for (int r = 0; r < tp; ++r)
        if (r != me) push4<kBf16>(peers.recv[r], slot_me * h, col0 + i, a);
    }
    __syncthreads();
    if ((int)threadIdx.x < tp && (int)threadIdx.x != me) st_release_sys_u32(flag_of((int)threadIdx.x, me), epoch);
  } else {
    if (threadIdx.x == 0) {
      st_release_sys_u32(flag_of(owner, me), epoch);
      wait_epoch(flag_of(me, owner), epoch);
    }
    __syncthreads();
  }

  // 2) sum in rank order from local memory, residual add, row statistics over the cluster, norm
  float* res = residual + (int64_t)row * h + col0;
  float sum = 0.f, sumsq = 0.f;
  for (int i = threadIdx.x * 4; i < cw; i += blockDim.x * 4) {
    float4 v = *reinterpret_cast<const float4*>(res + i);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!two_shot) {
      for (int r = 0; r < tp; ++r) {
        const float4 x = pull4<kBf16>(peers.recv[me], ((int64_t)(parity * tp + r) * peers.max_rows + row) * h, col0 + i);
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
    } else if (own) {
      a = *reinterpret_cast<const float4*>(srow + i);
    } else {
      a = pull4<kBf16>(peers.recv[me], ((int64_t)(parity * tp + owner) * peers.max_rows + row) * h, col0 + i);
    }
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    if (bias != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(bias + col0 + i);
      v.x += bf16_lo(b.x); v.y += bf16_hi(b.x); v.z += bf16_lo(b.y); v.w += bf16_hi(b.y);
    }
    *reinterpret_cast<float4*>(res + i) = v;
    *reinterpret_cast<float4*>(srow + i) = v;
    sum += (v.x + v.y) + (v.z + v.w);
    sumsq += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  float mean = 0.f, rstd;
  if (layer_norm) {
    mean = cluster_sum(block_sum(sum, red), &cl_slot[0], C) / (float)h;
    float var = 0.f;
    for (int i = threadIdx.x * 4; i < cw; i += blockDim.x * 4) {
      const float4 v = *reinterpret_cast<const float4*>(srow + i);
      const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
      var += (a * a + b * b) + (c * c + d * d);
    }
    var = cluster_sum(block_sum(var, red), &cl_slot[1], C) / (float)h;
    rstd = 1.0f / sqrtf(var + eps);
  } else {
    rstd = 1.0f / sqrtf(cluster_sum(block_sum(sumsq, red), &cl_slot[0], C) / (float)h + eps);
  }
  bf16* o = out + (int64_t)row * h + col0;
  for (int i = threadIdx.x * 4; i < cw; i += blockDim.x * 4) {
    const float4 v = *reinterpret_cast<const float4*>(srow + i);
    const uint2 g = *reinterpret_cast<const uint2*>(gamma + col0 + i);
    float y0 = (v.x - mean) * rstd * bf16_lo(g.x), y1 = (v.y - mean) * rstd * bf16_hi(g.x);
    float y2 = (v.z - mean) * rstd * bf16_lo(g.y), y3 = (v.w - mean) * rstd * bf16_hi(g.y);
    if (beta != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(beta + col0 + i);
      y0 += bf16_lo(b.x); y1 += bf16_hi(b.x); y2 += bf16_lo(b.y); y3 += bf16_hi(b.y);
    }
    *reinterpret_cast<uint2*>(o + i) = make_uint2(pack_bf16(y0, y1), pack_bf16(y2, y3));
  }
}

}  // namespace

// -----------------------------------------------------------------------------------------------
// Flag-free ("Lamport") variant.  The flag protocol above pays, per hop, for a release at system scope (all of
// the CTA's NVLink stores acknowledged) before the flag may leave, and then for the flag's own trip: ~9 us per
// hop measured at tp=8 (tools/bench_c1.py).  Here every 32-bit word validates itself: receive slots hold a
// SENTINEL (-0.0f, a value the senders never transmit: they turn -0.0 into +0.0, which changes no sum), a reader
// polls the very words it is going to add until none of them is the sentinel, and writes the sentinel back
// once it has consumed them.  No flags, no fences, no epoch counter: one NVLink store latency per hop.
// Why two parities are enough: a slot written at call k is cleared by its reader inside call k; the writer's
// next store to it belongs to call k+2, which it can only reach after finishing call k+1, which needed the
// reader's call-k+1 data, which the reader's kernel k+1 sent after its kernel k (clears included) completed.
// One CTA per row (splitting rows over more CTAs measured slower).  Same one-shot / two-shot modes, same
// rank-order sums, so results are bit-identical to the flag protocol.
// -----------------------------------------------------------------------------------------------
constexpr unsigned kSentinel = 0x80000000u;      // -0.0f

__device__ __forceinline__ float no_neg_zero(float v) { return __float_as_uint(v) == kSentinel ? 0.f : v; }
__device__ __forceinline__ float4 no_neg_zero4(float4 v) {
  return make_float4(no_neg_zero(v.x), no_neg_zero(v.y), no_neg_zero(v.z), no_neg_zero(v.w));
}
// wait until the four words at p have all arrived, return them and put the sentinel back
__device__ __forceinline__ float4 take4(float* p) {
  float4 v = ld_volatile_f4(p);
  auto missing = [](const float4& x) {
    return __float_as_uint(x.x) == kSentinel || __float_as_uint(x.y) == kSentinel || __float_as_uint(x.z) == kSentinel ||
           __float_as_uint(x.w) == kSentinel;
  };
  if (missing(v)) {
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    do {
      v = ld_volatile_f4(p);
      if ((++spins & 1023u) == 0u && globaltimer_ns() - t0 > 60000000000ull) __trap();   // a lost peer must not hang the GPU
    } while (missing(v));
  }
  const float s = __uint_as_float(kSentinel);
  *reinterpret_cast<float4*>(p) = make_float4(s, s, s, s);
  return v;
}

__global__ void __launch_bounds__(1024) ar_lamport_add_norm_kernel(
    float* __restrict__ residual, const float* __restrict__ partials, int n_partials, int64_t partial_stride,
    const bf16* __restrict__ bias, const bf16* __restrict__ gamma, const bf16* __restrict__ beta, float eps,
    int layer_norm, bf16* __restrict__ out, int h, ArPeers peers, int call_idx, int two_shot) {
  extern __shared__ __align__(16) float srow[];
  __shared__ float red[32];
  pdl_trigger();
  pdl_wait();
  const int row = blockIdx.x, tp = peers.tp, me = peers.rank;
  const int parity = call_idx & 1;
  auto slot = [&](int at_rank, int src) {
    return peers.recv[at_rank] + ((int64_t)(parity * tp + src) * peers.max_rows + row) * h;
  };
  const int owner = two_shot ? row % tp : -1;
  const bool own = owner == me;

  // 1) fold this rank's slices of the row; send it to every rank (one-shot) or to the row's owner
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    const float* pp = partials + (int64_t)row * h + i;
    float4 v = *reinterpret_cast<const float4*>(pp);
    int p = 1;
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
    v = no_neg_zero4(v);
    if (!two_shot) {
      for (int r = 0; r < tp; ++r)
        if (r != me) *reinterpret_cast<float4*>(slot(r, me) + i) = v;
      *reinterpret_cast<float4*>(srow + i) = v;          // this rank's own contribution never leaves the SM
    } else if (own) {
      *reinterpret_cast<float4*>(srow + i) = v;
    } else {
      *reinterpret_cast<float4*>(slot(owner, me) + i) = v;
    }
  }
  // 2) every thread gathers the words it owns (the same thread wrote srow[i..i+3]): rank-order sum
  float* res = residual + (int64_t)row * h;
  float sum = 0.f, sumsq = 0.f;
  for (int i = threadIdx.x * 4; i < h; i += blockDim.x * 4) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!two_shot || own) {
      for (int r = 0; r < tp; ++r) {
        const float4 x = r == me ? *reinterpret_cast<const float4*>(srow + i) : take4(slot(me, r) + i);
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
      if (two_shot) {                                    // the owner sends the reduced row back to everybody
        a = no_neg_zero4(a);
        for (int r = 0; r < tp; ++r)
          if (r != me) *reinterpret_cast<float4*>(slot(r, me) + i) = a;
      }
    } else {
      a = take4(slot(me, owner) + i);
    }
    float4 v = *reinterpret_cast<const float4*>(res + i);
    v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    if (bias != nullptr) {
      const uint2 b = *reinterpret_cast<const uint2*>(bias + i);
      v.x += bf16_lo(b.x); v.y += bf16_hi(b.x); v.z += bf16_lo(b.y); v.w += bf16_hi(b.y);
    }
    *reinterpret_cast<float4*>(res + i) = v;
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

__global__ void fill_u32_kernel(unsigned* p, size_t n, unsigned v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// Column chunks per row for the cluster variant: chunks of >= 512 columns (128 threads), a power of two <= 8.
int ar_chunks_for(int h) {
  int c = 1;
  while (c < kArMaxChunks && h % (c * 2 * 4) == 0 && h / (c * 2) >= 512) c *= 2;
  return c;
}

int launch_ar_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                       const bf16* bias, const bf16* gamma, const bf16* beta, float eps, int layer_norm, bf16* out,
                       int T, int h, const ArPeers& peers, int call_idx, const unsigned* step_ptr, int calls_per_step,
                       int two_shot, int chunks, int bf16_payload, int lamport, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && T <= peers.max_rows && h % 4 == 0 && h * 4 <= 160 * 1024, "ar_add_norm: bad shape");
  BFLY_CHECK(n_partials >= 1 && partials != nullptr && peers.tp >= 2 && peers.tp <= kMaxPeers, "ar_add_norm: bad peers");
  if (lamport) {           // flag-free protocol: the receive buffers must have been filled with ar_fill_sentinel()
    static PerDeviceOnce attr3;
    BFLY_TRY(attr3.run([] {
      BFLY_CUDA(cudaFuncSetAttribute(ar_lamport_add_norm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      return (int)BFLY_OK;
    }));
    const int threads = std::min(1024, std::max(128, ((h / 4 + 31) / 32) * 32));
    BFLY_CUDA(launch_pdl(ar_lamport_add_norm_kernel, dim3(T), dim3(threads), (size_t)h * sizeof(float), stream, residual,
                         partials, n_partials, partial_stride, bias, gamma, beta, eps, layer_norm, out, h, peers, call_idx,
                         two_shot));
    count_launch();
    return BFLY_OK;
  }
  BFLY_CHECK(chunks >= 0 && chunks <= kArMaxChunks && (chunks == 0 || h % (chunks * 4) == 0), "ar_add_norm: chunks");
  if (chunks == 0) {       // one CTA per row (round-1 kernel; fp32 payload only)
    const size_t smem = (size_t)h * sizeof(float);
    static PerDeviceOnce attr;
    BFLY_TRY(attr.run([] {
      BFLY_CUDA(cudaFuncSetAttribute(ar_add_norm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
      return (int)BFLY_OK;
    }));
    const int threads = std::min(1024, std::max(128, ((h / 4 + 31) / 32) * 32));
    BFLY_CUDA(launch_pdl(ar_add_norm_kernel, dim3(T), dim3(threads), smem, stream, residual, partials, n_partials,
                         partial_stride, bias, gamma, beta, eps, layer_norm, out, h, peers, call_idx, step_ptr,
                         calls_per_step, two_shot));
    count_launch();
    return BFLY_OK;
  }
  BFLY_CHECK(T <= 65535, "ar_add_norm: too many rows");
  const int cw = h / chunks;
  static PerDeviceOnce attr2;
  BFLY_TRY(attr2.run([] {
    BFLY_CUDA(cudaFuncSetAttribute(ar_add_norm_cluster_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    BFLY_CUDA(cudaFuncSetAttribute(ar_add_norm_cluster_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    return (int)BFLY_OK;
  }));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)chunks, (unsigned)T);
  cfg.blockDim = dim3((unsigned)std::min(1024, std::max(64, ((cw / 4 + 31) / 32) * 32)));
  cfg.dynamicSmemBytes = (size_t)cw * sizeof(float);
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  at[1].id = cudaLaunchAttributeClusterDimension;
  at[1].val.clusterDim.x = (unsigned)chunks;
  at[1].val.clusterDim.y = 1;
  at[1].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 2;
  if (bf16_payload)
    BFLY_CUDA(cudaLaunchKernelEx(&cfg, ar_add_norm_cluster_kernel<true>, residual, partials, n_partials, partial_stride, bias,
                                 gamma, beta, eps, layer_norm, out, h, peers, call_idx, step_ptr, calls_per_step, two_shot));
  else
    BFLY_CUDA(cudaLaunchKernelEx(&cfg, ar_add_norm_cluster_kernel<false>, residual, partials, n_partials, partial_stride, bias,
                                 gamma, beta, eps, layer_norm, out, h, peers, call_idx, step_ptr, calls_per_step, two_shot));
  count_launch();
  return BFLY_OK;
}

// Lamport protocol set-up: every word of a receive buffer starts as the sentinel.
int ar_fill_sentinel(float* recv, size_t n_floats, cudaStream_t stream) {
  fill_u32_kernel<<<512, 256, 0, stream>>>(reinterpret_cast<unsigned*>(recv), n_floats, kSentinel);
  BFLY_CUDA(cudaGetLastError());
  return BFLY_OK;
}

}  // namespace bfly
