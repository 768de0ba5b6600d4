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

#include "kernels.cuh"

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
  const unsigned epoch = (*step_ptr) * (unsigned)calls_per_step + (unsigned)call_idx + 1u;
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

}  // namespace

int launch_ar_add_norm(float* residual, const float* partials, int n_partials, int64_t partial_stride,
                       const bf16* bias, const bf16* gamma, const bf16* beta, float eps, int layer_norm, bf16* out,
                       int T, int h, const ArPeers& peers, int call_idx, const unsigned* step_ptr, int calls_per_step,
                       int two_shot, cudaStream_t stream) {
  BFLY_CHECK(T > 0 && T <= peers.max_rows && h % 4 == 0 && h * 4 <= 160 * 1024, "ar_add_norm: bad shape");
  BFLY_CHECK(n_partials >= 1 && partials != nullptr && peers.tp >= 2 && peers.tp <= kMaxPeers, "ar_add_norm: bad peers");
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

}  // namespace bfly
