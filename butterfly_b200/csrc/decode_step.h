// Fused decode step: the whole partitioned block forward of ONE decode step (embedding, every layer's
// norm / QKV / RoPE+KV append / paged attention / attention-out / [all-reduce] / norm / gate-up / SwiGLU /
// down / [all-reduce], final norm, LM head, greedy keys) as ONE persistent cooperative kernel, one CTA per
// SM, with grid-wide phase counters in place of kernel boundaries (decode_step.cu).  This header is what the
// engine (host) and the kernel share.
#pragma once

#include "kernels.cuh"

namespace bfly {

constexpr int kStepThreads = 256;       // warps 0..2: TMA producers, warp 3: MMA issuer, warps 4..7: epilogue / attention
constexpr int kStepMaxTokens = 64;      // rows of one decode step the fused kernel accepts (UMMA N <= 64)
constexpr int kStepRingBytes = 213504;  // operand ring (GEMM phases) = attention rings + merge buffers (attention phase)

// One GEMM of the step, partitioned stream-K: the m_tiles x nkb (128-row weight tile, 64-wide K block) units
// are dealt out in order, unit range [lo(c), lo(c+1)) to CTA c < ge, so every CTA streams the same number of
// weight bytes (+-1 block).  A CTA's units of one tile form a segment whose partial sum goes to fp32 slice
// (c - first[tile]) of `out`; the consumer folds the nsl[tile] slices of a tile in slice (= K) order.
struct StepGemm {
  int M, K, nkb, m_tiles, units, ge;
  int max_slices;
  float* out;                    // fp32 [max_slices][T][M]
  const unsigned short* first;   // [m_tiles] first CTA with units of the tile
  const unsigned char* nsl;      // [m_tiles] slices (CTAs) the tile is spread over
};
__host__ __device__ inline int sk_lo(const StepGemm& g, int c) { return (int)((int64_t)c * g.units / g.ge); }
__host__ __device__ inline int sk_owner(const StepGemm& g, int u) {
  int c = (int)((int64_t)u * g.ge / g.units);
  while (c + 1 < g.ge && sk_lo(g, c + 1) <= u) ++c;
  return c;
}

struct StepLayer {
  const bf16 *attn_norm_w, *ffn_norm_w;
  bf16* kv;                      // this layer's pages
};

enum StepGemmId : int { SG_QKV = 0, SG_WO = 1, SG_GU = 2, SG_DOWN = 3, SG_HEAD = 4, SG_COUNT = 5 };

struct StepParams {
  // shapes
  int T, bn, stages, n_layers, h, nq, nk, f_local, qkv_rows, v_local, v0, vocab;
  int chunks;                    // column chunks per row in the norm phases (T * chunks items)
  int tune;                      // experiment bits (BFLY_STEP_TUNE): 1 = default L2 policy for weight tiles
  int n_producers;               // TMA producer warps used by the GEMM phases (1..3)
  int attn_splits;
  float eps, attn_scale_log2;
  // GEMMs (same geometry in every layer) and their tensor maps: maps[layer * 4 + {qkv, wo, gate_up, down}],
  // then lm_head, then the three activation operands (xg, attn_out, act)
  StepGemm gemm[SG_COUNT];
  const void* maps;              // CUtensorMap[n_layers * 4 + 4], 128-byte aligned, device memory
  alignas(64) unsigned char map_head_param[128];   // experiment: the LM-head and xg maps as kernel parameters
  alignas(64) unsigned char map_xg_param[128];
  const StepLayer* layers;       // device array
  const bf16* final_norm_w;
  // activations
  float* residual;               // fp32 [T][h]
  bf16* xg;                      // bf16 [T][h]   residual * gamma (NOT yet scaled by rstd: consumers apply it)
  float* ssq;                    // fp32 [3][T][chunks] partial sums of squares (attn norm, ffn norm, final norm)
  bf16 *q, *attn_out, *act;
  float *part_o, *part_ml;
  // tokens / head
  const int* tokens;
  const bf16* embed;
  int* history;
  const int* step_counter;
  unsigned long long* keys;
  float* logits;                 // nullptr unless the caller wants them
  int64_t logits_ld;
  // paged KV
  const int* seq_lens;
  const int* page_table;
  int pages_per_seq, page_size;
  const float2* rope_table;
  // synchronisation
  unsigned* counters;            // one per phase, zeroed before the launch
  unsigned long long* trace;     // optional [phases + 2]: globaltimer when CTA 0 saw each phase complete (profiling),
                                 // followed by [phases][grid]: when each CTA arrived at each phase
  // tensor parallel: the all-reduce folded into the norm phases (same buffers and protocol as allreduce_norm.cu)
  ArPeers peers;
  const unsigned* ar_step;
  int ar_on;                     // 0: tp == 1 or collectives skipped (profiling)
};

int step_phase_count(int n_layers, int attn_splits);
int launch_decode_step(const StepParams& p, int grid, cudaStream_t stream);
int decode_step_init();                                       // kernel attributes (per device)
// gemm_tcgen05.cu: a 2D bf16 tensor map (box = 64 columns x box_rows rows, 128-byte swizzle) written to out128
int make_tensor_map_2d(void* out128, const bf16* ptr, int64_t rows, int cols, int64_t pitch, int box_rows);

}  // namespace bfly
