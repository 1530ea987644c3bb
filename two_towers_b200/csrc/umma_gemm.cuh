// Host-side interface of the tcgen05 bf16x3 GEMM engine (umma_gemm.cu).
//
// Every fp32 operand is carried as two bf16 planes (hi = bf16(x), lo = bf16(x - hi)); a product is evaluated as
// hi*hi + lo*hi + hi*lo with fp32 accumulation in TMEM, i.e. ~16 mantissa bits per operand (measured end-to-end
// error vs fp32 ~1e-5) at 3 tensor-core passes.  Single-pass bf16 (3.5e-3) and TF32 (5e-4 fwd, 3e-2
// on gradients through the margin loss) both miss the 1e-3 parity bound.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "common.cuh"

namespace tt {

typedef __nv_bfloat16 bf16;

// row-major bf16 plane pair; pitch in elements (multiple of 8 so rows are 16-byte aligned for TMA)
struct Planes {
  const bf16* hi;
  const bf16* lo;
  int64_t pitch;
};

static inline int64_t plane_pitch(int64_t cols) { return (cols + 7) / 8 * 8; }

// C[M,N] = alpha * A * B^T (+ beta*C) (+ bias[n]), batched over up to 2 independent problems of the same shape.
//   a_mn == false: A is K-major   : planes hold [M rows][K cols]
//   a_mn == true : A is MN-major  : planes hold [K rows][M cols]   (reduction over rows, e.g. tokens -> wgrad)
//   same for B with N.
struct UmmaGemm {
  Planes A[2], B[2];
  float* C[2];
  const float* bias[2];
  int64_t ldc;
  int32_t M, N, K, nbatch;
  bool a_mn, b_mn;
  float alpha, beta;
  int32_t splitk;      // 0 = choose; >1 accumulates with atomics (C is pre-scaled by beta)
  // ---- extensions (all default 0 = the plain bf16 x3 row-major GEMM)
  int32_t f16;         // planes hold IEEE fp16 instead of bf16
  int32_t npass;       // 0/3: hi*hi + lo*hi + hi*lo ; 1: hi*hi only (lo planes unused)
  int64_t a_rows, b_rows;   // rows of the plane arrays the tensor maps cover (0: derived from M/N/K)
  int32_t a_row0, b_row0;   // K-major operands: first plane row of the problem (e.g. t*Bp for one timestep)
  int32_t max_ctas;         // 0 = one CTA per SM; otherwise cap of the persistent grid (GEMMs co-scheduled with another kernel)
  int32_t a_seg2;           // optional second threshold of the remap below: i >= a_seg2 skips a_skip once more (K-major A)
  int32_t a_seg, a_skip;    // A column remap: logical index i >= a_seg reads plane column i + a_skip
                            //   (K-major A: i = k ; MN-major A: i = m).  Both multiples of 64 (128 for MN-major).
  const float* alpha_ptr;   // optional device scalar multiplied into alpha (loss-scale undo)
  int32_t epi;              // 0 row-major C ; 1 tiled C (see TiledOut) ; 2 no C: per-row candidate selection (below)
  int32_t rm_Bp, rm_B, rm_T;  // row-major epilogue: rows are time-major m = t*Bp + b -> C row b*T + t; b >= rm_B skipped
  int32_t tl_Bp, tl_G, tl_Hc; // tiled epilogue: C[(((t*MT + mt)*G + g)*(Hc/4) + u/4)*512 + row*4 + u%4],
                              //   t = m / Bp, mt = (m % Bp) / 128, row = m % 128, g = n / Hc, u = n % Hc, MT = Bp / 128
  // ---- epi == 2: fused row top-K2 selection (K-H / K-F retrieval scan).  Scores never leave the SM: each epilogue thread
  // (one accumulator row, half of the tile's columns) keeps the K2 best (score desc, column asc) of the columns it has
  // seen for the current 128-row tile; a CTA walks a contiguous range of column tiles of one row tile and writes its lists
  // (in no particular order) to sel_v / sel_i [M][sel_slots][32] (slot = 2 * part + column half; unused entries keep the
  // caller's initialisation).
  // Column sel_mask_off + row (if sel_mask_off >= 0) scores -1 (times 1/sel_invx[row] when sel_invx is given: the scores
  // are cosines scaled by |x|); NaN scores become +inf so the row fails the caller's acceptance test.
  // Needs K-major operands, splitk 1, nbatch 1; umma_sel_slots() gives sel_slots for a problem shape.
  float* sel_v; int* sel_i; const float* sel_invx;
  int64_t sel_mask_off;
  int32_t sel_k2, sel_slots;
  int* sel_thr;             // optional [M] ints, every word initialised to the bits of -inf (0xff800000) before the launch: the
                            //   partitions of a row share a lower bound of its K2-th best score through it
  int32_t sel_two;          // B's lo plane is not used (hi*hi + lo*hi only: |error| <= 2^-11 |a||b| more); needs umma_sel_two_ok(K)
};
bool umma_sel_two_ok(int32_t K);
int umma_sel_slots(int32_t M, int64_t N);
void umma_set_pairs(int on);            // 1 (default): large K-major GEMMs on 2-CTA clusters (cta_group::2)
void umma_set_sel_prefetch(int tiles); // epi == 2, resident A: L2 prefetch distance in column tiles (0 = off)
void umma_set_sel_resident(int on);   // epi == 2: keep the unit's A tile in shared memory (default on; off = stream it per tile)
int umma_gemm(const UmmaGemm& g, cudaStream_t s);
// encode the tensor maps / pick the configuration once, launch many times (per-timestep GEMMs only change the first
// plane row of A, given per batched problem)
struct UmmaPrepared { alignas(64) unsigned char blob[1408]; int kind, items, max_ctas; };
int umma_gemm_prepare(const UmmaGemm& g, UmmaPrepared* out);
int umma_gemm_launch(const UmmaPrepared& p, const int* a_row0, cudaStream_t s);

// fp32 [rows, cols] (row stride ld) -> hi/lo planes [rows, pitch]; optional inverted dropout (p, seed, offset over
// the logical index row*cols+col) fused in; columns [cols, pitch) are written as zeros.
int split_planes(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch, float p,
                 uint64_t seed, uint64_t offset, cudaStream_t s);
// transpose + split: x [rows, cols] -> planes [cols, pitch(rows)]   (weights for dgrad: W^T)
int split_planes_transposed(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch,
                            cudaStream_t s);

// 2-D tensor map over a 16-bit plane [rows][pitch] exposing `cols` columns, box {64, box_rows}, 128-byte swizzle
int make_plane_map(CUtensorMap* m, const void* base, int64_t rows, int64_t cols, int64_t pitch, int box_rows, int f16);

bool umma_available();   // false if the driver entry point for tensor maps is missing

}  // namespace tt
