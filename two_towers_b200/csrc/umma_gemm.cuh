// Host-side interface of the tcgen05 bf16x3 GEMM engine (umma_gemm.cu).
//
// Every fp32 operand is carried as two bf16 planes (hi = bf16(x), lo = bf16(x - hi)); a product is evaluated as
// hi*hi + lo*hi + hi*lo with fp32 accumulation in TMEM, i.e. ~16 mantissa bits per operand (measured end-to-end
// error vs fp32 ~1e-5) at 3 tensor-core passes.  Single-pass bf16 (3.5e-3) and TF32 (5e-4 fwd, 3e-2
// on gradients through the margin loss) both miss the 1e-3 parity bound.
#pragma once
#include <cuda_bf16.h>
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
};
int umma_gemm(const UmmaGemm& g, cudaStream_t s);

// fp32 [rows, cols] (row stride ld) -> hi/lo planes [rows, pitch]; optional inverted dropout (p, seed, offset over
// the logical index row*cols+col) fused in; columns [cols, pitch) are written as zeros.
int split_planes(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch, float p,
                 uint64_t seed, uint64_t offset, cudaStream_t s);
// transpose + split: x [rows, cols] -> planes [cols, pitch(rows)]   (weights for dgrad: W^T)
int split_planes_transposed(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch,
                            cudaStream_t s);

bool umma_available();   // false if the driver entry point for tensor maps is missing

}  // namespace tt
