// 128x128x16 SIMT tile machinery shared by the fp32 GEMM and the exact top-k scan.
#pragma once
#include "common.cuh"
namespace tt {
constexpr int BM = 128, BN = 128, BK = 16, NT = 256, LDS = BM + 4;

__device__ __forceinline__ int64_t seg_row(int r, int seg_len, int seg_stride, int off) {
  return seg_len > 0 ? (int64_t)(r / seg_len) * seg_stride + (r % seg_len) + off : (int64_t)r;
}

// load 8 elements (m fixed per thread, r = r0..r0+7) of a [rows x BK] operand tile into registers
template <bool R_CONTIG>
__device__ __forceinline__ void load_frag(float (&v)[8], const float* __restrict__ P, int64_t rs, int64_t cs,
                                          int row0, int nrows, int k0, int kend, int seg_len, int seg_stride,
                                          int seg_off, int t) {
  const int m = R_CONTIG ? (t >> 1) : (t & 127);
  const int r0 = R_CONTIG ? (t & 1) * 8 : (t >> 7) * 8;
  const int mg = row0 + m;
  const bool mok = mg < nrows;
  const float* base = P + (int64_t)mg * rs;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int r = k0 + r0 + i;
    float x = 0.f;
    if (mok && r < kend) x = __ldg(base + seg_row(r, seg_len, seg_stride, seg_off) * cs);
    v[i] = x;
  }
}
template <bool R_CONTIG>
__device__ __forceinline__ void store_frag(const float (&v)[8], float (*S)[LDS], int t) {
  const int m = R_CONTIG ? (t >> 1) : (t & 127);
  const int r0 = R_CONTIG ? (t & 1) * 8 : (t >> 7) * 8;
#pragma unroll
  for (int i = 0; i < 8; ++i) S[r0 + i][m] = v[i];
}


// acc[8][8] += outer products of one BK-deep smem tile pair (k ascending -> one FMA chain per element)
__device__ __forceinline__ void tile_fma(float (&acc)[8][8], const float (*As)[LDS], const float (*Bs)[LDS], int tx, int ty) {
#pragma unroll
  for (int kk = 0; kk < BK; ++kk) {
    float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
    float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
    float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
    float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
    float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}
// micro-tile coordinates: local row of acc[i][*] / local col of acc[*][j]
__device__ __forceinline__ int tile_row(int ty, int i) { return i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4); }
__device__ __forceinline__ int tile_col(int tx, int j) { return j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4); }
}  // namespace tt
