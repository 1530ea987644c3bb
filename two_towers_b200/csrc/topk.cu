// K-F / K-H exact-fp32 score + streaming row top-k, K-I shard merge, MRR rank kernel.
//
// Two routes to the same result.  Exact route (engine 0, and the fall-back of the other): the fused scan below.
// Tensor route: approximate scores from the tcgen05 engine (fp16 hi/lo planes, 3 passes, |error| <= 2^-13 |x||y|)
// only SELECT k+8 candidates per row; the candidates are then re-scored with the exact chain and ordered by the same
// total order, and a row is accepted only if its k-th exact score clears every non-candidate by more than the error
// bound - otherwise that row is redone by the exact scan.  Indices are therefore bit-exact on either route.
//
// score(i,j) is ONE fp32 FMA chain over d = 0..D-1 (ascending), optionally scaled to a cosine with
// pre-computed inverse norms: (dot * invx[i]) * invy[j].  The scan never stores scores: a 128x128 tile is
// reduced in registers, each value is compared against its row's current k-th best (score desc, index asc
// total order) and the few survivors are inserted into a per-row sorted list held in registers of one warp
// (lane r = rank r).  Because the filter and the insertion use the same total order, the result does not
// depend on the order in which threads report candidates: ties always resolve to the lowest index.
#include <limits.h>
#include <stdlib.h>
#include "common.cuh"
#include "simt_tile.cuh"
#include "umma_gemm.cuh"
#include "gru_umma.cuh"

namespace tt {

constexpr int ROUND_COLS = 16;   // candidate slots per row per filter round

__device__ __forceinline__ bool better(float v, int i, float v2, int i2) { return v > v2 || (v == v2 && i < i2); }

// insert (cv,ci) into the warp-resident sorted list (lane r holds rank r)
__device__ __forceinline__ void list_insert(float& lv, int& li, float cv, int ci, int lane) {
  const unsigned ahead = __ballot_sync(0xffffffffu, better(lv, li, cv, ci));
  const int p = __popc(ahead);                 // list is sorted, so these are lanes 0..p-1
  const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
  const int ui = __shfl_up_sync(0xffffffffu, li, 1);
  if (lane == p) { lv = cv; li = ci; }
  else if (lane > p) { lv = uv; li = ui; }
}

__global__ void inv_norm_kernel(const float* __restrict__ x, int64_t n, int D, float eps, float* __restrict__ inv) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= n) return;
  float s = 0.f;
  for (int c = lane; c < D; c += 32) { const float v = x[row * D + c]; s = fmaf(v, v, s); }
  s = warp_sum(s);
  if (lane == 0) inv[row] = 1.f / fmaxf(sqrtf(s), eps);
}

// grid: (ceil(Q/128), nsplit).  Split s scans columns [s*cols_per_split, ...) and writes list s.
__global__ void __launch_bounds__(NT) rows_topk_kernel(const float* __restrict__ X, int Q, const float* __restrict__ Y,
                                                      int64_t N, int D, int k, const float* __restrict__ invx,
                                                      const float* __restrict__ invy, int64_t mask_pos_offset,
                                                      int64_t index_base, int64_t cols_per_split,
                                                      float* __restrict__ out_val, int64_t* __restrict__ out_idx,
                                                      const int* __restrict__ need, const int* __restrict__ nfail, int nfail_min) {
  __shared__ __align__(16) float As[BK][LDS];
  __shared__ __align__(16) float Bs[BK][LDS];
  __shared__ float thr_v[BM];
  __shared__ int thr_i[BM];
  __shared__ int cnt[BM];
  __shared__ float cand_v[BM][ROUND_COLS];
  __shared__ int cand_i[BM][ROUND_COLS];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4, lane = t & 31, warp = t >> 5;
  const int bm = blockIdx.x * BM;
  if (nfail && *nfail <= nfail_min) return;            // few rejected rows: rescan_rows_kernel has them
  if (need && !__syncthreads_or(t < BM && bm + t < Q && need[bm + t] != 0)) return;   // no row of this tile was rejected
  const int64_t cbeg = (int64_t)blockIdx.y * cols_per_split;
  const int64_t cend = cbeg + cols_per_split < N ? cbeg + cols_per_split : N;

  // warp w owns rows w*16 .. w*16+15; lane r holds rank r of each
  float lv[16]; int li[16];
#pragma unroll
  for (int r = 0; r < 16; ++r) { lv[r] = -INFINITY; li[r] = INT_MAX; }
  if (t < BM) { thr_v[t] = -INFINITY; thr_i[t] = INT_MAX; cnt[t] = 0; }
  __syncthreads();

  float rinv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = bm + tile_row(ty, i);
    rinv[i] = (invx && m < Q) ? invx[m] : 1.f;
  }

  for (int64_t c0 = cbeg; c0 < cend; c0 += BN) {
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const float* Yt = Y + c0 * D;
    const int ncols = (int)((cend - c0) < BN ? (cend - c0) : BN);
    float ra[8], rb[8];
    load_frag<true>(ra, X, D, 1, bm, Q, 0, D, 0, 0, 0, t);
    load_frag<true>(rb, Yt, D, 1, 0, ncols, 0, D, 0, 0, 0, t);
    for (int k0 = 0; k0 < D; k0 += BK) {
      __syncthreads();
      store_frag<true>(ra, As, t);
      store_frag<true>(rb, Bs, t);
      __syncthreads();
      if (k0 + BK < D) {
        load_frag<true>(ra, X, D, 1, bm, Q, k0 + BK, D, 0, 0, 0, t);
        load_frag<true>(rb, Yt, D, 1, 0, ncols, k0 + BK, D, 0, 0, 0, t);
      }
      tile_fma(acc, As, Bs, tx, ty);
    }
    // finalise scores in place
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int cl = tile_col(tx, j);
      const int64_t cg = c0 + cl;
      const float yinv = (invy && cl < ncols) ? invy[cg] : 1.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int m = bm + tile_row(ty, i);
        float sc = invx ? (acc[i][j] * rinv[i]) * yinv : acc[i][j];
        if (mask_pos_offset >= 0 && cg == mask_pos_offset + m) sc = -1.f;
        if (cl >= ncols || m >= Q) sc = -INFINITY;
        acc[i][j] = sc;
      }
    }
    // 8 filter rounds of 16 columns: round r covers local columns [16r, 16r+16)
#pragma unroll
    for (int round = 0; round < 8; ++round) {
      int any = 0;
      const bool mine = (tx >> 2) == (round & 3);      // 4 consecutive tx own 16 consecutive columns of a 64-col half
      if (mine) {
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j = (round >> 2) * 4 + jj;
          const int cl = tile_col(tx, j);
          const int ci = (int)(c0 - cbeg) + cl;          // index relative to the split start (fits int)
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rl = tile_row(ty, i);
            const float sc = acc[i][j];
            if (sc > -INFINITY && better(sc, ci, thr_v[rl], thr_i[rl])) {
              const int slot = atomicAdd(&cnt[rl], 1);
              cand_v[rl][slot] = sc; cand_i[rl][slot] = ci;
              any = 1;
            }
          }
        }
      }
      if (__syncthreads_or(any)) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int rl = warp * 16 + r;
          const int n = cnt[rl];
          if (n > 0) {
            for (int c = 0; c < n; ++c) list_insert(lv[r], li[r], cand_v[rl][c], cand_i[rl][c], lane);
            if (lane == k - 1) { thr_v[rl] = lv[r]; thr_i[rl] = li[r]; }
            if (lane == 0) cnt[rl] = 0;
          }
        }
        __syncthreads();
      }
    }
  }
  // emit
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    const int m = bm + warp * 16 + r;
    if (m < Q && lane < k && !(need && need[m] == 0)) {
      const int64_t o = ((int64_t)blockIdx.y * Q + m) * k + lane;
      const bool empty = li[r] == INT_MAX;
      out_val[o] = empty ? -INFINITY : lv[r];
      out_idx[o] = empty ? -1 : index_base + cbeg + li[r];
    }
  }
}

// one warp per query: merge S lists of k (vals [S,Q,k], idx [S,Q,k]) by (score desc, idx asc)
__global__ void topk_merge_kernel(const float* __restrict__ vals, const int64_t* __restrict__ idx, int S, int Q, int k,
                                  float* __restrict__ out_val, int64_t* __restrict__ out_idx, const int* __restrict__ need,
                                  const int* __restrict__ nfail, int nfail_min) {
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (nfail && *nfail <= nfail_min) return;
  if (q >= Q || (need && need[q] == 0)) return;
  float lv = -INFINITY; long long li = LLONG_MAX;
  for (int s = 0; s < S; ++s)
    for (int e = 0; e < k; ++e) {
      const int64_t o = ((int64_t)s * Q + q) * k + e;
      const float cv = vals[o]; const long long ci = idx[o];
      if (ci < 0) continue;                                   // uniform across the warp
      const unsigned ahead = __ballot_sync(0xffffffffu, lv > cv || (lv == cv && li < ci));
      const int p = __popc(ahead);
      const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
      const long long ui = __shfl_up_sync(0xffffffffu, li, 1);
      if (lane == p) { lv = cv; li = ci; }
      else if (lane > p) { lv = uv; li = ui; }
    }
  if (lane < k) {
    const bool empty = li == LLONG_MAX;
    out_val[(int64_t)q * k + lane] = empty ? -INFINITY : lv;
    out_idx[(int64_t)q * k + lane] = empty ? -1 : li;
  }
}

__global__ void mrr_ranks_kernel(const int64_t* __restrict__ topk_idx, const int64_t* __restrict__ relevant, int Q, int k,
                                 int32_t* __restrict__ first_rank, float* __restrict__ rr) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  int rank = 0;
  const int64_t rel = relevant[q];
  for (int e = 0; e < k; ++e)
    if (topk_idx[(int64_t)q * k + e] == rel && rel >= 0) { rank = e + 1; break; }
  first_rank[q] = rank;
  rr[q] = rank ? 1.f / rank : 0.f;
}


// ------------------------------------------------------------------------------------------ sparse exact re-scan
// The acceptance test of the tensor route rejects a row now and then (the k-th exact score closer to the last candidate
// than the error bound: ~1 row in 10^4 on random data); the tile-shaped exact scan above would spend a 128-row tile's
// work (tens of ms at corpus scale) on it.  Up to kRescanCap rejected rows are instead gathered and re-scanned one row
// per warp: every CTA owns a contiguous chunk of the corpus, stages it through shared memory 64 documents at a time and
// each warp scores them against ITS row with the same fp32 FMA chain (d ascending) and score formula as the exact scan,
// so the result is bit-identical.  More rejected rows than kRescanCap: the tile-shaped scan takes all of them.
constexpr int kRescanCap = 256, kRescanTile = 64, kRescanCtas = 148;   // (4 CTAs per SM measured slower: 4x the lists to merge)
static inline uint32_t rescan_smem(int D) { return (uint32_t)((kRescanTile * (D + 1) + 8 * D) * sizeof(float)); }
static inline int rescan_cap(int D) { return rescan_smem(D) <= 200 * 1024 ? kRescanCap : 0; }

__global__ void compact_need_kernel(const int* __restrict__ need, int Q, int* __restrict__ rows, int* __restrict__ nfail) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q < Q && need[q] != 0) {
    const int slot = atomicAdd(nfail, 1);
    if (slot < kRescanCap) rows[slot] = q;
  }
}
// grid kRescanCtas x 256 threads.  part_v / part_i: [kRescanCtas][kRescanCap][k]
__global__ void __launch_bounds__(256) rescan_rows_kernel(const float* __restrict__ X, const float* __restrict__ Y, int64_t N, int D,
                                                          int k, int cosine, const float* __restrict__ invx,
                                                          const float* __restrict__ invy, int64_t mask_pos_offset,
                                                          int64_t index_base, const int* __restrict__ rows,
                                                          const int* __restrict__ nfail, float* __restrict__ part_v,
                                                          int64_t* __restrict__ part_i) {
  extern __shared__ float rs_smem[];
  const int nf = *nfail;
  if (nf <= 0 || nf > kRescanCap) return;
  float* ys = rs_smem;                                   // [kRescanTile][D + 1]
  float* xs = rs_smem + kRescanTile * (D + 1);           // [8][D]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t per = (N + gridDim.x - 1) / gridDim.x;
  const int64_t c0 = (int64_t)blockIdx.x * per, c1 = c0 + per < N ? c0 + per : N;
  for (int f0 = 0; f0 < nf; f0 += 8) {
    const int f = f0 + warp;
    const bool has = f < nf;
    const int m = has ? rows[f] : 0;
    __syncthreads();                                     // previous pass done with xs / ys
    if (has) for (int d = lane; d < D; d += 32) xs[warp * D + d] = X[(int64_t)m * D + d];
    const float xinv = (cosine && has) ? invx[m] : 1.f;
    const int64_t mcol = (mask_pos_offset >= 0 && has) ? mask_pos_offset + m : -1;
    float lv = -INFINITY; int li = INT_MAX;              // lane r holds rank r; li = document index relative to c0
    for (int64_t t0 = c0; t0 < c1; t0 += kRescanTile) {
      const int nd = (int)(c1 - t0 < kRescanTile ? c1 - t0 : kRescanTile);
      __syncthreads();
      if ((D & 3) == 0) {                                  // a warp per document row, 16-byte loads, 8 rows in flight per CTA
        for (int dl = warp; dl < nd; dl += 8) {
          const float4* src = reinterpret_cast<const float4*>(Y + (t0 + dl) * D);
          float* dst = ys + dl * (D + 1);
          for (int d4 = lane; d4 < (D >> 2); d4 += 32) {
            const float4 v = __ldg(src + d4);
            dst[4 * d4] = v.x; dst[4 * d4 + 1] = v.y; dst[4 * d4 + 2] = v.z; dst[4 * d4 + 3] = v.w;
          }
        }
      } else {
        for (int e = threadIdx.x; e < nd * D; e += 256) ys[(e / D) * (D + 1) + (e % D)] = Y[t0 * D + e];
      }
      __syncthreads();
      if (!has) continue;
      const float* xr = xs + warp * D;
      float a0 = 0.f, a1 = 0.f;
      const float* y0 = ys + lane * (D + 1);
      const float* y1 = ys + (lane + 32) * (D + 1);
      for (int d = 0; d < D; ++d) { const float xv = xr[d]; a0 = fmaf(xv, y0[d], a0); a1 = fmaf(xv, y1[d], a1); }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int dl = lane + 32 * h;
        const int64_t doc = t0 + dl;
        float sc = h ? a1 : a0;
        if (cosine) sc = (sc * xinv) * (dl < nd ? invy[doc] : 1.f);
        if (doc == mcol) sc = -1.f;
        if (dl >= nd) sc = -INFINITY;
        const int ci = (int)(doc - c0);
        const float tv = __shfl_sync(0xffffffffu, lv, k - 1);
        const int ti = __shfl_sync(0xffffffffu, li, k - 1);
        unsigned mm = __ballot_sync(0xffffffffu, sc > -INFINITY && better(sc, ci, tv, ti));
        while (mm) {
          const int b = __ffs(mm) - 1;
          mm &= mm - 1;
          const float cv = __shfl_sync(0xffffffffu, sc, b);
          const int cc = __shfl_sync(0xffffffffu, ci, b);
          if (better(cv, cc, __shfl_sync(0xffffffffu, lv, k - 1), __shfl_sync(0xffffffffu, li, k - 1))) list_insert(lv, li, cv, cc, lane);
        }
      }
    }
    if (has && lane < k) {
      const int64_t o = ((int64_t)blockIdx.x * kRescanCap + f) * k + lane;
      const bool empty = li == INT_MAX;
      part_v[o] = empty ? -INFINITY : lv;
      part_i[o] = empty ? -1 : index_base + c0 + li;
    }
  }
}
// one warp per rejected row: merge the per-CTA lists (score desc, index asc) into the row's output
__global__ void rescan_merge_kernel(const float* __restrict__ part_v, const int64_t* __restrict__ part_i, int nctas, int k,
                                    const int* __restrict__ rows, const int* __restrict__ nfail, float* __restrict__ out_val,
                                    int64_t* __restrict__ out_idx) {
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int nf = *nfail;
  if (nf <= 0 || nf > kRescanCap || f >= nf) return;
  float lv = -INFINITY; long long li = LLONG_MAX;
  for (int c = 0; c < nctas; ++c)
    for (int e = 0; e < k; ++e) {
      const int64_t o = ((int64_t)c * kRescanCap + f) * k + e;
      const float cv = part_v[o]; const long long ci = part_i[o];
      if (ci < 0) continue;                                   // uniform across the warp
      const unsigned ahead = __ballot_sync(0xffffffffu, lv > cv || (lv == cv && li < ci));
      const int p = __popc(ahead);
      const float uv = __shfl_up_sync(0xffffffffu, lv, 1);
      const long long ui = __shfl_up_sync(0xffffffffu, li, 1);
      if (lane == p) { lv = cv; li = ci; }
      else if (lane > p) { lv = uv; li = ui; }
    }
  if (lane < k) {
    const int m = rows[f];
    const bool empty = li == LLONG_MAX;
    out_val[(int64_t)m * k + lane] = empty ? -INFINITY : lv;
    out_idx[(int64_t)m * k + lane] = empty ? -1 : li;
  }
}

// ------------------------------------------------------------------------------------------ tensor route
constexpr int kCandSlack = 8;                                // candidates beyond k (fewer when k + 8 would exceed the engine's 16)
constexpr int kSelMax = 16;                                  // list entries per row the selection epilogue keeps (umma_gemm.cu)
static inline int cand_k2(int k) { return k + kCandSlack <= kSelMax ? k + kCandSlack : kSelMax; }
constexpr float kScoreErr = 1.220703125e-4f;                 // 2^-13: bound on |approx - exact| / (|x| |y|), 4x margin
// two-pass scan (x_hi y_hi + x_lo y_hi: the corpus contributes its fp16 hi plane only): rounding y to fp16 moves a score by
// at most |x| |y - y_hi| (Cauchy-Schwarz).  |y - y_hi| is not bounded a priori by 2^-11 |y| here but MEASURED: the index
// build takes the maximum over the corpus rows of the residual's 2-norm (resid_norm_kernel, ~0.4 * 2^-11 for normalised
// rows), and the acceptance test adds it to the bound above.
static int g_sel_passes = 2, g_sel_share = 1, g_rescan = 1;

// candidate lists empty; the first nthr words of thr = bits of -inf (the rows' shared lower bounds)
__global__ void list_init_kernel(float* __restrict__ lv, int* __restrict__ li, int64_t n, int* __restrict__ thr, int nthr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { lv[i] = -INFINITY; li[i] = INT_MAX; }
  if (i < nthr) thr[i] = __float_as_int(-INFINITY);
}
// one warp per plane row: max over rows of | v - hi(v) |_2, v = y[row][d] * (row_scale ? row_scale[row] : 1) as the planes
// were built (fp32 subtraction of a value and its own fp16 rounding is exact); scaled up by 2^-9 for the summation's rounding
__global__ void resid_norm_kernel(const float* __restrict__ y, int64_t N, int D, const float* __restrict__ row_scale,
                                  unsigned int* __restrict__ out_max) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  float s = 0.f;
  if (row < N) {
    const float sc = row_scale ? row_scale[row] : 1.f;
    for (int d = lane; d < D; d += 32) {
      float v = y[row * D + d];
      if (row_scale) v *= sc;
      const float r = v - __half2float(__float2half_rn(v));
      s = fmaf(r, r, s);
    }
  }
  s = warp_max(sqrtf(warp_sum(s)) * 1.001953125f);
  __shared__ float sh[8];
  if (lane == 0) sh[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) m = fmaxf(m, sh[i]);
    if (m == m) atomicMax(out_max, __float_as_uint(m)); else atomicMax(out_max, 0x7f800000u);   // NaN row: +inf (every row re-scanned)
  }
}
__global__ void min_pos_kernel(const float* __restrict__ v, int64_t n, unsigned int* __restrict__ out) {
  float m = INFINITY;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) m = fminf(m, v[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMin(out, __float_as_uint(m));       // positive floats order like their bit patterns
}
// one warp per row: fold the per-(CTA, column half) candidate lists the GEMM engine's selection epilogue wrote
// (part_v / part_i [Q][slots][32], unused entries = (-inf, INT_MAX)) into the row's K2 best (lane r = rank r).
// rscale (cosine: 1/max(|x|,eps) per row, the scores arrive scaled by |x|) turns the kept scores into the units of the
// acceptance test.
__global__ void merge_parts_kernel(const float* __restrict__ part_v, const int* __restrict__ part_i, int Q, int slots, int K2,
                                   const float* __restrict__ rscale, float* __restrict__ list_v, int* __restrict__ list_i) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= Q) return;
  float lv = -INFINITY, tv = -INFINITY;
  int li = INT_MAX, ti = INT_MAX;
  const float* pv = part_v + (int64_t)row * slots * 32;
  const int* pi = part_i + (int64_t)row * slots * 32;
  for (int sl = 0; sl < slots; ++sl) {
    const float cvl = pv[sl * 32 + lane];
    const int cil = pi[sl * 32 + lane];
    unsigned m = __ballot_sync(0xffffffffu, cil != INT_MAX && better(cvl, cil, tv, ti));
    while (m) {
      const int b = __ffs(m) - 1;
      m &= m - 1;
      const float cv = __shfl_sync(0xffffffffu, cvl, b);
      const int ci = __shfl_sync(0xffffffffu, cil, b);
      if (better(cv, ci, tv, ti)) {
        list_insert(lv, li, cv, ci, lane);
        tv = __shfl_sync(0xffffffffu, lv, K2 - 1);
        ti = __shfl_sync(0xffffffffu, li, K2 - 1);
      }
    }
  }
  list_v[(int64_t)row * 32 + lane] = rscale ? lv * rscale[row] : lv;
  list_i[(int64_t)row * 32 + lane] = li;
}
// one warp per row: exact chain for each candidate, exact order, acceptance test
__global__ void rescore_rows_kernel(const float* __restrict__ X, int Q, const float* __restrict__ Y, int D, int k, int K2,
                                    int cosine, const float* __restrict__ invx, const float* __restrict__ invy,
                                    const unsigned int* __restrict__ min_invy, int64_t mask_pos_offset, int64_t index_base,
                                    const float* __restrict__ list_v, const int* __restrict__ list_i, int two_pass,
                                    float* __restrict__ out_val, int64_t* __restrict__ out_idx, int* __restrict__ need) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= Q) return;
  const float av = list_v[(int64_t)row * 32 + lane];
  const int ci = list_i[(int64_t)row * 32 + lane];
  const bool valid = lane < K2 && ci != INT_MAX;
  float ev = -INFINITY;
  if (valid) {
    const float* x = X + (int64_t)row * D;
    const float* y = Y + (int64_t)ci * D;
    float acc = 0.f;
    for (int d = 0; d < D; ++d) acc = fmaf(x[d], y[d], acc);
    ev = cosine ? (acc * invx[row]) * invy[ci] : acc;
    if (mask_pos_offset >= 0 && ci == mask_pos_offset + row) ev = -1.f;
  }
  const int ei = valid ? ci : INT_MAX - 31 + lane;           // distinct keys: empty slots rank after every candidate, in lane order
  int rank = 0;
#pragma unroll 1
  for (int l = 0; l < 32; ++l) {
    const float ov = __shfl_sync(0xffffffffu, ev, l);
    const int oi = __shfl_sync(0xffffffffu, ei, l);
    if (l != lane && better(ov, oi, ev, ei)) ++rank;
  }
  // acceptance: every non-candidate has approx <= a_min, hence exact <= a_min + err; the k-th exact score must clear that
  const float a_min = __shfl_sync(0xffffffffu, av, K2 - 1);
  const bool full = __shfl_sync(0xffffffffu, ci, K2 - 1) != INT_MAX;       // list not full: every column is a candidate
  const unsigned at_k = __ballot_sync(0xffffffffu, rank == k - 1 && valid);
  float e_k = -INFINITY;
  if (at_k) e_k = __shfl_sync(0xffffffffu, ev, __ffs(at_k) - 1);
  // cosine: |x| invx and |y| invy are <= 1; dot: |x| max|y|  (invx = 1/max(|x|,1e-8))
  const float scale = cosine ? 1.f : (1.f / invx[row]) * (1.f / __uint_as_float(*min_invy));
  // two passes: plus |x| times the largest fp16 rounding residual of a corpus row (min_invy[1], see resid_norm_kernel)
  const float extra = two_pass ? __uint_as_float(min_invy[1]) * (cosine ? 1.f : 1.f / invx[row]) : 0.f;
  // the query's own hi + lo split is exact to ~2^-22 |x| only while its components stay above the fp16 subnormal range:
  // add the MEASURED |x - x_hi - x_lo|_2 times the largest corpus row norm (cosine: planes are normalised, times invx)
  float rx = 0.f;
  for (int d = lane; d < D; d += 32) {
    const float xv = X[(int64_t)row * D + d];
    const float h = __half2float(__float2half_rn(xv));
    const float l = __half2float(__float2half_rn(xv - h));
    const float r = (xv - h) - l;
    rx = fmaf(r, r, rx);
  }
  rx = sqrtf(warp_sum(rx)) * 1.001953125f;
  const float extra_x = rx * (cosine ? invx[row] : 1.f / __uint_as_float(*min_invy));
  const bool ok = !full || (e_k - a_min > kScoreErr * scale + extra + extra_x);     // false for NaN/inf as well
  if (lane == 0) need[row] = ok ? 0 : 1;
  if (ok && rank < k) {
    out_val[(int64_t)row * k + rank] = valid ? ev : -INFINITY;
    out_idx[(int64_t)row * k + rank] = valid ? index_base + ci : -1;
  }
}

static int topk_splits(int Q, int64_t N) {
  const int64_t row_tiles = ceil_div(Q, BM);
  int64_t s = ceil_div(2 * 148, row_tiles);
  const int64_t maxs = ceil_div(N, 4 * BN);
  if (s > maxs) s = maxs;
  if (s > 64) s = 64;
  return (int)(s < 1 ? 1 : s);
}

}  // namespace tt
using namespace tt;

extern "C" {

// tensor route: queries in chunks of QT rows; the approximate scores are consumed inside the GEMM engine's epilogue
static inline bool topk_tc(int64_t N, int k) { return engine_mode() >= 1 && umma_available() && N >= 256 && N < INT_MAX - 64 && k + 4 <= kSelMax; }
static inline int topk_qt(int Q) { return Q < 8192 ? Q : 8192; }
struct TopkWs {
  float *invx, *invy; unsigned int* ymin; float* pv; int64_t* pi;
  __half *xh, *xl, *yh, *yl; float* part_v; int* part_i; int64_t slots; float* lv; int* li; int* need; int64_t dp; int* thr; int* nfail; int* fail_rows; float* rs_v; int64_t* rs_i;
};
// The per-corpus part of the scan (tt_corpus_index_build): row norms, their minimum, and the hi/lo planes of the tensor route
struct CorpusIdx { float* invy; unsigned int* ymin; __half *yh, *yl; int64_t dp; bool planes; };
static inline bool corpus_planes(int64_t N) { return umma_available() && N >= 256 && N < INT_MAX - 64; }
static void carve_corpus(Workspace& w, int64_t N, int D, CorpusIdx* o) {
  o->dp = plane_pitch(D);
  o->invy = w.take<float>(N);
  o->ymin = w.take<unsigned int>(64);
  o->planes = corpus_planes(N);
  o->yh = o->yl = nullptr;
  if (o->planes) { o->yh = w.take<__half>(N * o->dp); o->yl = w.take<__half>(N * o->dp); }
}
static void carve_topk(Workspace& w, int Q, int64_t N, int D, int mode, int k, TopkWs* o, bool own_corpus = true) {
  const bool tc = topk_tc(N, k);
  o->invx = o->invy = nullptr;
  if (mode == 1 || tc) { o->invx = w.take<float>(Q); if (own_corpus) o->invy = w.take<float>(N); }
  const int S = topk_splits(Q, N);
  o->pv = nullptr; o->pi = nullptr;
  if (S > 1) { o->pv = w.take<float>((int64_t)S * Q * 32); o->pi = w.take<int64_t>((int64_t)S * Q * 32); }
  if (tc) {
    o->dp = plane_pitch(D);
    o->ymin = nullptr; o->yh = o->yl = nullptr;
    if (own_corpus) o->ymin = w.take<unsigned int>(64);
    o->xh = w.take<__half>((int64_t)Q * o->dp); o->xl = w.take<__half>((int64_t)Q * o->dp);
    if (own_corpus) { o->yh = w.take<__half>(N * o->dp); o->yl = w.take<__half>(N * o->dp); }
    // candidate lists of the selection epilogue: [rows of a query chunk][slots][32]; full chunks and the last, shorter one
    const int QT = topk_qt(Q), rem = Q % QT;
    int64_t rows_slots = (int64_t)QT * umma_sel_slots(QT, N);
    if (rem) { const int64_t r2 = (int64_t)rem * umma_sel_slots(rem, N); rows_slots = r2 > rows_slots ? r2 : rows_slots; }
    o->slots = rows_slots;                                      // rows x slots capacity (checked per chunk)
    o->part_v = w.take<float>(rows_slots * 32);
    o->part_i = w.take<int>(rows_slots * 32);
    o->lv = w.take<float>((int64_t)Q * 32); o->li = w.take<int>((int64_t)Q * 32); o->need = w.take<int>(Q);
    o->thr = w.take<int>(Q);
    o->nfail = w.take<int>(64); o->fail_rows = w.take<int>(kRescanCap);
    o->rs_v = w.take<float>((int64_t)kRescanCtas * kRescanCap * 32); o->rs_i = w.take<int64_t>((int64_t)kRescanCtas * kRescanCap * 32);
  }
}

uint64_t tt_rows_topk_workspace(int32_t Q, int64_t N, int32_t D, int32_t mode) {
  Workspace w(nullptr, 0);
  TopkWs t;
  carve_topk(w, Q, N, D, mode, 1, &t);          // k only gates the route (k + slack <= 32): size for the larger one
  return w.used + 256;
}

uint64_t tt_rows_topk_indexed_workspace(int32_t Q, int64_t N, int32_t D, int32_t mode) {
  Workspace w(nullptr, 0);
  TopkWs t;
  carve_topk(w, Q, N, D, mode, 1, &t, false);
  return w.used + 256;
}

uint64_t tt_corpus_index_bytes(int64_t N, int32_t D) {
  Workspace w(nullptr, 0);
  CorpusIdx c;
  carve_corpus(w, N > 0 ? N : 0, D, &c);
  return w.used + 256;
}

// per-corpus work of the scan: row norms (+ their minimum: the error bound of the dot mode) and the planes of the tensor route
static int corpus_prepare(const float* y, int64_t N, int D, int mode, const CorpusIdx& c, cudaStream_t s) {
  if (N <= 0) return TT_OK;
  note_launch(); inv_norm_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, s>>>(y, N, D, 1e-8f, c.invy);
  TT_CUDA(cudaMemsetAsync(c.ymin, 0x7f, sizeof(unsigned int), s));
  note_launch(); min_pos_kernel<<<296, 256, 0, s>>>(c.invy, N, c.ymin);
  TT_CUDA(cudaMemsetAsync(c.ymin + 1, 0, sizeof(unsigned int), s));
  if (c.planes) { note_launch(); resid_norm_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, s>>>(y, N, D, mode == 1 ? c.invy : nullptr, c.ymin + 1); }
  // cosine: the corpus rows go into the planes already divided by their norm (the exact re-score uses the fp32 rows)
  if (c.planes) TT_TRY(split_rows_f16_scaled(y, N, D, c.dp, c.yh, c.yl, mode == 1 ? c.invy : nullptr, s));
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_corpus_index_build(const float* y, int64_t N, int32_t D, int32_t mode, void* index, uint64_t index_bytes, void* stream) {
  TT_CHECK_ARG(y && index && N >= 0 && D > 0, "corpus_index_build: bad args");
  TT_CHECK_ARG(mode == 0 || mode == 1, "corpus_index_build: mode must be 0 (dot) or 1 (cosine)");
  Workspace w(index, index_bytes);
  CorpusIdx c;
  carve_corpus(w, N, D, &c);
  if (!w.ok()) { set_error("corpus_index_build: index buffer too small"); return TT_ERR_WORKSPACE; }
  return corpus_prepare(y, N, D, mode, c, (cudaStream_t)stream);
}

static int rows_topk_impl(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                          int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx, const void* index,
                          uint64_t index_bytes, void* workspace, uint64_t workspace_bytes, void* stream) {
  TT_CHECK_ARG(x && y && out_val && out_idx && Q > 0 && N >= 0 && D > 0, "rows_topk: bad args");
  TT_CHECK_ARG(k >= 1 && k <= 32, "rows_topk: k=%d must be in [1,32]", k);
  TT_CHECK_ARG(mode == 0 || mode == 1, "rows_topk: mode must be 0 (dot) or 1 (cosine)");
  cudaStream_t s = (cudaStream_t)stream;
  Workspace w(workspace, workspace_bytes);
  TopkWs t;
  carve_topk(w, Q, N, D, mode, k, &t, index == nullptr);
  if (!w.ok()) { set_error("rows_topk: workspace too small"); return TT_ERR_WORKSPACE; }
  const bool tc = topk_tc(N, k);
  CorpusIdx c;
  if (index) {
    Workspace wi(const_cast<void*>(index), index_bytes);
    carve_corpus(wi, N, D, &c);
    if (!wi.ok()) { set_error("rows_topk: corpus index too small for N=%lld D=%d", (long long)N, D); return TT_ERR_WORKSPACE; }
    TT_CHECK_ARG(!tc || c.planes, "rows_topk: corpus index has no planes");
  } else {
    c.invy = t.invy; c.ymin = t.ymin; c.yh = t.yh; c.yl = t.yl; c.dp = plane_pitch(D); c.planes = tc;
  }
  float* invx = t.invx; float* invy = c.invy;
  if (invx) { tt::note_launch(); inv_norm_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, s>>>(x, Q, D, 1e-8f, invx); }
  if (!index && invx && N > 0) {
    if (tc) TT_TRY(corpus_prepare(y, N, D, mode, c, s));
    else { note_launch(); inv_norm_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, s>>>(y, N, D, 1e-8f, invy); }
  }
  const int* need = nullptr;
  int rs_cap = 0;
  if (tc) {
    // ---- tensor route: approximate tile scores -> candidates -> exact re-score -> acceptance
    const int K2 = cand_k2(k), QT = topk_qt(Q);
    const bool two = g_sel_passes == 2 && umma_sel_two_ok(D);
    TT_TRY(split_rows_f16(x, Q, D, t.dp, t.xh, t.xl, s));
    for (int q0 = 0; q0 < Q; q0 += QT) {
      const int nq = Q - q0 < QT ? Q - q0 : QT;
      const int use_slots = umma_sel_slots(nq, N);
      TT_CHECK_ARG((int64_t)nq * use_slots <= t.slots, "rows_topk: part buffer too small");
      note_launch(); list_init_kernel<<<(unsigned)ceil_div((int64_t)nq * use_slots * 32, 256), 256, 0, s>>>(t.part_v, t.part_i, (int64_t)nq * use_slots * 32, t.thr + q0, nq);
      UmmaGemm g{};
      g.A[0] = Planes{(const bf16*)(t.xh + (int64_t)q0 * t.dp), (const bf16*)(t.xl + (int64_t)q0 * t.dp), t.dp};
      g.B[0] = Planes{(const bf16*)c.yh, (const bf16*)c.yl, c.dp};
      g.M = nq; g.N = (int)N; g.K = D; g.nbatch = 1;
      g.alpha = 1.f; g.beta = 0.f; g.splitk = 1; g.f16 = 1; g.npass = 3;
      g.sel_two = two ? 1 : 0;
      g.sel_thr = g_sel_share ? t.thr + q0 : nullptr;
      g.epi = 2; g.sel_v = t.part_v; g.sel_i = t.part_i; g.sel_k2 = K2; g.sel_slots = use_slots;
      g.sel_invx = mode == 1 ? invx + q0 : nullptr;
      g.sel_mask_off = mask_pos_offset >= 0 ? mask_pos_offset + q0 : -1;
      TT_TRY(umma_gemm(g, s));
      note_launch();
      merge_parts_kernel<<<(unsigned)ceil_div(nq, 8), 256, 0, s>>>(t.part_v, t.part_i, nq, use_slots, K2, mode == 1 ? invx + q0 : nullptr,
                                                                  t.lv + (int64_t)q0 * 32, t.li + (int64_t)q0 * 32);
    }
    note_launch();
    rescore_rows_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, s>>>(x, Q, y, D, k, K2, mode, invx, invy, c.ymin, mask_pos_offset,
                                                                index_base, t.lv, t.li, two ? 1 : 0, out_val, out_idx, t.need);
    TT_LAUNCH_CHECK();
    need = t.need;
    // rejected rows: a few -> gathered, one warp per row; many -> the tile-shaped exact scan below (gated on the count)
    rs_cap = g_rescan ? rescan_cap(D) : 0;
    if (rs_cap > 0) {
      static DeviceOnce attr_once;
      const uint32_t smem = rescan_smem(D);
      if (attr_once.need()) { TT_CUDA(cudaFuncSetAttribute(rescan_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); attr_once.done(); }
      TT_CUDA(cudaMemsetAsync(t.nfail, 0, sizeof(int), s));
      note_launch(); compact_need_kernel<<<(unsigned)ceil_div(Q, 256), 256, 0, s>>>(t.need, Q, t.fail_rows, t.nfail);
      note_launch();
      rescan_rows_kernel<<<kRescanCtas, 256, smem, s>>>(x, y, N, D, k, mode, invx, invy, mask_pos_offset, index_base, t.fail_rows,
                                                       t.nfail, t.rs_v, t.rs_i);
      note_launch();
      rescan_merge_kernel<<<kRescanCap / 8, 256, 0, s>>>(t.rs_v, t.rs_i, kRescanCtas, k, t.fail_rows, t.nfail, out_val, out_idx);
      TT_LAUNCH_CHECK();
    }
    if (mode == 0) { invx = nullptr; invy = nullptr; }      // the exact scan applies norms only for the cosine
  }
  // ---- exact scan: everything (engine 0) or only the rows the acceptance test rejected
  const int S = topk_splits(Q, N);
  float* pv = t.pv; int64_t* pi = t.pi;
  int64_t cps = ceil_div(N > 0 ? N : 1, S);
  cps = ceil_div(cps, BN) * BN;
  dim3 grid((unsigned)ceil_div(Q, BM), (unsigned)S);
  tt::note_launch(); rows_topk_kernel<<<grid, NT, 0, s>>>(x, Q, y, N, D, k, invx, invy, mask_pos_offset, index_base, cps,
                                      S > 1 ? pv : out_val, S > 1 ? pi : out_idx, need, rs_cap > 0 ? t.nfail : nullptr, rs_cap);
  if (S > 1) { note_launch(); topk_merge_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, s>>>(pv, pi, S, Q, k, out_val, out_idx, need, rs_cap > 0 ? t.nfail : nullptr, rs_cap); }
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_rows_topk(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                 int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx, void* workspace,
                 uint64_t workspace_bytes, void* stream) {
  return rows_topk_impl(x, Q, y, N, D, k, mode, mask_pos_offset, index_base, out_val, out_idx, nullptr, 0, workspace,
                        workspace_bytes, stream);
}

int tt_rows_topk_indexed(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                         int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx, const void* index,
                         uint64_t index_bytes, void* workspace, uint64_t workspace_bytes, void* stream) {
  TT_CHECK_ARG(index != nullptr, "rows_topk_indexed: null index");
  return rows_topk_impl(x, Q, y, N, D, k, mode, mask_pos_offset, index_base, out_val, out_idx, index, index_bytes, workspace,
                        workspace_bytes, stream);
}

int tt_debug_sel_resident(int32_t on) {
  tt::umma_set_sel_resident(on);
  return TT_OK;
}
// byte offset, inside the workspace of a tt_rows_topk (indexed == 0) / tt_rows_topk_indexed (indexed != 0) call of this shape,
// of the int32[Q] flags "row failed the acceptance test of the tensor route and was re-done by the exact scan";
// -1 when the call does not take the tensor route
int64_t tt_debug_topk_fallback_flags(int32_t Q, int64_t N, int32_t D, int32_t mode, int32_t k, int32_t indexed) {
  if (!topk_tc(N, k)) return -1;
  Workspace w(nullptr, 0);
  TopkWs t;
  carve_topk(w, Q, N, D, mode, k, &t, indexed == 0);
  return (int64_t)(reinterpret_cast<char*>(t.need) - (char*)nullptr);
}
int tt_debug_sel_passes(int32_t passes) {
  const int p = passes % 10, f = passes / 10;
  TT_CHECK_ARG((p == 2 || p == 3) && f >= 0 && f <= 3, "tt_debug_sel_passes: 2 or 3 (+10: no shared row bound, +20: no sparse re-scan)");
  tt::g_sel_passes = p;
  tt::g_sel_share = (f & 1) == 0;
  tt::g_rescan = (f & 2) == 0;
  return TT_OK;
}
int tt_debug_gemm_pairs(int32_t on) {
  tt::umma_set_pairs(on);
  return TT_OK;
}
int tt_debug_sel_prefetch(int32_t tiles) {
  TT_CHECK_ARG(tiles >= 0 && tiles <= 64, "tt_debug_sel_prefetch: 0..64 column tiles");
  tt::umma_set_sel_prefetch(tiles);
  return TT_OK;
}

int tt_topk_merge(const float* vals, const int64_t* idx, int32_t S, int32_t Q, int32_t k, float* out_val,
                  int64_t* out_idx, void* stream) {
  TT_CHECK_ARG(vals && idx && out_val && out_idx && S > 0 && Q > 0 && k >= 1 && k <= 32, "topk_merge: bad args");
  tt::note_launch(); topk_merge_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, (cudaStream_t)stream>>>(vals, idx, S, Q, k, out_val, out_idx, nullptr, nullptr, 0);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_mrr_ranks(const int64_t* topk_idx, const int64_t* relevant, int32_t Q, int32_t k, int32_t* first_rank, float* rr,
                 void* stream) {
  TT_CHECK_ARG(topk_idx && relevant && first_rank && rr && Q > 0 && k > 0, "mrr_ranks: bad args");
  tt::note_launch(); mrr_ranks_kernel<<<(unsigned)ceil_div(Q, 256), 256, 0, (cudaStream_t)stream>>>(topk_idx, relevant, Q, k, first_rank, rr);
  TT_LAUNCH_CHECK();
  return TT_OK;
}



}  // extern "C"
