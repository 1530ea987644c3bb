// K-F / K-H exact-fp32 score + streaming row top-k, K-I shard merge, MRR rank kernel.
//
// score(i,j) is ONE fp32 FMA chain over d = 0..D-1 (ascending), optionally scaled to a cosine with
// pre-computed inverse norms: (dot * invx[i]) * invy[j].  The scan never stores scores: a 128x128 tile is
// reduced in registers, each value is compared against its row's current k-th best (score desc, index asc
// total order) and the few survivors are inserted into a per-row sorted list held in registers of one warp
// (lane r = rank r).  Because the filter and the insertion use the same total order, the result does not
// depend on the order in which threads report candidates: ties always resolve to the lowest index.
#include <limits.h>
#include "common.cuh"
#include "simt_tile.cuh"

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
                                                      float* __restrict__ out_val, int64_t* __restrict__ out_idx) {
  __shared__ __align__(16) float As[BK][LDS];
  __shared__ __align__(16) float Bs[BK][LDS];
  __shared__ float thr_v[BM];
  __shared__ int thr_i[BM];
  __shared__ int cnt[BM];
  __shared__ float cand_v[BM][ROUND_COLS];
  __shared__ int cand_i[BM][ROUND_COLS];

  const int t = threadIdx.x, tx = t & 15, ty = t >> 4, lane = t & 31, warp = t >> 5;
  const int bm = blockIdx.x * BM;
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
    if (m < Q && lane < k) {
      const int64_t o = ((int64_t)blockIdx.y * Q + m) * k + lane;
      const bool empty = li[r] == INT_MAX;
      out_val[o] = empty ? -INFINITY : lv[r];
      out_idx[o] = empty ? -1 : index_base + cbeg + li[r];
    }
  }
}

// one warp per query: merge S lists of k (vals [S,Q,k], idx [S,Q,k]) by (score desc, idx asc)
__global__ void topk_merge_kernel(const float* __restrict__ vals, const int64_t* __restrict__ idx, int S, int Q, int k,
                                  float* __restrict__ out_val, int64_t* __restrict__ out_idx) {
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (q >= Q) return;
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

uint64_t tt_rows_topk_workspace(int32_t Q, int64_t N, int32_t D, int32_t mode) {
  (void)D;
  Workspace w(nullptr, 0);
  if (mode == 1) { w.take<float>(Q); w.take<float>(N); }
  const int S = topk_splits(Q, N);
  if (S > 1) { w.take<float>((int64_t)S * Q * 32); w.take<int64_t>((int64_t)S * Q * 32); }
  return w.used + 256;
}

int tt_rows_topk(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                 int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx, void* workspace,
                 uint64_t workspace_bytes, void* stream) {
  TT_CHECK_ARG(x && y && out_val && out_idx && Q > 0 && N >= 0 && D > 0, "rows_topk: bad args");
  TT_CHECK_ARG(k >= 1 && k <= 32, "rows_topk: k=%d must be in [1,32]", k);
  TT_CHECK_ARG(mode == 0 || mode == 1, "rows_topk: mode must be 0 (dot) or 1 (cosine)");
  cudaStream_t s = (cudaStream_t)stream;
  Workspace w(workspace, workspace_bytes);
  float *invx = nullptr, *invy = nullptr;
  if (mode == 1) { invx = w.take<float>(Q); invy = w.take<float>(N); }
  const int S = topk_splits(Q, N);
  float* pv = nullptr; int64_t* pi = nullptr;
  if (S > 1) { pv = w.take<float>((int64_t)S * Q * 32); pi = w.take<int64_t>((int64_t)S * Q * 32); }
  if (!w.ok()) { set_error("rows_topk: workspace too small"); return TT_ERR_WORKSPACE; }
  if (mode == 1) {
    tt::note_launch(); inv_norm_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, s>>>(x, Q, D, 1e-8f, invx);
    if (N > 0) { note_launch(); inv_norm_kernel<<<(unsigned)ceil_div(N, 8), 256, 0, s>>>(y, N, D, 1e-8f, invy); }
  }
  int64_t cps = ceil_div(N > 0 ? N : 1, S);
  cps = ceil_div(cps, BN) * BN;
  dim3 grid((unsigned)ceil_div(Q, BM), (unsigned)S);
  tt::note_launch(); rows_topk_kernel<<<grid, NT, 0, s>>>(x, Q, y, N, D, k, invx, invy, mask_pos_offset, index_base, cps,
                                      S > 1 ? pv : out_val, S > 1 ? pi : out_idx);
  if (S > 1) { note_launch(); topk_merge_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, s>>>(pv, pi, S, Q, k, out_val, out_idx); }
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_topk_merge(const float* vals, const int64_t* idx, int32_t S, int32_t Q, int32_t k, float* out_val,
                  int64_t* out_idx, void* stream) {
  TT_CHECK_ARG(vals && idx && out_val && out_idx && S > 0 && Q > 0 && k >= 1 && k <= 32, "topk_merge: bad args");
  tt::note_launch(); topk_merge_kernel<<<(unsigned)ceil_div(Q, 8), 256, 0, (cudaStream_t)stream>>>(vals, idx, S, Q, k, out_val, out_idx);
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
