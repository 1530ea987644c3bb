// tcgen05 / TMA / TMEM GEMM engine with bf16x3 operands (see umma_gemm.cuh).
//
// Persistent, warp-specialised: warp 0 = TMA producer (one elected lane), warp 1 = tcgen05.mma issuer (one lane) and
// TMEM owner, warps 2-5 = epilogue (TMEM -> registers -> global).  Tile 128 x BN x 64, operands staged in 128-byte
// swizzled shared memory by TMA (hi and lo planes of A and B per stage), fp32 accumulators double-buffered in TMEM
// so the epilogue of tile i overlaps the main loop of tile i+1.  Per k-block the issuer runs
// 3 (hi*hi, lo*hi, hi*lo) x 4 (UMMA_K = 16) tcgen05.mma into the same accumulator.
#include <limits.h>
#include <string.h>
#include "umma.cuh"
#include "umma_gemm.cuh"

namespace tt {
using namespace umma;

constexpr int UBM = 128, UBK = 64;
constexpr int kThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue (two per TMEM lane quarter, half the columns each)

struct alignas(64) UmmaParams {
  CUtensorMap a_hi[2], a_lo[2], b_hi[2], b_lo[2];
  float* C[2];
  const float* bias[2];
  const float* alpha_ptr;
  int64_t ldc;
  int32_t M, N, K, nbatch, splitk;
  float alpha, beta;
  uint32_t idesc;
  int32_t a_row0[2], b_row0, a_seg, a_skip, a_seg2;
  int32_t epi, rm_Bp, rm_B, rm_T, tl_Bp, tl_G, tl_Hc;
  float* sel_v; int* sel_i; const float* sel_invx;
  int64_t sel_mask_off;
  int32_t sel_k2, sel_slots, sel_S, sel_cpr;   // epi == 2: S column ranges of sel_cpr column tiles per row tile
  int* sel_thr;    // epi == 2: [M] per-row lower bound of the final K2-th best score (float bits; -inf at launch), shared by all
                   //   partitions of a row
  int32_t sel_two; // epi == 2, resident A: B contributes its hi plane only (hi*hi + lo*hi): 2 of the 3 passes, half the B bytes
  int32_t sel_pf;  // epi == 2, resident A: column tiles ahead that one CTA of each lockstep group prefetches into L2 (0 = off)
  int32_t a_res;   // epi == 2, K <= 128: the unit's A tile (every k-block, both planes) stays in shared memory for the whole
                   //   run of column tiles; the stage ring carries B only (L2 -> SM traffic per tile 192 -> 128 KB)
};
__device__ __forceinline__ bool sel_better(float v, int i, float v2, int i2) { return v > v2 || (v == v2 && i < i2); }
constexpr int kSelStride = 256;          // epilogue threads
constexpr int kSelCap = 16;              // list entries per thread held in shared memory (K2 <= kSelCap)
// cold path of the selection epilogue, entered by the whole (converged) warp when any lane has a column that is not below
// its row's current K2-th best: bit j of a lane's `m` = column gn0 + j of its row.  The flagged scores are re-read from
// tensor memory one column at a time (a warp-wide tcgen05.ld per column flagged by ANY lane - one or two per call) instead
// of being passed in an indexable array: that array would live in local memory, i.e. in L2 with ~200 KB of the SM carved
// out as shared memory.  Kept out of line: inlined into the filter it made the hot loop stream its code through the
// instruction cache.
struct SelState { float tv; int ti; int wpos; };   // the WORST entry of the thread's list (score, column) and its position
__device__ __forceinline__ float lds_f32(uint32_t a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ int lds_s32(uint32_t a) { int v; asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void sts_s32(uint32_t a, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
// lv_s / li_s: shared-space addresses of this thread's K2 candidates, entry r at + r * kSelStride * 4, in NO particular order:
// a new candidate replaces the worst entry and the new worst is found by one pass over the entries - 2 * kSelCap independent
// shared loads and a 4-level compare tree (keeping the list sorted cost a chain of ~K2/2 dependent load-compare-store
// steps per insertion).  State goes in and out by value (registers): by reference it lived in local memory = L2.
__device__ __noinline__ SelState sel_insert_warp(uint32_t m, uint32_t taddr, int gn0, int N, int64_t mcol, float mval, int K2,
                                                 uint32_t lv_s, uint32_t li_s, SelState st, int* thr) {
  bool changed = false;
  uint32_t wm = __reduce_or_sync(0xffffffffu, m);
  while (wm) {
    const int j = __ffs(wm) - 1;
    wm &= wm - 1;
    const int col = gn0 + j;
    if (col >= N) break;                                     // bits are in column order (warp-uniform)
    float sc = tmem_ld1(taddr + (uint32_t)j);
    if (!((m >> j) & 1u)) continue;
    if (col == mcol) sc = mval;
    if (sc != sc) sc = INFINITY;
    if (sel_better(sc, col, st.tv, st.ti)) {
      sts_f32(lv_s + st.wpos * (kSelStride * 4), sc);
      sts_s32(li_s + st.wpos * (kSelStride * 4), col);
      float pv[kSelCap];
      int pi[kSelCap], pp[kSelCap];
#pragma unroll
      for (int r = 0; r < kSelCap; ++r) {
        // entries past K2 count as better than anything, so they are never the worst
        pv[r] = r < K2 ? lds_f32(lv_s + r * (kSelStride * 4)) : INFINITY;
        pi[r] = r < K2 ? lds_s32(li_s + r * (kSelStride * 4)) : INT_MIN;
        pp[r] = r;
      }
#pragma unroll
      for (int w = kSelCap / 2; w >= 1; w >>= 1) {
#pragma unroll
        for (int r = 0; r < w; ++r) {
          if (sel_better(pv[r], pi[r], pv[r + w], pi[r + w])) { pv[r] = pv[r + w]; pi[r] = pi[r + w]; pp[r] = pp[r + w]; }
        }
      }
      st.tv = pv[0]; st.ti = pi[0]; st.wpos = pp[0];
      changed = true;
    }
  }
  // publish the row's K2-th best so far: every other partition of the row (the other column half in this CTA, other column
  // ranges in other CTAs) may skip scores below it.  Positive floats order like their bit patterns; the word starts at -inf.
  if (changed && thr && st.tv > 0.f) atomicMax(thr, __float_as_int(st.tv));
  return st;
}

template <int BN, int NPASS>
struct Cfg {
  static constexpr int A_BYTES = UBM * UBK * 2;              // one plane of the A tile
  static constexpr int B_BYTES = BN * UBK * 2;
  static constexpr int NPLANES = NPASS == 3 ? 2 : 1;
  static constexpr int STAGE_BYTES = NPLANES * (A_BYTES + B_BYTES);
  static constexpr int STAGES = (200 * 1024) / STAGE_BYTES > 6 ? 6 : (200 * 1024) / STAGE_BYTES;
  static constexpr int TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
  static constexpr int TR_BYTES = 8 * 32 * 32 * 4;           // one 32x32 fp32 transpose tile per epilogue warp
  // selection epilogue: per-thread candidate lists (kSelCap x (fp32 + int32) x 256 threads = 32 KB) share the transpose
  // area and extend it where the stage ring leaves room
  static constexpr int SEL_BYTES = kSelCap * kSelStride * 8;
  static constexpr bool SEL_OK = STAGES * STAGE_BYTES + 1024 + 256 + SEL_BYTES <= 232448;
  static constexpr int TAIL_BYTES = SEL_OK ? SEL_BYTES : TR_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + TAIL_BYTES;
};

// Drain one accumulator tile (this thread: row `row` of the 128, columns [chalf*BN/2, (chalf+1)*BN/2)) to C: plain row-major
// (transposed through shared memory where a full 32-column block allows), time-major -> batch-first remap, tiled gi layout,
// split-K atomics.  Shared by the single-CTA and the CTA-pair kernels.
template <int BN>
__device__ __forceinline__ void epi_store_tile(const UmmaParams& p, uint32_t tmem_acc, float* tr_smem, int warp, int lane, int quarter,
                                               int chalf, int gm, int m_blk, int n_blk, int sp, int bz, bool has_k, float alpha) {
  const int kb0 = 0, kb1 = has_k ? 1 : 0;
  const float* bias = p.bias[bz];
  bool row_ok = gm < p.M;
  float* crow;
  int64_t tl_base = 0;
  if (p.epi == 1) {                        // tiled: 128-row tile == one batch tile of one timestep
    const int t = gm / p.tl_Bp, mt = (gm % p.tl_Bp) >> 7;
    tl_base = ((int64_t)t * (p.tl_Bp >> 7) + mt) * p.tl_G;
    crow = p.C[bz] + (gm & 127) * 4;
  } else if (p.rm_Bp > 0) {                // time-major rows -> batch-first C rows
    const int t = gm / p.rm_Bp, b = gm % p.rm_Bp;
    row_ok = row_ok && b < p.rm_B;
    crow = p.C[bz] + ((int64_t)b * p.rm_T + t) * p.ldc;
  } else {
    crow = p.C[bz] + (int64_t)gm * p.ldc;
  }
  const bool vec_ok = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C[bz]) & 15) == 0);
#pragma unroll 1
  for (int c0 = chalf * (BN / 2); c0 < (chalf + 1) * (BN / 2); c0 += 32) {
    float v[32];
    tmem_ld32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
    const int gn0 = n_blk * BN + c0;
    // plain row-major store of a full 32-column block: transpose the warp's 32x32 block through shared memory so
    // that 8 lanes write one 128-byte row segment (a TMEM lane is a row: direct stores would put 32 rows, 16 bytes
    // each, into every instruction).  All conditions are warp-uniform.
    if (p.epi == 0 && p.splitk == 1 && p.beta == 0.f && p.rm_Bp == 0 && vec_ok && gn0 + 32 <= p.N && kb0 < kb1) {
      float* T = tr_smem + (warp - 2) * 1024;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 o = make_float4(alpha * v[4 * j], alpha * v[4 * j + 1], alpha * v[4 * j + 2], alpha * v[4 * j + 3]);
        if (bias) { o.x += bias[gn0 + 4 * j]; o.y += bias[gn0 + 4 * j + 1]; o.z += bias[gn0 + 4 * j + 2]; o.w += bias[gn0 + 4 * j + 3]; }
        *reinterpret_cast<float4*>(T + lane * 32 + ((j ^ (lane & 7)) << 2)) = o;
      }
      __syncwarp();
      const int c4 = lane & 7, rsub = lane >> 3;
      float* cbase = p.C[bz] + (int64_t)(m_blk * UBM + quarter * 32) * p.ldc + gn0 + 4 * c4;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int rr = 4 * i + rsub;
        const float4 o = *reinterpret_cast<const float4*>(T + rr * 32 + ((c4 ^ (rr & 7)) << 2));
        if (m_blk * UBM + quarter * 32 + rr < p.M) *reinterpret_cast<float4*>(cbase + (int64_t)rr * p.ldc) = o;
      }
      __syncwarp();
      continue;
    }
    if (row_ok && gn0 < p.N && kb0 < kb1) {
      if (p.epi == 1) {
        const int hc4 = p.tl_Hc >> 2;
        float4* dst[8];
        float4 cv[8], bv[8];
        // (gate, unit) of column n = gn0 + 4 j: ONE division per 32-column block, then steps of 4 with a wrap (eight runtime
        // divisions per block made this loop longer than the MMAs of a K = 300 tile: ncu, tensor pipe 50 % active)
        int g = gn0 / p.tl_Hc, u = gn0 - g * p.tl_Hc;
        const bool bias4 = bias && (reinterpret_cast<uintptr_t>(bias) & 15) == 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {          // all loads first (read-modify-write values, bias): one L2 round trip, not eight
          const int n = gn0 + 4 * j;
          dst[j] = reinterpret_cast<float4*>(crow + ((tl_base + g) * hc4 + (u >> 2)) * 512);
          cv[j] = (p.beta != 0.f && n < p.N) ? *dst[j] : make_float4(0.f, 0.f, 0.f, 0.f);
          bv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias && n < p.N) {
            if (bias4) bv[j] = __ldg(reinterpret_cast<const float4*>(bias + n));
            else bv[j] = make_float4(bias[n], bias[n + 1], bias[n + 2], bias[n + 3]);
          }
          u += 4;
          while (u >= p.tl_Hc) { u -= p.tl_Hc; ++g; }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int n = gn0 + 4 * j;
          if (n < p.N) {
            float4 o = make_float4(alpha * v[4 * j] + p.beta * cv[j].x, alpha * v[4 * j + 1] + p.beta * cv[j].y,
                                   alpha * v[4 * j + 2] + p.beta * cv[j].z, alpha * v[4 * j + 3] + p.beta * cv[j].w);
            o.x += bv[j].x; o.y += bv[j].y; o.z += bv[j].z; o.w += bv[j].w;
            *dst[j] = o;
          }
        }
      } else if (p.splitk > 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (gn0 + j < p.N) {
            float x = alpha * v[j];
            if (sp == 0 && bias) x += bias[gn0 + j];
            atomicAdd(crow + gn0 + j, x);
          }
      } else if (vec_ok && gn0 + 32 <= p.N) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float4 o = make_float4(alpha * v[j], alpha * v[j + 1], alpha * v[j + 2], alpha * v[j + 3]);
          if (p.beta != 0.f) {
            const float4 c = *reinterpret_cast<const float4*>(crow + gn0 + j);
            o.x += p.beta * c.x; o.y += p.beta * c.y; o.z += p.beta * c.z; o.w += p.beta * c.w;
          }
          if (bias) { o.x += bias[gn0 + j]; o.y += bias[gn0 + j + 1]; o.z += bias[gn0 + j + 2]; o.w += bias[gn0 + j + 3]; }
          *reinterpret_cast<float4*>(crow + gn0 + j) = o;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (gn0 + j < p.N) {
            float x = alpha * v[j];
            if (p.beta != 0.f) x += p.beta * crow[gn0 + j];
            if (bias) x += bias[gn0 + j];
            crow[gn0 + j] = x;
          }
      }
    }
  }
}

template <int BN, bool A_MN, bool B_MN, int NPASS>
__global__ void __launch_bounds__(kThreads, 1) umma_gemm_kernel(const __grid_constant__ UmmaParams p) {
  using C_ = Cfg<BN, NPASS>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C_::STAGES * C_::STAGE_BYTES);
  uint64_t* full = bars;                       // [STAGES]
  uint64_t* empty = bars + C_::STAGES;         // [STAGES]
  uint64_t* tfull = bars + 2 * C_::STAGES;     // [4] (2 used unless nacc == 4)
  uint64_t* tempty = tfull + 4;                // [4]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 4);      // <= bars + 20
  uint64_t* afull = bars + 30;                 // resident-A mode: A tile of the unit loaded / no longer read
  uint64_t* aempty = bars + 31;
  uint64_t* rfull = bars + 22;                 // resident-A mode: its own B ring barriers [4] (the ring may be deeper than STAGES)
  uint64_t* rempty = bars + 26;                // [4]
  // accumulators in flight: 2, or - selection epilogue on 128-column tiles - 4 (all 512 TMEM columns): the time the epilogue
  // spends on a tile varies (candidate insertions), and with two accumulators every slow tile stalled the MMA issuer
  const int nacc = (p.epi == 2 && BN == 128) ? 4 : 2;
  const uint32_t tmem_cols = nacc == 4 ? 512 : C_::TMEM_COLS;
  // resident-A layout: [k-block][hi 16 KB | lo 16 KB] at the start of the ring area, then B stages of NPLANES * B_BYTES
  constexpr int RES_A_BYTES = 2 * C_::NPLANES * C_::A_BYTES;
  const int res_stage_bytes = (p.sel_two ? 1 : C_::NPLANES) * C_::B_BYTES;
  const int res_stages_ = (C_::STAGES * C_::STAGE_BYTES - RES_A_BYTES) / res_stage_bytes;
  const int res_stages = res_stages_ > 4 ? 4 : res_stages_;
  float* tr_smem = reinterpret_cast<float*>(smem + C_::STAGES * C_::STAGE_BYTES + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int MT = (p.M + UBM - 1) / UBM, NT = (p.N + BN - 1) / BN, KB = (p.K + UBK - 1) / UBK;
  const int kb_per = (KB + p.splitk - 1) / p.splitk;
  const int items = p.nbatch * p.splitk * MT * NT;
  // work walk: items strided over the persistent grid; epi == 2: units = (row tile, column range) strided over the grid,
  // each unit a run of consecutive column tiles of ONE row tile (the per-row candidate lists live across the run).  Unit
  // u = range * MT + row tile, so the CTAs running at the same time sweep the same columns and share them through L2.
  const bool epi2 = p.epi == 2;
  const int units = epi2 ? MT * p.sel_S : items;
#define TT_WALK_BEGIN                                                                      \
  for (int u_ = blockIdx.x; u_ < units; u_ += gridDim.x) {                                 \
    const int nb0_ = epi2 ? (u_ / MT) * p.sel_cpr : 0;                                     \
    const int nb1_ = epi2 ? min(NT, nb0_ + p.sel_cpr) : 1;                                 \
    for (int nb_ = nb0_; nb_ < nb1_; ++nb_) {                                              \
      const int it = epi2 ? (u_ % MT) * NT + nb_ : u_;
#define TT_WALK_END }}

  if (threadIdx.x == 0) {
    for (int i = 0; i < C_::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 256); }
    mbar_init(afull, 1); mbar_init(aempty, 1);
    for (int i = 0; i < 4; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rempty[i], 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      for (int b = 0; b < p.nbatch; ++b) {
        tma_prefetch_desc(&p.a_hi[b]); tma_prefetch_desc(&p.b_hi[b]);
        if (NPASS == 3) { tma_prefetch_desc(&p.a_lo[b]); tma_prefetch_desc(&p.b_lo[b]); }
      }
      int stage = 0, phase = 0, unit_n = 0;
      TT_WALK_BEGIN
        const int n_blk = it % NT, m_blk = (it / NT) % MT, sp = (it / (NT * MT)) % p.splitk, bz = it / (NT * MT * p.splitk);
        const int kb0 = sp * kb_per, kb1 = min(KB, kb0 + kb_per);
        if (!A_MN && !B_MN && p.a_res) {
          if (nb_ == nb0_) {                   // new unit: its A tile, once the MMAs of the previous unit have read theirs
            mbar_wait(aempty, (unit_n & 1) ^ 1);
            ++unit_n;
            mbar_arrive_expect_tx(afull, KB * C_::NPLANES * C_::A_BYTES);
            for (int kb = 0; kb < KB; ++kb) {
              uint8_t* sa = smem + kb * C_::NPLANES * C_::A_BYTES;
              tma_load_2d(&p.a_hi[0], afull, sa, kb * UBK, m_blk * UBM + p.a_row0[0]);
              if (NPASS == 3) tma_load_2d(&p.a_lo[0], afull, sa + C_::A_BYTES, kb * UBK, m_blk * UBM + p.a_row0[0]);
            }
          }
          // the corpus streams from HBM (it is several times the L2) and the ring holds one tile: the CTAs of a column range
          // take turns to pull the tile sel_pf steps ahead into L2, so the loads below are L2 hits
          if (p.sel_pf > 0 && nb_ + p.sel_pf < nb1_ && (nb_ + p.sel_pf) % MT == m_blk) {
            for (int kb = 0; kb < KB; ++kb) {
              tma_prefetch_l2_2d(&p.b_hi[0], kb * UBK, (nb_ + p.sel_pf) * BN + p.b_row0);
              if (NPASS == 3 && !p.sel_two) tma_prefetch_l2_2d(&p.b_lo[0], kb * UBK, (nb_ + p.sel_pf) * BN + p.b_row0);
            }
          }
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&rempty[stage], phase ^ 1);
            uint8_t* sb = smem + RES_A_BYTES + stage * res_stage_bytes;
            mbar_arrive_expect_tx(&rfull[stage], res_stage_bytes);
            tma_load_2d(&p.b_hi[0], &rfull[stage], sb, kb * UBK, n_blk * BN + p.b_row0);
            if (NPASS == 3 && !p.sel_two) tma_load_2d(&p.b_lo[0], &rfull[stage], sb + C_::B_BYTES, kb * UBK, n_blk * BN + p.b_row0);
            if (++stage == res_stages) { stage = 0; phase ^= 1; }
          }
          continue;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* st = smem + stage * C_::STAGE_BYTES;
          uint8_t* sa_hi = st, *sb_hi = st + C_::A_BYTES;
          uint8_t* sa_lo = st + C_::A_BYTES + C_::B_BYTES, *sb_lo = sa_lo + C_::A_BYTES;
          mbar_arrive_expect_tx(&full[stage], C_::STAGE_BYTES);
          if (!A_MN) {
            int ka = kb * UBK;
            if (p.a_seg > 0 && ka >= p.a_seg) ka += p.a_skip;
            if (p.a_seg2 > 0 && kb * UBK >= p.a_seg2) ka += p.a_skip;
            tma_load_2d(&p.a_hi[bz], &full[stage], sa_hi, ka, m_blk * UBM + p.a_row0[bz]);
            if (NPASS == 3) tma_load_2d(&p.a_lo[bz], &full[stage], sa_lo, ka, m_blk * UBM + p.a_row0[bz]);
          } else {
            int ma = m_blk * UBM;
            if (p.a_seg > 0 && ma >= p.a_seg) ma += p.a_skip;
#pragma unroll
            for (int h = 0; h < UBM / 64; ++h) {
              tma_load_2d(&p.a_hi[bz], &full[stage], sa_hi + h * 8192, ma + h * 64, kb * UBK);
              if (NPASS == 3) tma_load_2d(&p.a_lo[bz], &full[stage], sa_lo + h * 8192, ma + h * 64, kb * UBK);
            }
          }
          if (!B_MN) {
            tma_load_2d(&p.b_hi[bz], &full[stage], sb_hi, kb * UBK, n_blk * BN + p.b_row0);
            if (NPASS == 3) tma_load_2d(&p.b_lo[bz], &full[stage], sb_lo, kb * UBK, n_blk * BN + p.b_row0);
          } else {
#pragma unroll
            for (int h = 0; h < BN / 64; ++h) {
              tma_load_2d(&p.b_hi[bz], &full[stage], sb_hi + h * 8192, n_blk * BN + h * 64, kb * UBK);
              if (NPASS == 3) tma_load_2d(&p.b_lo[bz], &full[stage], sb_lo + h * 8192, n_blk * BN + h * 64, kb * UBK);
            }
          }
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1; }
        }
      TT_WALK_END
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = p.idesc;
      // K-major: rows of 128 B, 8-row groups 1024 B apart.  MN-major: 64-element groups 8192 B apart (LBO),
      // 8-k groups 1024 B apart (SBO).  One UMMA_K=16 step advances 32 B (K-major) or 2048 B (MN-major).
      constexpr uint32_t a_lbo = A_MN ? 8192 : 16, a_step = A_MN ? 2048 : 32;
      constexpr uint32_t b_lbo = B_MN ? 8192 : 16, b_step = B_MN ? 2048 : 32;
      int stage = 0, phase = 0, tile = 0, unit_n = 0;
      TT_WALK_BEGIN
        const int sp = (it / (NT * MT)) % p.splitk;
        const int kb0 = sp * kb_per, kb1 = min(KB, kb0 + kb_per);
        const int acc = nacc == 4 ? (tile & 3) : (tile & 1), acc_phase = (nacc == 4 ? (tile >> 2) : (tile >> 1)) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        if (!A_MN && !B_MN && p.a_res) {
          if (nb_ == nb0_) { mbar_wait(afull, unit_n & 1); ++unit_n; }
          const int nterms = p.sel_two ? 2 : NPASS;
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&rfull[stage], phase);
            tc_fence_after();
            const uint32_t sa_hi = smem_u32(smem + kb * C_::NPLANES * C_::A_BYTES), sa_lo = sa_hi + C_::A_BYTES;
            const uint32_t sb_hi = smem_u32(smem + RES_A_BYTES + stage * res_stage_bytes), sb_lo = sb_hi + C_::B_BYTES;
#pragma unroll
            for (int term = 0; term < NPASS; ++term) {
              if (term >= nterms) break;
              const uint32_t sa = term == 1 ? sa_lo : sa_hi;
              const uint32_t sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
              for (int k = 0; k < UBK / 16; ++k) {
                const uint64_t ad = make_smem_desc(sa + k * a_step, a_lbo, 1024);
                const uint64_t bd = make_smem_desc(sb + k * b_step, b_lbo, 1024);
                mma_bf16_ss(d_tmem, ad, bd, idesc, (kb > 0 || term > 0 || k > 0) ? 1u : 0u);
              }
            }
            mma_commit(&rempty[stage]);
            if (++stage == res_stages) { stage = 0; phase ^= 1; }
          }
          mma_commit(&tfull[acc]);
          if (nb_ == nb1_ - 1) mma_commit(aempty);     // the unit's A tile may be replaced once everything issued so far has read it
          ++tile;
          continue;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + stage * C_::STAGE_BYTES);
          const uint32_t sa_hi = st, sb_hi = st + C_::A_BYTES;
          const uint32_t sa_lo = st + C_::A_BYTES + C_::B_BYTES, sb_lo = sa_lo + C_::A_BYTES;
#pragma unroll
          for (int term = 0; term < NPASS; ++term) {
            const uint32_t sa = term == 1 ? sa_lo : sa_hi;
            const uint32_t sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
            for (int k = 0; k < UBK / 16; ++k) {
              const uint64_t ad = make_smem_desc(sa + k * a_step, a_lbo, 1024);
              const uint64_t bd = make_smem_desc(sb + k * b_step, b_lbo, 1024);
              mma_bf16_ss(d_tmem, ad, bd, idesc, (kb > kb0 || term > 0 || k > 0) ? 1u : 0u);
            }
          }
          mma_commit(&empty[stage]);          // smem slot reusable once these MMAs have read it
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit(&tfull[acc]);              // accumulator complete -> epilogue
        ++tile;
      TT_WALK_END
    }
  } else {
    // ================================================================= epilogue (warps 2..5)
    const int quarter = warp & 3;             // TMEM lanes this warp may touch: [32*quarter, 32*quarter+32)
    const int chalf = (warp - 2) >> 2;        // which half of the tile's columns this warp drains
    const int row = quarter * 32 + lane;
    const float alpha = p.alpha * (p.alpha_ptr ? __ldg(p.alpha_ptr) : 1.f);
    int tile = 0;
    // epi == 2 state: this thread's K2 best of (row, its half of the columns) for the current row tile, sorted
    float* lv = tr_smem + (threadIdx.x - 64);                                 // [r * kSelStride]
    int* li = reinterpret_cast<int*>(tr_smem + kSelCap * kSelStride) + (threadIdx.x - 64);
    const uint32_t lv_s = smem_u32(lv), li_s = smem_u32(li);
    SelState st{-INFINITY, INT_MAX, 0};
    float sel_mval = -1.f, gthr_next = -INFINITY;
    int cur_u = -1;
    auto sel_flush = [&](int u_done) {                       // unit = (column range, row tile)
      if (u_done < 0) return;
      const int gmf = (u_done % MT) * UBM + row;
      if (gmf >= p.M) return;
      const int part = u_done / MT;
      const int64_t o = ((int64_t)gmf * p.sel_slots + 2 * part + chalf) * 32;
      for (int r = 0; r < p.sel_k2; ++r) { p.sel_v[o + r] = lv[r * kSelStride]; p.sel_i[o + r] = li[r * kSelStride]; }
    };
    TT_WALK_BEGIN
      const int n_blk = it % NT, m_blk = (it / NT) % MT, sp = (it / (NT * MT)) % p.splitk, bz = it / (NT * MT * p.splitk);
      const int kb0 = sp * kb_per, kb1 = min(KB, kb0 + kb_per);
      const int acc = nacc == 4 ? (tile & 3) : (tile & 1), acc_phase = (nacc == 4 ? (tile >> 2) : (tile >> 1)) & 1;
      ++tile;
      if (epi2 && u_ != cur_u) {
        sel_flush(cur_u);
        cur_u = u_;
        for (int r = 0; r < kSelCap; ++r) { lv[r * kSelStride] = -INFINITY; li[r * kSelStride] = INT_MAX; }
        st.tv = -INFINITY; st.ti = INT_MAX; st.wpos = 0;
        // per unit, not per tile (both are global loads whose L2 latency - ~1 us per tile - sat on the critical path)
        const int gmu = m_blk * UBM + row;
        sel_mval = (p.sel_invx && gmu < p.M) ? -1.f / __ldg(p.sel_invx + gmu) : -1.f;
        gthr_next = -INFINITY;
      }
      const int gm = m_blk * UBM + row;
      // the row's shared lower bound: the value used for this tile was loaded while the previous tile was filtered, the load
      // issued here is consumed by the next tile (a stale value is still a valid bound, it only filters less)
      const float gthr = gthr_next;
      if (epi2 && p.sel_thr && gm < p.M) gthr_next = __int_as_float(__ldcg(p.sel_thr + gm));
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      if (p.epi == 2) {
        const bool rok = gm < p.M;
        const int64_t mcol = p.sel_mask_off >= 0 ? p.sel_mask_off + gm : -1;
        const float mval = sel_mval;
        const int K2 = p.sel_k2;
#pragma unroll 1
        for (int c0 = chalf * (BN / 2); c0 < (chalf + 1) * (BN / 2); c0 += 32) {
          float v[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + c0);
          tmem_ld32(taddr, v);
          const int gn0 = n_blk * BN + c0;
          if (gn0 >= p.N) continue;
          // hot path: one 3-input max tree over the 32 scores and ONE comparison with the row's K2-th best (NaN propagates
          // and compares as "not below")
          float t[12];
#pragma unroll
          for (int j = 0; j < 10; ++j) t[j] = max3_nan(v[3 * j], v[3 * j + 1], v[3 * j + 2]);
          t[10] = v[30]; t[11] = v[31];
#pragma unroll
          for (int j = 0; j < 4; ++j) t[j] = max3_nan(t[3 * j], t[3 * j + 1], t[3 * j + 2]);
          const float mx = max3_nan(max3_nan(t[0], t[1], t[2]), t[3], t[3]);
          const bool inmask = gn0 <= mcol && mcol < gn0 + 32;     // the masked column is judged by its mask value
          const float tv = fmaxf(st.tv, gthr);
          const bool hit = rok && (!(mx < tv) || inmask);
          if (__any_sync(0xffffffffu, hit)) {
            uint32_t m = 0;
            if (hit) {
#pragma unroll
              for (int j = 0; j < 32; ++j) m |= !(v[j] < tv) ? (1u << j) : 0u;      // also set for NaN
              if (inmask) m |= 1u << (int)(mcol - gn0);
            }
            st = sel_insert_warp(m, taddr, gn0, p.N, mcol, mval, K2, lv_s, li_s, st, (rok && p.sel_thr) ? p.sel_thr + gm : nullptr);
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty[acc]);
        continue;
      }
      epi_store_tile<BN>(p, tmem_base + (uint32_t)(acc * BN), tr_smem, warp, lane, quarter, chalf, gm, m_blk, n_blk, sp, bz, kb0 < kb1, alpha);
      tc_fence_before();
      mbar_arrive(&tempty[acc]);
    TT_WALK_END
    if (epi2) sel_flush(cur_u);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}


// ------------------------------------------------------------------------------------------ CTA-pair variant (cta_group::2)
// 256 x 256 x 64 tiles on 2-CTA clusters, K-major operands: each CTA stages its own 128 rows of A and HALF of the B tile
// (128 of the 256 columns), the leader's single thread issues M = 256 tcgen05.mma.cta_group::2 that read both CTAs' shared
// memory and write both CTAs' tensor memory; every CTA drains its own 128 x 256 accumulator with the shared epilogue.  Why:
// the 3-pass GEMMs are bound by operand feed per SM (96 KB of planes per k-block at ~67 GB/s; ncu: tensor pipe 50-76 %
// active, DRAM 35 %) - a pair pulls 64 KB per CTA and k-block for the same MMAs.  Both batched problems of a row tile run
// back to back (they share A), so A crosses DRAM once.
template <int NPASS>
struct PairCfg {
  static constexpr int BN = 256;
  static constexpr int A_BYTES = UBM * UBK * 2;
  static constexpr int B_BYTES = (BN / 2) * UBK * 2;
  static constexpr int NPLANES = NPASS == 3 ? 2 : 1;
  static constexpr int STAGE_BYTES = NPLANES * (A_BYTES + B_BYTES);
  static constexpr int STAGES = (192 * 1024) / STAGE_BYTES > 6 ? 6 : (192 * 1024) / STAGE_BYTES;
  static constexpr int TR_BYTES = 8 * 32 * 32 * 4;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + TR_BYTES;
};

template <int NPASS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) umma_gemm_pair_kernel(const __grid_constant__ UmmaParams p) {
  using C_ = PairCfg<NPASS>;
  constexpr int BN = C_::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C_::STAGES * C_::STAGE_BYTES);
  uint64_t* full = bars;                       // [STAGES]  (the LEADER's are used: bytes of both CTAs' loads)
  uint64_t* empty = bars + C_::STAGES;         // [STAGES]  (own: multicast commit of the leader's MMAs)
  uint64_t* tfull = bars + 2 * C_::STAGES;     // [2]       (own: multicast commit)
  uint64_t* tempty = tfull + 2;                // [2]       (the LEADER's: one arrival per CTA)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  float* tr_smem = reinterpret_cast<float*>(smem + C_::STAGES * C_::STAGE_BYTES + 256);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int MT = (p.M + UBM - 1) / UBM, NT = (p.N + BN - 1) / BN, KB = (p.K + UBK - 1) / UBK;
  const int MTP = MT >> 1;                     // row-tile pairs (the host guarantees an even MT)
  const int items = p.nbatch * MTP * NT;
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  // item -> column tile fastest, then the batched problem, then the row-tile pair: consecutive items share A
#define TT_PAIR_ITEM                                                                             \
  const int n_blk = it % NT, bz = (it / NT) % p.nbatch, m_blk = 2 * (it / (NT * p.nbatch)) + (int)rank;

  if (threadIdx.x == 0) {
    for (int i = 0; i < C_::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 2); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // both CTAs' barriers exist before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer (both CTAs: own A rows, own half of B)
    if (lane == 0) {
      for (int b = 0; b < p.nbatch; ++b) {
        tma_prefetch_desc(&p.a_hi[b]); tma_prefetch_desc(&p.b_hi[b]);
        if (NPASS == 3) { tma_prefetch_desc(&p.a_lo[b]); tma_prefetch_desc(&p.b_lo[b]); }
      }
      int stage = 0, phase = 0;
      for (int it = pair; it < items; it += npairs) {
        TT_PAIR_ITEM
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* st = smem + stage * C_::STAGE_BYTES;
          uint8_t* sa_hi = st, *sb_hi = st + C_::A_BYTES;
          uint8_t* sa_lo = st + C_::A_BYTES + C_::B_BYTES, *sb_lo = sa_lo + C_::A_BYTES;
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C_::STAGE_BYTES);
          int ka = kb * UBK;
          if (p.a_seg > 0 && ka >= p.a_seg) ka += p.a_skip;
          if (p.a_seg2 > 0 && kb * UBK >= p.a_seg2) ka += p.a_skip;
          tma_load_2d_pair(&p.a_hi[bz], &full[stage], sa_hi, ka, m_blk * UBM + p.a_row0[bz]);
          if (NPASS == 3) tma_load_2d_pair(&p.a_lo[bz], &full[stage], sa_lo, ka, m_blk * UBM + p.a_row0[bz]);
          tma_load_2d_pair(&p.b_hi[bz], &full[stage], sb_hi, kb * UBK, n_blk * BN + (int)rank * (BN / 2) + p.b_row0);
          if (NPASS == 3) tma_load_2d_pair(&p.b_lo[bz], &full[stage], sb_lo, kb * UBK, n_blk * BN + (int)rank * (BN / 2) + p.b_row0);
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (one thread of the pair)
    if (lane == 0 && leader) {
      const uint32_t idesc = p.idesc;            // M = 256, N = 256
      int stage = 0, phase = 0, tile = 0;
      for (int it = pair; it < items; it += npairs, ++tile) {
        const int acc = tile & 1, acc_phase = (tile >> 1) & 1;
        mbar_wait_cluster(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t st = smem_u32(smem + stage * C_::STAGE_BYTES);
          const uint32_t sa_hi = st, sb_hi = st + C_::A_BYTES;
          const uint32_t sa_lo = st + C_::A_BYTES + C_::B_BYTES, sb_lo = sa_lo + C_::A_BYTES;
#pragma unroll
          for (int term = 0; term < NPASS; ++term) {
            const uint32_t sa = term == 1 ? sa_lo : sa_hi;
            const uint32_t sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
            for (int k = 0; k < UBK / 16; ++k)
              mma_f16_ss_pair(d_tmem, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                              (kb > 0 || term > 0 || k > 0) ? 1u : 0u);
          }
          mma_commit_pair(&empty[stage], 3);     // both CTAs' stage is reusable once these MMAs have read it
          if (++stage == C_::STAGES) { stage = 0; phase ^= 1; }
        }
        mma_commit_pair(&tfull[acc], 3);         // accumulator complete in both CTAs -> their epilogues
      }
    }
  } else {
    // ================================================================= epilogue (warps 2..9 of both CTAs: own 128 rows)
    const int quarter = warp & 3, chalf = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const float alpha = p.alpha * (p.alpha_ptr ? __ldg(p.alpha_ptr) : 1.f);
    int tile = 0;
    for (int it = pair; it < items; it += npairs, ++tile) {
      TT_PAIR_ITEM
      const int acc = tile & 1, acc_phase = (tile >> 1) & 1;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      epi_store_tile<BN>(p, tmem_base + (uint32_t)(acc * BN), tr_smem, warp, lane, quarter, chalf, m_blk * UBM + row, m_blk, n_blk, 0, bz,
                         true, alpha);
      tc_fence_before();
      named_bar_sync(1, 256);
      if (threadIdx.x == 64) {
        if (leader) mbar_arrive(&tempty[acc]);
        else mbar_arrive_leader(&tempty[acc]);
      }
    }
  }
#undef TT_PAIR_ITEM
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                          // nobody leaves (or frees tensor memory) while the pair's MMAs / commits may touch it
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}


// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}
bool umma_available() { return encode_fn() != nullptr; }

// 2-D 16-bit map over a row-major [rows][pitch] plane exposing `cols` logical columns; box = {64 cols, box_rows};
// 128-byte swizzle; out-of-bounds elements read as zero (this is what makes ragged M/N/K tiles correct).
int make_plane_map(CUtensorMap* m, const void* base, int64_t rows, int64_t cols, int64_t pitch, int box_rows, int f16) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled not available"); return TT_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  if ((reinterpret_cast<uintptr_t>(base) & 15) || (strides[0] & 15)) {
    set_error("umma: plane base/pitch not 16-byte aligned");
    return TT_ERR_INVALID;
  }
  CUresult r = fn(m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base),
                  dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld pitch=%lld", (int)r, (long long)rows, (long long)cols, (long long)pitch); return TT_ERR_CUDA; }
  return TT_OK;
}

__global__ void prescale_kernel(float* C, int64_t ldc, int M, int N, float beta) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * N) return;
  float* q = C + (i / N) * ldc + (i % N);
  *q = beta == 0.f ? 0.f : beta * *q;
}

template <int BN, bool A_MN, bool B_MN, int NPASS>
static int launch_cfg(const UmmaParams& p, int items, int max_ctas, cudaStream_t s) {
  auto kern = umma_gemm_kernel<BN, A_MN, B_MN, NPASS>;
  static DeviceOnce attr_once;                  // one per template instantiation
  if (attr_once.need()) {
    TT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN, NPASS>::SMEM_BYTES));
    attr_once.done();
  }
  const int num_sms = device_sm_count();
  const int cap = (max_ctas > 0 && max_ctas < num_sms && p.epi != 2) ? max_ctas : num_sms;
  const int grid = items < cap ? items : cap;
  note_launch();
  prof_begin_gemm(s, p.M, p.N, p.K, p.nbatch, NPASS, (A_MN ? 1 : 0) | (B_MN ? 2 : 0) | (p.epi << 4) | (p.splitk > 1 ? 256 : 0));
  ProfScope prof(PROF_GEMM_TC, s);
  kern<<<grid, kThreads, Cfg<BN, NPASS>::SMEM_BYTES, s>>>(p);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

template <int NPASS>
static int launch_pair(const UmmaParams& p, int items, int max_ctas, cudaStream_t s) {
  auto kern = umma_gemm_pair_kernel<NPASS>;
  static DeviceOnce attr_once;
  if (attr_once.need()) {
    TT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfg<NPASS>::SMEM_BYTES));
    attr_once.done();
  }
  const int num_sms = device_sm_count();
  const int cap = (max_ctas > 0 && max_ctas < num_sms) ? max_ctas : num_sms;
  int pairs = cap / 2;
  if (items < pairs) pairs = items;
  if (pairs < 1) pairs = 1;
  note_launch();
  prof_begin_gemm(s, p.M, p.N, p.K, p.nbatch, NPASS, (p.epi << 4) | 512);       // flag 512: CTA pairs
  ProfScope prof(PROF_GEMM_TC, s);
  kern<<<2 * pairs, kThreads, PairCfg<NPASS>::SMEM_BYTES, s>>>(p);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

static_assert(sizeof(UmmaParams) <= sizeof(((UmmaPrepared*)0)->blob), "UmmaPrepared blob too small");

static int sm_count() {
  return device_sm_count();
}
// slots per row of the selection epilogue's output for an [M, N] problem (must mirror umma_gemm_prepare)
static int g_sel_resident = 1, g_sel_prefetch = 4, g_sel_bn128 = 0;   // 128-column tiles x 4 accumulators measured slower (7.8 vs 5.65 ms)
void umma_set_sel_resident(int on) { g_sel_resident = on ? 1 : 0; g_sel_bn128 = (on & 4) ? 1 : 0; }
void umma_set_sel_prefetch(int tiles) { g_sel_prefetch = tiles < 0 ? 0 : tiles; }
static int g_gemm_pairs = 1;             // 1: large K-major GEMMs run on 2-CTA clusters (umma_gemm_pair_kernel)
void umma_set_pairs(int on) { g_gemm_pairs = on ? 1 : 0; }

bool umma_sel_two_ok(int32_t K) { return g_sel_resident && ceil_div(K, UBK) <= 2; }

int umma_sel_slots(int32_t M, int64_t N) {
  // sized for 128-column tiles (the most parts a row can have); with 256-column tiles the extra slots stay empty
  const int64_t MT = ceil_div(M, UBM), BN = 128, NT = ceil_div(N, BN);
  int64_t S = sm_count() / MT < 1 ? 1 : sm_count() / MT;
  if (S > NT) S = NT;
  return (int)(2 * S);
}

int umma_gemm_prepare(const UmmaGemm& g, UmmaPrepared* out) {
  memset(out, 0, sizeof(*out));
  if (g.M <= 0 || g.N <= 0) { out->kind = -1; return TT_OK; }
  TT_CHECK_ARG(g.nbatch >= 1 && g.nbatch <= 2 && g.K > 0, "umma_gemm: bad nbatch/K");
  const int npass = g.npass == 1 ? 1 : 3;
  const int MT = (int)ceil_div(g.M, UBM), KB = (int)ceil_div(g.K, UBK);
  // BN = 128 when N is small, or when 256-wide tiles would leave most SMs idle
  int BN = g.N <= 128 ? 128 : 256;
  if (g.epi == 2 && g_sel_bn128 && g_sel_resident && KB <= 2) BN = 128;      // 4 accumulators in flight (see the kernel)
  if (BN == 256 && g.splitk == 1 && MT * ceil_div(g.N, 256) * g.nbatch < 100 && (g.N % 256) == 0 && g.epi != 2) BN = 128;
  // CTA pairs: K-major operands, 256-column tiles, an even number of row tiles and enough of them to fill the GPU
  const bool pair = g_gemm_pairs && BN == 256 && !g.a_mn && !g.b_mn && g.epi != 2 && g.splitk == 1 && (MT % 2) == 0 &&
                    (int64_t)MT * ceil_div(g.N, 256) * g.nbatch >= 2 * 148;
  UmmaParams& p = *reinterpret_cast<UmmaParams*>(out->blob);
  const int f16 = g.f16;
  const int64_t skipA = g.a_seg > 0 ? (int64_t)g.a_skip * (g.a_seg2 > 0 ? 2 : 1) : 0;
  for (int b = 0; b < g.nbatch; ++b) {
    // K-major: plane [MN rows][K cols], box rows = tile MN extent.  MN-major: plane [K rows][MN cols], box 64x64.
    if (!g.a_mn) {
      const int64_t rows = g.a_rows > 0 ? g.a_rows : g.M;
      TT_TRY(make_plane_map(&p.a_hi[b], g.A[b].hi, rows, g.K + skipA, g.A[b].pitch, UBM, f16));
      if (npass == 3) TT_TRY(make_plane_map(&p.a_lo[b], g.A[b].lo, rows, g.K + skipA, g.A[b].pitch, UBM, f16));
    } else {
      const int64_t rows = g.a_rows > 0 ? g.a_rows : g.K;
      TT_TRY(make_plane_map(&p.a_hi[b], g.A[b].hi, rows, g.M + skipA, g.A[b].pitch, 64, f16));
      if (npass == 3) TT_TRY(make_plane_map(&p.a_lo[b], g.A[b].lo, rows, g.M + skipA, g.A[b].pitch, 64, f16));
    }
    if (!g.b_mn) {
      const int64_t rows = g.b_rows > 0 ? g.b_rows : g.N;
      TT_TRY(make_plane_map(&p.b_hi[b], g.B[b].hi, rows, g.K, g.B[b].pitch, pair ? BN / 2 : BN, f16));
      if (npass == 3) TT_TRY(make_plane_map(&p.b_lo[b], g.B[b].lo, rows, g.K, g.B[b].pitch, pair ? BN / 2 : BN, f16));
    } else {
      const int64_t rows = g.b_rows > 0 ? g.b_rows : g.K;
      TT_TRY(make_plane_map(&p.b_hi[b], g.B[b].hi, rows, g.N, g.B[b].pitch, 64, f16));
      if (npass == 3) TT_TRY(make_plane_map(&p.b_lo[b], g.B[b].lo, rows, g.N, g.B[b].pitch, 64, f16));
    }
    p.C[b] = g.C[b]; p.bias[b] = g.bias[b]; p.a_row0[b] = g.a_row0;
  }
  p.ldc = g.ldc; p.M = g.M; p.N = g.N; p.K = g.K; p.nbatch = g.nbatch; p.alpha = g.alpha; p.beta = g.beta;
  p.alpha_ptr = g.alpha_ptr; p.b_row0 = g.b_row0; p.a_seg = g.a_seg; p.a_skip = g.a_skip; p.a_seg2 = g.a_seg2;
  p.epi = g.epi; p.rm_Bp = g.rm_Bp; p.rm_B = g.rm_B; p.rm_T = g.rm_T; p.tl_Bp = g.tl_Bp; p.tl_G = g.tl_G; p.tl_Hc = g.tl_Hc;
  p.idesc = f16 ? make_idesc_f16(pair ? 2 * UBM : UBM, BN, g.a_mn ? 1 : 0, g.b_mn ? 1 : 0)
                : make_idesc_bf16(pair ? 2 * UBM : UBM, BN, g.a_mn ? 1 : 0, g.b_mn ? 1 : 0);
  if (g.epi == 1)
    TT_CHECK_ARG(g.splitk == 1 && (g.tl_Bp % 128) == 0 && (g.tl_Hc % 4) == 0 && (g.N % 4) == 0 && (g.M % 128) == 0,
                 "umma_gemm: tiled epilogue needs splitk 1, Bp %% 128 == 0, Hc %% 4 == 0");
  const int NT = (int)ceil_div(g.N, BN);
  if (g.epi == 2) {
    TT_CHECK_ARG(g.splitk == 1 && g.nbatch == 1 && !g.a_mn && !g.b_mn && g.sel_v && g.sel_i && g.sel_k2 >= 1 && g.sel_k2 <= kSelCap &&
                     npass == 3,
                 "umma_gemm: selection epilogue needs K-major operands, 3 passes, splitk 1, nbatch 1, 1 <= K2 <= 16");
    p.sel_S = sm_count() / MT < 1 ? 1 : sm_count() / MT;
    if (p.sel_S > NT) p.sel_S = NT;
    p.sel_cpr = (int)ceil_div(NT, p.sel_S);
    TT_CHECK_ARG(g.sel_slots >= 2 * p.sel_S, "umma_gemm: sel_slots too small");
    p.sel_v = g.sel_v; p.sel_i = g.sel_i; p.sel_invx = g.sel_invx; p.sel_mask_off = g.sel_mask_off;
    p.sel_k2 = g.sel_k2; p.sel_slots = g.sel_slots;
    p.a_res = (g_sel_resident && KB <= 2) ? 1 : 0;
    p.sel_pf = g_sel_prefetch;
    TT_CHECK_ARG(!g.sel_two || p.a_res, "umma_gemm: two-pass selection needs the resident-A mode (K <= 128)");
    p.sel_two = g.sel_two ? 1 : 0;
    p.sel_thr = g.sel_thr;
  }
  int splitk = g.splitk;
  if (splitk <= 0) {
    splitk = 1;
    const int tiles = MT * NT * g.nbatch;
    if (tiles < 74 && KB >= 32) {
      // ONE wave of the persistent grid: tiles * splitk <= CTAs (rounding up gave e.g. 6 tiles x 25 splits = 150 items on
      // 148 CTAs - a second wave of two items that took as long as the first)
      const int ctas = (g.max_ctas > 0 && g.max_ctas < 148) ? g.max_ctas : 148;
      splitk = ctas / tiles;
      if (splitk > KB / 8) splitk = KB / 8;
      if (splitk < 1) splitk = 1;
    }
  }
  if (splitk > KB) splitk = KB;
  // no empty splits: shrink until ceil(KB/kb_per) == splitk
  while (splitk > 1 && (int)ceil_div(KB, ceil_div(KB, splitk)) < splitk) --splitk;
  p.splitk = splitk;
  out->items = MT * NT * g.nbatch * splitk;
  if (g.epi == 2) out->items = MT * p.sel_S;                 // grid = work units (row tile, column range)
  const int major = (!g.a_mn && !g.b_mn) ? 0 : (g.a_mn && g.b_mn) ? 1 : (!g.a_mn && g.b_mn) ? 2 : 3;
  if (major == 3) {
    set_error("umma_gemm: operand major / pass combination not instantiated");
    return TT_ERR_INVALID;
  }
  out->kind = (BN == 128 ? 0 : 8) + (npass == 1 ? 4 : 0) + major;
  if (pair) {
    out->kind = npass == 1 ? 17 : 16;
    out->items = (MT / 2) * NT * g.nbatch;
  }
  out->max_ctas = g.max_ctas;
  return TT_OK;
}

int umma_gemm_launch(const UmmaPrepared& prep, const int* a_row0, cudaStream_t s) {
  if (prep.kind < 0) return TT_OK;
  UmmaParams p = *reinterpret_cast<const UmmaParams*>(prep.blob);
  if (a_row0)
    for (int b = 0; b < p.nbatch; ++b) p.a_row0[b] = a_row0[b];
  if (p.splitk > 1 && p.beta != 1.f) {
    for (int b = 0; b < p.nbatch; ++b) {
      const int64_t n = (int64_t)p.M * p.N;
      note_launch();
      prescale_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(p.C[b], p.ldc, p.M, p.N, p.beta);
    }
    TT_LAUNCH_CHECK();
  }
  const int items = prep.items;
  switch (prep.kind) {
    case 0: return launch_cfg<128, false, false, 3>(p, items, prep.max_ctas, s);
    case 1: return launch_cfg<128, true, true, 3>(p, items, prep.max_ctas, s);
    case 2: return launch_cfg<128, false, true, 3>(p, items, prep.max_ctas, s);
    case 4: return launch_cfg<128, false, false, 1>(p, items, prep.max_ctas, s);
    case 5: return launch_cfg<128, true, true, 1>(p, items, prep.max_ctas, s);
    case 6: return launch_cfg<128, false, true, 1>(p, items, prep.max_ctas, s);
    case 8: return launch_cfg<256, false, false, 3>(p, items, prep.max_ctas, s);
    case 9: return launch_cfg<256, true, true, 3>(p, items, prep.max_ctas, s);
    case 10: return launch_cfg<256, false, true, 3>(p, items, prep.max_ctas, s);
    case 12: return launch_cfg<256, false, false, 1>(p, items, prep.max_ctas, s);
    case 13: return launch_cfg<256, true, true, 1>(p, items, prep.max_ctas, s);
    case 14: return launch_cfg<256, false, true, 1>(p, items, prep.max_ctas, s);
    case 16: return launch_pair<3>(p, items, prep.max_ctas, s);
    case 17: return launch_pair<1>(p, items, prep.max_ctas, s);
  }
  set_error("umma_gemm: bad prepared kind %d", prep.kind);
  return TT_ERR_INVALID;
}

int umma_gemm(const UmmaGemm& g, cudaStream_t s) {
  UmmaPrepared prep;
  TT_TRY(umma_gemm_prepare(g, &prep));
  return umma_gemm_launch(prep, nullptr, s);
}

// ------------------------------------------------------------------------------------------ plane producers
__global__ void split_planes_kernel(const float* __restrict__ x, int64_t ld, int64_t rows, int cols, bf16* __restrict__ hi,
                                    bf16* __restrict__ lo, int64_t pitch, float p, float inv_keep, uint64_t seed,
                                    uint64_t offset) {
  const int64_t pairs_per_row = pitch >> 1;
  const int64_t total = rows * pairs_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / pairs_per_row;
    const int c = (int)(i % pairs_per_row) * 2;
    float v0 = c < cols ? x[r * ld + c] : 0.f;
    float v1 = c + 1 < cols ? x[r * ld + c + 1] : 0.f;
    if (p > 0.f) {
      if (c < cols) v0 *= dropout_scale(seed, offset, (uint64_t)(r * cols + c), p, inv_keep);
      if (c + 1 < cols) v1 *= dropout_scale(seed, offset, (uint64_t)(r * cols + c + 1), p, inv_keep);
    }
    bf16 h0, l0, h1, l1;
    split_bf16(v0, h0, l0);
    split_bf16(v1, h1, l1);
    *reinterpret_cast<__nv_bfloat162*>(hi + r * pitch + c) = __halves2bfloat162(h0, h1);
    *reinterpret_cast<__nv_bfloat162*>(lo + r * pitch + c) = __halves2bfloat162(l0, l1);
  }
}
int split_planes(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch, float p,
                 uint64_t seed, uint64_t offset, cudaStream_t s) {
  if (rows <= 0) return TT_OK;
  int64_t n = rows * (pitch >> 1), nb = ceil_div(n, 256);
  if (nb > 148 * 32) nb = 148 * 32;
  note_launch();
  split_planes_kernel<<<(unsigned)nb, 256, 0, s>>>(x, ld, rows, cols, hi, lo, pitch, p, p > 0.f ? 1.f / (1.f - p) : 1.f,
                                                 seed, offset);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

// out planes [cols][pitch >= rows] = x^T ; 32x32 smem transpose
__global__ void split_planes_t_kernel(const float* __restrict__ x, int64_t ld, int rows, int cols, bf16* __restrict__ hi,
                                      bf16* __restrict__ lo, int64_t pitch) {
  __shared__ float t[32][33];
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    t[i][threadIdx.x] = (r < rows && c < cols) ? x[(int64_t)r * ld + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + threadIdx.x;      // output row = c, output col = r
    if (c < cols && r < pitch) {
      bf16 h, l;
      split_bf16(t[threadIdx.x][i], h, l);
      hi[(int64_t)c * pitch + r] = h;
      lo[(int64_t)c * pitch + r] = l;
    }
  }
}
int split_planes_transposed(const float* x, int64_t ld, int64_t rows, int32_t cols, bf16* hi, bf16* lo, int64_t pitch,
                            cudaStream_t s) {
  if (rows <= 0 || cols <= 0) return TT_OK;
  note_launch();
  split_planes_t_kernel<<<dim3((unsigned)ceil_div(cols, 32), (unsigned)ceil_div(pitch, 32)), dim3(32, 8), 0, s>>>(
      x, ld, (int)rows, cols, hi, lo, pitch);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

}  // namespace tt

// test hook: C = alpha * A B^T + bias through the tcgen05 engine (A [M,K], B [N,K] fp32, split on the fly);
// trans != 0 evaluates the same product with both operands given transposed ([K,M], [K,N]) -> MN-major path.
extern "C" int tt_gemm_x3(const float* A, const float* B, float* C, const float* bias, int32_t M, int32_t N, int32_t K,
                          float alpha, float beta, int32_t trans, int32_t splitk, void* workspace,
                          uint64_t workspace_bytes, void* stream) {
  using namespace tt;
  TT_CHECK_ARG(A && B && C && M > 0 && N > 0 && K > 0, "gemm_x3: bad args");
  cudaStream_t s = (cudaStream_t)stream;
  Workspace w(workspace, workspace_bytes);
  const int64_t ar = trans ? K : M, ac = trans ? M : K, br = trans ? K : N, bc = trans ? N : K;
  const int64_t pa = plane_pitch(ac), pb = plane_pitch(bc);
  bf16* ahi = w.take<bf16>(ar * pa); bf16* alo = w.take<bf16>(ar * pa);
  bf16* bhi = w.take<bf16>(br * pb); bf16* blo = w.take<bf16>(br * pb);
  if (!w.ok()) { set_error("gemm_x3: workspace too small (%llu needed)", (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
  TT_TRY(split_planes(A, ac, ar, (int)ac, ahi, alo, pa, 0.f, 0, 0, s));
  TT_TRY(split_planes(B, bc, br, (int)bc, bhi, blo, pb, 0.f, 0, 0, s));
  UmmaGemm g{};
  g.A[0] = {ahi, alo, pa}; g.B[0] = {bhi, blo, pb}; g.C[0] = C; g.bias[0] = bias; g.ldc = N;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.a_mn = g.b_mn = trans != 0; g.alpha = alpha; g.beta = beta; g.splitk = splitk;
  return umma_gemm(g, s);
}
// test hook for the v2 operand formats: flags bit0 = fp16 planes, bit1 = single pass, bit2 = A given as [K,M]
// (MN-major), bit3 = B given as [K,N] (MN-major).  C = alpha * A B^T (+ beta C).
__global__ void split_f16_kernel(const float* __restrict__ x, int64_t rows, int cols, __half* __restrict__ hi,
                                 __half* __restrict__ lo, int64_t pitch) {
  const int64_t total = rows * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % pitch);
    const float v = c < cols ? x[(i / pitch) * cols + c] : 0.f;
    __half a, b;
    tt::umma::split_f16(v, a, b);
    hi[i] = a; lo[i] = b;
  }
}
extern "C" int tt_gemm_test(const float* A, const float* B, float* C, int32_t M, int32_t N, int32_t K, float alpha,
                            float beta, int32_t flags, void* workspace, uint64_t workspace_bytes, void* stream) {
  using namespace tt;
  TT_CHECK_ARG(A && B && C && M > 0 && N > 0 && K > 0, "gemm_test: bad args");
  cudaStream_t s = (cudaStream_t)stream;
  const bool f16 = flags & 1, one = flags & 2, amn = flags & 4, bmn = flags & 8;
  Workspace w(workspace, workspace_bytes);
  const int64_t ar = amn ? K : M, ac = amn ? M : K, br = bmn ? K : N, bc = bmn ? N : K;
  const int64_t pa = plane_pitch(ac), pb = plane_pitch(bc);
  bf16* ahi = w.take<bf16>(ar * pa); bf16* alo = w.take<bf16>(ar * pa);
  bf16* bhi = w.take<bf16>(br * pb); bf16* blo = w.take<bf16>(br * pb);
  if (!w.ok()) { set_error("gemm_test: workspace too small"); return TT_ERR_WORKSPACE; }
  if (f16) {
    note_launch(); split_f16_kernel<<<592, 256, 0, s>>>(A, ar, (int)ac, (__half*)ahi, (__half*)alo, pa);
    note_launch(); split_f16_kernel<<<592, 256, 0, s>>>(B, br, (int)bc, (__half*)bhi, (__half*)blo, pb);
    TT_LAUNCH_CHECK();
  } else {
    TT_TRY(split_planes(A, ac, ar, (int)ac, ahi, alo, pa, 0.f, 0, 0, s));
    TT_TRY(split_planes(B, bc, br, (int)bc, bhi, blo, pb, 0.f, 0, 0, s));
  }
  UmmaGemm g{};
  g.A[0] = {ahi, alo, pa}; g.B[0] = {bhi, blo, pb}; g.C[0] = C; g.ldc = N;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.a_mn = amn; g.b_mn = bmn; g.alpha = alpha; g.beta = beta; g.splitk = 1;
  g.f16 = f16; g.npass = one ? 1 : 3;
  return umma_gemm(g, s);
}

extern "C" uint64_t tt_gemm_x3_workspace(int32_t M, int32_t N, int32_t K) {
  return 4ull * ((uint64_t)(M + 8) * (K + 8) + (uint64_t)(N + 8) * (K + 8)) + 4096;
}
