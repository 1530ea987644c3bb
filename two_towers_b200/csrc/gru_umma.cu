// GRU layer on the tcgen05 bf16x3 engine.
//   K-B   input projection  gi = x W_ih^T + b_ih          : umma_gemm (all timesteps, both directions batched)
//   K-C   per step          gh = h W_hh^T fused with the GRU cell in the epilogue (this file, gru_step kernel)
//   K-C'  per step          gate gradients (elementwise) + dh += dGH W_hh on umma_gemm; then the three big
//                           reductions over tokens (dW_ih, dW_hh, dX) on umma_gemm (MN-major / K-major)
// h is carried between steps as bf16 hi/lo planes (the MMA operand) plus an fp32 copy (the cell's h_prev).
#include "umma.cuh"
#include "umma_gemm.cuh"

namespace tt {
using namespace umma;

constexpr int GBM = 128, GBK = 64, GUN = 64 /*hidden units per tile*/, GBN = 3 * GUN;
constexpr int G_A_BYTES = GBM * GBK * 2, G_B_BYTES = GBN * GBK * 2;
constexpr int G_STAGE = 2 * G_A_BYTES + 2 * G_B_BYTES;          // 81920
constexpr int G_STAGES = 2;
constexpr int G_SMEM = G_STAGES * G_STAGE + 1024 + 128;
constexpr int G_TMEM_COLS = 256;

struct alignas(64) GruStepParams {
  CUtensorMap h_hi[2], h_lo[2];      // per direction: current h planes [B][Hp]  (box 64 x 128)
  CUtensorMap w_hi[2], w_lo[2];      // per direction: W_hh planes [3H][Hp]      (box 64 x 64)
  const float* gi;                   // [B*T, ndir*3H]  (b_ih folded in)
  const float* b_hh[2];
  float* hs;                         // [ndir][B][H] fp32 running state (in place)
  float* out;                        // [B,T,ndir*H]
  float* gates;                      // [ndir][B][T][4H] or null
  bf16* hn_hi[2];                    // next-step h planes [B][Hp]
  bf16* hn_lo[2];
  const int32_t* lengths;
  int32_t B, T, H, ndir, step;
  int64_t hpitch;
};

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// grid: (ceil(H/64), ceil(B/128), ndir).  One 128 x (3 x 64) tile per CTA: columns [0,64) = r, [64,128) = z,
// [128,192) = n pre-activations (h-part) of hidden units [u0, u0+64).
__global__ void __launch_bounds__(192, 1) gru_step_kernel(const __grid_constant__ GruStepParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G_STAGES * G_STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + G_STAGES;
  uint64_t* tfull = bars + 2 * G_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.z, m0 = blockIdx.y * GBM, u0 = blockIdx.x * GUN;
  const int H = p.H, KB = (H + GBK - 1) / GBK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < G_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, G_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* st = smem + stage * G_STAGE;
        uint8_t* sa_hi = st, *sa_lo = st + G_A_BYTES, *sb_hi = st + 2 * G_A_BYTES, *sb_lo = sb_hi + G_B_BYTES;
        mbar_arrive_expect_tx(&full[stage], G_STAGE);
        tma_load_2d(&p.h_hi[dir], &full[stage], sa_hi, kb * GBK, m0);
        tma_load_2d(&p.h_lo[dir], &full[stage], sa_lo, kb * GBK, m0);
#pragma unroll
        for (int g = 0; g < 3; ++g) {       // gate g rows [g*H + u0, +64) of W_hh -> tile columns [64 g, 64 g + 64)
          tma_load_2d(&p.w_hi[dir], &full[stage], sb_hi + g * (GUN * 128), kb * GBK, g * H + u0);
          tma_load_2d(&p.w_lo[dir], &full[stage], sb_lo + g * (GUN * 128), kb * GBK, g * H + u0);
        }
        if (++stage == G_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(GBM, GBN, 0, 0);
      int stage = 0, phase = 0;
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t st = smem_u32(smem + stage * G_STAGE);
        const uint32_t sa_hi = st, sa_lo = st + G_A_BYTES, sb_hi = st + 2 * G_A_BYTES, sb_lo = sb_hi + G_B_BYTES;
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t sa = term == 1 ? sa_lo : sa_hi, sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
          for (int k = 0; k < GBK / 16; ++k)
            mma_bf16_ss(tmem_base, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                        (kb > 0 || term > 0 || k > 0) ? 1u : 0u);
        }
        mma_commit(&empty[stage]);
        if (++stage == G_STAGES) { stage = 0; phase ^= 1; }
      }
      mma_commit(tfull);
    }
  } else {
    // ------------------------------------------------------------------ fused GRU cell (one thread = one batch row)
    const int quarter = warp & 3;
    const int b = m0 + quarter * 32 + lane;
    const int T = p.T, nd = p.ndir;
    const int t = dir == 0 ? p.step : T - 1 - p.step;
    const bool row_ok = b < p.B;
    const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
    const float* gi = p.gi + ((int64_t)(row_ok ? b : 0) * T + t) * (nd * 3 * H) + dir * 3 * H;
    const float* bh = p.b_hh[dir];
    float* hs = p.hs + ((int64_t)dir * p.B + (row_ok ? b : 0)) * H;
    float* out = p.out + ((int64_t)(row_ok ? b : 0) * T + t) * (nd * H) + dir * H;
    float* gs = p.gates ? p.gates + (((int64_t)dir * p.B + (row_ok ? b : 0)) * T + t) * (4 * H) : nullptr;
    bf16* nhi = p.hn_hi[dir] + (int64_t)(row_ok ? b : 0) * p.hpitch;
    bf16* nlo = p.hn_lo[dir] + (int64_t)(row_ok ? b : 0) * p.hpitch;
    mbar_wait(tfull, 0);
    tc_fence_after();
    const uint32_t tl = tmem_base + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
    for (int c0 = 0; c0 < GUN; c0 += 16) {
      float ar[16], az[16], an[16];
      tmem_ld16(tl + c0, ar);
      tmem_ld16(tl + GUN + c0, az);
      tmem_ld16(tl + 2 * GUN + c0, an);
      tmem_ld_wait();
      if (!row_ok) continue;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int u = u0 + c0 + j;
        if (u >= H) break;
        float h;
        if (live) {
          const float r = sigmoidf_(gi[u] + ar[j] + bh[u]);
          const float z = sigmoidf_(gi[H + u] + az[j] + bh[H + u]);
          const float hn = an[j] + bh[2 * H + u];
          const float n = tanhf(gi[2 * H + u] + r * hn);
          const float hp = hs[u];
          h = (1.f - z) * n + z * hp;
          hs[u] = h;
          out[u] = h;
          if (gs) { gs[u] = r; gs[H + u] = z; gs[2 * H + u] = n; gs[3 * H + u] = hn; }
        } else {
          h = hs[u];                 // padded step: state carries, output is zero
          out[u] = 0.f;
          if (gs) { gs[u] = 0.f; gs[H + u] = 1.f; gs[2 * H + u] = 0.f; gs[3 * H + u] = 0.f; }
        }
        bf16 a, c;
        split_bf16(h, a, c);
        nhi[u] = a; nlo[u] = c;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, G_TMEM_COLS);
}


// ------------------------------------------------------------------------------------------ helper kernels
// token ids -> bf16 hi/lo planes of the (optionally dropped-out) embedding rows; one warp per token
template <typename IdT>
__global__ void gather_split_kernel(const IdT* __restrict__ ids, int64_t n_tokens, const float* __restrict__ table,
                                    int64_t vocab, int E, int64_t pad_id, bf16* __restrict__ hi, bf16* __restrict__ lo,
                                    int64_t pitch, float p, float inv_keep, uint64_t seed, uint64_t offset) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t tok = warp; tok < n_tokens; tok += nwarps) {
    const int64_t id = (int64_t)ids[tok];
    const bool zero = id < 0 || id >= vocab || id == pad_id;
    const float* src = table + (zero ? 0 : id) * E;
    for (int c = lane * 2; c < pitch; c += 64) {
      float v0 = (!zero && c < E) ? __ldg(src + c) : 0.f;
      float v1 = (!zero && c + 1 < E) ? __ldg(src + c + 1) : 0.f;
      if (p > 0.f) {
        v0 *= dropout_scale(seed, offset, (uint64_t)tok * E + c, p, inv_keep);
        v1 *= dropout_scale(seed, offset, (uint64_t)tok * E + c + 1, p, inv_keep);
      }
      bf16 h0, l0, h1, l1;
      split_bf16(v0, h0, l0);
      split_bf16(v1, h1, l1);
      *reinterpret_cast<__nv_bfloat162*>(hi + tok * pitch + c) = __halves2bfloat162(h0, h1);
      *reinterpret_cast<__nv_bfloat162*>(lo + tok * pitch + c) = __halves2bfloat162(l0, l1);
    }
  }
}

// h_{prev}[dir][b,t] = out[b, t-1 (fwd) | t+1 (rev), dir*H : dir*H+H] (zero at the sequence start) -> planes
__global__ void hprev_planes_kernel(const float* __restrict__ out, bf16* __restrict__ hi, bf16* __restrict__ lo, int B, int T,
                                    int H, int nd, int64_t pitch) {
  const int64_t total = (int64_t)nd * B * T * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int u = (int)(i % pitch);
    const int64_t row = i / pitch;              // dir*B*T + b*T + t
    const int t = (int)(row % T);
    const int b = (int)((row / T) % B);
    const int dir = (int)(row / ((int64_t)T * B));
    const int tp = dir == 0 ? t - 1 : t + 1;
    const float v = (u < H && tp >= 0 && tp < T) ? out[((int64_t)b * T + tp) * (nd * H) + dir * H + u] : 0.f;
    bf16 a, c;
    split_bf16(v, a, c);
    hi[i] = a; lo[i] = c;
  }
}

// gate gradients of one step, written as planes dG[dir][b*T+t][4H] = (dr_pre, dz_pre, dn_pre*r, dn_pre)
__global__ void gru_cell_bwd_planes_kernel(float* __restrict__ dhs, const float* __restrict__ d_out,
                                           const float* __restrict__ gates, const float* __restrict__ out,
                                           bf16* __restrict__ dg_hi, bf16* __restrict__ dg_lo,
                                           const int32_t* __restrict__ lengths, int B, int T, int H, int ndir, int step) {
  const int64_t total = (int64_t)ndir * B * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % H);
    const int b = (int)((i / H) % B);
    const int dir = (int)(i / ((int64_t)H * B));
    const int t = dir == 0 ? step : T - 1 - step;
    const int64_t o = (((int64_t)dir * B + b) * T + t) * (4 * H);
    float g4[4] = {0.f, 0.f, 0.f, 0.f};
    if (!(lengths && t >= lengths[b])) {
      float dh = dhs[i];
      if (d_out) dh += d_out[((int64_t)b * T + t) * (ndir * H) + dir * H + j];
      const float* g = gates + o;
      const float r = g[j], z = g[H + j], n = g[2 * H + j], hn = g[3 * H + j];
      const int tp = dir == 0 ? t - 1 : t + 1;
      const float hp = (tp >= 0 && tp < T) ? out[((int64_t)b * T + tp) * (ndir * H) + dir * H + j] : 0.f;
      const float dn_pre = dh * (1.f - z) * (1.f - n * n);
      g4[0] = dn_pre * hn * r * (1.f - r);
      g4[1] = dh * (hp - n) * z * (1.f - z);
      g4[2] = dn_pre * r;
      g4[3] = dn_pre;
      dhs[i] = dh * z;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      bf16 a, c;
      split_bf16(g4[q], a, c);
      dg_hi[o + q * H + j] = a; dg_lo[o + q * H + j] = c;
    }
  }
}

// column sums of a plane pair: out[n] = sum_rows (hi + lo)
__global__ void colsum_planes_kernel(const bf16* __restrict__ hi, const bf16* __restrict__ lo, int64_t pitch, int64_t rows,
                                     int N, float* out) {
  __shared__ float red[8][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  float acc = 0.f;
  if (n < N)
    for (int64_t m = (int64_t)blockIdx.y * 8 + threadIdx.y; m < rows; m += (int64_t)gridDim.y * 8)
      acc += __bfloat162float(hi[m * pitch + n]) + __bfloat162float(lo[m * pitch + n]);
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) v += red[i][threadIdx.x];
    atomicAdd(out + n, v);
  }
}

static inline unsigned eblocks(int64_t n, int cap = 148 * 16) {
  int64_t b = ceil_div(n, 256);
  return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int make_map2(CUtensorMap* m, const bf16* base, int64_t rows, int64_t cols, int64_t pitch, int box_rows) {
  static EncodeTiledFn2 fn = []() -> EncodeTiledFn2 {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &q, cudaEnableDefault, &r) != cudaSuccess) return nullptr;
    return (EncodeTiledFn2)q;
  }();
  if (!fn) { set_error("cuTensorMapEncodeTiled not available"); return TT_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch * sizeof(bf16)};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("gru_umma: cuTensorMapEncodeTiled failed (%d)", (int)r); return TT_ERR_CUDA; }
  return TT_OK;
}

struct PlaneBuf { bf16* hi; bf16* lo; int64_t pitch; Planes view() const { return Planes{hi, lo, pitch}; } };
static PlaneBuf take_planes(Workspace& w, int64_t rows, int64_t cols) {
  PlaneBuf b;
  b.pitch = plane_pitch(cols);
  b.hi = w.take<bf16>(rows * b.pitch);
  b.lo = w.take<bf16>(rows * b.pitch);
  return b;
}

// effective layer input (x, dropped-out x, or gathered table rows) as planes [ntok][Inp]
static int make_x_planes(const tt_gru_layer* a, const PlaneBuf& xp, cudaStream_t s) {
  const int64_t ntok = (int64_t)a->B * a->T;
  if (a->ids) {
    int64_t blocks = ceil_div(ntok, 8);
    if (blocks > 148 * 16) blocks = 148 * 16;
    const float p = a->dropout_p, ik = p > 0.f ? 1.f / (1.f - p) : 1.f;
    note_launch();
    if (a->ids_are_i64)
      gather_split_kernel<int64_t><<<(unsigned)blocks, 256, 0, s>>>((const int64_t*)a->ids, ntok, a->table, a->vocab, a->In,
                                                                 a->pad_id, xp.hi, xp.lo, xp.pitch, p, ik, a->seed, a->offset);
    else
      gather_split_kernel<int32_t><<<(unsigned)blocks, 256, 0, s>>>((const int32_t*)a->ids, ntok, a->table, a->vocab, a->In,
                                                                 a->pad_id, xp.hi, xp.lo, xp.pitch, p, ik, a->seed, a->offset);
    TT_LAUNCH_CHECK();
    return TT_OK;
  }
  return split_planes(a->x, a->In, ntok, a->In, xp.hi, xp.lo, xp.pitch, a->dropout_p, a->seed, a->offset, s);
}

// ------------------------------------------------------------------------------------------ forward
uint64_t gru_fwd_workspace_umma(const tt_gru_layer* a) {
  Workspace w(nullptr, 0);
  const int64_t ntok = (int64_t)a->B * a->T;
  take_planes(w, ntok, a->In);
  for (int d = 0; d < a->ndir; ++d) { take_planes(w, 3 * a->H, a->In); take_planes(w, 3 * a->H, a->H); }
  for (int i = 0; i < 2 * a->ndir; ++i) take_planes(w, a->B, a->H);
  w.take<float>(ntok * a->ndir * 3 * a->H);
  return w.used + 1024;
}

int gru_layer_fwd_umma(const tt_gru_layer* a, cudaStream_t s) {
  const int B = a->B, T = a->T, In = a->In, H = a->H, nd = a->ndir;
  const int64_t ntok = (int64_t)B * T;
  Workspace w(a->workspace, a->workspace_bytes);
  PlaneBuf xp = take_planes(w, ntok, In);
  PlaneBuf wih[2], whh[2], hbuf[2][2];
  for (int d = 0; d < nd; ++d) { wih[d] = take_planes(w, 3 * H, In); whh[d] = take_planes(w, 3 * H, H); }
  for (int i = 0; i < 2; ++i) for (int d = 0; d < nd; ++d) hbuf[i][d] = take_planes(w, B, H);
  float* gi = w.take<float>(ntok * nd * 3 * H);
  if (!w.ok()) { set_error("gru fwd (umma): workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }

  TT_TRY(make_x_planes(a, xp, s));
  for (int d = 0; d < nd; ++d) {
    TT_TRY(split_planes(a->w_ih[d], In, 3 * H, In, wih[d].hi, wih[d].lo, wih[d].pitch, 0.f, 0, 0, s));
    TT_TRY(split_planes(a->w_hh[d], H, 3 * H, H, whh[d].hi, whh[d].lo, whh[d].pitch, 0.f, 0, 0, s));
  }
  // K-B: gi[:, d*3H:(d+1)*3H] = x W_ih[d]^T + b_ih[d]
  UmmaGemm g{};
  for (int d = 0; d < nd; ++d) {
    g.A[d] = xp.view(); g.B[d] = wih[d].view(); g.C[d] = gi + (int64_t)d * 3 * H; g.bias[d] = a->b_ih[d];
  }
  g.ldc = (int64_t)nd * 3 * H; g.M = (int)ntok; g.N = 3 * H; g.K = In; g.nbatch = nd; g.alpha = 1.f; g.beta = 0.f; g.splitk = 1;
  TT_TRY(umma_gemm(g, s));

  // zero initial state: fp32 copy and both plane buffers (pad columns stay zero forever)
  TT_CUDA(cudaMemsetAsync(a->h_n, 0, sizeof(float) * nd * B * H, s));
  for (int i = 0; i < 2; ++i)
    for (int d = 0; d < nd; ++d) {
      TT_CUDA(cudaMemsetAsync(hbuf[i][d].hi, 0, sizeof(bf16) * B * hbuf[i][d].pitch, s));
      TT_CUDA(cudaMemsetAsync(hbuf[i][d].lo, 0, sizeof(bf16) * B * hbuf[i][d].pitch, s));
    }
  static DeviceOnce attr_once;
  if (attr_once.need()) {
    TT_CUDA(cudaFuncSetAttribute(gru_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    attr_once.done();
  }
  GruStepParams prm[2];
  for (int i = 0; i < 2; ++i) {
    GruStepParams& p = prm[i];
    memset(&p, 0, sizeof(p));
    for (int d = 0; d < nd; ++d) {
      TT_TRY(make_map2(&p.h_hi[d], hbuf[i][d].hi, B, H, hbuf[i][d].pitch, GBM));
      TT_TRY(make_map2(&p.h_lo[d], hbuf[i][d].lo, B, H, hbuf[i][d].pitch, GBM));
      TT_TRY(make_map2(&p.w_hi[d], whh[d].hi, 3 * H, H, whh[d].pitch, GUN));
      TT_TRY(make_map2(&p.w_lo[d], whh[d].lo, 3 * H, H, whh[d].pitch, GUN));
      p.b_hh[d] = a->b_hh[d];
      p.hn_hi[d] = hbuf[i ^ 1][d].hi; p.hn_lo[d] = hbuf[i ^ 1][d].lo;
    }
    p.gi = gi; p.hs = a->h_n; p.out = a->out; p.gates = a->save_for_backward ? a->gates : nullptr;
    p.lengths = a->lengths; p.B = B; p.T = T; p.H = H; p.ndir = nd; p.hpitch = hbuf[0][0].pitch;
  }
  dim3 grid((unsigned)ceil_div(H, GUN), (unsigned)ceil_div(B, GBM), (unsigned)nd);
  for (int step = 0; step < T; ++step) {
    GruStepParams& p = prm[step & 1];
    p.step = step;
    note_launch();
    gru_step_kernel<<<grid, 192, G_SMEM, s>>>(p);
  }
  TT_LAUNCH_CHECK();
  return TT_OK;
}


// ------------------------------------------------------------------------------------------ backward
uint64_t gru_bwd_workspace_umma(const tt_gru_layer_grad* a) {
  const tt_gru_layer* f = &a->fwd;
  Workspace w(nullptr, 0);
  const int64_t ntok = (int64_t)f->B * f->T;
  take_planes(w, ntok, f->In);
  take_planes(w, (int64_t)f->ndir * ntok, 4 * f->H);
  take_planes(w, (int64_t)f->ndir * ntok, f->H);
  for (int d = 0; d < f->ndir; ++d) { take_planes(w, f->H, 3 * f->H); take_planes(w, f->In, 3 * f->H); }
  w.take<float>((int64_t)f->ndir * f->B * f->H);
  w.take<float>((int64_t)f->ndir * 4 * f->H);
  return w.used + 1024;
}

int gru_layer_bwd_umma(const tt_gru_layer_grad* a, cudaStream_t s) {
  const tt_gru_layer* f = &a->fwd;
  const int B = f->B, T = f->T, In = f->In, H = f->H, nd = f->ndir;
  const int64_t ntok = (int64_t)B * T;
  Workspace w(a->workspace, a->workspace_bytes);
  PlaneBuf xp = take_planes(w, ntok, In);
  PlaneBuf dg = take_planes(w, (int64_t)nd * ntok, 4 * H);        // pitch == 4H (H % 8 == 0)
  PlaneBuf hp = take_planes(w, (int64_t)nd * ntok, H);
  PlaneBuf whhT[2], wihT[2];
  for (int d = 0; d < nd; ++d) { whhT[d] = take_planes(w, H, 3 * H); wihT[d] = take_planes(w, In, 3 * H); }
  float* dhs = w.take<float>((int64_t)nd * B * H);
  float* csum = w.take<float>((int64_t)nd * 4 * H);
  if (!w.ok()) { set_error("gru bwd (umma): workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
  const int64_t G4 = dg.pitch;   // 4H

  TT_TRY(make_x_planes(f, xp, s));
  for (int d = 0; d < nd; ++d) {
    TT_TRY(split_planes_transposed(f->w_hh[d], H, 3 * H, H, whhT[d].hi, whhT[d].lo, whhT[d].pitch, s));
    if (a->d_x) TT_TRY(split_planes_transposed(f->w_ih[d], In, 3 * H, In, wihT[d].hi, wihT[d].lo, wihT[d].pitch, s));
  }
  note_launch();
  hprev_planes_kernel<<<eblocks((int64_t)nd * ntok * hp.pitch, 148 * 32), 256, 0, s>>>(f->out, hp.hi, hp.lo, B, T, H, nd, hp.pitch);
  if (a->d_h_n) TT_CUDA(cudaMemcpyAsync(dhs, a->d_h_n, sizeof(float) * nd * B * H, cudaMemcpyDeviceToDevice, s));
  else TT_CUDA(cudaMemsetAsync(dhs, 0, sizeof(float) * nd * B * H, s));

  // per step (reverse): gate gradients -> planes, then dhs += dGH_t W_hh   (A rows = batch, row pitch T*4H)
  UmmaGemm r{};
  r.ldc = H; r.M = B; r.N = H; r.K = 3 * H; r.nbatch = nd; r.alpha = 1.f; r.beta = 1.f; r.splitk = 1;
  for (int d = 0; d < nd; ++d) { r.B[d] = whhT[d].view(); r.C[d] = dhs + (int64_t)d * B * H; r.bias[d] = nullptr; }
  const unsigned cb = eblocks((int64_t)nd * B * H);
  for (int step = T - 1; step >= 0; --step) {
    note_launch();
    gru_cell_bwd_planes_kernel<<<cb, 256, 0, s>>>(dhs, a->d_out, f->gates, f->out, dg.hi, dg.lo, f->lengths, B, T, H, nd, step);
    if (step == 0) break;
    for (int d = 0; d < nd; ++d) {
      const int t = d == 0 ? step : T - 1 - step;
      const int64_t off = ((int64_t)d * ntok + t) * G4;
      r.A[d] = Planes{dg.hi + off, dg.lo + off, (int64_t)T * G4};
    }
    TT_TRY(umma_gemm(r, s));
  }
  TT_LAUNCH_CHECK();

  for (int d = 0; d < nd; ++d) {
    const bf16* dhi = dg.hi + (int64_t)d * ntok * G4;
    const bf16* dlo = dg.lo + (int64_t)d * ntok * G4;
    float* cs = csum + (int64_t)d * 4 * H;
    TT_CUDA(cudaMemsetAsync(cs, 0, sizeof(float) * 4 * H, s));
    int64_t gy = ceil_div(ntok, 64);
    if (gy > 128) gy = 128;
    note_launch();
    colsum_planes_kernel<<<dim3((unsigned)ceil_div(4 * H, 32), (unsigned)gy), dim3(32, 8), 0, s>>>(dhi, dlo, G4, ntok, 4 * H, cs);
    TT_CUDA(cudaMemcpyAsync(a->d_b_hh[d], cs, sizeof(float) * 3 * H, cudaMemcpyDeviceToDevice, s));
    TT_CUDA(cudaMemcpyAsync(a->d_b_ih[d], cs, sizeof(float) * 2 * H, cudaMemcpyDeviceToDevice, s));
    TT_CUDA(cudaMemcpyAsync(a->d_b_ih[d] + 2 * H, cs + 3 * H, sizeof(float) * H, cudaMemcpyDeviceToDevice, s));

    // dW_hh[3H,H] = dGH^T h_prev  (reduction over all tokens; both operands MN-major)
    UmmaGemm g{};
    g.A[0] = Planes{dhi, dlo, G4};
    g.B[0] = Planes{hp.hi + (int64_t)d * ntok * hp.pitch, hp.lo + (int64_t)d * ntok * hp.pitch, hp.pitch};
    g.C[0] = a->d_w_hh[d]; g.ldc = H; g.M = 3 * H; g.N = H; g.K = (int)ntok; g.nbatch = 1; g.a_mn = g.b_mn = true;
    g.alpha = 1.f; g.beta = 0.f; g.splitk = 0;
    TT_TRY(umma_gemm(g, s));
    // dW_ih rows [0,2H) <- dG cols [0,2H); rows [2H,3H) <- dG cols [3H,4H)
    for (int piece = 0; piece < 2; ++piece) {
      UmmaGemm q{};
      const int64_t co = piece ? 3 * H : 0;
      q.A[0] = Planes{dhi + co, dlo + co, G4};
      q.B[0] = xp.view();
      q.C[0] = a->d_w_ih[d] + (int64_t)(piece ? 2 * H : 0) * In; q.ldc = In;
      q.M = piece ? H : 2 * H; q.N = In; q.K = (int)ntok; q.nbatch = 1; q.a_mn = q.b_mn = true;
      q.alpha = 1.f; q.beta = 0.f; q.splitk = 0;
      TT_TRY(umma_gemm(q, s));
    }
    // dX (+)= dGI W_ih : K segments [0,2H) and [3H,4H) of dG against columns [0,2H) / [2H,3H) of W_ih^T
    if (a->d_x) {
      for (int piece = 0; piece < 2; ++piece) {
        UmmaGemm q{};
        const int64_t ca = piece ? 3 * H : 0, cbk = piece ? 2 * H : 0;
        q.A[0] = Planes{dhi + ca, dlo + ca, G4};
        q.B[0] = Planes{wihT[d].hi + cbk, wihT[d].lo + cbk, wihT[d].pitch};
        q.C[0] = a->d_x; q.ldc = In; q.M = (int)ntok; q.N = In; q.K = piece ? H : 2 * H; q.nbatch = 1;
        q.alpha = 1.f; q.beta = (d == 0 && piece == 0) ? 0.f : 1.f; q.splitk = 1;
        TT_TRY(umma_gemm(q, s));
      }
    }
  }
  return TT_OK;
}

}  // namespace tt
