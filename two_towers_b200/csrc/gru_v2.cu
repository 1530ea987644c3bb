// GRU stack, "v2" pipeline (engine 2): the whole multi-layer bidirectional GRU behind tt_gru_stack_fwd/bwd with
// library-private activation layouts chosen for the hardware:
//
//   * tokens are TIME-MAJOR inside the library: row m = t*Bp + b (Bp = B rounded up to 128), so a 128-row GEMM /
//     TMEM tile is one batch tile of ONE timestep and the recurrence reads/writes contiguous slabs per step;
//   * GEMM operands are fp16 hi/lo plane pairs.  Forward products use 3 tensor passes (hi*hi + lo*hi + hi*lo, ~22
//     mantissa bits); backward GEMMs (dh, dX, dW) use ONE fp16 pass on loss-scaled gradients;
//   * everything a one-thread-per-row TMEM epilogue touches per element (gi, running h, saved gates, dh, dX) is
//     stored TILED as [..][unit/4][128 rows][4] so that the 32 lanes of a warp (32 consecutive rows) move 512
//     contiguous bytes per 16-byte access;
//   * 16-bit operand planes produced by an epilogue go through a swizzled smem tile and one TMA store.
//
//   K-A  gather_planes / x_planes      ids or x -> layer-0 operand planes
//   K-B  umma_gemm (3 passes, tiled)   gi = X W_ih^T + b_ih (+ b_hr, b_hz folded)          all timesteps, both dirs
//   K-C  gru_step2_kernel              per step: h W_hh^T on tcgen05, GRU cell fused in the epilogue
//   K-C' cell_bwd2_kernel + umma_gemm  per step: gate gradients -> fp16 planes; dh += dGH W_hh (1 pass)
//        umma_gemm (1 pass)            dW_hh, dW_ih (MN-major, split-K), dX (tiled, feeds the layer below)
#include "umma.cuh"
#include "umma_gemm.cuh"
#include "gru_v2.cuh"
#include <mutex>

namespace tt {
using namespace umma;
typedef __half f16;

// ------------------------------------------------------------------------------------------ small device helpers
// MUFU-only transcendental paths (ex2.approx.ftz / rcp.approx.ftz: 2 ulp / 1 ulp, no denormal fix-up code):
// sigmoid(x) = 1 / (1 + 2^(-x log2 e)),  tanh(x) = 1 - 2 / (1 + 2^(2 x log2 e));  saturate correctly at +-inf.
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_ftz(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_sigmoid(float x) { return rcp_ftz(1.f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float fast_tanh(float x) { return fmaf(-2.f, rcp_ftz(1.f + ex2_ftz(2.8853900817779268f * x)), 1.f); }

// keep-mask scales of the 8 consecutive logical elements 8*e8 .. 8*e8+7: one Philox4x32-10 call, 16 random bits per
// element (keep iff bits >= round(p * 65536)); forward (layer output) and backward (d_above) use the same function.
__device__ __forceinline__ void dropout_scale8(uint64_t seed, uint64_t e8, float p, float inv_keep, float (&s)[8]) {
  uint32_t r[4];
  philox4x32(seed, e8, r);
  const uint32_t thr = (uint32_t)__float2uint_rn(p * 65536.f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    s[2 * i] = (r[i] & 0xFFFFu) >= thr ? inv_keep : 0.f;
    s[2 * i + 1] = (r[i] >> 16) >= thr ? inv_keep : 0.f;
  }
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// 8 floats -> 8 fp16 hi (uint4) and 8 fp16 lo (uint4)
__device__ __forceinline__ void split8(const float (&x)[8], uint4& hi, uint4& lo) {
  float l[8];
  uint32_t h[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half a = __float2half_rn(x[2 * i]), b = __float2half_rn(x[2 * i + 1]);
    l[2 * i] = x[2 * i] - __half2float(a);
    l[2 * i + 1] = x[2 * i + 1] - __half2float(b);
    __half2 hh = __halves2half2(a, b);
    h[i] = *reinterpret_cast<uint32_t*>(&hh);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(pack_h2(l[0], l[1]), pack_h2(l[2], l[3]), pack_h2(l[4], l[5]), pack_h2(l[6], l[7]));
}

// ---- saved gate records (what the BPTT needs of a forward step), five 16-bit planes of 8 units per 16-byte word:
//   Z = z, R = r                       u16, x 65535
//   A = (1 - z)(1 - n^2)               u16, x 65535          dn_pre       = dh * A
//   C = (h_t - n)(1 - z)               s16, x 65000 (|C| <= 1/2: h_t - n = z (h_prev - n))    dz_pre = dh * C
//   B = (h W_hn + b_hn) r (1 - r)      fp16                   dr_pre       = dn_pre * B
// Everything that does not depend on the running gradient is formed ONCE, in the forward epilogue (which has the
// operands in registers), instead of per BPTT step from r, z, n, q and an fp32 hn plane: the BPTT kernel runs on one CTA
// per (row tile, direction) and is bound by the instructions of its gate-gradient loop (40 -> ~25 per unit), and the
// saved state shrinks from 12 H to 10 H bytes per (token, direction).  Padded steps / pad rows store z = 1 and zeros:
// the gradient passes through and every gate gradient is zero without any masking in the BPTT.
// Word index of (dir, t, row tile mt, unit group u8 = unit / 8, row): ((((dir T + t) MT + mt) H/8 + u8) 128 + row); plane
// k starts k * g16_plane words further on.
enum { GP_Z = 0, GP_R = 1, GP_A = 2, GP_C = 3, GP_B = 4, GP_PLANES = 5 };
constexpr float kCScale = 65000.f;
// 16-bit fixed-point codes through the mantissa of 2^23 (magic-number FMA, PRMT to unpack) on the FMA / ALU pipes: F2I / I2F
// issue at 16 per clock and SM on the XU pipe, which also carries the ex2 / rcp of the gate activations.
// quantisers without clamps (the value ranges are guaranteed by construction); the result's LOW 16 bits are the code
__device__ __forceinline__ uint32_t qf_u16(float x) { return __float_as_uint(fmaf(x, 65535.f, 8388608.f)); }
__device__ __forceinline__ uint32_t qf_s16(float x, float scale) { return __float_as_uint(fmaf(x, scale, 8388608.f + 32768.f)); }
__device__ __forceinline__ uint32_t pk_lo16(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x5410); }   // a.lo16 | b.lo16 << 16
// code in the low (HALF = 0) / high (HALF = 1) 16 bits of w -> float, one PRMT + one FFMA
template <int HALF> __device__ __forceinline__ float dqf_u16(uint32_t w) {
  constexpr float c = 1.f / 65535.f;
  return fmaf(__uint_as_float(__byte_perm(w, 0x4B000000u, HALF ? 0x7632 : 0x7610)), c, -8388608.f * c);
}
template <int HALF> __device__ __forceinline__ float dqf_s16(uint32_t w, float inv_scale) {
  return fmaf(__uint_as_float(__byte_perm(w, 0x4B000000u, HALF ? 0x7632 : 0x7610)), inv_scale, -(8388608.f + 32768.f) * inv_scale);
}
// forward: records of 8 consecutive units of one row.  dst = plane-Z word of (.., u8, row).
__device__ __forceinline__ void store_rec8(uint4* __restrict__ dst, int64_t plane, const float (&r)[8], const float (&z)[8],
                                           const float (&n)[8], const float (&q)[8], const float (&hn)[8]) {
  uint32_t cz[8], cr[8], ca[8], cc[8];
  float bv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float omz = 1.f - z[j];
    cz[j] = qf_u16(z[j]);
    cr[j] = qf_u16(r[j]);
    ca[j] = qf_u16(omz * fmaf(-n[j], n[j], 1.f));
    cc[j] = qf_s16(q[j] * omz, kCScale);
    const float t = fmaf(-r[j], r[j], r[j]);                       // r (1 - r)
    bv[j] = hn[j] * t;
  }
  dst[GP_Z * plane] = make_uint4(pk_lo16(cz[0], cz[1]), pk_lo16(cz[2], cz[3]), pk_lo16(cz[4], cz[5]), pk_lo16(cz[6], cz[7]));
  dst[GP_R * plane] = make_uint4(pk_lo16(cr[0], cr[1]), pk_lo16(cr[2], cr[3]), pk_lo16(cr[4], cr[5]), pk_lo16(cr[6], cr[7]));
  dst[GP_A * plane] = make_uint4(pk_lo16(ca[0], ca[1]), pk_lo16(ca[2], ca[3]), pk_lo16(ca[4], ca[5]), pk_lo16(ca[6], ca[7]));
  dst[GP_C * plane] = make_uint4(pk_lo16(cc[0], cc[1]), pk_lo16(cc[2], cc[3]), pk_lo16(cc[4], cc[5]), pk_lo16(cc[6], cc[7]));
  dst[GP_B * plane] = make_uint4(pack_h2(bv[0], bv[1]), pack_h2(bv[2], bv[3]), pack_h2(bv[4], bv[5]), pack_h2(bv[6], bv[7]));
}
// backward: the four planes the gate gradients of 8 units need (Z is consumed separately, by the dh * z pass)
struct GateRec8 { uint4 r, a, c, b; };
__device__ __forceinline__ void load_rec8(GateRec8& g, const uint4* __restrict__ src, int64_t plane) {
  g.r = __ldg(src + GP_R * plane);
  g.a = __ldg(src + GP_A * plane);
  g.c = __ldg(src + GP_C * plane);
  g.b = __ldg(src + GP_B * plane);
}
// gate gradients of 8 units from dL/dh: packed fp16 (dr | dz | dn*r | dn), 16 bytes each
__device__ __forceinline__ void grad_rec8(const GateRec8& g, const float (&dh)[8], uint4& o_dr, uint4& o_dz, uint4& o_dnr, uint4& o_dn) {
  const uint32_t rw[4] = {g.r.x, g.r.y, g.r.z, g.r.w}, aw[4] = {g.a.x, g.a.y, g.a.z, g.a.w};
  const uint32_t cw[4] = {g.c.x, g.c.y, g.c.z, g.c.w}, bw[4] = {g.b.x, g.b.y, g.b.z, g.b.w};
  uint32_t pr[4], pz[4], pnr[4], pn[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 bb = __half22float2(*reinterpret_cast<const __half2*>(&bw[i]));
    const float dn0 = dh[2 * i] * dqf_u16<0>(aw[i]), dn1 = dh[2 * i + 1] * dqf_u16<1>(aw[i]);
    pn[i] = pack_h2(dn0, dn1);
    pnr[i] = pack_h2(dn0 * dqf_u16<0>(rw[i]), dn1 * dqf_u16<1>(rw[i]));
    pr[i] = pack_h2(dn0 * bb.x, dn1 * bb.y);
    pz[i] = pack_h2(dh[2 * i] * dqf_s16<0>(cw[i], 1.f / kCScale), dh[2 * i + 1] * dqf_s16<1>(cw[i], 1.f / kCScale));
  }
  o_dr = make_uint4(pr[0], pr[1], pr[2], pr[3]);
  o_dz = make_uint4(pz[0], pz[1], pz[2], pz[3]);
  o_dnr = make_uint4(pnr[0], pnr[1], pnr[2], pnr[3]);
  o_dn = make_uint4(pn[0], pn[1], pn[2], pn[3]);
}
// dh[0..7] *= z of 8 units (one Z word)
__device__ __forceinline__ void mul_z8(float* dh, const uint4& zw) {
  const uint32_t w[4] = {zw.x, zw.y, zw.z, zw.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) { dh[2 * i] *= dqf_u16<0>(w[i]); dh[2 * i + 1] *= dqf_u16<1>(w[i]); }
}

// ------------------------------------------------------------------------------------------ operand producers
// x fp32 [B,T,In] (batch-first) -> fp16 hi/lo planes [T][Bp][Kp] (time-major); pad rows / pad columns are zero.
__global__ void x_planes_kernel(const float* __restrict__ x, int B, int Bp, int T, int In, int Kp, f16* __restrict__ hi,
                                f16* __restrict__ lo) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t m = warp; m < (int64_t)T * Bp; m += nwarps) {
    const int t = (int)(m / Bp), b = (int)(m % Bp);
    const float* src = x + ((int64_t)b * T + t) * In;
    for (int c = lane * 2; c < Kp; c += 64) {
      const float v0 = (b < B && c < In) ? __ldg(src + c) : 0.f;
      const float v1 = (b < B && c + 1 < In) ? __ldg(src + c + 1) : 0.f;
      f16 h0, l0, h1, l1;
      split_f16(v0, h0, l0);
      split_f16(v1, h1, l1);
      *reinterpret_cast<__half2*>(hi + m * Kp + c) = __halves2half2(h0, h1);
      *reinterpret_cast<__half2*>(lo + m * Kp + c) = __halves2half2(l0, l1);
    }
  }
}

// token ids [B,T] -> planes of the gathered table rows (K-A fused into the operand producer); one warp per token
template <typename IdT>
__global__ void gather_planes_kernel(const IdT* __restrict__ ids, int B, int Bp, int T, const float* __restrict__ table,
                                     int64_t vocab, int E, int Kp, int64_t pad_id, f16* __restrict__ hi, f16* __restrict__ lo) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t m = warp; m < (int64_t)T * Bp; m += nwarps) {
    const int t = (int)(m / Bp), b = (int)(m % Bp);
    int64_t id = -1;
    if (b < B) id = (int64_t)ids[(int64_t)b * T + t];
    const bool zero = id < 0 || id >= vocab || id == pad_id;
    const float* src = table + (zero ? 0 : id) * E;
    for (int c = lane * 2; c < Kp; c += 64) {
      const float v0 = (!zero && c < E) ? __ldg(src + c) : 0.f;
      const float v1 = (!zero && c + 1 < E) ? __ldg(src + c + 1) : 0.f;
      f16 h0, l0, h1, l1;
      split_f16(v0, h0, l0);
      split_f16(v1, h1, l1);
      *reinterpret_cast<__half2*>(hi + m * Kp + c) = __halves2half2(h0, h1);
      *reinterpret_cast<__half2*>(lo + m * Kp + c) = __halves2half2(l0, l1);
    }
  }
}

// weights fp32 [rows][cols] -> planes [rows][pitch]
__global__ void w_planes_kernel(const float* __restrict__ w, int rows, int cols, int pitch, f16* __restrict__ hi,
                                f16* __restrict__ lo) {
  const int64_t total = (int64_t)rows * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % pitch);
    const float v = c < cols ? w[(i / pitch) * cols + c] : 0.f;
    f16 a, b;
    split_f16(v, a, b);
    hi[i] = a;
    if (lo) lo[i] = b;
  }
}

// folded projection bias: b_ih + (b_hr, b_hz, 0)   (b_hn stays inside r * (...))
__global__ void fold_bias_kernel(const float* __restrict__ b_ih, const float* __restrict__ b_hh, int H, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < 3 * H) out[n] = b_ih[n] + (n < 2 * H ? b_hh[n] : 0.f);
}

// ------------------------------------------------------------------------------------------ K-C forward step
unsigned long long* g_step_dbg = nullptr;
static int g_seq_mode = 4;                    // 0: one launch per timestep; 2 / 4: persistent kernel, epilogue warps per quarter
static int seq_mode() { return g_seq_mode; }
static int g_wgrad_overlap = 1;               // 1: weight-gradient GEMMs of layer l overlap the BPTT of layer l-1 (helper stream)
static int g_bwd_seq_mode = 2;                // 2: fused persistent BPTT, on unit-split CTA pairs when H == 256 and the pairs fit in half the SMs;
                                              // 3: unit-split CTA pairs whenever H == 256; 4: tcgen05 CTA pairs (cta_group::2, two row tiles of one
                                              //    direction per cluster) whenever H == 256 and the row tiles pair up;
                                              // 1: always one independent CTA per (row tile, direction) (H <= 256); 0: two launches per timestep
static int g_bwd_segments = 1;                // time segments per fused BPTT (weight gradients of a finished segment overlap the next one)
static int bwd_seq_mode() { return g_bwd_seq_mode; }    // tt_debug_step_buffer(): phase timestamps of the LAST launched step
constexpr int S_BM = 128, S_BK = 64, S_UN = 64, S_BN = 3 * S_UN;
constexpr int S_A_BYTES = S_BM * S_BK * 2, S_B_BYTES = S_BN * S_BK * 2;
constexpr int S_STAGE = 2 * S_A_BYTES + 2 * S_B_BYTES;   // 81920
constexpr int S_STAGES = 2;
constexpr int S_SMEM = S_STAGES * S_STAGE + 1024 + 512;   // + barriers, tmem slot, b_hn slice
constexpr int S_TMEM_COLS = 256;
constexpr int S_THREADS = 320;                           // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue

struct alignas(64) Step2Params {
  CUtensorMap h_hi[2], h_lo[2];      // per direction: H planes [(T+1)*Bp][H], box 64 x 128 (load h_{t-1}, store h_t)
  CUtensorMap w_hi[2], w_lo[2];      // per direction: W_hh planes [3H][H], box 64 x 64
  CUtensorMap x_hi, x_lo;            // next layer's input planes [T*Bp][ndir*H], box 64 x 128 (store)
  const float* gi;                   // tiled [ndir][T][MT][3][H/4][128][4], biases folded
  const float* b_hh[2];              // only the n-gate third is read
  uint4* g16;                        // saved gate records or null: GP_PLANES planes [ndir][T][MT][H/8][128] of 16-byte words
  int64_t g16_plane;                 // words per plane (see store_rec8)
  float* hn;                         // unused (the records carry hn r (1 - r))
  float* out;                        // fp32 [B,T,ndir*H] or null
  float* h_n;                        // fp32 [ndir][B][H] (written at the last step)
  const int32_t* lengths;
  int32_t B, Bp, T, H, ndir, step, has_next;
  float drop_p, inv_keep;
  uint64_t seed;
  const uint64_t* seed_dev;          // optional device scalar added to seed (CUDA-graph replay draws a new mask)
  unsigned long long* dbg;           // optional phase timestamps, 8 per CTA (tools/step_phases.py)
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define DBG_MARK(slot)                                                                                   \
  do {                                                                                                   \
    if (p.dbg) p.dbg[((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (slot)] = gtimer(); \
  } while (0)

__global__ void __launch_bounds__(S_THREADS, 1) gru_step2_kernel(const __grid_constant__ Step2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S_STAGES * S_STAGE);
  uint64_t* full = bars;
  uint64_t* empty = bars + S_STAGES;
  uint64_t* tfull = bars + 2 * S_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  float* bhn_s = reinterpret_cast<float*>(bars + 8);                 // [64] b_hn of this CTA's units
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.z, mt = blockIdx.y, m0 = mt * S_BM, u0 = blockIdx.x * S_UN;
  const int H = p.H, T = p.T, KB = H / S_BK, MT = p.Bp >> 7;
  // k-blocks are consumed in rotated order so that this CTA's OWN 64 units (= h_{t-1} of the units it updates) are
  // the last A tile to land: it is still in shared memory when the epilogue needs h_prev (no fp32 state in HBM)
  const int own_kb = blockIdx.x, own_stage = (KB - 1) % S_STAGES;
  const int t = dir == 0 ? p.step : T - 1 - p.step;
  const int rslot = dir == 0 ? t : t + 1, wslot = dir == 0 ? t + 1 : t;

  if (threadIdx.x == 0) {
    for (int i = 0; i < S_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, S_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* st = smem + stage * S_STAGE;
        uint8_t* sa_hi = st, *sa_lo = st + S_A_BYTES, *sb_hi = st + 2 * S_A_BYTES, *sb_lo = sb_hi + S_B_BYTES;
        mbar_arrive_expect_tx(&full[stage], S_STAGE);
        const int kk = (own_kb + 1 + kb) % KB;
        tma_load_2d(&p.h_hi[dir], &full[stage], sa_hi, kk * S_BK, rslot * p.Bp + m0);
        tma_load_2d(&p.h_lo[dir], &full[stage], sa_lo, kk * S_BK, rslot * p.Bp + m0);
#pragma unroll
        for (int g = 0; g < 3; ++g) {       // gate g rows [g*H + u0, +64) of W_hh -> tile columns [64 g, 64 g + 64)
          tma_load_2d(&p.w_hi[dir], &full[stage], sb_hi + g * (S_UN * 128), kk * S_BK, g * H + u0);
          tma_load_2d(&p.w_lo[dir], &full[stage], sb_lo + g * (S_UN * 128), kk * S_BK, g * H + u0);
        }
        if (++stage == S_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(S_BM, S_BN, 0, 0);
      int stage = 0, phase = 0;
      DBG_MARK(1);
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait(&full[stage], phase);
        if (kb == 0) DBG_MARK(2);
        tc_fence_after();
        const uint32_t st = smem_u32(smem + stage * S_STAGE);
        const uint32_t sa_hi = st, sa_lo = st + S_A_BYTES, sb_hi = st + 2 * S_A_BYTES, sb_lo = sb_hi + S_B_BYTES;
#pragma unroll
        for (int term = 0; term < 3; ++term) {
          const uint32_t sa = term == 1 ? sa_lo : sa_hi, sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
          for (int k = 0; k < S_BK / 16; ++k)
            mma_bf16_ss(tmem_base, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                        (kb > 0 || term > 0 || k > 0) ? 1u : 0u);
        }
        mma_commit(&empty[stage]);
        if (++stage == S_STAGES) { stage = 0; phase ^= 1; }
      }
      mma_commit(tfull);
      DBG_MARK(3);
    }
  } else {
    // ------------------------------------------------------------------ fused GRU cell: thread = (batch row, 32 units)
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane, b = m0 + row;
    const int uh = u0 + 32 * half;                            // first hidden unit of this thread
    const bool row_ok = b < p.B;
    const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
    const int h4 = H >> 2;
    // prefetch this row's input-projection terms (do not depend on the MMA): 3 gates x 32 units
    float4 gi_r[8], gi_z[8], gi_n[8];
    {
      const float4* g = reinterpret_cast<const float4*>(p.gi) +
                        ((((int64_t)dir * T + t) * MT + mt) * 3 * h4 + (uh >> 2)) * 128 + row;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        gi_r[i] = __ldg(g + (int64_t)i * 128);
        gi_z[i] = __ldg(g + ((int64_t)h4 + i) * 128);
        gi_n[i] = __ldg(g + ((int64_t)2 * h4 + i) * 128);
      }
    }
    {
      const int e = threadIdx.x - 64;
      if (e < S_UN) bhn_s[e] = __ldg(p.b_hh[dir] + 2 * H + u0 + e);
    }
    const int64_t gbase = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + (uh >> 3)) * 128 + row;
    // 4 x 16 KB staging tiles in the operand stage that does NOT hold the own A tile (free once tfull has fired)
    uint8_t* stg = smem + (own_stage == 0 ? S_STAGE : 0);
    const uint8_t* own_hi = smem + own_stage * S_STAGE, *own_lo = own_hi + S_A_BYTES;
    named_bar_sync(2, 256);                                    // bhn_s visible
    mbar_wait(tfull, 0);
    mbar_wait(&full[own_stage], ((KB - 1) / S_STAGES) & 1);   // already complete: orders the TMA-written tile before our reads
    tc_fence_after();
    if (threadIdx.x == 64) DBG_MARK(4);
    const uint32_t tl = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(32 * half);
#pragma unroll
    for (int c = 0; c < 4; ++c) {                              // 8 units per chunk
      float ar[8], az[8], an[8];
      tmem_ld8(tl + 8 * c, ar);
      tmem_ld8(tl + S_UN + 8 * c, az);
      tmem_ld8(tl + 2 * S_UN + 8 * c, an);
      const uint32_t off = sw128_off(row, 4 * half + c);
      const uint4 ph = *reinterpret_cast<const uint4*>(own_hi + off), pl = *reinterpret_cast<const uint4*>(own_lo + off);
      const float4 bn0 = *reinterpret_cast<const float4*>(bhn_s + 32 * half + 8 * c);
      const float4 bn1 = *reinterpret_cast<const float4*>(bhn_s + 32 * half + 8 * c + 4);
      tmem_wait_ld();
      float hp[8];
      {
        const uint32_t hw[4] = {ph.x, ph.y, ph.z, ph.w}, lw[4] = {pl.x, pl.y, pl.z, pl.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&hw[i]));
          const float2 b2 = __half22float2(*reinterpret_cast<const __half2*>(&lw[i]));
          hp[2 * i] = a.x + b2.x; hp[2 * i + 1] = a.y + b2.y;
        }
      }
      const float bn[8] = {bn0.x, bn0.y, bn0.z, bn0.w, bn1.x, bn1.y, bn1.z, bn1.w};
      const float gr[8] = {gi_r[2 * c].x, gi_r[2 * c].y, gi_r[2 * c].z, gi_r[2 * c].w,
                           gi_r[2 * c + 1].x, gi_r[2 * c + 1].y, gi_r[2 * c + 1].z, gi_r[2 * c + 1].w};
      const float gz[8] = {gi_z[2 * c].x, gi_z[2 * c].y, gi_z[2 * c].z, gi_z[2 * c].w,
                           gi_z[2 * c + 1].x, gi_z[2 * c + 1].y, gi_z[2 * c + 1].z, gi_z[2 * c + 1].w};
      const float gn[8] = {gi_n[2 * c].x, gi_n[2 * c].y, gi_n[2 * c].z, gi_n[2 * c].w,
                           gi_n[2 * c + 1].x, gi_n[2 * c + 1].y, gi_n[2 * c + 1].z, gi_n[2 * c + 1].w};
      float h[8], xo[8], r[8], z[8], n[8], q[8], hn[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (live) {
          r[j] = fast_sigmoid(gr[j] + ar[j]);
          z[j] = fast_sigmoid(gz[j] + az[j]);
          hn[j] = an[j] + bn[j];
          n[j] = fast_tanh(gn[j] + r[j] * hn[j]);
          h[j] = (1.f - z[j]) * n[j] + z[j] * hp[j];
          q[j] = h[j] - n[j];
          xo[j] = h[j];
        } else {                                             // padded step / pad row: state carries, output is zero
          r[j] = 0.f; z[j] = 1.f; hn[j] = 0.f; n[j] = 0.f; q[j] = 0.f;
          h[j] = row_ok ? hp[j] : 0.f;
          xo[j] = 0.f;
        }
      }
      if (p.g16) store_rec8(p.g16 + gbase + (int64_t)c * 128, p.g16_plane, r, z, n, q, hn);
      // operand planes of h_t (next step's A tile, backward's h_prev) -> swizzled staging tiles
      {
        uint4 hi, lo;
        split8(h, hi, lo);
        *reinterpret_cast<uint4*>(stg + off) = hi;
        *reinterpret_cast<uint4*>(stg + 16384 + off) = lo;
        if (p.has_next) {
          if (p.drop_p > 0.f && row_ok) {
            const uint64_t e = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + uh + 8 * c);
            float s8[8];
            dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), e >> 3, p.drop_p, p.inv_keep, s8);
#pragma unroll
            for (int j = 0; j < 8; ++j) xo[j] *= s8[j];
          }
          split8(xo, hi, lo);
          *reinterpret_cast<uint4*>(stg + 32768 + off) = hi;
          *reinterpret_cast<uint4*>(stg + 49152 + off) = lo;
        }
      }
      if (row_ok) {
        if (p.out) {
          float4* o = reinterpret_cast<float4*>(p.out + ((int64_t)b * T + t) * (p.ndir * H) + dir * H + uh + 8 * c);
          o[0] = live ? make_float4(h[0], h[1], h[2], h[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
          o[1] = live ? make_float4(h[4], h[5], h[6], h[7]) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (p.step == T - 1) {
          float4* o = reinterpret_cast<float4*>(p.h_n + ((int64_t)dir * p.B + b) * H + uh + 8 * c);
          o[0] = make_float4(h[0], h[1], h[2], h[3]);
          o[1] = make_float4(h[4], h[5], h[6], h[7]);
        }
      }
    }
    fence_proxy_async();
    named_bar_sync(1, 256);
    if (warp == 2 && lane == 0) {
      DBG_MARK(5);
      tma_store_2d(&p.h_hi[dir], stg, u0, wslot * p.Bp + m0);
      tma_store_2d(&p.h_lo[dir], stg + 16384, u0, wslot * p.Bp + m0);
      if (p.has_next) {
        tma_store_2d(&p.x_hi, stg + 32768, dir * H + u0, t * p.Bp + m0);
        tma_store_2d(&p.x_lo, stg + 49152, dir * H + u0, t * p.Bp + m0);
      }
      bulk_commit();
      bulk_wait_read0();
      DBG_MARK(6);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, S_TMEM_COLS);
  if (threadIdx.x == 0) DBG_MARK(7);
}

// ------------------------------------------------------------------------------------------ K-C persistent recurrence
// One cooperative launch runs ALL T timesteps of one layer (both directions).  Every CTA owns a fixed set of
// (64-unit slice, batch tile, direction) tiles and walks them step by step; the step-s tile of (mt, dir) may start
// once the H/64 CTAs that produce h_{s-1} of that batch tile have published it (release counter per (dir, mt),
// written after their TMA stores completed).  H-plane slots, X planes and gates are distinct per timestep, so RAW
// through the counters is the only inter-CTA dependency.  Inside a CTA three roles run a software pipeline over
// (step, tile) items: warp 0 = TMA producer (operand ring), warp 1 = tcgen05.mma issuer (two TMEM accumulators),
// warps 2.. = fused GRU-cell epilogue, so the cell math of one tile overlaps the operand traffic of the next.
constexpr int Q_STAGING = 4 * 16384;
constexpr int Q_SMEM = S_STAGES * S_STAGE + Q_STAGING + 1024 /*align*/ + 1024 /*barriers, b_hn*/;
constexpr int Q_TMEM_COLS = 512;

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

template <int N> __device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

struct alignas(64) SeqParams {
  Step2Params s;
  int* ready;                        // [ndir][MT] counters, zero before the launch
  int32_t tiles;                     // (H/64) * MT * ndir
};

// In-kernel %globaltimer phase marks of the two persistent kernels (tools/step_phases.py, tools/bwd_phases.py) are compiled
// in only with -DTT_PHASE_MARKS (TT_PHASE_MARKS=1 python -m two_towers_b200.build --force): even predicated off they sat in
// front of the named barriers of the per-item loop (ncu: the thread-0 warp stalled on them, every other warp on the barrier).
#ifdef TT_PHASE_MARKS
#define SEQ_MARK(item, slot)                                                              \
  do {                                                                                    \
    if (p.dbg && blockIdx.x == 0 && (item) < 256) p.dbg[(item) * 8 + (slot)] = gtimer(); \
  } while (0)
#else
#define SEQ_MARK(item, slot) do { } while (0)
#endif

template <int EWQ>                                              // epilogue warps per TMEM lane quarter (2 or 4)
__global__ void __launch_bounds__(128 * EWQ + 128, 1) gru_seq_kernel(const __grid_constant__ SeqParams sp) {
  constexpr int UT = S_UN / EWQ;                                // hidden units per epilogue thread (32 or 16)
  constexpr int NCH = UT / 8;                                   // 8-unit chunks per thread
  constexpr int EPI_THREADS = 128 * EWQ;
  constexpr int W_TMA = 4 * EWQ, W_MMA = 4 * EWQ + 1;           // control warpgroup follows the epilogue warpgroups
  const Step2Params& p = sp.s;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stg = smem + S_STAGES * S_STAGE;                     // 4 x 16 KB staging tiles (epilogue -> TMA store)
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + Q_STAGING);
  uint64_t* full = bars;                                        // [S_STAGES]
  uint64_t* empty = bars + S_STAGES;                            // [S_STAGES], 2 arrivals per use
  uint64_t* tfull = bars + 2 * S_STAGES;                        // [2]
  uint64_t* tempty = tfull + 2;                                 // [2]
  uint64_t* sready = tempty + 2;                                // staging tiles written (epilogue -> store warp)
  uint64_t* sfree = sready + 1;                                 // staging tiles read out by the TMA store (store warp -> epilogue)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sfree + 1);
  float* bhn_s = reinterpret_cast<float*>(bars + 16);           // [2][64]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int H = p.H, T = p.T, KB = H / S_BK, MT = p.Bp >> 7, US = H / S_UN;
  constexpr int W_STORE = 4 * EWQ + 2;

  if (threadIdx.x == 0) {
    for (int i = 0; i < S_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 2); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 1); }
    mbar_init(sready, 1);
    mbar_init(sfree, 1);
    fence_barrier_init();
  }
  if (warp == W_MMA) tmem_alloc(tmem_slot, Q_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 4 * EWQ) {
    if (EWQ == 4) setmaxnreg_dec<32>();                          // 512 * 112 + 128 * 32 == 640 * 96: what the launch allocated
  if (warp == W_TMA) {
    // ================================================================= TMA producer
    if (lane == 0) {
      int stage = 0, phase = 0, pitem = 0;
      for (int step = 0; step < T; ++step) {
        for (int tile = blockIdx.x; tile < sp.tiles; tile += gridDim.x, ++pitem) {
          const int us = tile % US, mt = (tile / US) % MT, dir = tile / (US * MT);
          const int t = dir == 0 ? step : T - 1 - step;
          const int rslot = dir == 0 ? t : t + 1;
          const int m0 = mt * S_BM, u0 = us * S_UN;
          for (int kb = 0; kb < KB; ++kb) {
            const int kk = (us + 1 + kb) % KB;                  // own k-block last (the epilogue reads h_prev from it)
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* st = smem + stage * S_STAGE;
            uint8_t* sa_hi = st, *sa_lo = st + S_A_BYTES, *sb_hi = st + 2 * S_A_BYTES, *sb_lo = sb_hi + S_B_BYTES;
            mbar_arrive_expect_tx(&full[stage], S_STAGE);
#pragma unroll
            for (int g = 0; g < 3; ++g) {
              tma_load_2d(&p.w_hi[dir], &full[stage], sb_hi + g * (S_UN * 128), kk * S_BK, g * H + u0);
              tma_load_2d(&p.w_lo[dir], &full[stage], sb_lo + g * (S_UN * 128), kk * S_BK, g * H + u0);
            }
            if (kb == 0) SEQ_MARK(pitem, 0);
            if (kb == 0 && step > 0) {                           // h_{step-1} of this batch tile published by all US producers?
              const int* flag = sp.ready + dir * MT + mt;
              const int need = US * step;
              while (ld_acquire_gpu(flag) < need) __nanosleep(40);
              fence_proxy_async_global();
            }
            if (kb == 0) SEQ_MARK(pitem, 1);
            if (kb == KB - 1) SEQ_MARK(pitem, 2);
            tma_load_2d(&p.h_hi[dir], &full[stage], sa_hi, kk * S_BK, rslot * p.Bp + m0);
            tma_load_2d(&p.h_lo[dir], &full[stage], sa_lo, kk * S_BK, rslot * p.Bp + m0);
            if (++stage == S_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == W_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(S_BM, S_BN, 0, 0);
      int stage = 0, phase = 0, item = 0;
      for (int step = 0; step < T; ++step) {
        for (int tile = blockIdx.x; tile < sp.tiles; tile += gridDim.x, ++item) {
          const int acc = item & 1, acc_phase = (item >> 1) & 1;
          mbar_wait(&tempty[acc], acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * 256;
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&full[stage], phase);
            tc_fence_after();
            const uint32_t st = smem_u32(smem + stage * S_STAGE);
            const uint32_t sa_hi = st, sa_lo = st + S_A_BYTES, sb_hi = st + 2 * S_A_BYTES, sb_lo = sb_hi + S_B_BYTES;
#pragma unroll
            for (int term = 0; term < 3; ++term) {
              const uint32_t sa = term == 1 ? sa_lo : sa_hi, sb = term == 2 ? sb_lo : sb_hi;
#pragma unroll
              for (int k = 0; k < S_BK / 16; ++k)
                mma_bf16_ss(d_tmem, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sb + k * 32, 16, 1024), idesc,
                            (kb > 0 || term > 0 || k > 0) ? 1u : 0u);
            }
            mma_commit(&empty[stage]);                           // arrival 1 of 2: the MMAs have read the stage
            if (kb != KB - 1) mbar_arrive(&empty[stage]);        // arrival 2 of 2 (the own tile is released by the epilogue)
            if (++stage == S_STAGES) { stage = 0; phase ^= 1; }
          }
          mma_commit(&tfull[acc]);
        }
      }
    }
  } else if (warp == W_STORE) {
    // ================================================================= h_t writer + publisher (off the epilogue threads' loop:
    // the epilogue goes on with the CTA's next item while the stores drain and the release counter is bumped)
    if (lane == 0) {
      int item = 0;
      for (int step = 0; step < T; ++step) {
        for (int tile = blockIdx.x; tile < sp.tiles; tile += gridDim.x, ++item) {
          const int us = tile % US, mt = (tile / US) % MT, dir = tile / (US * MT);
          const int t = dir == 0 ? step : T - 1 - step;
          const int wslot = dir == 0 ? t + 1 : t;
          const int m0 = mt * S_BM, u0 = us * S_UN;
          mbar_wait(sready, item & 1);
          tma_store_2d(&p.h_hi[dir], stg, u0, wslot * p.Bp + m0);
          tma_store_2d(&p.h_lo[dir], stg + 16384, u0, wslot * p.Bp + m0);
          if (p.has_next) {
            tma_store_2d(&p.x_hi, stg + 32768, dir * H + u0, t * p.Bp + m0);
            tma_store_2d(&p.x_lo, stg + 49152, dir * H + u0, t * p.Bp + m0);
          }
          bulk_commit();
          bulk_wait_read0();                                   // staging read out: the epilogue may overwrite it
          mbar_arrive(sfree);
          bulk_wait0();                                        // writes complete (not only the smem reads)
          fence_proxy_async_global();
          __threadfence();
          atomicAdd(sp.ready + dir * MT + mt, 1);              // publish h_t of (us, mt, dir)
          SEQ_MARK(item, 6);
        }
      }
    }
  }
  } else {
    if (EWQ == 4) setmaxnreg_inc<112>();
    // ================================================================= fused GRU cell: thread = (batch row, UT units)
    const int e = threadIdx.x;
    const int quarter = warp & 3, sub = warp >> 2;
    const int row = quarter * 32 + lane;
    const int h4 = H >> 2;
    int item = 0;
    // b_hn of the CTA's tiles: constant over the timesteps.  With at most two tiles per CTA (slot = item & 1 = position of
    // the tile in the CTA's round) it is loaded ONCE; it used to be a global load in front of named barrier 1 of every item.
    const bool bhn_once = (int)blockIdx.x + 2 * (int)gridDim.x >= sp.tiles;
    if (bhn_once) {
      const int own = (int)blockIdx.x + (int)gridDim.x < sp.tiles ? 2 : 1;       // tiles of this CTA
      for (int k = 0; k < 2; ++k) {                                              // one tile: both slots hold it
        const int tile = (int)blockIdx.x + (k % own) * (int)gridDim.x;
        if (e < S_UN && tile < sp.tiles) bhn_s[k * S_UN + e] = __ldg(p.b_hh[tile / (US * MT)] + 2 * H + (tile % US) * S_UN + e);
      }
      named_bar_sync(1, EPI_THREADS);
    }
    for (int step = 0; step < T; ++step) {
      for (int tile = blockIdx.x; tile < sp.tiles; tile += gridDim.x, ++item) {
        const int us = tile % US, mt = (tile / US) % MT, dir = tile / (US * MT);
        const int t = dir == 0 ? step : T - 1 - step;
        const int m0 = mt * S_BM, u0 = us * S_UN, b = m0 + row;
        const int uh = u0 + UT * sub;
        const bool row_ok = b < p.B;
        const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
        const int acc = item & 1, acc_phase = (item >> 1) & 1;
        // input-projection terms of this (step, tile): independent of the recurrence -> in flight during the MMA
        float4 gi_r[UT / 4], gi_z[UT / 4], gi_n[UT / 4];
        {
          const float4* g = reinterpret_cast<const float4*>(p.gi) +
                            ((((int64_t)dir * T + t) * MT + mt) * 3 * h4 + (uh >> 2)) * 128 + row;
#pragma unroll
          for (int i = 0; i < UT / 4; ++i) {
            gi_r[i] = __ldg(g + (int64_t)i * 128);
            gi_z[i] = __ldg(g + ((int64_t)h4 + i) * 128);
            gi_n[i] = __ldg(g + ((int64_t)2 * h4 + i) * 128);
          }
        }
        float* bh = bhn_s + (item & 1) * S_UN;
        if (!bhn_once && e < S_UN) bh[e] = __ldg(p.b_hh[dir] + 2 * H + u0 + e);
        const int64_t gbase = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + (uh >> 3)) * 128 + row;
        // ring position of the own A tile (last k-block of this item)
        const int own_pos = (item + 1) * KB - 1, own_stage = own_pos % S_STAGES;
        const uint8_t* own_hi = smem + own_stage * S_STAGE, *own_lo = own_hi + S_A_BYTES;
        if (e == 0) SEQ_MARK(item, 3);
        mbar_wait(&tfull[acc], acc_phase);
        if (e == 0) SEQ_MARK(item, 4);
        mbar_wait(&full[own_stage], (own_pos / S_STAGES) & 1);  // already complete: orders the TMA-written tile before our reads
        tc_fence_after();
        // h_prev of this thread's units out of the operand tile, then hand the ring stage back to the producer
        uint4 ph[NCH], pl[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          const uint32_t off = sw128_off(row, NCH * sub + c);
          ph[c] = *reinterpret_cast<const uint4*>(own_hi + off);
          pl[c] = *reinterpret_cast<const uint4*>(own_lo + off);
        }
        if (e == 0 && item > 0) mbar_wait(sfree, (item - 1) & 1);   // previous item's staging tiles have been read out
        named_bar_sync(1, EPI_THREADS);      // all h_prev reads done, bhn_s written, staging free
        if (e == 0) mbar_arrive(&empty[own_stage]);
        const uint32_t tl = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 256 + UT * sub);
#pragma unroll
        for (int c = 0; c < NCH; ++c) {                          // 8 units per chunk
          float ar[8], az[8], an[8];
          tmem_ld8(tl + 8 * c, ar);
          tmem_ld8(tl + S_UN + 8 * c, az);
          tmem_ld8(tl + 2 * S_UN + 8 * c, an);
          const float4 bn0 = *reinterpret_cast<const float4*>(bh + UT * sub + 8 * c);
          const float4 bn1 = *reinterpret_cast<const float4*>(bh + UT * sub + 8 * c + 4);
          tmem_wait_ld();
          float hp[8];
          {
            const uint32_t hw[4] = {ph[c].x, ph[c].y, ph[c].z, ph[c].w}, lw[4] = {pl[c].x, pl[c].y, pl[c].z, pl[c].w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&hw[i]));
              const float2 b2 = __half22float2(*reinterpret_cast<const __half2*>(&lw[i]));
              hp[2 * i] = a.x + b2.x; hp[2 * i + 1] = a.y + b2.y;
            }
          }
          const float bn[8] = {bn0.x, bn0.y, bn0.z, bn0.w, bn1.x, bn1.y, bn1.z, bn1.w};
          const float gr[8] = {gi_r[2 * c].x, gi_r[2 * c].y, gi_r[2 * c].z, gi_r[2 * c].w,
                               gi_r[2 * c + 1].x, gi_r[2 * c + 1].y, gi_r[2 * c + 1].z, gi_r[2 * c + 1].w};
          const float gz[8] = {gi_z[2 * c].x, gi_z[2 * c].y, gi_z[2 * c].z, gi_z[2 * c].w,
                               gi_z[2 * c + 1].x, gi_z[2 * c + 1].y, gi_z[2 * c + 1].z, gi_z[2 * c + 1].w};
          const float gn[8] = {gi_n[2 * c].x, gi_n[2 * c].y, gi_n[2 * c].z, gi_n[2 * c].w,
                               gi_n[2 * c + 1].x, gi_n[2 * c + 1].y, gi_n[2 * c + 1].z, gi_n[2 * c + 1].w};
          float h[8], xo[8], r[8], z[8], n[8], q[8], hn[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (live) {
              r[j] = fast_sigmoid(gr[j] + ar[j]);
              z[j] = fast_sigmoid(gz[j] + az[j]);
              hn[j] = an[j] + bn[j];
              n[j] = fast_tanh(gn[j] + r[j] * hn[j]);
              h[j] = (1.f - z[j]) * n[j] + z[j] * hp[j];
              q[j] = h[j] - n[j];
              xo[j] = h[j];
            } else {                                           // padded step / pad row: state carries, output is zero
              r[j] = 0.f; z[j] = 1.f; hn[j] = 0.f; n[j] = 0.f; q[j] = 0.f;
              h[j] = row_ok ? hp[j] : 0.f;
              xo[j] = 0.f;
            }
          }
          if (p.g16) store_rec8(p.g16 + gbase + (int64_t)c * 128, p.g16_plane, r, z, n, q, hn);
          {
            const uint32_t off = sw128_off(row, NCH * sub + c);
            uint4 hi, lo;
            split8(h, hi, lo);
            *reinterpret_cast<uint4*>(stg + off) = hi;
            *reinterpret_cast<uint4*>(stg + 16384 + off) = lo;
            if (p.has_next) {
              if (p.drop_p > 0.f && row_ok) {
                const uint64_t el = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + uh + 8 * c);
                float s8[8];
                dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), el >> 3, p.drop_p, p.inv_keep, s8);
#pragma unroll
                for (int j = 0; j < 8; ++j) xo[j] *= s8[j];
              }
              split8(xo, hi, lo);
              *reinterpret_cast<uint4*>(stg + 32768 + off) = hi;
              *reinterpret_cast<uint4*>(stg + 49152 + off) = lo;
            }
          }
          if (row_ok) {
            if (p.out) {
              float4* o = reinterpret_cast<float4*>(p.out + ((int64_t)b * T + t) * (p.ndir * H) + dir * H + uh + 8 * c);
              o[0] = live ? make_float4(h[0], h[1], h[2], h[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
              o[1] = live ? make_float4(h[4], h[5], h[6], h[7]) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (step == T - 1) {
              float4* o = reinterpret_cast<float4*>(p.h_n + ((int64_t)dir * p.B + b) * H + uh + 8 * c);
              o[0] = make_float4(h[0], h[1], h[2], h[3]);
              o[1] = make_float4(h[4], h[5], h[6], h[7]);
            }
          }
        }
        // accumulator drained -> the MMA warp may start the item after next; staging -> global through TMA
        tc_fence_before();
        fence_proxy_async();
        named_bar_sync(2, EPI_THREADS);
        if (e == 0) {
          SEQ_MARK(item, 5);
          mbar_arrive(&tempty[acc]);
          mbar_arrive(sready);                                 // the store warp writes h_t out and publishes it
        }
        // (the staging tiles are reused by the next item only after named barrier 1, which thread e == 0 reaches
        //  after the store warp has reported them read out)
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == W_MMA) tmem_dealloc(tmem_base, Q_TMEM_COLS);
}

// ------------------------------------------------------------------------------------------ K-C' gate gradients
// dG planes per direction: [T*Bp][4H] fp16, columns (dr_pre | dz_pre | dn_pre * r | dn_pre), in loss-scaled units.
struct alignas(64) CellBwdParams {
  CUtensorMap dg[2];                 // store maps, box 64 x 128
  const uint4* g16;
  int64_t g16_plane;
  const float* hn;
  float* dhs;                        // tiled running dL/dh [ndir][MT][H/4][128][4] (scaled)
  const float* d_above;              // tiled [T][MT][ndir*H/4][128][4] (scaled) or null
  const int32_t* lengths;
  int32_t B, Bp, T, H, ndir, step;
  float drop_p, inv_keep;            // dropout that was applied to this layer's OUTPUT (the next layer's input)
  uint64_t seed;
  const uint64_t* seed_dev;
};
constexpr int CB_SMEM = 4 * 16384 + 1024;

__global__ void __launch_bounds__(512) cell_bwd2_kernel(const __grid_constant__ CellBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stg = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int dir = blockIdx.z, mt = blockIdx.y, m0 = mt * 128, u0 = blockIdx.x * 64;
  const int row = threadIdx.x & 127, qtr = threadIdx.x >> 7, b = m0 + row;      // thread = (batch row, 16 units)
  const int H = p.H, T = p.T, MT = p.Bp >> 7, h4 = H >> 2;
  const int t = dir == 0 ? p.step : T - 1 - p.step;
  const int uh = u0 + 16 * qtr;
  const bool row_ok = b < p.B;
  const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
  const int64_t gbase = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + (uh >> 3)) * 128 + row;
  float4* dhs = reinterpret_cast<float4*>(p.dhs) + (((int64_t)dir * MT + mt) * h4 + (uh >> 2)) * 128 + row;
  const float4* dab = p.d_above ? reinterpret_cast<const float4*>(p.d_above) +
                                      (((int64_t)t * MT + mt) * (p.ndir * h4) + ((dir * H + uh) >> 2)) * 128 + row
                                : nullptr;
  // records of padded steps / pad rows hold z = 1 and zeros (store_rec8): no masking needed, except that the upstream
  // gradient of a padded step's (all-zero) output must not enter dh
  GateRec8 g[2];
  uint4 zw[2];
  float4 dh4[4], da4[4];
#pragma unroll
  for (int hf = 0; hf < 2; ++hf) {
    load_rec8(g[hf], p.g16 + gbase + (int64_t)hf * 128, p.g16_plane);
    zw[hf] = __ldg(p.g16 + GP_Z * p.g16_plane + gbase + (int64_t)hf * 128);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    dh4[i] = dhs[i * 128];
    da4[i] = (dab && live) ? __ldg(dab + (int64_t)i * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int hf = 0; hf < 2; ++hf) {
    float dh[8];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = 2 * hf + k;
      float4 da = da4[i];
      if (dab && p.drop_p > 0.f) {
        const uint64_t e = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + uh + 4 * i);
        float s8[8];
        dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), e >> 3, p.drop_p, p.inv_keep, s8);
        const int o8 = (i & 1) * 4;
        da.x *= s8[o8]; da.y *= s8[o8 + 1]; da.z *= s8[o8 + 2]; da.w *= s8[o8 + 3];
      }
      dh[4 * k] = dh4[i].x + da.x; dh[4 * k + 1] = dh4[i].y + da.y; dh[4 * k + 2] = dh4[i].z + da.z; dh[4 * k + 3] = dh4[i].w + da.w;
    }
    uint4 o_dr, o_dz, o_dnr, o_dn;
    grad_rec8(g[hf], dh, o_dr, o_dz, o_dnr, o_dn);
    const uint32_t off = sw128_off(row, 2 * qtr + hf);
    *reinterpret_cast<uint4*>(stg + off) = o_dr;
    *reinterpret_cast<uint4*>(stg + 16384 + off) = o_dz;
    *reinterpret_cast<uint4*>(stg + 32768 + off) = o_dnr;
    *reinterpret_cast<uint4*>(stg + 49152 + off) = o_dn;
    mul_z8(dh, zw[hf]);                                           // dh_prev (before the recurrent product) = dh * z
    dhs[(2 * hf) * 128] = make_float4(dh[0], dh[1], dh[2], dh[3]);
    dhs[(2 * hf + 1) * 128] = make_float4(dh[4], dh[5], dh[6], dh[7]);
  }
  fence_proxy_async();
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int g = 0; g < 4; ++g) tma_store_2d(&p.dg[dir], stg + g * 16384, g * H + u0, t * p.Bp + m0);
    bulk_commit();
    bulk_wait_read0();
  }
}

// ------------------------------------------------------------------------------------------ K-C' persistent BPTT
// One launch runs the whole reverse-time recurrence of one layer.  A CTA owns ALL hidden units of one
// (128-row batch tile, direction), so consecutive steps only depend on the CTA itself: no inter-CTA traffic at all.
// The running gradient dL/dh lives in TENSOR MEMORY for the entire sequence: two 256-column buffers alternate,
//   buf[cur] = dL/dh_t  ->  pass 1: dht = buf[cur] (+ d_above[t]);  buf[cur] <- dht;  buf[nxt] <- dht * z_t
//                           pass 2: per 64-unit chunk, gate gradients (dr, dz, dn*r, dn) -> fp16 swizzled smem tiles
//                                   -> (a) TMA store to the dG planes (wgrad / dX GEMMs later), (b) A operand of
//                                   buf[nxt] += dGH_chunk W_hh[chunk rows]   (tcgen05, single fp16 pass, W_hh streamed
//                                   from L2 through a TMA ring; the operand never touches HBM on its way to the MMA)
// so after the 12 k-blocks buf[nxt] = dL/dh_{t-1} and the roles of the buffers swap.
constexpr int R_WSTAGE = 64 * 256 * 2;             // one k-block of W_hh: [64 k][256 n] fp16 = 32 KB
constexpr int R_WSTAGES = 3;
constexpr int R_STG = 4 * 16384;                   // one staging buffer: tiles dr | dz | dn*r | dn, [128][64] fp16 each
constexpr int R_SMEM = R_WSTAGES * R_WSTAGE + 2 * R_STG + 1024 + 512;
constexpr int R_THREADS = 512 + 128;              // warps 0-15 gate gradients; control warpgroup: 16 TMA, 17 MMA, 18 dG writer, 19 idle
constexpr int R_WARP_TMA = 16, R_WARP_MMA = 17, R_WARP_STORE = 18;


struct alignas(64) BwdSeqParams {
  CUtensorMap dg[2];                 // store maps over the dG planes [T*Bp][4H], box 64 x 128
  CUtensorMap w[2];                  // load maps over W_hh hi planes [3H][H], box 64 x 64
  const uint4* g16;
  int64_t g16_plane;
  const float* hn;
  const float* dhs0;                 // tiled initial dL/dh [ndir][MT][H/4][128][4] (scaled)
  const float* d_above;              // tiled [T][MT][ndir*H/4][128][4] (scaled) or null
  const int32_t* lengths;
  int32_t B, Bp, T, H, ndir, tiles;
  float drop_p, inv_keep;
  uint64_t seed;
  const uint64_t* seed_dev;
  unsigned long long* dbg;
  // time segment of this launch: BPTT steps step_begin-1, step_begin-2, ..., step_end (step s is timestep s of the
  // forward direction, T-1-s of the reverse one).  A whole sequence is (T, 0).  dhs_out (may alias dhs0, may be null):
  // dL/dh after the segment's last step, in the dhs0 layout - the next segment's launch starts from it.
  int32_t step_begin, step_end;
  float* dhs_out;
};
#ifdef TT_PHASE_MARKS
#define BWD_MARK(itv, slot)                                                              \
  do {                                                                                   \
    if (p.dbg && blockIdx.x == 0 && (itv) < 64) p.dbg[(itv) * 16 + (slot)] = gtimer(); \
  } while (0)
#else
#define BWD_MARK(itv, slot) do { } while (0)
#endif

__device__ __forceinline__ void tmem_ld16v(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2_bulk(const void* gptr, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gptr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

// requires H in {64, 128, 256} (one <=256-column accumulator holds every hidden unit of the tile); NCH = H / 64.
// The gate-gradient threads are latency-bound on the saved-gate records, so every global load is issued one
// pipeline stage ahead of its use: all z words of a step before the accumulator wait, d_above one chunk ahead,
// the (r,z,n,q,hn) records one half-chunk (8 units) ahead - across the chunk barriers.
template <int NCH, bool HAS_DAB>
__global__ void __launch_bounds__(R_THREADS, 1) gru_bwd_seq_kernel(const __grid_constant__ BwdSeqParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wring = smem;
  uint8_t* stg = smem + R_WSTAGES * R_WSTAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + 2 * R_STG);
  uint64_t* wfull = bars;                 // [3]
  uint64_t* wempty = bars + 3;            // [3]
  uint64_t* sfull = bars + 6;             // [2] staging written (cell threads -> MMA)
  uint64_t* sempty = bars + 8;            // [2] staging consumed (MMA commit -> cell threads)
  uint64_t* init_done = bars + 10;        // buf[nxt] holds dht * z
  uint64_t* acc_full = bars + 11;         // all MMAs of the step complete
  uint64_t* sstored = bars + 12;          // [2] TMA store has read the staging buffer (store warp -> cell threads)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int H = 64 * NCH, h4 = H >> 2;
  const int T = p.T, MT = p.Bp >> 7;

  if (threadIdx.x == 0) {
    for (int i = 0; i < R_WSTAGES; ++i) { mbar_init(&wfull[i], 1); mbar_init(&wempty[i], 1); }
    // sfull / init_done: one arrival per gate-gradient WARP (16) - no block-wide barrier in the per-chunk loop, so a warp
    // that is ahead goes on into the other staging buffer instead of waiting for the slowest one at every hand-over
    for (int i = 0; i < 2; ++i) { mbar_init(&sfull[i], 16); mbar_init(&sempty[i], 1); mbar_init(&sstored[i], 1); }
    mbar_init(init_done, 16);
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == R_WARP_MMA) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // register budget: the 16 gate-gradient warps (warpgroups 0-3) take what the control warpgroup gives up
  if (warp >= 16) {
    setmaxnreg_dec<32>();   // 512 * 112 + 128 * 32 == 640 * 96 registers: exactly what the launch allocated
  if (warp == R_WARP_TMA) {
    // ================================================================= W_hh producer (same 3*NCH k-blocks every step)
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        const int dir = tile / MT;
        for (int step = p.step_end; step < p.step_begin; ++step)
          for (int c = 0; c < NCH; ++c)
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wempty[stage], phase ^ 1);
              uint8_t* sb = wring + stage * R_WSTAGE;
              mbar_arrive_expect_tx(&wfull[stage], 64 * H * 2);
#pragma unroll
              for (int hh = 0; hh < NCH; ++hh) tma_load_2d(&p.w[dir], &wfull[stage], sb + hh * 8192, hh * 64, g * H + c * 64);
              if (++stage == R_WSTAGES) { stage = 0; phase ^= 1; }
            }
      }
    }
  } else if (warp == R_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(128, H, 0, 1);
      int stage = 0, phase = 0, it = 0, use = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        for (int step = p.step_end; step < p.step_begin; ++step, ++it) {
          const uint32_t d_tmem = tmem_base + ((it & 1) ^ 1) * 256;        // buf[nxt]
          mbar_wait(init_done, it & 1);
          tc_fence_after();
          for (int c = 0; c < NCH; ++c, ++use) {
            const int sb = use & 1;
            mbar_wait(&sfull[sb], (use >> 1) & 1);
            tc_fence_after();
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wfull[stage], phase);
              tc_fence_after();
              const uint32_t sa = smem_u32(stg + sb * R_STG + g * 16384), sw = smem_u32(wring + stage * R_WSTAGE);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_bf16_ss(d_tmem, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sw + k * 2048, 8192, 1024), idesc, 1u);
              mma_commit(&wempty[stage]);
              if (++stage == R_WSTAGES) { stage = 0; phase ^= 1; }
            }
            mma_commit(&sempty[sb]);
          }
          mma_commit(acc_full);
          BWD_MARK(it, 15);
        }
      }
    }
  } else if (warp == R_WARP_STORE) {
    // ================================================================= dG plane writer (off the cell threads' critical path)
    if (lane == 0) {
      int use = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        const int mt = tile % MT, dir = tile / MT, m0 = mt * 128;
        for (int step = p.step_begin - 1; step >= p.step_end; --step) {
          const int t = dir == 0 ? step : T - 1 - step;
          if (step > p.step_end) {
            // pull the NEXT step's saved-gate records (contiguous per (dir, t, row tile)) into L2 one whole step ahead:
            // the cell threads' loads then see L2 latency instead of HBM latency
            const int tn = dir == 0 ? step - 1 : T - step;
            // each plane's words of (dir, tn, mt) are contiguous: H/8 groups x 128 rows x 16 B = NCH x 16 KB
            const int64_t w0 = (((int64_t)dir * T + tn) * MT + mt) * (H >> 3) * 128;
#pragma unroll 1
            for (int k = 0; k < GP_PLANES; ++k)
              prefetch_l2_bulk(reinterpret_cast<const char*>(p.g16 + k * p.g16_plane + w0), NCH * 16384);
            if (HAS_DAB) {
              const char* dp = reinterpret_cast<const char*>(p.d_above) +
                               (((int64_t)tn * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2)) * 128 * 16;
#pragma unroll 1
              for (int i = 0; i < (NCH + 1) / 2; ++i) prefetch_l2_bulk(dp + i * 65536, NCH == 1 ? 32768 : 65536);
            }
          }
          for (int c = 0; c < NCH; ++c, ++use) {
            const int sb = use & 1;
            mbar_wait(&sfull[sb], (use >> 1) & 1);
            const uint8_t* sbuf = stg + sb * R_STG;
#pragma unroll
            for (int g = 0; g < 4; ++g) tma_store_2d(&p.dg[dir], sbuf + g * 16384, g * H + 64 * c, t * p.Bp + m0);
            bulk_commit();
            bulk_wait_read0();
            mbar_arrive(&sstored[sb]);
          }
        }
      }
    }
  }
  } else {
    setmaxnreg_inc<112>();
    // ================================================================= gate gradients: thread = (batch row, 16 units per chunk)
    const int e = threadIdx.x;
    const int quarter = warp & 3, sub = warp >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int it = 0, use = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      const int mt = tile % MT, dir = tile / MT, m0 = mt * 128, b = m0 + row;
      const bool row_ok = b < p.B;
      // ---- dL/dh at the end of the sequence -> buf[cur of this tile's first step]
      if (it > 0) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
      {
        const float4* src = reinterpret_cast<const float4*>(p.dhs0) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = __ldg(src + (int64_t)(16 * c + 4 * sub + i) * 128);
            v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
          }
          tmem_st16(lane_base + (it & 1) * 256 + 64 * c + 16 * sub, v);
        }
        tmem_wait_st();
      }
      for (int step = p.step_begin - 1; step >= p.step_end; --step, ++it) {
        if (e == 0) BWD_MARK(it, 14);
        const int t = dir == 0 ? step : T - 1 - step;
        const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
        const uint32_t cur = lane_base + (it & 1) * 256, nxt = lane_base + ((it & 1) ^ 1) * 256;
        // word index (plane Z) of this thread's first 8-unit group; group (chunk c, half hf) is at rec0 + (8 c + hf) * 128
        const int64_t rec0 = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + 2 * sub) * 128 + row;
        const float4* dab0 = HAS_DAB ? reinterpret_cast<const float4*>(p.d_above) +
                                           (((int64_t)t * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2) + 4 * sub) * 128 + row
                                     : nullptr;
        // ---------------- all z words of the step: in flight while the previous step's MMAs finish
        uint4 zw[NCH * 2];                        // (padded steps / pad rows hold z = 1: the gradient passes through)
#pragma unroll
        for (int g = 0; g < NCH * 2; ++g)
          zw[g] = __ldg(p.g16 + GP_Z * p.g16_plane + rec0 + (int64_t)((g >> 1) * 8 + (g & 1)) * 128);
        float4 da[4];
        if (HAS_DAB) {
#pragma unroll
          for (int i = 0; i < 4; ++i) da[i] = live ? __ldg(dab0 + (int64_t)i * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (e == 0) BWD_MARK(it, 0);
        if (step != p.step_begin - 1) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
        if (e == 0) BWD_MARK(it, 1);
        // ---------------- pass 1: dht = dh (+ d_above); buf[cur] <- dht; buf[nxt] <- dht * z
        // (tcgen05.ld / .st are warp-collective: every lane executes them; only the VALUES depend on `live`)
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float dh[16];
          tmem_ld16v(cur + 64 * c + 16 * sub, dh);
          if (HAS_DAB) {
            float4* d4 = da;
            if (p.drop_p > 0.f && live) {
              const uint64_t el = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + 64 * c + 16 * sub);
              float s8[8];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), el >> 3, p.drop_p, p.inv_keep, s8);
              d4[0].x *= s8[0]; d4[0].y *= s8[1]; d4[0].z *= s8[2]; d4[0].w *= s8[3];
              d4[1].x *= s8[4]; d4[1].y *= s8[5]; d4[1].z *= s8[6]; d4[1].w *= s8[7];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), (el >> 3) + 1, p.drop_p, p.inv_keep, s8);
              d4[2].x *= s8[0]; d4[2].y *= s8[1]; d4[2].z *= s8[2]; d4[2].w *= s8[3];
              d4[3].x *= s8[4]; d4[3].y *= s8[5]; d4[3].z *= s8[6]; d4[3].w *= s8[7];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) { dh[4 * i] += d4[i].x; dh[4 * i + 1] += d4[i].y; dh[4 * i + 2] += d4[i].z; dh[4 * i + 3] += d4[i].w; }
            if (c + 1 < NCH) {                                   // next chunk's d_above (L2-resident: prefetched a step ahead)
#pragma unroll
              for (int i = 0; i < 4; ++i)
                da[i] = live ? __ldg(dab0 + (int64_t)(16 * (c + 1) + i) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            tmem_st16(cur + 64 * c + 16 * sub, dh);
          }
          mul_z8(dh, zw[2 * c]);
          mul_z8(dh + 8, zw[2 * c + 1]);
          tmem_st16(nxt + 64 * c + 16 * sub, dh);
        }
        // first half-chunk of gate records for pass 2 (in flight across the barrier)
        GateRec8 g0, g1;
#define TT_LOAD_HALF(G, HCV) load_rec8(G, p.g16 + rec0 + (int64_t)(((HCV) >> 1) * 8 + ((HCV) & 1)) * 128, p.g16_plane);
#define TT_GRAD_HALF(G, CV, CCV, SBUF)                                                                         \
  {                                                                                                            \
    float dh[8];                                                                                               \
    tmem_ld8(cur + 64 * (CV) + 16 * sub + 8 * (CCV), dh);                                                      \
    tmem_wait_ld();                                                                                            \
    uint4 o_dr, o_dz, o_dnr, o_dn;                                                                             \
    grad_rec8(G, dh, o_dr, o_dz, o_dnr, o_dn);                                                                 \
    const uint32_t off = sw128_off(row, 2 * sub + (CCV));                                                      \
    *reinterpret_cast<uint4*>(SBUF + off) = o_dr;                                                              \
    *reinterpret_cast<uint4*>(SBUF + 16384 + off) = o_dz;                                                      \
    *reinterpret_cast<uint4*>(SBUF + 32768 + off) = o_dnr;                                                     \
    *reinterpret_cast<uint4*>(SBUF + 49152 + off) = o_dn;                                                      \
  }
        TT_LOAD_HALF(g0, 0)
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(init_done);
        if (e == 0) BWD_MARK(it, 2);
        // ---------------- pass 2: gate gradients, half-chunk (8 units) at a time, loads one half-chunk ahead.
        // NOT unrolled over chunks: the loop body must stay inside the instruction cache (the unrolled version spent
        // ~6 us per timestep refetching 65 KB of straight-line code)
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int sb = (use + c) & 1;
          uint8_t* sbuf = stg + sb * R_STG;
          TT_LOAD_HALF(g1, 2 * c + 1)
          if (use + c >= 2) {                                    // staging buffer free: MMAs and the TMA store of its last use are done
            mbar_wait(&sempty[sb], (((use + c) >> 1) - 1) & 1);  // (every thread waits itself: the phase cannot advance twice before
            mbar_wait(&sstored[sb], (((use + c) >> 1) - 1) & 1); //  this warp has arrived on sfull for the use in between)
          }
          if (e == 0) BWD_MARK(it, 3 + 3 * c);
          TT_GRAD_HALF(g0, c, 0, sbuf)
          if (c + 1 < NCH) { TT_LOAD_HALF(g0, 2 * c + 2) }
          TT_GRAD_HALF(g1, c, 1, sbuf)
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sfull[sb]);
          if (e == 0) BWD_MARK(it, 4 + 3 * c);
        }
#undef TT_LOAD_HALF
#undef TT_GRAD_HALF
        use += NCH;
      }
      // ---- hand dL/dh to the next time segment: after the last step it lives in buf[it & 1] (that step's buf[nxt])
      if (p.dhs_out && p.step_begin > p.step_end) {
        mbar_wait(acc_full, (it - 1) & 1);
        tc_fence_after();
        float4* dst = reinterpret_cast<float4*>(p.dhs_out) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float v[16];
          tmem_ld16v(lane_base + (it & 1) * 256 + 64 * c + 16 * sub, v);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[(int64_t)(16 * c + 4 * sub + i) * 128] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == R_WARP_MMA) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------ fused BPTT, CTA pairs
// H = 256 only.  The (row tile, direction) work item of gru_bwd_seq_kernel is split over a 2-CTA cluster by hidden
// units: CTA r owns units [128 r, 128 r + 128) - their dL/dh accumulators (2 x 128 TMEM columns), their gate
// gradients (2 chunks of 64 units per step instead of 4) and their 128 output columns of dh_prev = dG W_hh.  That
// product reduces over ALL 768 gate units, so every staged chunk of gate gradients (dr | dz | dn*r tiles, 48 KB) is
// also copied into the peer's landing slot with one shared::cta -> shared::cluster bulk copy that completes on the
// peer's mbarrier; the peer's MMA thread consumes it like a TMA load and reports the slot free again through a
// multicast tcgen05.commit.  Halves the per-SM gate-gradient work (the bound of the single-CTA kernel: instruction
// issue / latency of 32 Ki gate units per timestep on one SM) and doubles the SMs in use.
constexpr int P_WSTAGE = 64 * 128 * 2;             // one k-block of W_hh for this CTA's outputs: [64 k][128 n] fp16 = 16 KB
constexpr int P_WSTAGES = 3;
constexpr int P_LAND = 3 * 16384;                  // landing slot: the peer's dr | dz | dn*r tiles of one chunk
constexpr int P_SMEM = P_WSTAGES * P_WSTAGE + 2 * R_STG + P_LAND + 1024 + 512;

__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// shared::cta -> peer shared memory, completion (bytes) on the PEER's mbarrier
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}
// arrive on the mbarrier at the same offset in every CTA of `mask` once this thread's earlier tcgen05.mma have completed
__device__ __forceinline__ void mma_commit_multicast(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(mask)
               : "memory");
}

template <bool HAS_DAB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(R_THREADS, 1) gru_bwd_pair_kernel(const __grid_constant__ BwdSeqParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wring = smem;
  uint8_t* stg = smem + P_WSTAGES * P_WSTAGE;
  uint8_t* land = stg + 2 * R_STG;
  uint64_t* bars = reinterpret_cast<uint64_t*>(land + P_LAND);
  uint64_t* wfull = bars;                 // [3]
  uint64_t* wempty = bars + 3;            // [3]
  uint64_t* sfull = bars + 6;             // [2] staging written (cell threads -> MMA, store warp)
  uint64_t* sempty = bars + 8;            // [2] staging consumed by the local MMAs
  uint64_t* init_done = bars + 10;        // buf[nxt] holds dht * z
  uint64_t* acc_full = bars + 11;         // all MMAs of the step complete
  uint64_t* sstored = bars + 12;          // [2] TMA store AND the peer are done with the staging buffer
  uint64_t* pfull = bars + 14;            // landing slot holds the peer's chunk (tx bytes from the peer's bulk copy)
  uint64_t* pfree = bars + 15;            // the peer's MMAs have consumed the chunk this CTA sent (multicast commit)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int H = 256, h4 = H >> 2, NCL = 2;          // NCL: chunks of 64 units this CTA owns
  const int T = p.T, MT = p.Bp >> 7;
  const uint32_t rank = cluster_ctarank(), peer = rank ^ 1u;
  const int cg0 = 2 * (int)rank;                         // first global chunk (of 4) of this CTA
  const int first_tile = blockIdx.x >> 1, tile_stride = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P_WSTAGES; ++i) { mbar_init(&wfull[i], 1); mbar_init(&wempty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sfull[i], 1); mbar_init(&sempty[i], 1); mbar_init(&sstored[i], 1); }
    mbar_init(init_done, 1);
    mbar_init(acc_full, 1);
    mbar_init(pfull, 1);
    mbar_init(pfree, 1);
    fence_barrier_init();
  }
  if (warp == R_WARP_MMA) tmem_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                     // both CTAs' barriers are initialised before any remote copy / commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 16) {
    setmaxnreg_dec<32>();
  if (warp == R_WARP_TMA) {
    // ================================================================= W_hh producer: k-blocks in MMA order
    // (own chunk 0, peer chunk 0, own chunk 1, peer chunk 1) x (3 gates); only this CTA's 128 output columns
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int tile = first_tile; tile < p.tiles; tile += tile_stride) {
        const int dir = tile / MT;
        for (int step = p.step_end; step < p.step_begin; ++step)
          for (int src = 0; src < 4; ++src) {
            const int cg = (src & 1) ? 2 * (int)peer + (src >> 1) : cg0 + (src >> 1);
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wempty[stage], phase ^ 1);
              uint8_t* sb = wring + stage * P_WSTAGE;
              mbar_arrive_expect_tx(&wfull[stage], P_WSTAGE);
#pragma unroll
              for (int hh = 0; hh < 2; ++hh) tma_load_2d(&p.w[dir], &wfull[stage], sb + hh * 8192, (cg0 + hh) * 64, g * H + cg * 64);
              if (++stage == P_WSTAGES) { stage = 0; phase ^= 1; }
            }
          }
      }
    }
  } else if (warp == R_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16(128, 128, 0, 1);
      int stage = 0, phase = 0, it = 0, use = 0, pu = 0;
      mbar_arrive_expect_tx(pfull, P_LAND);                                // arm the landing slot for the first chunk
      for (int tile = first_tile; tile < p.tiles; tile += tile_stride) {
        for (int step = p.step_end; step < p.step_begin; ++step, ++it) {
          const uint32_t d_tmem = tmem_base + ((it & 1) ^ 1) * 128;        // buf[nxt]
          mbar_wait(init_done, it & 1);
          tc_fence_after();
          for (int src = 0; src < 4; ++src) {
            uint32_t sa0;
            int sb = 0;
            if (src & 1) {
              mbar_wait(pfull, pu & 1);
              mbar_arrive_expect_tx(pfull, P_LAND);                        // arm the next phase (its bytes cannot arrive before our commit)
              sa0 = smem_u32(land);
            } else {
              sb = use & 1;
              mbar_wait(&sfull[sb], (use >> 1) & 1);
              sa0 = smem_u32(stg + sb * R_STG);
            }
            tc_fence_after();
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wfull[stage], phase);
              tc_fence_after();
              const uint32_t sa = sa0 + g * 16384, sw = smem_u32(wring + stage * P_WSTAGE);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_bf16_ss(d_tmem, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sw + k * 2048, 8192, 1024), idesc, 1u);
              mma_commit(&wempty[stage]);
              if (++stage == P_WSTAGES) { stage = 0; phase ^= 1; }
            }
            if (src & 1) { mma_commit_multicast(pfree, (uint16_t)(1u << peer)); ++pu; }
            else { mma_commit(&sempty[sb]); ++use; }
          }
          mma_commit(acc_full);
        }
      }
    }
  } else if (warp == R_WARP_STORE) {
    // ================================================================= dG plane writer + sender of the staged chunks
    if (lane == 0) {
      int use = 0;
      const uint32_t peer_land = mapa_u32(smem_u32(land), peer), peer_pfull = mapa_u32(smem_u32(pfull), peer);
      for (int tile = first_tile; tile < p.tiles; tile += tile_stride) {
        const int mt = tile % MT, dir = tile / MT, m0 = mt * 128;
        for (int step = p.step_begin - 1; step >= p.step_end; --step) {
          const int t = dir == 0 ? step : T - 1 - step;
          if (step > p.step_end) {
            // next step's saved-gate records of THIS CTA's units (16 contiguous 8-unit groups = 32 KB per plane) into L2
            const int tn = dir == 0 ? step - 1 : T - step;
            const int64_t w0 = ((((int64_t)dir * T + tn) * MT + mt) * (H >> 3) + 16 * (int)rank) * 128;
#pragma unroll 1
            for (int k = 0; k < GP_PLANES; ++k)
              prefetch_l2_bulk(reinterpret_cast<const char*>(p.g16 + k * p.g16_plane + w0), 32768);
            if (HAS_DAB) {
              const char* dp = reinterpret_cast<const char*>(p.d_above) +
                               (((int64_t)tn * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2) + 32 * (int)rank) * 128 * 16;
              prefetch_l2_bulk(dp, 65536);
            }
          }
          for (int c = 0; c < NCL; ++c, ++use) {
            const int sb = use & 1;
            mbar_wait(&sfull[sb], (use >> 1) & 1);
            const uint8_t* sbuf = stg + sb * R_STG;
#pragma unroll
            for (int g = 0; g < 4; ++g) tma_store_2d(&p.dg[dir], sbuf + g * 16384, g * H + 64 * (cg0 + c), t * p.Bp + m0);
            bulk_commit();
            if (use > 0) {
              // the peer has consumed the previous chunk: its landing slot is free again and the previous staging buffer
              // has been read completely by the copy engine (its TMA store was drained when it was issued)
              mbar_wait(pfree, (use - 1) & 1);
              mbar_arrive(&sstored[sb ^ 1]);
            }
            bulk_copy_to_peer(peer_land, smem_u32(sbuf), P_LAND, peer_pfull);
            bulk_wait_read0();
          }
        }
      }
      if (use > 0) mbar_wait(pfree, (use - 1) & 1);           // the peer is done with everything this CTA sent
    }
  }
  } else {
    setmaxnreg_inc<112>();
    // ================================================================= gate gradients: thread = (batch row, 16 units per chunk)
    const int e = threadIdx.x;
    const int quarter = warp & 3, sub = warp >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int it = 0, use = 0;
    for (int tile = first_tile; tile < p.tiles; tile += tile_stride) {
      const int mt = tile % MT, dir = tile / MT, m0 = mt * 128, b = m0 + row;
      const bool row_ok = b < p.B;
      if (it > 0) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
      {
        const float4* src = reinterpret_cast<const float4*>(p.dhs0) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCL; ++c) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = __ldg(src + (int64_t)(16 * (cg0 + c) + 4 * sub + i) * 128);
            v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
          }
          tmem_st16(lane_base + (it & 1) * 128 + 64 * c + 16 * sub, v);
        }
        tmem_wait_st();
      }
      for (int step = p.step_begin - 1; step >= p.step_end; --step, ++it) {
        const int t = dir == 0 ? step : T - 1 - step;
        const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
        const uint32_t cur = lane_base + (it & 1) * 128, nxt = lane_base + ((it & 1) ^ 1) * 128;
        // word index (plane Z) of this thread's first 8-unit group; LOCAL chunk c, half hf is at rec0 + (8 c + hf) * 128
        const int64_t rec0 = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + 8 * cg0 + 2 * sub) * 128 + row;
        const float4* dab0 = HAS_DAB ? reinterpret_cast<const float4*>(p.d_above) +
                                           (((int64_t)t * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2) + 16 * cg0 + 4 * sub) * 128 + row
                                     : nullptr;
        uint4 zw[NCL * 2];
#pragma unroll
        for (int g = 0; g < NCL * 2; ++g)
          zw[g] = __ldg(p.g16 + GP_Z * p.g16_plane + rec0 + (int64_t)((g >> 1) * 8 + (g & 1)) * 128);
        float4 da[4];
        if (HAS_DAB) {
#pragma unroll
          for (int i = 0; i < 4; ++i) da[i] = live ? __ldg(dab0 + (int64_t)i * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (step != p.step_begin - 1) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
        // ---------------- pass 1: dht = dh (+ d_above); buf[cur] <- dht; buf[nxt] <- dht * z
#pragma unroll
        for (int c = 0; c < NCL; ++c) {
          float dh[16];
          tmem_ld16v(cur + 64 * c + 16 * sub, dh);
          if (HAS_DAB) {
            float4* d4 = da;
            if (p.drop_p > 0.f && live) {
              const uint64_t el = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + 64 * (cg0 + c) + 16 * sub);
              float s8[8];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), el >> 3, p.drop_p, p.inv_keep, s8);
              d4[0].x *= s8[0]; d4[0].y *= s8[1]; d4[0].z *= s8[2]; d4[0].w *= s8[3];
              d4[1].x *= s8[4]; d4[1].y *= s8[5]; d4[1].z *= s8[6]; d4[1].w *= s8[7];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), (el >> 3) + 1, p.drop_p, p.inv_keep, s8);
              d4[2].x *= s8[0]; d4[2].y *= s8[1]; d4[2].z *= s8[2]; d4[2].w *= s8[3];
              d4[3].x *= s8[4]; d4[3].y *= s8[5]; d4[3].z *= s8[6]; d4[3].w *= s8[7];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) { dh[4 * i] += d4[i].x; dh[4 * i + 1] += d4[i].y; dh[4 * i + 2] += d4[i].z; dh[4 * i + 3] += d4[i].w; }
            if (c + 1 < NCL) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                da[i] = live ? __ldg(dab0 + (int64_t)(16 * (c + 1) + i) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            tmem_st16(cur + 64 * c + 16 * sub, dh);
          }
          mul_z8(dh, zw[2 * c]);
          mul_z8(dh + 8, zw[2 * c + 1]);
          tmem_st16(nxt + 64 * c + 16 * sub, dh);
        }
        GateRec8 g0, g1;
#define TT_LOAD_HALF(G, HCV) load_rec8(G, p.g16 + rec0 + (int64_t)(((HCV) >> 1) * 8 + ((HCV) & 1)) * 128, p.g16_plane);
#define TT_GRAD_HALF(G, CV, CCV, SBUF)                                                                         \
  {                                                                                                            \
    float dh[8];                                                                                               \
    tmem_ld8(cur + 64 * (CV) + 16 * sub + 8 * (CCV), dh);                                                      \
    tmem_wait_ld();                                                                                            \
    uint4 o_dr, o_dz, o_dnr, o_dn;                                                                             \
    grad_rec8(G, dh, o_dr, o_dz, o_dnr, o_dn);                                                                 \
    const uint32_t off = sw128_off(row, 2 * sub + (CCV));                                                      \
    *reinterpret_cast<uint4*>(SBUF + off) = o_dr;                                                              \
    *reinterpret_cast<uint4*>(SBUF + 16384 + off) = o_dz;                                                      \
    *reinterpret_cast<uint4*>(SBUF + 32768 + off) = o_dnr;                                                     \
    *reinterpret_cast<uint4*>(SBUF + 49152 + off) = o_dn;                                                      \
  }
        TT_LOAD_HALF(g0, 0)
        tmem_wait_st();
        tc_fence_before();
        named_bar_sync(1, 512);
        if (e == 0) mbar_arrive(init_done);
        // ---------------- pass 2: gate gradients of this CTA's two chunks
#pragma unroll 1
        for (int c = 0; c < NCL; ++c) {
          const int sb = (use + c) & 1;
          uint8_t* sbuf = stg + sb * R_STG;
          TT_LOAD_HALF(g1, 2 * c + 1)
          if (e == 0 && use + c >= 2) {                          // staging buffer free: local MMAs, TMA store and the peer are done with it
            mbar_wait(&sempty[sb], (((use + c) >> 1) - 1) & 1);
            mbar_wait(&sstored[sb], (((use + c) >> 1) - 1) & 1);
          }
          named_bar_sync(2, 512);
          TT_GRAD_HALF(g0, c, 0, sbuf)
          if (c + 1 < NCL) { TT_LOAD_HALF(g0, 2 * c + 2) }
          TT_GRAD_HALF(g1, c, 1, sbuf)
          fence_proxy_async();
          named_bar_sync(3, 512);
          if (e == 0) mbar_arrive(&sfull[sb]);
        }
#undef TT_LOAD_HALF
#undef TT_GRAD_HALF
        use += NCL;
      }
      if (p.dhs_out && p.step_begin > p.step_end) {
        mbar_wait(acc_full, (it - 1) & 1);
        tc_fence_after();
        float4* dst = reinterpret_cast<float4*>(p.dhs_out) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCL; ++c) {
          float v[16];
          tmem_ld16v(lane_base + (it & 1) * 128 + 64 * c + 16 * sub, v);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[(int64_t)(16 * (cg0 + c) + 4 * sub + i) * 128] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                     // nobody leaves while the peer may still copy into / read from this CTA's shared memory
  if (warp == R_WARP_MMA) tmem_dealloc(tmem_base, 256);
}

// ------------------------------------------------------------------------------------------ fused BPTT, tcgen05 CTA pairs
// H = 256.  Two row tiles of the SAME direction run as a 2-CTA cluster whose recurrent product dh_prev += dG W_hh is ONE
// tcgen05.mma.cta_group::2 per k-step: M = 256 (each CTA's 128 rows, accumulators in each CTA's own tensor memory),
// N = 256 with the W_hh k-block split by output column over the two CTAs' shared memory.  Each CTA therefore streams
// HALF of W_hh per timestep (196 KB instead of 393 KB) through a ring of twice as many stages: the single-CTA kernel's
// timestep was bound by that stream (latency x bytes in flight: 96 KB per ~1.2 us against 393 KB per step; in-kernel
// marks: the MMAs of a step trailed the gate-gradient threads by 3.3 us).  Everything else - gate gradients, dG stores,
// tensor-memory-resident dh - is the single-CTA kernel's, per CTA; the pair only shares the MMA issue (leader CTA) and
// its barriers: the peer's warps arrive on the leader's `init_done` / `sfull` (mbarrier.arrive.release.cluster), and
// every tcgen05.commit is multicast to both CTAs.  Same SM count as the single-CTA kernel (one CTA per (row tile, dir)).
constexpr int D_WSTAGE = 64 * 128 * 2;             // this CTA's half of a k-block: [64 k][128 n] fp16 = 16 KB
constexpr int D_WSTAGES = 6;
static_assert(D_WSTAGES * D_WSTAGE == R_WSTAGES * R_WSTAGE, "same shared-memory budget as the single-CTA kernel");

template <bool HAS_DAB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(R_THREADS, 1) gru_bwd_duo_kernel(const __grid_constant__ BwdSeqParams p) {
  constexpr int NCH = 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wring = smem;
  uint8_t* stg = smem + D_WSTAGES * D_WSTAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + 2 * R_STG);
  uint64_t* wfull = bars;                 // [6] LEADER's: both CTAs' halves of a W k-block have landed (tx bytes of both)
  uint64_t* wempty = bars + 6;            // [6] per CTA, arrived by the leader's multicast commit
  uint64_t* sfull = bars + 12;            // [2] LEADER's: both CTAs' staging written (32 warp arrivals) -> MMA
  uint64_t* sfull_l = bars + 14;          // [2] this CTA's staging written (16 warps) -> its dG store warp
  uint64_t* sempty = bars + 16;           // [2] staging consumed (multicast commit -> both CTAs' cell threads)
  uint64_t* init_done = bars + 18;        // LEADER's: buf[nxt] holds dht * z in both CTAs (32 warp arrivals)
  uint64_t* acc_full = bars + 19;         // all MMAs of the step complete (multicast commit)
  uint64_t* sstored = bars + 20;          // [2] TMA store has read the staging buffer (store warp -> cell threads)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 22);
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int H = 64 * NCH, h4 = H >> 2;
  const int T = p.T, MT = p.Bp >> 7;

  if (threadIdx.x == 0) {
    for (int i = 0; i < D_WSTAGES; ++i) { mbar_init(&wfull[i], 1); mbar_init(&wempty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sfull[i], 32); mbar_init(&sfull_l[i], 16); mbar_init(&sempty[i], 1); mbar_init(&sstored[i], 1); }
    mbar_init(init_done, 32);
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == R_WARP_MMA) tmem_alloc_pair(tmem_slot, 512);      // the same warp of BOTH CTAs
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                     // both CTAs' barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // register budget: the 16 gate-gradient warps (warpgroups 0-3) take what the control warpgroup gives up
  if (warp >= 16) {
    setmaxnreg_dec<32>();   // 512 * 112 + 128 * 32 == 640 * 96 registers: exactly what the launch allocated
  if (warp == R_WARP_TMA) {
    // ================================================================= W_hh producer: THIS CTA's half (128 of the 256 output
    // columns) of every k-block; the bytes of both halves are counted on the leader's barrier
    if (lane == 0) {
      int stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        const int dir = tile / MT;
        for (int step = p.step_end; step < p.step_begin; ++step)
          for (int c = 0; c < NCH; ++c)
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wempty[stage], phase ^ 1);
              uint8_t* sb = wring + stage * D_WSTAGE;
              if (leader) mbar_arrive_expect_tx(&wfull[stage], 2 * D_WSTAGE);
#pragma unroll
              for (int hh = 0; hh < 2; ++hh)
                tma_load_2d_pair(&p.w[dir], &wfull[stage], sb + hh * 8192, (2 * (int)rank + hh) * 64, g * H + c * 64);
              if (++stage == D_WSTAGES) { stage = 0; phase ^= 1; }
            }
      }
    }
  } else if (warp == R_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0 && leader) {           // one thread of the pair issues the M = 256 MMAs of both row tiles
      constexpr uint32_t idesc = make_idesc_f16(256, H, 0, 1);
      int stage = 0, phase = 0, it = 0, use = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        for (int step = p.step_end; step < p.step_begin; ++step, ++it) {
          const uint32_t d_tmem = tmem_base + ((it & 1) ^ 1) * 256;        // buf[nxt]
          mbar_wait_cluster(init_done, it & 1);
          tc_fence_after();
          for (int c = 0; c < NCH; ++c, ++use) {
            const int sb = use & 1;
            mbar_wait_cluster(&sfull[sb], (use >> 1) & 1);
            tc_fence_after();
            for (int g = 0; g < 3; ++g) {
              mbar_wait(&wfull[stage], phase);
              tc_fence_after();
              const uint32_t sa = smem_u32(stg + sb * R_STG + g * 16384), sw = smem_u32(wring + stage * D_WSTAGE);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                mma_f16_ss_pair(d_tmem, make_smem_desc(sa + k * 32, 16, 1024), make_smem_desc(sw + k * 2048, 8192, 1024), idesc, 1u);
              mma_commit_pair(&wempty[stage], 3);
              if (++stage == D_WSTAGES) { stage = 0; phase ^= 1; }
            }
            mma_commit_pair(&sempty[sb], 3);
          }
          mma_commit_pair(acc_full, 3);
          BWD_MARK(it, 15);
        }
      }
    }
  } else if (warp == R_WARP_STORE) {
    // ================================================================= dG plane writer (off the cell threads' critical path)
    if (lane == 0) {
      int use = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        const int mt = tile % MT, dir = tile / MT, m0 = mt * 128;
        for (int step = p.step_begin - 1; step >= p.step_end; --step) {
          const int t = dir == 0 ? step : T - 1 - step;
          if (step > p.step_end) {
            // pull the NEXT step's saved-gate records (contiguous per (dir, t, row tile)) into L2 one whole step ahead:
            // the cell threads' loads then see L2 latency instead of HBM latency
            const int tn = dir == 0 ? step - 1 : T - step;
            // each plane's words of (dir, tn, mt) are contiguous: H/8 groups x 128 rows x 16 B = NCH x 16 KB
            const int64_t w0 = (((int64_t)dir * T + tn) * MT + mt) * (H >> 3) * 128;
#pragma unroll 1
            for (int k = 0; k < GP_PLANES; ++k)
              prefetch_l2_bulk(reinterpret_cast<const char*>(p.g16 + k * p.g16_plane + w0), NCH * 16384);
            if (HAS_DAB) {
              const char* dp = reinterpret_cast<const char*>(p.d_above) +
                               (((int64_t)tn * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2)) * 128 * 16;
#pragma unroll 1
              for (int i = 0; i < (NCH + 1) / 2; ++i) prefetch_l2_bulk(dp + i * 65536, NCH == 1 ? 32768 : 65536);
            }
          }
          for (int c = 0; c < NCH; ++c, ++use) {
            const int sb = use & 1;
            mbar_wait(&sfull_l[sb], (use >> 1) & 1);
            const uint8_t* sbuf = stg + sb * R_STG;
#pragma unroll
            for (int g = 0; g < 4; ++g) tma_store_2d(&p.dg[dir], sbuf + g * 16384, g * H + 64 * c, t * p.Bp + m0);
            bulk_commit();
            bulk_wait_read0();
            mbar_arrive(&sstored[sb]);
          }
        }
      }
    }
  }
  } else {
    setmaxnreg_inc<112>();
    // ================================================================= gate gradients: thread = (batch row, 16 units per chunk)
    const int e = threadIdx.x;
    const int quarter = warp & 3, sub = warp >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
    int it = 0, use = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      const int mt = tile % MT, dir = tile / MT, m0 = mt * 128, b = m0 + row;
      const bool row_ok = b < p.B;
      // ---- dL/dh at the end of the sequence -> buf[cur of this tile's first step]
      if (it > 0) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
      {
        const float4* src = reinterpret_cast<const float4*>(p.dhs0) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = __ldg(src + (int64_t)(16 * c + 4 * sub + i) * 128);
            v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
          }
          tmem_st16(lane_base + (it & 1) * 256 + 64 * c + 16 * sub, v);
        }
        tmem_wait_st();
      }
      for (int step = p.step_begin - 1; step >= p.step_end; --step, ++it) {
        if (e == 0) BWD_MARK(it, 14);
        const int t = dir == 0 ? step : T - 1 - step;
        const bool live = row_ok && !(p.lengths && t >= p.lengths[row_ok ? b : 0]);
        const uint32_t cur = lane_base + (it & 1) * 256, nxt = lane_base + ((it & 1) ^ 1) * 256;
        // word index (plane Z) of this thread's first 8-unit group; group (chunk c, half hf) is at rec0 + (8 c + hf) * 128
        const int64_t rec0 = ((((int64_t)dir * T + t) * MT + mt) * (H >> 3) + 2 * sub) * 128 + row;
        const float4* dab0 = HAS_DAB ? reinterpret_cast<const float4*>(p.d_above) +
                                           (((int64_t)t * MT + mt) * (p.ndir * h4) + ((dir * H) >> 2) + 4 * sub) * 128 + row
                                     : nullptr;
        // ---------------- all z words of the step: in flight while the previous step's MMAs finish
        uint4 zw[NCH * 2];                        // (padded steps / pad rows hold z = 1: the gradient passes through)
#pragma unroll
        for (int g = 0; g < NCH * 2; ++g)
          zw[g] = __ldg(p.g16 + GP_Z * p.g16_plane + rec0 + (int64_t)((g >> 1) * 8 + (g & 1)) * 128);
        float4 da[4];
        if (HAS_DAB) {
#pragma unroll
          for (int i = 0; i < 4; ++i) da[i] = live ? __ldg(dab0 + (int64_t)i * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (e == 0) BWD_MARK(it, 0);
        if (step != p.step_begin - 1) { mbar_wait(acc_full, (it - 1) & 1); tc_fence_after(); }
        if (e == 0) BWD_MARK(it, 1);
        // ---------------- pass 1: dht = dh (+ d_above); buf[cur] <- dht; buf[nxt] <- dht * z
        // (tcgen05.ld / .st are warp-collective: every lane executes them; only the VALUES depend on `live`)
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float dh[16];
          tmem_ld16v(cur + 64 * c + 16 * sub, dh);
          if (HAS_DAB) {
            float4* d4 = da;
            if (p.drop_p > 0.f && live) {
              const uint64_t el = ((uint64_t)b * T + t) * (uint64_t)(p.ndir * H) + (uint64_t)(dir * H + 64 * c + 16 * sub);
              float s8[8];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), el >> 3, p.drop_p, p.inv_keep, s8);
              d4[0].x *= s8[0]; d4[0].y *= s8[1]; d4[0].z *= s8[2]; d4[0].w *= s8[3];
              d4[1].x *= s8[4]; d4[1].y *= s8[5]; d4[1].z *= s8[6]; d4[1].w *= s8[7];
              dropout_scale8(p.seed + (p.seed_dev ? __ldg(p.seed_dev) : 0ull), (el >> 3) + 1, p.drop_p, p.inv_keep, s8);
              d4[2].x *= s8[0]; d4[2].y *= s8[1]; d4[2].z *= s8[2]; d4[2].w *= s8[3];
              d4[3].x *= s8[4]; d4[3].y *= s8[5]; d4[3].z *= s8[6]; d4[3].w *= s8[7];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) { dh[4 * i] += d4[i].x; dh[4 * i + 1] += d4[i].y; dh[4 * i + 2] += d4[i].z; dh[4 * i + 3] += d4[i].w; }
            if (c + 1 < NCH) {                                   // next chunk's d_above (L2-resident: prefetched a step ahead)
#pragma unroll
              for (int i = 0; i < 4; ++i)
                da[i] = live ? __ldg(dab0 + (int64_t)(16 * (c + 1) + i) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            tmem_st16(cur + 64 * c + 16 * sub, dh);
          }
          mul_z8(dh, zw[2 * c]);
          mul_z8(dh + 8, zw[2 * c + 1]);
          tmem_st16(nxt + 64 * c + 16 * sub, dh);
        }
        // first half-chunk of gate records for pass 2 (in flight across the barrier)
        GateRec8 g0, g1;
#define TT_LOAD_HALF(G, HCV) load_rec8(G, p.g16 + rec0 + (int64_t)(((HCV) >> 1) * 8 + ((HCV) & 1)) * 128, p.g16_plane);
#define TT_GRAD_HALF(G, CV, CCV, SBUF)                                                                         \
  {                                                                                                            \
    float dh[8];                                                                                               \
    tmem_ld8(cur + 64 * (CV) + 16 * sub + 8 * (CCV), dh);                                                      \
    tmem_wait_ld();                                                                                            \
    uint4 o_dr, o_dz, o_dnr, o_dn;                                                                             \
    grad_rec8(G, dh, o_dr, o_dz, o_dnr, o_dn);                                                                 \
    const uint32_t off = sw128_off(row, 2 * sub + (CCV));                                                      \
    *reinterpret_cast<uint4*>(SBUF + off) = o_dr;                                                              \
    *reinterpret_cast<uint4*>(SBUF + 16384 + off) = o_dz;                                                      \
    *reinterpret_cast<uint4*>(SBUF + 32768 + off) = o_dnr;                                                     \
    *reinterpret_cast<uint4*>(SBUF + 49152 + off) = o_dn;                                                      \
  }
        TT_LOAD_HALF(g0, 0)
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(init_done);
        if (e == 0) BWD_MARK(it, 2);
        // ---------------- pass 2: gate gradients, half-chunk (8 units) at a time, loads one half-chunk ahead.
        // NOT unrolled over chunks: the loop body must stay inside the instruction cache (the unrolled version spent
        // ~6 us per timestep refetching 65 KB of straight-line code)
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int sb = (use + c) & 1;
          uint8_t* sbuf = stg + sb * R_STG;
          TT_LOAD_HALF(g1, 2 * c + 1)
          if (use + c >= 2) {                                    // staging buffer free: MMAs and the TMA store of its last use are done
            mbar_wait(&sempty[sb], (((use + c) >> 1) - 1) & 1);  // (every thread waits itself: the phase cannot advance twice before
            mbar_wait(&sstored[sb], (((use + c) >> 1) - 1) & 1); //  this warp has arrived on sfull for the use in between)
          }
          if (e == 0) BWD_MARK(it, 3 + 3 * c);
          TT_GRAD_HALF(g0, c, 0, sbuf)
          if (c + 1 < NCH) { TT_LOAD_HALF(g0, 2 * c + 2) }
          TT_GRAD_HALF(g1, c, 1, sbuf)
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) { mbar_arrive(&sfull_l[sb]); mbar_arrive_leader(&sfull[sb]); }
          if (e == 0) BWD_MARK(it, 4 + 3 * c);
        }
#undef TT_LOAD_HALF
#undef TT_GRAD_HALF
        use += NCH;
      }
      // ---- hand dL/dh to the next time segment: after the last step it lives in buf[it & 1] (that step's buf[nxt])
      if (p.dhs_out && p.step_begin > p.step_end) {
        mbar_wait(acc_full, (it - 1) & 1);
        tc_fence_after();
        float4* dst = reinterpret_cast<float4*>(p.dhs_out) + (((int64_t)dir * MT + mt) * h4) * 128 + row;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          float v[16];
          tmem_ld16v(lane_base + (it & 1) * 256 + 64 * c + 16 * sub, v);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[(int64_t)(16 * c + 4 * sub + i) * 128] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                     // nobody leaves (or frees tensor memory) while the pair's MMAs / commits may still touch it
  if (warp == R_WARP_MMA) tmem_dealloc_pair(tmem_base, 512);
}


// ------------------------------------------------------------------------------------------ backward helpers
// loss scale: s = 2^k with max|g| * s ~ 4  (fp16 planes of the gate gradients then keep >= 11 bits for everything
// within 2^-16 of the largest incoming gradient and have 2^14 of headroom).  scale[0] = s, scale[1] = 1/s.
__global__ void absmax_kernel(const float* __restrict__ x, int64_t n, unsigned int* __restrict__ out) {
  float m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = fabsf(x[i]);
    if (v < 3.0e38f) m = fmaxf(m, v);                 // ignore inf / nan
  }
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));
}
__global__ void finish_scale_kernel(const unsigned int* __restrict__ mx, float* __restrict__ scale) {
  const float m = __uint_as_float(*mx);
  float s = 1.f;
  if (m > 0.f) {
    int e;
    frexpf(m, &e);                                     // m = f * 2^e, f in [0.5, 1)
    int k = 3 - e;                                     // m * 2^k in [4, 8)
    k = max(-100, min(100, k));
    s = ldexpf(1.f, k);
  }
  scale[0] = s;
  scale[1] = 1.f / s;
}
// d_h_n rows of one layer [ndir][B][H] fp32 -> tiled running state (scaled); null source -> zeros
__global__ void init_dhs_kernel(const float* __restrict__ d_h_n, const float* __restrict__ scale, int B, int Bp, int H, int ndir,
                                float* __restrict__ dhs) {
  const int64_t total = (int64_t)ndir * Bp * (H >> 2);
  const float s = scale[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i & 127);
    const int u4 = (int)((i >> 7) % (H >> 2));
    const int mt = (int)((i >> 7) / (H >> 2) % (Bp >> 7));
    const int d = (int)((i >> 7) / (H >> 2) / (Bp >> 7));
    const int b = mt * 128 + row;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (d_h_n && b < B) {
      const float4 g = *reinterpret_cast<const float4*>(d_h_n + ((int64_t)d * B + b) * H + 4 * u4);
      v = make_float4(g.x * s, g.y * s, g.z * s, g.w * s);
    }
    reinterpret_cast<float4*>(dhs)[i] = v;
  }
}
// API d_out [B,T,C] fp32 -> tiled [T][MT][C/4][128][4] (scaled)
__global__ void dout_tiled_kernel(const float* __restrict__ d_out, const float* __restrict__ scale, int B, int Bp, int T, int C,
                                  float* __restrict__ out) {
  const int64_t total = (int64_t)T * Bp * (C >> 2);
  const float s = scale[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i & 127);
    const int c4 = (int)((i >> 7) % (C >> 2));
    const int64_t tm = (i >> 7) / (C >> 2);
    const int mt = (int)(tm % (Bp >> 7)), t = (int)(tm / (Bp >> 7));
    const int b = mt * 128 + row;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b < B) {
      const float4 g = *reinterpret_cast<const float4*>(d_out + ((int64_t)b * T + t) * C + 4 * c4);
      v = make_float4(g.x * s, g.y * s, g.z * s, g.w * s);
    }
    reinterpret_cast<float4*>(out)[i] = v;
  }
}
constexpr int kColsumBlocks = 592;
// column sums of the dG planes of one direction -> partial sums cs[block * rpi + phase][4H] (fp32, still scaled)
__global__ void __launch_bounds__(256) colsum_dg_kernel(const f16* __restrict__ dg, int64_t rows, int C, float* __restrict__ cs) {
  // thread = (8 consecutive columns, row phase): 16-byte loads, 4 rows in flight per thread
  const int cg = C >> 3;
  for (int g0 = 0; g0 < cg; g0 += 256) {
    const int lanes = min(cg - g0, 256), rpi = 256 / lanes;
    const int g = g0 + (int)threadIdx.x % lanes, rr = (int)threadIdx.x / lanes;
    if (rr >= rpi) continue;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    const uint4* src = reinterpret_cast<const uint4*>(dg) + g;
    const int64_t stride = (int64_t)gridDim.x * rpi;
    int64_t r = (int64_t)blockIdx.x * rpi + rr;
    for (; r + 3 * stride < rows; r += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(src + (r + u * stride) * cg);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
          acc[2 * i] += x.x; acc[2 * i + 1] += x.y;
        }
      }
    }
    for (; r < rows; r += stride) {
      const uint4 v = __ldg(src + r * cg);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        acc[2 * i] += x.x; acc[2 * i + 1] += x.y;
      }
    }
    // per-(block, row phase) partial sums, reduced by bias_grads_kernel (same-address atomics from every block
    // serialise in L2 and cost as much as the streaming read itself)
    float* out = cs + ((int64_t)blockIdx.x * rpi + rr) * C + 8 * g;
    *reinterpret_cast<float4*>(out) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(out + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}
// one block per 32 columns of the dG plane: sums the `parts` partial rows, then scatters to the two bias gradients
__global__ void __launch_bounds__(256) bias_grads_kernel(const float* __restrict__ cs, int pitch, int parts, const float* __restrict__ scale, int H,
                                                         float* __restrict__ d_b_ih, float* __restrict__ d_b_hh) {
  __shared__ float red[8][33];
  const int col = blockIdx.x * 32 + (threadIdx.x & 31), ph = threadIdx.x >> 5;   // cs already points at this direction's columns
  float v = 0.f;
  for (int pi = ph; pi < parts; pi += 8) v += cs[(int64_t)pi * pitch + col];
  red[ph][threadIdx.x & 31] = v;
  __syncthreads();
  if (ph == 0) {
#pragma unroll
    for (int i = 1; i < 8; ++i) v += red[i][threadIdx.x & 31];
    v *= scale[1];
    const int gate = col / H, u = col - gate * H;         // plane columns: dr | dz | dn*r | dn
    if (gate < 3) d_b_hh[col] = v;                        // (dr, dz, dn*r)
    if (gate < 2) d_b_ih[col] = v;                        // (dr, dz, dn)
    else if (gate == 3) d_b_ih[2 * H + u] = v;
  }
}

static inline unsigned gblocks(int64_t n, int per = 256, int cap = 148 * 16) {
  int64_t b = ceil_div(n, per);
  return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
}

// ------------------------------------------------------------------------------------------ buffers
struct V2Dims {
  int B, Bp, MT, T, In, H, L, nd;
  int64_t ntok;                       // T * Bp
  int in_l(int l) const { return l == 0 ? In : nd * H; }
  int kp_l(int l) const { return (int)plane_pitch(in_l(l)); }
};
static V2Dims dims_of(const tt_gru_stack* a) {
  V2Dims d;
  d.B = a->B; d.Bp = (a->B + 127) / 128 * 128; d.MT = d.Bp / 128; d.T = a->T; d.In = a->In; d.H = a->H; d.L = a->L; d.nd = a->ndir;
  d.ntok = (int64_t)d.T * d.Bp;
  return d;
}
// saved for backward, per layer: X hi plane, H hi planes (per direction, T+1 slots), 16-bit gates, hn
struct V2Saved {
  f16* xhi[8]; f16* hhi[8][2]; uint4* g16[8]; float* hn[8];
};
static uint64_t carve_saved(const V2Dims& d, void* base, uint64_t size, V2Saved* out, bool with_gates) {
  Workspace w(base, size);
  for (int l = 0; l < d.L; ++l) {
    f16* x = w.take<f16>(d.ntok * d.kp_l(l));
    f16* h0 = w.take<f16>((int64_t)(d.T + 1) * d.Bp * d.H);
    f16* h1 = d.nd > 1 ? w.take<f16>((int64_t)(d.T + 1) * d.Bp * d.H) : nullptr;
    uint4* g = nullptr; float* hn = nullptr;
    if (with_gates) {
      g = w.take<uint4>((int64_t)d.nd * d.ntok * (d.H >> 3) * GP_PLANES);       // 10 H bytes per (token, direction)
    }
    if (out) { out->xhi[l] = x; out->hhi[l][0] = h0; out->hhi[l][1] = h1; out->g16[l] = g; out->hn[l] = hn; }
  }
  return w.used + 256;
}

bool v2_supported(const tt_gru_stack* a) {
  return engine_mode() == 2 && a->H % 64 == 0 && a->H <= 1024 && a->L >= 1 && a->L <= 8 && (a->ndir == 1 || a->ndir == 2) &&
         umma_available();
}

uint64_t gru_stack_saved_bytes_v2(const tt_gru_stack* a) {
  return carve_saved(dims_of(a), nullptr, 0, nullptr, true);
}

struct V2FwdWs {
  f16* xlo[2]; f16* hlo[2]; float* gi; float* hs; float* bias; int* ready; f16* wih_hi[2]; f16* wih_lo[2]; f16* whh_hi[2]; f16* whh_lo[2];
  V2Saved sv;                          // only used when the caller gave no saved buffer
};
static uint64_t carve_fwd_ws(const V2Dims& d, void* base, uint64_t size, V2FwdWs* out, bool own_saved) {
  Workspace w(base, size);
  int kpmax = 0;
  for (int l = 0; l < d.L; ++l) kpmax = d.kp_l(l) > kpmax ? d.kp_l(l) : kpmax;
  V2FwdWs f{};
  for (int i = 0; i < 2; ++i) f.xlo[i] = w.take<f16>(d.ntok * kpmax);
  for (int i = 0; i < d.nd; ++i) f.hlo[i] = w.take<f16>((int64_t)(d.T + 1) * d.Bp * d.H);
  f.gi = w.take<float>((int64_t)d.nd * d.ntok * 3 * d.H);
  f.hs = w.take<float>((int64_t)d.nd * d.Bp * d.H);
  f.bias = w.take<float>((int64_t)d.nd * 3 * d.H);
  f.ready = w.take<int>((int64_t)d.nd * d.MT + 4);
  for (int i = 0; i < d.nd; ++i) {
    f.wih_hi[i] = w.take<f16>((int64_t)3 * d.H * kpmax); f.wih_lo[i] = w.take<f16>((int64_t)3 * d.H * kpmax);
    f.whh_hi[i] = w.take<f16>((int64_t)3 * d.H * d.H); f.whh_lo[i] = w.take<f16>((int64_t)3 * d.H * d.H);
  }
  uint64_t used = w.used;
  if (own_saved) {
    const uint64_t off = align_up(used, 256);
    used = off + carve_saved(d, base ? (char*)base + off : nullptr, size > off ? size - off : 0, out ? &f.sv : nullptr, false);
  }
  if (out) *out = f;
  return used + 256;
}
uint64_t gru_stack_fwd_workspace_v2(const tt_gru_stack* a) {
  return carve_fwd_ws(dims_of(a), nullptr, 0, nullptr, !(a->save_for_backward && a->saved));
}

static const float* wptr(const tt_gru_stack* a, int l, int d, int k) { return a->weights[(l * a->ndir + d) * 4 + k]; }

// ------------------------------------------------------------------------------------------ forward
int gru_stack_fwd_v2(const tt_gru_stack* a, cudaStream_t s) {
  const V2Dims d = dims_of(a);
  const int B = d.B, Bp = d.Bp, T = d.T, H = d.H, nd = d.nd, L = d.L, MT = d.MT;
  const bool save = a->save_for_backward && a->saved;
  V2FwdWs f;
  const uint64_t need = carve_fwd_ws(d, a->workspace, a->workspace_bytes, &f, !save);
  if (need > a->workspace_bytes || !a->workspace) { set_error("gru stack fwd (v2): workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)need); return TT_ERR_WORKSPACE; }
  V2Saved sv = f.sv;
  if (save) {
    const uint64_t sneed = carve_saved(d, a->saved, a->saved_bytes, &sv, true);
    if (sneed > a->saved_bytes) { set_error("gru stack fwd (v2): saved buffer too small (%llu < %llu)", (unsigned long long)a->saved_bytes, (unsigned long long)sneed); return TT_ERR_WORKSPACE; }
  }
  static DeviceOnce attr_once;
  if (attr_once.need()) {
    TT_CUDA(cudaFuncSetAttribute(gru_step2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, S_SMEM));
    attr_once.done();
  }
  // layer-0 operand planes
  {
    const int kp = d.kp_l(0);
    note_launch();
    if (a->ids) {
      if (a->ids_are_i64)
        gather_planes_kernel<int64_t><<<gblocks(d.ntok, 8), 256, 0, s>>>((const int64_t*)a->ids, B, Bp, T, a->table, a->vocab, d.In, kp, a->pad_id, sv.xhi[0], f.xlo[0]);
      else
        gather_planes_kernel<int32_t><<<gblocks(d.ntok, 8), 256, 0, s>>>((const int32_t*)a->ids, B, Bp, T, a->table, a->vocab, d.In, kp, a->pad_id, sv.xhi[0], f.xlo[0]);
    } else {
      x_planes_kernel<<<gblocks(d.ntok, 8), 256, 0, s>>>(a->x, B, Bp, T, d.In, kp, sv.xhi[0], f.xlo[0]);
    }
    TT_LAUNCH_CHECK();
  }
  for (int l = 0; l < L; ++l) {
    const int in = d.in_l(l), kp = d.kp_l(l);
    const bool has_next = l + 1 < L;
    f16* xlo = f.xlo[l & 1];
    for (int dd = 0; dd < nd; ++dd) {
      note_launch(); w_planes_kernel<<<gblocks((int64_t)3 * H * kp), 256, 0, s>>>(wptr(a, l, dd, 0), 3 * H, in, kp, f.wih_hi[dd], f.wih_lo[dd]);
      note_launch(); w_planes_kernel<<<gblocks((int64_t)3 * H * H), 256, 0, s>>>(wptr(a, l, dd, 1), 3 * H, H, H, f.whh_hi[dd], f.whh_lo[dd]);
      note_launch(); fold_bias_kernel<<<(unsigned)ceil_div(3 * H, 256), 256, 0, s>>>(wptr(a, l, dd, 2), wptr(a, l, dd, 3), H, f.bias + (int64_t)dd * 3 * H);
    }
    TT_LAUNCH_CHECK();
    // K-B: gi (tiled) = X W_ih^T + folded bias, all timesteps, both directions, 3 passes
    {
      UmmaGemm g{};
      for (int dd = 0; dd < nd; ++dd) {
        g.A[dd] = Planes{(const bf16*)sv.xhi[l], (const bf16*)xlo, kp};
        g.B[dd] = Planes{(const bf16*)f.wih_hi[dd], (const bf16*)f.wih_lo[dd], kp};
        g.C[dd] = f.gi + (int64_t)dd * d.ntok * 3 * H;
        g.bias[dd] = f.bias + (int64_t)dd * 3 * H;
      }
      g.M = (int)d.ntok; g.N = 3 * H; g.K = in; g.nbatch = nd; g.alpha = 1.f; g.beta = 0.f; g.splitk = 1;
      g.f16 = 1; g.npass = 3; g.epi = 1; g.tl_Bp = Bp; g.tl_G = 3; g.tl_Hc = H;
      TT_TRY(umma_gemm(g, s));
    }
    // zero initial state: the h_{-1} slot of each direction's planes
    for (int dd = 0; dd < nd; ++dd) {
      const int64_t slot = dd == 0 ? 0 : T;
      TT_CUDA(cudaMemsetAsync(sv.hhi[l][dd] + slot * Bp * H, 0, sizeof(f16) * Bp * H, s));
      TT_CUDA(cudaMemsetAsync(f.hlo[dd] + slot * Bp * H, 0, sizeof(f16) * Bp * H, s));
    }
    Step2Params p;
    memset(&p, 0, sizeof(p));
    for (int dd = 0; dd < nd; ++dd) {
      TT_TRY(make_plane_map(&p.h_hi[dd], sv.hhi[l][dd], (int64_t)(T + 1) * Bp, H, H, S_BM, 1));
      TT_TRY(make_plane_map(&p.h_lo[dd], f.hlo[dd], (int64_t)(T + 1) * Bp, H, H, S_BM, 1));
      TT_TRY(make_plane_map(&p.w_hi[dd], f.whh_hi[dd], 3 * H, H, H, S_UN, 1));
      TT_TRY(make_plane_map(&p.w_lo[dd], f.whh_lo[dd], 3 * H, H, H, S_UN, 1));
      p.b_hh[dd] = wptr(a, l, dd, 3);
    }
    if (has_next) {
      TT_TRY(make_plane_map(&p.x_hi, sv.xhi[l + 1], d.ntok, nd * H, d.kp_l(l + 1), S_BM, 1));
      TT_TRY(make_plane_map(&p.x_lo, f.xlo[(l + 1) & 1], d.ntok, nd * H, d.kp_l(l + 1), S_BM, 1));
    }
    p.gi = f.gi;
    p.g16 = save ? sv.g16[l] : nullptr; p.hn = save ? sv.hn[l] : nullptr;
    p.g16_plane = (int64_t)nd * T * MT * (H >> 3) * 128;
    p.out = (a->out && !has_next) ? a->out : nullptr;
    p.h_n = a->h_n + (int64_t)l * nd * B * H;
    p.lengths = a->lengths;
    p.B = B; p.Bp = Bp; p.T = T; p.H = H; p.ndir = nd; p.has_next = has_next ? 1 : 0;
    p.drop_p = has_next ? a->dropout_p : 0.f;
    p.inv_keep = p.drop_p > 0.f ? 1.f / (1.f - p.drop_p) : 1.f;
    p.seed = a->seed + (uint64_t)(l + 1) * 0x9E3779B97F4A7C15ull;     // per-layer stream (layer l+1's input mask)
    p.seed_dev = a->seed_dev;
    p.dbg = g_step_dbg;
    const int tiles = (H / S_UN) * MT * nd;
    if (seq_mode() != 0) {
      // persistent cooperative launch: all T steps, one CTA per SM at most
      const int num_sms = device_sm_count();
      TT_CHECK_ARG(device_cooperative() == 1, "gru stack fwd (v2): device lacks cooperative launch");
      const int tpc = (int)ceil_div(tiles, num_sms);
      const int ncta = (int)ceil_div(tiles, tpc);
      SeqParams sp;
      memset(&sp, 0, sizeof(sp));
      sp.s = p; sp.ready = f.ready; sp.tiles = tiles;
      TT_CUDA(cudaMemsetAsync(f.ready, 0, sizeof(int) * nd * MT, s));
      void* args[] = {&sp};
      note_launch();
      ProfScope prof(PROF_GRU_SEQ_FWD, s);
      if (seq_mode() == 2) {
        static DeviceOnce a2;
        if (a2.need()) { TT_CUDA(cudaFuncSetAttribute(gru_seq_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Q_SMEM)); a2.done(); }
        TT_CUDA(cudaLaunchCooperativeKernel((const void*)gru_seq_kernel<2>, dim3(ncta), dim3(256 + 128), args, Q_SMEM, s));
      } else {
        static DeviceOnce a4;
        if (a4.need()) { TT_CUDA(cudaFuncSetAttribute(gru_seq_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, Q_SMEM)); a4.done(); }
        TT_CUDA(cudaLaunchCooperativeKernel((const void*)gru_seq_kernel<4>, dim3(ncta), dim3(512 + 128), args, Q_SMEM, s));
      }
    } else {
      dim3 grid((unsigned)(H / S_UN), (unsigned)MT, (unsigned)nd);
      for (int step = 0; step < T; ++step) {
        p.step = step;
        note_launch();
        gru_step2_kernel<<<grid, S_THREADS, S_SMEM, s>>>(p);
      }
    }
    TT_LAUNCH_CHECK();
  }
  return TT_OK;
}

// ------------------------------------------------------------------------------------------ backward
struct V2BwdWs {
  f16* dgbuf[2]; float* csbuf[2];   // [layer & 1]: the weight-gradient GEMMs of layer l overlap the BPTT of layer l-1
  f16* dg[2]; float* dhs; float* dab[2]; float* cs; float* scale; unsigned int* mx; f16* whh_hi[2]; f16* wih_hi[2];
};
static uint64_t carve_bwd_ws(const V2Dims& d, void* base, uint64_t size, V2BwdWs* out) {
  Workspace w(base, size);
  V2BwdWs f{};
  int kpmax = 0;
  for (int l = 0; l < d.L; ++l) kpmax = d.kp_l(l) > kpmax ? d.kp_l(l) : kpmax;
  // dG: ONE plane [T*Bp][ndir*4H] per buffer, direction dd = columns [4H*dd, 4H*dd + 4H)
  f.dgbuf[0] = w.take<f16>(d.ntok * d.nd * 4 * d.H);
  f.dgbuf[1] = d.L > 1 ? w.take<f16>(d.ntok * d.nd * 4 * d.H) : f.dgbuf[0];
  f.dhs = w.take<float>((int64_t)d.nd * d.Bp * d.H);
  for (int i = 0; i < 2; ++i) f.dab[i] = w.take<float>(d.ntok * d.nd * d.H);
  {
    const int cg = d.nd * d.H / 2, rpi = cg >= 256 ? 1 : 256 / cg;      // as colsum_dg_kernel splits its 256 threads
    f.csbuf[0] = w.take<float>((int64_t)kColsumBlocks * rpi * d.nd * 4 * d.H);   // per-(block, row phase) partial column sums
    f.csbuf[1] = d.L > 1 ? w.take<float>((int64_t)kColsumBlocks * rpi * d.nd * 4 * d.H) : f.csbuf[0];
  }
  f.scale = w.take<float>(4);
  f.mx = w.take<unsigned int>(4);
  for (int i = 0; i < d.nd; ++i) f.whh_hi[i] = w.take<f16>((int64_t)3 * d.H * d.H);
  f.wih_hi[0] = w.take<f16>((int64_t)d.nd * 3 * d.H * kpmax);   // both directions stacked along the gate rows: [ndir*3H][kp]
  f.wih_hi[1] = nullptr;
  if (out) *out = f;
  return w.used + 256;
}
uint64_t gru_stack_bwd_workspace_v2(const tt_gru_stack_grad* a) { return carve_bwd_ws(dims_of(&a->fwd), nullptr, 0, nullptr); }

// Helper stream per caller stream: the weight-gradient GEMMs + bias sums of layer l run here, concurrently with the
// BPTT kernel of layer l-1 on the caller's stream (that kernel occupies one CTA per (row tile, direction), the GEMMs
// get the remaining SMs).  Forked / joined with events, so to the caller all work stays ordered on its stream.
struct SideStream { int dev; cudaStream_t caller, side; cudaEvent_t fork, done[2]; };
static SideStream* side_for(cudaStream_t caller) {
  // keyed by (device, caller stream): stream handle 0 is the same value on every device.  Entries are never freed (a
  // process uses a handful of streams); past 64 of them the caller simply gets no overlap.
  static SideStream table[64];
  static int n = 0;
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < n; ++i) if (table[i].caller == caller && table[i].dev == dev) return &table[i];
  if (n == 64) return nullptr;
  SideStream& t = table[n];
  if (cudaStreamCreateWithFlags(&t.side, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
  if (cudaEventCreateWithFlags(&t.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
  for (int i = 0; i < 2; ++i) if (cudaEventCreateWithFlags(&t.done[i], cudaEventDisableTiming) != cudaSuccess) return nullptr;
  t.caller = caller; t.dev = dev;
  return &table[n++];
}

int gru_stack_bwd_v2(const tt_gru_stack_grad* a, cudaStream_t s) {
  const tt_gru_stack* fw = &a->fwd;
  const V2Dims d = dims_of(fw);
  const int B = d.B, Bp = d.Bp, T = d.T, H = d.H, nd = d.nd, L = d.L, MT = d.MT;
  TT_CHECK_ARG(fw->saved != nullptr, "gru stack bwd (v2): forward state missing");
  V2Saved sv;
  const uint64_t sneed = carve_saved(d, fw->saved, fw->saved_bytes, &sv, true);
  if (sneed > fw->saved_bytes) { set_error("gru stack bwd (v2): saved buffer too small"); return TT_ERR_WORKSPACE; }
  V2BwdWs f;
  const uint64_t need = carve_bwd_ws(d, a->workspace, a->workspace_bytes, &f);
  if (need > a->workspace_bytes || !a->workspace) { set_error("gru stack bwd (v2): workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)need); return TT_ERR_WORKSPACE; }
  static DeviceOnce attr_once;
  if (attr_once.need()) {
    TT_CUDA(cudaFuncSetAttribute(cell_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CB_SMEM));
    attr_once.done();
  }
  // loss scale from the incoming gradients
  TT_CUDA(cudaMemsetAsync(f.mx, 0, 16, s));
  if (a->d_h_n) { note_launch(); absmax_kernel<<<gblocks((int64_t)L * nd * B * H), 256, 0, s>>>(a->d_h_n, (int64_t)L * nd * B * H, f.mx); }
  if (a->d_out) { note_launch(); absmax_kernel<<<gblocks((int64_t)B * T * nd * H), 256, 0, s>>>(a->d_out, (int64_t)B * T * nd * H, f.mx); }
  note_launch(); finish_scale_kernel<<<1, 1, 0, s>>>(f.mx, f.scale);
  TT_LAUNCH_CHECK();

  const float* d_above = nullptr;                       // tiled gradient wrt this layer's output sequence
  if (a->d_out) {
    note_launch();
    dout_tiled_kernel<<<gblocks(d.ntok * (nd * H / 4)), 256, 0, s>>>(a->d_out, f.scale, B, Bp, T, nd * H, f.dab[(L - 1) & 1]);
    TT_LAUNCH_CHECK();
    d_above = f.dab[(L - 1) & 1];
  }
  const int num_sms_b = device_sm_count();
  // CTA pairs halve the latency of a timestep but cost more SM-time (5.2 ms on 128 SMs vs 6.6 ms on 64 at B = 4096, where the
  // idle SMs are busy with the weight-gradient GEMMs and the other tower anyway): used when the pairs fit in half the GPU
  // (measured: 12.6 vs 13.8 ms/step at B = 2048, 9.4 vs 11.4 at B = 1024); mode 3 forces them.
  const bool pair_bptt = H == 256 && (bwd_seq_mode() == 3 || (bwd_seq_mode() == 2 && 4 * MT * nd <= num_sms_b));
  // tcgen05 CTA pairs: two row tiles of one direction per cluster (MT even), every cluster resident (MT * nd CTAs <= SMs)
  // Opt-in (mode 4): measured at B = 4096 it halves the wait for the previous step's MMAs (2.5 -> 0.8 us per timestep) but the gate
  // threads slow down by as much once the MMAs really overlap them (shared-memory bandwidth: 12.5 vs 11.8 us per timestep,
  // profiles/r2_summary.md), so the independent one-CTA kernel stays the default.
  const bool duo_bptt = !pair_bptt && H == 256 && (MT % 2) == 0 && MT * nd <= (num_sms_b & ~1) && bwd_seq_mode() == 4;
  const int bptt_want = MT * nd * (pair_bptt ? 2 : 1);
  const int bptt_ctas = bptt_want < num_sms_b ? bptt_want : num_sms_b;
  const int free_ctas = num_sms_b - bptt_ctas;           // SMs the BPTT kernel of the next layer leaves idle
  const bool fused_bptt = bwd_seq_mode() != 0 && (H == 64 || H == 128 || H == 256);
  // Time segments: the persistent BPTT kernel runs the sequence as `nseg` launches (dL/dh handed over through f.dhs);
  // the weight-gradient GEMMs over a finished segment's tokens run on the helper stream under the next segment's
  // launch, so they leave the critical path for EVERY layer (not only for layers that have another layer below them).
  int nseg = (g_wgrad_overlap && fused_bptt && free_ctas >= 32) ? g_bwd_segments : 1;
  if (nseg > T / 16) nseg = T / 16;
  if (nseg < 1) nseg = 1;
  SideStream* side = (g_wgrad_overlap && (L > 1 || nseg > 1) && free_ctas >= 32 && fused_bptt) ? side_for(s) : nullptr;
  if (!side) nseg = 1;
  bool pending[2] = {false, false};                      // side work still reading dgbuf[i] / csbuf[i]
  for (int l = L - 1; l >= 0; --l) {
    const int in = d.in_l(l), kp = d.kp_l(l);
    const int pb = l & 1;
    f.dg[0] = f.dgbuf[pb]; f.dg[1] = f.dgbuf[pb] + 4 * H; f.cs = f.csbuf[pb];
    if (side && pending[pb]) { TT_CUDA(cudaStreamWaitEvent(s, side->done[pb], 0)); pending[pb] = false; }
    const bool has_next = l + 1 < L;
    for (int dd = 0; dd < nd; ++dd) {
      note_launch(); w_planes_kernel<<<gblocks((int64_t)3 * H * H), 256, 0, s>>>(wptr(fw, l, dd, 1), 3 * H, H, H, f.whh_hi[dd], nullptr);
      note_launch(); w_planes_kernel<<<gblocks((int64_t)3 * H * kp), 256, 0, s>>>(wptr(fw, l, dd, 0), 3 * H, in, kp, f.wih_hi[0] + (int64_t)dd * 3 * H * kp, nullptr);
    }
    note_launch();
    init_dhs_kernel<<<gblocks((int64_t)nd * Bp * (H / 4)), 256, 0, s>>>(a->d_h_n ? a->d_h_n + (int64_t)l * nd * B * H : nullptr, f.scale, B, Bp, H, nd, f.dhs);
    TT_LAUNCH_CHECK();

    CellBwdParams cp;
    memset(&cp, 0, sizeof(cp));
    const int dgp_pitch = nd * 4 * H;
    for (int dd = 0; dd < nd; ++dd) TT_TRY(make_plane_map(&cp.dg[dd], f.dg[dd], d.ntok, 4 * H, dgp_pitch, 128, 1));
    // weight gradients over the token rows [row0[dd], row0[dd] + rows) of direction dd (time-major rows: a time segment):
    //   dW_hh[3H,H] (+)= dGH^T h_prev   (h_prev(t) = H-plane slot t (fwd) / t+1 (rev))
    //   dW_ih[3H,In] (+)= dGI^T X       (dGI = columns [0,2H) and [3H,4H) of the direction's dG columns)
    auto issue_wgrads = [&](const int64_t* row0, int64_t rows, float beta, cudaStream_t ws, int cap) -> int {
      for (int dd = 0; dd < nd; ++dd) {
        const Planes dgp{(const bf16*)(f.dg[dd] + row0[dd] * dgp_pitch), nullptr, dgp_pitch};
        {
          UmmaGemm g{};
          g.A[0] = dgp;
          g.B[0] = Planes{(const bf16*)(sv.hhi[l][dd] + (dd == 0 ? 0 : (int64_t)Bp * H) + row0[dd] * H), nullptr, H};
          g.C[0] = a->d_weights[(l * nd + dd) * 4 + 1]; g.ldc = H; g.M = 3 * H; g.N = H; g.K = (int)rows; g.nbatch = 1;
          g.a_mn = g.b_mn = true; g.alpha = 1.f; g.alpha_ptr = f.scale + 1; g.beta = beta; g.splitk = 0; g.f16 = 1; g.npass = 1;
          g.max_ctas = cap;
          TT_TRY(umma_gemm(g, ws));
        }
        {
          UmmaGemm g{};
          g.A[0] = dgp; g.a_seg = 2 * H; g.a_skip = H;
          g.B[0] = Planes{(const bf16*)(sv.xhi[l] + row0[dd] * kp), nullptr, kp};
          g.C[0] = a->d_weights[(l * nd + dd) * 4 + 0]; g.ldc = in; g.M = 3 * H; g.N = in; g.K = (int)rows; g.nbatch = 1;
          g.a_mn = g.b_mn = true; g.alpha = 1.f; g.alpha_ptr = f.scale + 1; g.beta = beta; g.splitk = 0; g.f16 = 1; g.npass = 1;
          g.max_ctas = cap;
          TT_TRY(umma_gemm(g, ws));
        }
      }
      return TT_OK;
    };
    cp.g16 = sv.g16[l]; cp.g16_plane = (int64_t)nd * T * MT * (H >> 3) * 128; cp.hn = sv.hn[l]; cp.dhs = f.dhs; cp.d_above = d_above; cp.lengths = fw->lengths;
    cp.B = B; cp.Bp = Bp; cp.T = T; cp.H = H; cp.ndir = nd;
    cp.drop_p = (has_next && d_above) ? fw->dropout_p : 0.f;
    cp.inv_keep = cp.drop_p > 0.f ? 1.f / (1.f - cp.drop_p) : 1.f;
    cp.seed = fw->seed + (uint64_t)(l + 1) * 0x9E3779B97F4A7C15ull;
    cp.seed_dev = fw->seed_dev;
    // per step: dhs (tiled) += dGH_t W_hh     (A rows t*Bp.., K = 3H; B = W_hh as stored = [K][N] MN-major)
    UmmaGemm r{};
    for (int dd = 0; dd < nd; ++dd) {
      r.A[dd] = Planes{(const bf16*)f.dg[dd], nullptr, dgp_pitch};
      r.B[dd] = Planes{(const bf16*)f.whh_hi[dd], nullptr, H};
      r.C[dd] = f.dhs + (int64_t)dd * Bp * H;
    }
    r.M = Bp; r.N = H; r.K = 3 * H; r.nbatch = nd; r.a_mn = false; r.b_mn = true; r.alpha = 1.f; r.beta = 1.f; r.splitk = 1;
    r.f16 = 1; r.npass = 1; r.a_rows = d.ntok; r.epi = 1; r.tl_Bp = Bp; r.tl_G = 1; r.tl_Hc = H;
    if ((H == 64 || H == 128 || H == 256) && bwd_seq_mode() != 0) {
      // fused persistent BPTT: dL/dh stays in tensor memory for the whole sequence
      const int num_sms = num_sms_b;
      BwdSeqParams bp;
      memset(&bp, 0, sizeof(bp));
      for (int dd = 0; dd < nd; ++dd) {
        bp.dg[dd] = cp.dg[dd];
        TT_TRY(make_plane_map(&bp.w[dd], f.whh_hi[dd], 3 * H, H, H, 64, 1));
      }
      bp.g16 = sv.g16[l]; bp.g16_plane = (int64_t)nd * T * MT * (H >> 3) * 128; bp.hn = sv.hn[l]; bp.dhs0 = f.dhs; bp.d_above = d_above; bp.lengths = fw->lengths;
      bp.B = B; bp.Bp = Bp; bp.T = T; bp.H = H; bp.ndir = nd; bp.tiles = MT * nd;
      bp.drop_p = cp.drop_p; bp.inv_keep = cp.inv_keep; bp.seed = cp.seed; bp.seed_dev = cp.seed_dev; bp.dbg = g_step_dbg;
      const int grid = bp.tiles < num_sms ? bp.tiles : num_sms;
      const int pair_grid = 2 * bp.tiles <= num_sms ? 2 * bp.tiles : (num_sms & ~1);
#define TT_BWD_PAIR(DABV)                                                                                             \
  do {                                                                                                                \
    static DeviceOnce attr;                                                                                           \
    if (attr.need()) {                                                                                                \
      TT_CUDA(cudaFuncSetAttribute(gru_bwd_pair_kernel<DABV>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));  \
      attr.done();                                                                                                    \
    }                                                                                                                 \
    gru_bwd_pair_kernel<DABV><<<pair_grid, R_THREADS, P_SMEM, s>>>(bp);                                               \
  } while (0)
#define TT_BWD_DUO(DABV)                                                                                              \
  do {                                                                                                                \
    static DeviceOnce attr;                                                                                           \
    if (attr.need()) {                                                                                                \
      TT_CUDA(cudaFuncSetAttribute(gru_bwd_duo_kernel<DABV>, cudaFuncAttributeMaxDynamicSharedMemorySize, R_SMEM));   \
      attr.done();                                                                                                    \
    }                                                                                                                 \
    gru_bwd_duo_kernel<DABV><<<bp.tiles, R_THREADS, R_SMEM, s>>>(bp);                                                 \
  } while (0)
#define TT_BWD_SEQ(NCHV, DABV)                                                                                        \
  do {                                                                                                                \
    static DeviceOnce attr;                                                                                           \
    if (attr.need()) {                                                                                                \
      TT_CUDA(cudaFuncSetAttribute(gru_bwd_seq_kernel<NCHV, DABV>, cudaFuncAttributeMaxDynamicSharedMemorySize, R_SMEM)); \
      attr.done();                                                                                                    \
    }                                                                                                                 \
    gru_bwd_seq_kernel<NCHV, DABV><<<grid, R_THREADS, R_SMEM, s>>>(bp);                                               \
  } while (0)
      const bool dab = d_above != nullptr;
      for (int seg = 0; seg < nseg; ++seg) {
        bp.step_begin = T - (int)((int64_t)T * seg / nseg);
        bp.step_end = T - (int)((int64_t)T * (seg + 1) / nseg);
        bp.dhs_out = nseg > 1 && seg + 1 < nseg ? f.dhs : nullptr;
        {
          note_launch();
          ProfScope prof(PROF_GRU_SEQ_BWD, s);
          if (pair_bptt) { if (dab) TT_BWD_PAIR(true); else TT_BWD_PAIR(false); }
          else if (duo_bptt) { if (dab) TT_BWD_DUO(true); else TT_BWD_DUO(false); }
          else if (H == 256) { if (dab) TT_BWD_SEQ(4, true); else TT_BWD_SEQ(4, false); }
          else if (H == 128) { if (dab) TT_BWD_SEQ(2, true); else TT_BWD_SEQ(2, false); }
          else { if (dab) TT_BWD_SEQ(1, true); else TT_BWD_SEQ(1, false); }
          TT_LAUNCH_CHECK();
        }
        if (nseg > 1) {
          // this segment's tokens: forward direction t in [step_end, step_begin), reverse direction t in [T-step_begin, T-step_end)
          TT_CUDA(cudaEventRecord(side->fork, s));
          TT_CUDA(cudaStreamWaitEvent(side->side, side->fork, 0));
          const int64_t row0[2] = {(int64_t)bp.step_end * Bp, (int64_t)(T - bp.step_begin) * Bp};
          TT_TRY(issue_wgrads(row0, (int64_t)(bp.step_begin - bp.step_end) * Bp, seg == 0 ? 0.f : 1.f, side->side, free_ctas));
        }
      }
#undef TT_BWD_SEQ
#undef TT_BWD_DUO
#undef TT_BWD_PAIR
    } else {
    UmmaPrepared prep;
    TT_TRY(umma_gemm_prepare(r, &prep));
    dim3 cgrid((unsigned)(H / 64), (unsigned)MT, (unsigned)nd);
    for (int step = T - 1; step >= 0; --step) {
      cp.step = step;
      note_launch();
      cell_bwd2_kernel<<<cgrid, 512, CB_SMEM, s>>>(cp);
      if (step == 0) break;
      const int row0[2] = {step * Bp, (T - 1 - step) * Bp};
      TT_TRY(umma_gemm_launch(prep, row0, s));
    }
    TT_LAUNCH_CHECK();
    }

    float* d_below = nullptr;                            // tiled dX for the layer below
    // dX = sum over directions of dGI W_ih, as ONE GEMM with the directions concatenated along the reduction:
    // A = plane columns {dir0: dr dz dn, dir1: dr dz dn} (two skipped dn*r blocks), B = W_ih rows stacked [ndir*3H][kp].
    // Tiled for the layer below, or the API's batch-first fp32 d_x.  First, because the next layer's BPTT waits for it.
    if (l > 0 || a->d_x) {
      UmmaGemm g{};
      g.A[0] = Planes{(const bf16*)f.dg[0], nullptr, dgp_pitch}; g.a_seg = 2 * H; g.a_skip = H; g.a_seg2 = nd == 2 ? 5 * H : 0;
      g.a_rows = d.ntok;
      g.B[0] = Planes{(const bf16*)f.wih_hi[0], nullptr, kp};
      g.M = (int)d.ntok; g.N = in; g.K = nd * 3 * H; g.nbatch = 1; g.a_mn = false; g.b_mn = true;
      g.alpha = 1.f; g.beta = 0.f; g.splitk = 1; g.f16 = 1; g.npass = 1;
      if (l > 0) {
        d_below = f.dab[(l - 1) & 1];
        g.C[0] = d_below; g.epi = 1; g.tl_Bp = Bp; g.tl_G = 1; g.tl_Hc = in;
      } else {
        g.C[0] = a->d_x; g.ldc = in; g.alpha_ptr = f.scale + 1; g.rm_Bp = Bp; g.rm_B = B; g.rm_T = T;
      }
      TT_TRY(umma_gemm(g, s));
    }
    // weight and bias gradients of this layer: off the critical path.  With a layer below still to do they go to the
    // helper stream (capped to the SMs its BPTT kernel leaves idle), otherwise they stay on the caller's stream.
    const bool segmented = nseg > 1;                       // weight gradients already queued per segment on the helper stream
    const bool overlap = side != nullptr && (l > 0 || segmented);
    cudaStream_t ws = s;
    if (overlap) {
      TT_CUDA(cudaEventRecord(side->fork, s));
      TT_CUDA(cudaStreamWaitEvent(side->side, side->fork, 0));
      ws = side->side;
    }
    const int cap = overlap ? free_ctas : 0;
    {
      // bias gradients: column sums of the whole dG plane (both directions in one pass)
      const int C = dgp_pitch, cgs = C >> 3, rpi = cgs >= 256 ? 1 : 256 / cgs;
      const int cblocks = overlap ? (free_ctas * 4 < kColsumBlocks ? free_ctas * 4 : kColsumBlocks) : kColsumBlocks;
      note_launch(); colsum_dg_kernel<<<cblocks, 256, 0, ws>>>(f.dg[0], d.ntok, C, f.cs);
      for (int dd = 0; dd < nd; ++dd) {
        note_launch();
        bias_grads_kernel<<<(unsigned)(4 * H / 32), 256, 0, ws>>>(f.cs + dd * 4 * H, C, cblocks * rpi, f.scale, H,
                                                               a->d_weights[(l * nd + dd) * 4 + 2], a->d_weights[(l * nd + dd) * 4 + 3]);
      }
      TT_LAUNCH_CHECK();
    }
    if (!segmented) {
      const int64_t row0[2] = {0, 0};
      TT_TRY(issue_wgrads(row0, d.ntok, 0.f, ws, cap));
    }
    if (overlap) { TT_CUDA(cudaEventRecord(side->done[pb], side->side)); pending[pb] = true; }
    d_above = d_below;
  }
  for (int i = 0; i < 2; ++i)
    if (side && pending[i]) TT_CUDA(cudaStreamWaitEvent(s, side->done[i], 0));   // join: the caller's stream sees every gradient
  return TT_OK;
}

}  // namespace tt

extern "C" int tt_debug_wgrad_overlap(int32_t on) {
  tt::g_wgrad_overlap = on ? 1 : 0;
  return TT_OK;
}
extern "C" int tt_debug_seq_mode(int32_t mode) {
  if (mode != 0 && mode != 2 && mode != 4) { tt::set_error("tt_debug_seq_mode: 0, 2 or 4"); return TT_ERR_INVALID; }
  tt::g_seq_mode = mode;
  return TT_OK;
}
extern "C" int tt_debug_bwd_segments(int32_t n) {
  if (n < 1 || n > 16) { tt::set_error("tt_debug_bwd_segments: 1..16"); return TT_ERR_INVALID; }
  tt::g_bwd_segments = n;
  return TT_OK;
}
extern "C" int tt_debug_bwd_seq_mode(int32_t mode) {
  if (mode < 0 || mode > 4) { tt::set_error("tt_debug_bwd_seq_mode: 0..4"); return TT_ERR_INVALID; }
  tt::g_bwd_seq_mode = mode;
  return TT_OK;
}

extern "C" int tt_debug_step_buffer(void* buf) {
  tt::g_step_dbg = (unsigned long long*)buf;
  return TT_OK;
}
