// K-D projection head: Linear -> LayerNorm -> ReLU [-> Linear] [-> L2 normalise], forward and backward.
// Engine 2 (default): the Linear layers and all their gradients run on the tcgen05 GEMM engine - forward products as 3
// passes over fp16 hi/lo planes (the precision of the GRU forward), gradient products as 3 passes over bf16 hi/lo planes
// (gradients live in fp16's subnormal range; bf16 x3 carries 16 mantissa bits at fp32's exponent range).  The pooled
// state cat(h_n[-2], h_n[-1]) is formed once, as operand planes (never as an fp32 tensor).  Engine 0: the exact-fp32
// SIMT GEMM.  LayerNorm + ReLU and the normalisation are one-warp-per-row kernels.
#include "common.cuh"
#include "umma.cuh"
#include "umma_gemm.cuh"
#include "gru_v2.cuh"

namespace tt {

// y2 = relu(LN(y1)) ; stats = (mean, rstd).  One warp per row.
__global__ void ln_relu_fwd_kernel(const float* __restrict__ y1, const float* __restrict__ w, const float* __restrict__ b,
                                   float* __restrict__ y2, float* __restrict__ stats, int B, int P, float eps, float p,
                                   float inv_keep, uint64_t seed, uint64_t offset, const uint64_t* __restrict__ seed_dev) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= B) return;
  if (seed_dev) seed += *seed_dev;
  const float* x = y1 + (int64_t)row * P;
  float s = 0.f;
  for (int c = lane; c < P; c += 32) s += x[c];
  const float mean = warp_sum(s) / P;
  float v = 0.f;
  for (int c = lane; c < P; c += 32) { float d = x[c] - mean; v += d * d; }
  const float rstd = rsqrtf(warp_sum(v) / P + eps);
  for (int c = lane; c < P; c += 32) {
    float v = fmaxf((x[c] - mean) * rstd * w[c] + b[c], 0.f);
    if (p > 0.f) v *= dropout_scale(seed, offset, (uint64_t)row * P + c, p, inv_keep);
    y2[(int64_t)row * P + c] = v;
  }
  if (lane == 0) { stats[2 * row] = mean; stats[2 * row + 1] = rstd; }
}

// in: d_y2 ; out: d_y1, gm = d_y2*mask, gx = gm*xhat   (column sums of gm/gx give d_ln_b/d_ln_w)
__global__ void ln_relu_bwd_kernel(const float* __restrict__ y1, const float* __restrict__ y2,
                                   const float* __restrict__ stats, const float* __restrict__ w,
                                   const float* __restrict__ d_y2, float* __restrict__ d_y1, float* __restrict__ gm,
                                   float* __restrict__ gx, int B, int P, float inv_keep) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= B) return;
  const int64_t o = (int64_t)row * P;
  const float mean = stats[2 * row], rstd = stats[2 * row + 1];
  float s1 = 0.f, s2 = 0.f;
  for (int c = lane; c < P; c += 32) {
    const float g = y2[o + c] > 0.f ? d_y2[o + c] * inv_keep : 0.f;   // stored y2 > 0 <=> relu active AND kept
    const float xh = (y1[o + c] - mean) * rstd;
    gm[o + c] = g; gx[o + c] = g * xh;
    const float dxh = g * w[c];
    s1 += dxh; s2 += dxh * xh;
  }
  s1 = warp_sum(s1) / P; s2 = warp_sum(s2) / P;
  for (int c = lane; c < P; c += 32) {
    const float xh = (y1[o + c] - mean) * rstd;
    d_y1[o + c] = rstd * (gm[o + c] * w[c] - s1 - xh * s2);
  }
}

__global__ void l2norm_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, float* __restrict__ norm, int B,
                                  int D, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= B) return;
  float s = 0.f;
  for (int c = lane; c < D; c += 32) { float v = x[(int64_t)row * D + c]; s += v * v; }
  const float n = sqrtf(warp_sum(s));
  const float inv = 1.f / fmaxf(n, eps);
  for (int c = lane; c < D; c += 32) y[(int64_t)row * D + c] = x[(int64_t)row * D + c] * inv;
  if (lane == 0 && norm) norm[row] = n;
}
// y = x / max(|x|, eps):  |x| > eps -> dx = (dy - y <y,dy>) / |x| ;  else dx = dy / eps
__global__ void l2norm_bwd_kernel(const float* __restrict__ y, const float* __restrict__ norm,
                                  const float* __restrict__ dy, float* __restrict__ dx, int B, int D, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= B) return;
  const int64_t o = (int64_t)row * D;
  const float n = norm[row];
  if (n > eps) {
    float s = 0.f;
    for (int c = lane; c < D; c += 32) s += y[o + c] * dy[o + c];
    s = warp_sum(s);
    const float inv = 1.f / n;
    for (int c = lane; c < D; c += 32) dx[o + c] = (dy[o + c] - y[o + c] * s) * inv;
  } else {
    for (int c = lane; c < D; c += 32) dx[o + c] = dy[o + c] / eps;
  }
}

static inline GemmArgs nt_gemm(const float* A, int64_t lda, const float* Bm, int64_t ldb, float* C, int64_t ldc,
                               const float* bias, int M, int N, int K, float beta) {
  GemmArgs g{};
  g.A[0] = A; g.B[0] = Bm; g.C[0] = C; g.bias[0] = bias;
  g.a_rs = lda; g.a_cs = 1; g.b_rs = ldb; g.b_cs = 1; g.ldc = ldc;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.splitk = 1; g.alpha = 1.f; g.beta = beta;
  return g;
}
// C[M,N] = sum_r A[r,M]^T B[r,N]  (reduction over rows r < R)
static inline GemmArgs tn_gemm(const float* A, int64_t lda, const float* Bm, int64_t ldb, float* C, int64_t ldc, int M,
                               int N, int R) {
  GemmArgs g{};
  g.A[0] = A; g.B[0] = Bm; g.C[0] = C;
  g.a_rs = 1; g.a_cs = lda; g.b_rs = 1; g.b_cs = ldb; g.ldc = ldc;
  g.M = M; g.N = N; g.K = R; g.nbatch = 1; g.splitk = 0; g.alpha = 1.f; g.beta = 0.f;
  return g;
}
// C[M,N] (+)= A[M,K] B[K,N]
static inline GemmArgs nn_gemm(const float* A, int64_t lda, const float* Bm, int64_t ldb, float* C, int64_t ldc, int M,
                               int N, int K, float beta) {
  GemmArgs g{};
  g.A[0] = A; g.B[0] = Bm; g.C[0] = C;
  g.a_rs = lda; g.a_cs = 1; g.b_rs = 1; g.b_cs = ldb; g.ldc = ldc;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.splitk = 1; g.alpha = 1.f; g.beta = beta;
  return g;
}

// second problem of the same shape in the same launch
static inline void pair_with(GemmArgs& g, const float* A, const float* Bm, float* C) {
  g.A[1] = A; g.B[1] = Bm; g.C[1] = C; g.bias[1] = nullptr; g.nbatch = 2;
}

// ---- tensor-engine route
// (operand planes need 16-byte rows: every dimension a multiple of 8 - otherwise the exact SIMT GEMMs run)
static bool head_dims_tc(const tt_head* a) { return (a->In % 16) == 0 && (a->P1 % 8) == 0 && (a->D2 % 8) == 0; }
static bool head_tc(const tt_head* a) { return engine_mode() == 2 && umma_available() && a->workspace != nullptr && head_dims_tc(a); }
// fp32 [rows, cols] (row stride ld_in) -> fp16 hi/lo planes at row stride ld_out (no padding written: the caller owns the
// columns outside [0, cols) - used to lay the two halves of the pooled state side by side)
__global__ void split_f16_ld_kernel(const float* __restrict__ x, int64_t ld_in, int64_t rows, int cols, __half* __restrict__ hi,
                                    __half* __restrict__ lo, int64_t ld_out) {
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols;
    const int c = (int)(i - r * cols);
    __half a, b;
    tt::umma::split_f16(x[r * ld_in + c], a, b);
    hi[r * ld_out + c] = a; lo[r * ld_out + c] = b;
  }
}
__global__ void split_bf16_ld_kernel(const float* __restrict__ x, int64_t ld_in, int64_t rows, int cols, bf16* __restrict__ hi,
                                     bf16* __restrict__ lo, int64_t ld_out) {
  const int64_t total = rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / cols;
    const int c = (int)(i - r * cols);
    bf16 a, b;
    tt::umma::split_bf16(x[r * ld_in + c], a, b);
    hi[r * ld_out + c] = a; lo[r * ld_out + c] = b;
  }
}
struct HeadPlanes { bf16* hi; bf16* lo; int64_t pitch; };
static HeadPlanes take_planes(Workspace& w, int64_t rows, int64_t cols) {
  HeadPlanes p;
  p.pitch = plane_pitch(cols);
  p.hi = w.take<bf16>(rows * p.pitch);
  p.lo = w.take<bf16>(rows * p.pitch);
  return p;
}
static int to_planes(const float* x, int64_t ld_in, int64_t rows, int cols, const HeadPlanes& p, int col0, bool f16, cudaStream_t s) {
  const unsigned blocks = (unsigned)(ceil_div(rows * cols, 256) < 592 ? (ceil_div(rows * cols, 256) < 1 ? 1 : ceil_div(rows * cols, 256)) : 592);
  note_launch();
  if (f16) split_f16_ld_kernel<<<blocks, 256, 0, s>>>(x, ld_in, rows, cols, (__half*)p.hi + col0, (__half*)p.lo + col0, p.pitch);
  else split_bf16_ld_kernel<<<blocks, 256, 0, s>>>(x, ld_in, rows, cols, p.hi + col0, p.lo + col0, p.pitch);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
// C[M,N] = A B^T (+ bias) with A / B K-major ([M][K], [N][K]) or MN-major ([K][M], [K][N]) plane pairs, 3 passes
static int tc_gemm(const HeadPlanes& A, int a_col0, bool a_mn, const HeadPlanes& B, int b_col0, bool b_mn, float* C, int64_t ldc,
                   const float* bias, int M, int N, int K, bool f16, cudaStream_t s) {
  UmmaGemm g{};
  g.A[0] = Planes{A.hi + a_col0, A.lo + a_col0, A.pitch};
  g.B[0] = Planes{B.hi + b_col0, B.lo + b_col0, B.pitch};
  g.C[0] = C; g.bias[0] = bias; g.ldc = ldc;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.a_mn = a_mn; g.b_mn = b_mn; g.alpha = 1.f; g.beta = 0.f;
  g.splitk = (a_mn && b_mn) ? 0 : 1;              // reductions over the batch rows: split-K to fill the machine
  g.f16 = f16 ? 1 : 0; g.npass = 3;
  return umma_gemm(g, s);
}
static uint64_t planes_bytes(int64_t rows, int64_t cols) { return 2 * (align_up((uint64_t)rows * plane_pitch(cols) * 2, 256) + 256); }

static int check_head(const tt_head* a) {
  TT_CHECK_ARG(a && a->B > 0 && a->In > 0 && (a->In % 2) == 0 && a->P1 > 0 && a->D2 >= 0, "head: bad dims");
  TT_CHECK_ARG(a->v_lo && a->v_hi && a->w1 && a->b1 && a->ln_w && a->ln_b && a->y1 && a->stats && a->y2, "head: null buffer");
  TT_CHECK_ARG(a->D2 == 0 || (a->w2 && a->b2 && a->y3), "head: second linear buffers missing");
  TT_CHECK_ARG(!a->normalize || (a->out && a->norm), "head: normalize needs out/norm");
  TT_CHECK_ARG(a->dropout_p >= 0.f && a->dropout_p < 1.f, "head: dropout_p out of range");
  return TT_OK;
}

}  // namespace tt
using namespace tt;

extern "C" {

int tt_l2norm_fwd(const float* x, int32_t B, int32_t D, float eps, float* y, float* norm, void* stream) {
  TT_CHECK_ARG(x && y && B >= 0 && D > 0, "l2norm_fwd: bad args");
  if (B == 0) return TT_OK;
  tt::note_launch(); l2norm_fwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, (cudaStream_t)stream>>>(x, y, norm, B, D, eps);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
int tt_l2norm_bwd(const float* y, const float* norm, const float* dy, int32_t B, int32_t D, float eps, float* dx,
                  void* stream) {
  TT_CHECK_ARG(y && norm && dy && dx && B >= 0 && D > 0, "l2norm_bwd: bad args");
  if (B == 0) return TT_OK;
  tt::note_launch(); l2norm_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, (cudaStream_t)stream>>>(y, norm, dy, dx, B, D, eps);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_head_fwd(const tt_head* a, void* stream) {
  TT_TRY(check_head(a));
  cudaStream_t s = (cudaStream_t)stream;
  const int B = a->B, In = a->In, P = a->P1, D2 = a->D2, Hh = In / 2;
  const bool tc = head_tc(a);
  Workspace w(a->workspace, a->workspace_bytes);
  HeadPlanes vcat{}, w1p{}, y2p{}, w2p{};
  if (tc) {
    vcat = take_planes(w, B, In); w1p = take_planes(w, P, In);
    if (D2 > 0) { y2p = take_planes(w, B, P); w2p = take_planes(w, D2, P); }
    if (!w.ok()) { set_error("head fwd: workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
    // y1 = cat(v_lo, v_hi) W1^T + b1: the pooled state goes side by side into ONE operand plane pair
    TT_TRY(to_planes(a->v_lo, Hh, B, Hh, vcat, 0, true, s));
    TT_TRY(to_planes(a->v_hi, Hh, B, Hh, vcat, Hh, true, s));
    TT_TRY(to_planes(a->w1, In, P, In, w1p, 0, true, s));
    TT_TRY(tc_gemm(vcat, 0, false, w1p, 0, false, a->y1, P, a->b1, B, P, In, true, s));
  } else {
    // y1 = v_lo W1[:, :Hh]^T + v_hi W1[:, Hh:]^T + b1: the two halves of the concatenated input as one launch
    GemmArgs g = nt_gemm(a->v_lo, Hh, a->w1, In, a->y1, P, a->b1, B, P, Hh, 0.f);
    pair_with(g, a->v_hi, a->w1 + Hh, a->y1);
    TT_TRY(gemm_f32(g, s));
  }
  const float pdrop = a->dropout_p;
  tt::note_launch(); ln_relu_fwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(a->y1, a->ln_w, a->ln_b, a->y2, a->stats, B, P, 1e-5f, pdrop,
                                                              pdrop > 0.f ? 1.f / (1.f - pdrop) : 1.f, a->seed, a->offset, a->seed_dev);
  TT_LAUNCH_CHECK();
  const float* last = a->y2;
  int Dl = P;
  if (D2 > 0) {
    if (tc) {
      TT_TRY(to_planes(a->y2, P, B, P, y2p, 0, true, s));
      TT_TRY(to_planes(a->w2, P, D2, P, w2p, 0, true, s));
      TT_TRY(tc_gemm(y2p, 0, false, w2p, 0, false, a->y3, D2, a->b2, B, D2, P, true, s));
    } else {
      TT_TRY(gemm_f32(nt_gemm(a->y2, P, a->w2, P, a->y3, D2, a->b2, B, D2, P, 0.f), s));
    }
    last = a->y3; Dl = D2;
  }
  if (a->normalize) {
    tt::note_launch(); l2norm_fwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(last, a->out, a->norm, B, Dl, 1e-12f);
    TT_LAUNCH_CHECK();
  }
  return TT_OK;
}

uint64_t tt_head_fwd_workspace(const tt_head* a) {
  if (!a) return 0;
  uint64_t n = planes_bytes(a->B, a->In) + planes_bytes(a->P1, a->In);
  if (a->D2 > 0) n += planes_bytes(a->B, a->P1) + planes_bytes(a->D2, a->P1);
  return n + 1024;
}

uint64_t tt_head_bwd_workspace(const tt_head_grad* a) {
  if (!a) return 0;
  const tt_head* f = &a->fwd;
  Workspace w(nullptr, 0);
  const int Dl = f->D2 > 0 ? f->D2 : f->P1;
  w.take<float>((int64_t)f->B * Dl);
  for (int i = 0; i < 4; ++i) w.take<float>((int64_t)f->B * f->P1);
  // tensor route: bf16 hi/lo planes of dlast, y2, W2, d_y1, cat(v), W1
  uint64_t n = w.used + 256 + planes_bytes(f->B, f->P1) + planes_bytes(f->B, f->In) + planes_bytes(f->P1, f->In);
  if (f->D2 > 0) n += planes_bytes(f->B, f->D2) + planes_bytes(f->B, f->P1) + planes_bytes(f->D2, f->P1);
  return n + 1024;
}

int tt_head_bwd(const tt_head_grad* a, void* stream) {
  TT_CHECK_ARG(a != nullptr, "head bwd: null args");
  const tt_head* f = &a->fwd;
  TT_TRY(check_head(f));
  TT_CHECK_ARG(a->d_out && a->d_v_lo && a->d_v_hi && a->d_w1 && a->d_b1 && a->d_ln_w && a->d_ln_b, "head bwd: null buffer");
  TT_CHECK_ARG(f->D2 == 0 || (a->d_w2 && a->d_b2), "head bwd: null w2 grads");
  cudaStream_t s = (cudaStream_t)stream;
  const int B = f->B, In = f->In, P = f->P1, D2 = f->D2, Hh = In / 2;
  const int Dl = D2 > 0 ? D2 : P;
  Workspace w(a->workspace, a->workspace_bytes);
  float* d_y = w.take<float>((int64_t)B * Dl);
  float* d_y2 = w.take<float>((int64_t)B * P);
  float* d_y1 = w.take<float>((int64_t)B * P);
  float* gm = w.take<float>((int64_t)B * P);
  float* gx = w.take<float>((int64_t)B * P);
  if (!w.ok()) { set_error("head bwd: workspace too small"); return TT_ERR_WORKSPACE; }
  const bool tc = engine_mode() == 2 && umma_available() && head_dims_tc(f);
  HeadPlanes dlp{}, y2p{}, w2p{}, dy1p{}, vcat{}, w1p{};
  if (tc) {
    if (D2 > 0) { dlp = take_planes(w, B, D2); y2p = take_planes(w, B, P); w2p = take_planes(w, D2, P); }
    dy1p = take_planes(w, B, P); vcat = take_planes(w, B, In); w1p = take_planes(w, P, In);
    if (!w.ok()) { set_error("head bwd: workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
  }

  const float* dlast = a->d_out;
  if (f->normalize) {
    tt::note_launch(); l2norm_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(f->out, f->norm, a->d_out, d_y, B, Dl, 1e-12f);
    TT_LAUNCH_CHECK();
    dlast = d_y;
  }
  const float* dy2 = dlast;
  if (D2 > 0) {
    if (tc) {
      TT_TRY(to_planes(dlast, D2, B, D2, dlp, 0, false, s));
      TT_TRY(to_planes(f->y2, P, B, P, y2p, 0, false, s));
      TT_TRY(to_planes(f->w2, P, D2, P, w2p, 0, false, s));
      // d_w2[D2,P] = dlast^T y2 (reduction over the batch rows: both operands MN-major); d_y2[B,P] = dlast W2
      TT_TRY(tc_gemm(dlp, 0, true, y2p, 0, true, a->d_w2, P, nullptr, D2, P, B, false, s));
      TT_TRY(tc_gemm(dlp, 0, false, w2p, 0, true, d_y2, P, nullptr, B, P, D2, false, s));
    } else {
      TT_TRY(gemm_f32(tn_gemm(dlast, D2, f->y2, P, a->d_w2, P, D2, P, B), s));
      TT_TRY(gemm_f32(nn_gemm(dlast, D2, f->w2, P, d_y2, P, B, P, D2, 0.f), s));
    }
    TT_TRY(colsum_f32(dlast, D2, B, D2, a->d_b2, s));
    dy2 = d_y2;
  }
  tt::note_launch(); ln_relu_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(f->y1, f->y2, f->stats, f->ln_w, dy2, d_y1, gm, gx, B, P,
                                                              f->dropout_p > 0.f ? 1.f / (1.f - f->dropout_p) : 1.f);
  TT_LAUNCH_CHECK();
  TT_TRY(colsum_f32(gm, P, B, P, a->d_ln_b, s));
  TT_TRY(colsum_f32(gx, P, B, P, a->d_ln_w, s));
  TT_TRY(colsum_f32(d_y1, P, B, P, a->d_b1, s));
  if (tc) {
    TT_TRY(to_planes(d_y1, P, B, P, dy1p, 0, false, s));
    TT_TRY(to_planes(f->v_lo, Hh, B, Hh, vcat, 0, false, s));
    TT_TRY(to_planes(f->v_hi, Hh, B, Hh, vcat, Hh, false, s));
    TT_TRY(to_planes(f->w1, In, P, In, w1p, 0, false, s));
    // d_w1[P,In] = d_y1^T cat(v) ; d_v_lo / d_v_hi [B,Hh] = d_y1 W1[:, :Hh] / W1[:, Hh:]
    TT_TRY(tc_gemm(dy1p, 0, true, vcat, 0, true, a->d_w1, In, nullptr, P, In, B, false, s));
    TT_TRY(tc_gemm(dy1p, 0, false, w1p, 0, true, a->d_v_lo, Hh, nullptr, B, Hh, P, false, s));
    TT_TRY(tc_gemm(dy1p, 0, false, w1p, Hh, true, a->d_v_hi, Hh, nullptr, B, Hh, P, false, s));
  } else {
    // the two column halves of d_w1, and the two halves of d_v, as one launch each (disjoint outputs)
    GemmArgs g = tn_gemm(d_y1, P, f->v_lo, Hh, a->d_w1, In, P, Hh, B);
    pair_with(g, d_y1, f->v_hi, a->d_w1 + Hh);
    TT_TRY(gemm_f32(g, s));
    g = nn_gemm(d_y1, P, f->w1, In, a->d_v_lo, Hh, B, Hh, P, 0.f);
    pair_with(g, d_y1, f->w1 + Hh, a->d_v_hi);
    TT_TRY(gemm_f32(g, s));
  }
  return TT_OK;
}

}  // extern "C"
