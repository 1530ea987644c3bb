// K-D projection head: Linear -> LayerNorm -> ReLU [-> Linear] [-> L2 normalise], forward and backward.
// The two Linear layers are tiny ([B,512]x[512,256] at the headline config, <0.1 % of a step) and run on the
// exact-fp32 GEMM; LayerNorm+ReLU and the normalisation are one-warp-per-row kernels.
#include "common.cuh"

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
  {
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
    TT_TRY(gemm_f32(nt_gemm(a->y2, P, a->w2, P, a->y3, D2, a->b2, B, D2, P, 0.f), s));
    last = a->y3; Dl = D2;
  }
  if (a->normalize) {
    tt::note_launch(); l2norm_fwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(last, a->out, a->norm, B, Dl, 1e-12f);
    TT_LAUNCH_CHECK();
  }
  return TT_OK;
}

uint64_t tt_head_bwd_workspace(const tt_head_grad* a) {
  if (!a) return 0;
  const tt_head* f = &a->fwd;
  Workspace w(nullptr, 0);
  const int Dl = f->D2 > 0 ? f->D2 : f->P1;
  w.take<float>((int64_t)f->B * Dl);
  for (int i = 0; i < 4; ++i) w.take<float>((int64_t)f->B * f->P1);
  return w.used + 256;
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

  const float* dlast = a->d_out;
  if (f->normalize) {
    tt::note_launch(); l2norm_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(f->out, f->norm, a->d_out, d_y, B, Dl, 1e-12f);
    TT_LAUNCH_CHECK();
    dlast = d_y;
  }
  const float* dy2 = dlast;
  if (D2 > 0) {
    TT_TRY(gemm_f32(tn_gemm(dlast, D2, f->y2, P, a->d_w2, P, D2, P, B), s));
    TT_TRY(colsum_f32(dlast, D2, B, D2, a->d_b2, s));
    TT_TRY(gemm_f32(nn_gemm(dlast, D2, f->w2, P, d_y2, P, B, P, D2, 0.f), s));
    dy2 = d_y2;
  }
  tt::note_launch(); ln_relu_bwd_kernel<<<(unsigned)ceil_div(B, 8), 256, 0, s>>>(f->y1, f->y2, f->stats, f->ln_w, dy2, d_y1, gm, gx, B, P,
                                                              f->dropout_p > 0.f ? 1.f / (1.f - f->dropout_p) : 1.f);
  TT_LAUNCH_CHECK();
  TT_TRY(colsum_f32(gm, P, B, P, a->d_ln_b, s));
  TT_TRY(colsum_f32(gx, P, B, P, a->d_ln_w, s));
  TT_TRY(colsum_f32(d_y1, P, B, P, a->d_b1, s));
  {
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
