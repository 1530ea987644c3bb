// K-E in-batch score + cross-entropy (fwd / bwd) and K-G explicit-negative hinge.
// The [Bl,Bg] logit matrix is never held whole: it is produced in column chunks sized to stay L2-resident
// (<= 32 MiB of the 126 MB L2), consumed by a row-statistics (fwd) or softmax-gradient (bwd) kernel, and
// overwritten by the next chunk.  Scores are exact fp32 (one FMA chain per element).
#include <cuda_fp16.h>
#include "common.cuh"
#include "umma.cuh"
#include "umma_gemm.cuh"
#include "gru_umma.cuh"

namespace tt {

constexpr int64_t kChunkBytes = 32ll << 20;

static inline int chunk_cols(int Bl, int Bg) {
  int64_t nc = kChunkBytes / (4ll * (Bl > 0 ? Bl : 1));
  nc = nc / 128 * 128;
  if (nc < 128) nc = 128;
  return (int)(nc < Bg ? nc : Bg);
}

// one warp per row: fold chunk S[:, c0:c0+nc] (already scaled by inv_temp) into running (max, sumexp)
__global__ void ce_rowstats_kernel(const float* __restrict__ S, int ld, int Bl, int c0, int nc, int pos_offset,
                                   float margin, float* __restrict__ run_max, float* __restrict__ run_sum,
                                   float* __restrict__ pos_logit, int first) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= Bl) return;
  const float* s = S + (int64_t)row * ld;
  const int pos = pos_offset + row;
  float m = -INFINITY;
  for (int c = lane; c < nc; c += 32) {
    const float v = s[c] - ((c0 + c) != pos ? margin : 0.f);
    m = fmaxf(m, v);
  }
  m = warp_max(m);
  float l = 0.f;
  for (int c = lane; c < nc; c += 32) {
    const float v = s[c] - ((c0 + c) != pos ? margin : 0.f);
    l += __expf(v - m);
  }
  l = warp_sum(l);
  if (lane == 0) {
    if (pos >= c0 && pos < c0 + nc) pos_logit[row] = s[pos - c0];
    if (first) { run_max[row] = m; run_sum[row] = l; }
    else {
      const float M = run_max[row], L = run_sum[row], nm = fmaxf(M, m);
      run_sum[row] = L * __expf(M - nm) + l * __expf(m - nm);
      run_max[row] = nm;
    }
  }
}
__global__ void ce_finish_kernel(const float* run_max, const float* run_sum, const float* pos_logit, float* row_loss,
                                 float* row_lse, int Bl) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Bl) return;
  const float lse = run_max[i] + logf(run_sum[i]);
  row_lse[i] = lse;
  row_loss[i] = lse - pos_logit[i];
}
// in place: S <- (softmax - onehot) * coef
__global__ void ce_grad_kernel(float* __restrict__ S, int ld, int Bl, int c0, int nc, int pos_offset, float margin,
                               const float* __restrict__ row_lse, float coef, const float* __restrict__ coef_dev) {
  if (coef_dev) coef *= __ldg(coef_dev);
  const int64_t total = (int64_t)Bl * nc;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / nc), c = (int)(i % nc);
    const bool is_pos = (c0 + c) == pos_offset + row;
    float* p = S + (int64_t)row * ld + c;
    const float v = *p - (is_pos ? 0.f : margin);
    *p = (__expf(v - row_lse[row]) - (is_pos ? 1.f : 0.f)) * coef;
  }
}

// fp32 [rows][cols] -> fp16 hi/lo planes [rows][pitch] (pad columns zero)
__global__ void split_rows_f16_kernel(const float* __restrict__ x, int64_t rows, int cols, int64_t pitch,
                                      __half* __restrict__ hi, __half* __restrict__ lo, const float* __restrict__ row_scale = nullptr) {
  const int64_t total = rows * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % pitch);
    float v = c < cols ? x[(i / pitch) * cols + c] : 0.f;
    if (row_scale) v *= row_scale[i / pitch];
    __half a, b;
    umma::split_f16(v, a, b);
    hi[i] = a; lo[i] = b;
  }
}
// G = (softmax - onehot) * coef as fp16 hi/lo planes [Bl][pitch]; columns [nc, pitch) are zero.  coef carries a power
// of two that lifts the tiny probabilities out of the fp16 subnormal range (undone by alpha of the gradient GEMMs).
__global__ void ce_grad_planes_kernel(const float* __restrict__ S, int ld, int Bl, int c0, int nc, int pos_offset, float margin,
                                      const float* __restrict__ row_lse, float coef, __half* __restrict__ gh,
                                      __half* __restrict__ gl, int pitch) {
  const int64_t total = (int64_t)Bl * pitch;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / pitch), c = (int)(i % pitch);
    float g = 0.f;
    if (c < nc) {
      const bool is_pos = (c0 + c) == pos_offset + row;
      const float v = S[(int64_t)row * ld + c] - (is_pos ? 0.f : margin);
      g = (__expf(v - row_lse[row]) - (is_pos ? 1.f : 0.f)) * coef;
    }
    __half a, b;
    umma::split_f16(g, a, b);
    gh[i] = a; gl[i] = b;
  }
}
int split_rows_f16(const float* x, int64_t rows, int cols, int64_t pitch, void* hi, void* lo, cudaStream_t s) {
  if (rows <= 0) return TT_OK;
  note_launch(); split_rows_f16_kernel<<<592, 256, 0, s>>>(x, rows, cols, pitch, (__half*)hi, (__half*)lo);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
// same with every row multiplied by row_scale[row] first (row_scale may be null)
int split_rows_f16_scaled(const float* x, int64_t rows, int cols, int64_t pitch, void* hi, void* lo, const float* row_scale,
                          cudaStream_t s) {
  if (rows <= 0) return TT_OK;
  note_launch(); split_rows_f16_kernel<<<592, 256, 0, s>>>(x, rows, cols, pitch, (__half*)hi, (__half*)lo, row_scale);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
static inline bool score_tc() { return engine_mode() >= 1 && umma_available(); }
struct ScorePlanes { __half *xh, *xl, *yh, *yl, *gh, *gl; int64_t dp, ncp; };
static void carve_score_planes(Workspace& w, int Bl, int Bg, int D, int NC, bool grad, ScorePlanes* o) {
  o->dp = plane_pitch(D); o->ncp = plane_pitch(NC);
  o->xh = w.take<__half>((int64_t)Bl * o->dp); o->xl = w.take<__half>((int64_t)Bl * o->dp);
  o->yh = w.take<__half>((int64_t)Bg * o->dp); o->yl = w.take<__half>((int64_t)Bg * o->dp);
  o->gh = o->gl = nullptr;
  if (grad) { o->gh = w.take<__half>((int64_t)Bl * o->ncp); o->gl = w.take<__half>((int64_t)Bl * o->ncp); }
}
static int split_xy(const float* x, int Bl, const float* y, int Bg, int D, const ScorePlanes& sp, cudaStream_t s) {
  note_launch(); split_rows_f16_kernel<<<592, 256, 0, s>>>(x, Bl, D, sp.dp, sp.xh, sp.xl);
  note_launch(); split_rows_f16_kernel<<<592, 256, 0, s>>>(y, Bg, D, sp.dp, sp.yh, sp.yl);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
// S[Bl, nc] = inv_temp * x y[c0:c0+nc]^T on the tensor engine (x3)
static int score_chunk_tc(const ScorePlanes& sp, int Bl, int Bg, int D, int c0, int nc, int NC, float inv_temp, float* S,
                          cudaStream_t s) {
  UmmaGemm g{};
  g.A[0] = Planes{(const bf16*)sp.xh, (const bf16*)sp.xl, sp.dp};
  g.B[0] = Planes{(const bf16*)sp.yh, (const bf16*)sp.yl, sp.dp};
  g.C[0] = S; g.ldc = NC; g.M = Bl; g.N = nc; g.K = D; g.nbatch = 1;
  g.b_rows = Bg; g.b_row0 = c0; g.alpha = inv_temp; g.beta = 0.f; g.splitk = 1; g.f16 = 1; g.npass = 3;
  return umma_gemm(g, s);
}

// ---------------------------------------------------------------------------------- K-G hinge
// one warp per query row.  cos(a,b) = <a,b> / (max(|a|,eps) max(|b|,eps)), eps = 1e-8
__device__ __forceinline__ void dot3(const float* a, const float* b, int D, int lane, float& ab, float& bb) {
  float s = 0.f, t = 0.f;
  for (int c = lane; c < D; c += 32) { const float x = a[c], y = b[c]; s += x * y; t += y * y; }
  ab = warp_sum(s); bb = warp_sum(t);
}
template <bool BWD>
__global__ void hinge_kernel(const float* __restrict__ q, const float* __restrict__ pos, const float* __restrict__ neg,
                             int B, int k, int D, float margin, float gscale, const float* __restrict__ gscale_dev,
                             float* __restrict__ row_loss, float* __restrict__ dq, float* __restrict__ dpos,
                             float* __restrict__ dneg) {
  if (BWD && gscale_dev) gscale *= __ldg(gscale_dev);
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= B) return;
  const float eps = 1e-8f;
  const float* qi = q + (int64_t)row * D;
  const float* pi = pos + (int64_t)row * D;
  float qq = 0.f;
  for (int c = lane; c < D; c += 32) qq += qi[c] * qi[c];
  qq = warp_sum(qq);
  const float nq = sqrtf(qq), nqc = fmaxf(nq, eps);
  float qp, pp;
  dot3(qi, pi, D, lane, qp, pp);
  const float np_ = sqrtf(pp), npc = fmaxf(np_, eps);
  const float cpos = qp / (nqc * npc);
  float cneg_sum = 0.f;
  for (int j = 0; j < k; ++j) {
    float qn, nn;
    dot3(qi, neg + ((int64_t)row * k + j) * D, D, lane, qn, nn);
    cneg_sum += qn / (nqc * fmaxf(sqrtf(nn), eps));
  }
  const float loss = margin - cpos + cneg_sum / k;
  if (!BWD) {
    if (lane == 0) row_loss[row] = fmaxf(loss, 0.f);
    return;
  }
  const float g = loss > 0.f ? gscale : 0.f;     // clamp(min=0) subgradient 0 at the kink, like torch
  // d cos(a,b)/da = b/(na_c nb_c) - [na>eps] cos * a / na^2
  const float gp = -g, gn = g / k;
  const float qs_pos = nq > eps ? cpos / (nq * nq) : 0.f;
  const float ps = np_ > eps ? cpos / (np_ * np_) : 0.f;
  for (int c = lane; c < D; c += 32) {
    dq[(int64_t)row * D + c] = gp * (pi[c] / (nqc * npc) - qs_pos * qi[c]);
    dpos[(int64_t)row * D + c] = gp * (qi[c] / (nqc * npc) - ps * pi[c]);
  }
  for (int j = 0; j < k; ++j) {
    const float* nj = neg + ((int64_t)row * k + j) * D;
    float qn, nn;
    dot3(qi, nj, D, lane, qn, nn);
    const float nnr = sqrtf(nn), nnc = fmaxf(nnr, eps);
    const float cn = qn / (nqc * nnc);
    const float qs = nq > eps ? cn / (nq * nq) : 0.f;
    const float ns = nnr > eps ? cn / (nnr * nnr) : 0.f;
    for (int c = lane; c < D; c += 32) {
      dq[(int64_t)row * D + c] += gn * (nj[c] / (nqc * nnc) - qs * qi[c]);
      dneg[((int64_t)row * k + j) * D + c] = gn * (qi[c] / (nqc * nnc) - ns * nj[c]);
    }
  }
}

}  // namespace tt
using namespace tt;

extern "C" {

uint64_t tt_score_ce_workspace(int32_t Bl, int32_t Bg, int32_t D) {
  Workspace w(nullptr, 0);
  w.take<float>((int64_t)Bl * chunk_cols(Bl, Bg));
  for (int i = 0; i < 3; ++i) w.take<float>(Bl);
  ScorePlanes sp;
  carve_score_planes(w, Bl, Bg, D, chunk_cols(Bl, Bg), true, &sp);   // tensor-engine operand planes (fwd uses a prefix)
  return w.used + 256;
}

int tt_score_ce_fwd(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                    float inv_temp, float margin, float* row_loss, float* row_lse, void* workspace,
                    uint64_t workspace_bytes, void* stream) {
  TT_CHECK_ARG(x && y && row_loss && row_lse && Bl > 0 && Bg > 0 && D > 0, "score_ce_fwd: bad args");
  TT_CHECK_ARG(pos_offset >= 0 && pos_offset + Bl <= Bg, "score_ce_fwd: positives [%d,%d) outside %d columns", pos_offset, pos_offset + Bl, Bg);
  cudaStream_t s = (cudaStream_t)stream;
  const int NC = chunk_cols(Bl, Bg);
  Workspace w(workspace, workspace_bytes);
  float* S = w.take<float>((int64_t)Bl * NC);
  float* rmax = w.take<float>(Bl); float* rsum = w.take<float>(Bl); float* plog = w.take<float>(Bl);
  const bool tc = score_tc();
  ScorePlanes sp{};
  if (tc) carve_score_planes(w, Bl, Bg, D, NC, false, &sp);
  if (!w.ok()) { set_error("score_ce_fwd: workspace too small"); return TT_ERR_WORKSPACE; }
  if (tc) TT_TRY(split_xy(x, Bl, y, Bg, D, sp, s));
  for (int c0 = 0; c0 < Bg; c0 += NC) {
    const int nc = (Bg - c0) < NC ? (Bg - c0) : NC;
    if (tc) {
      TT_TRY(score_chunk_tc(sp, Bl, Bg, D, c0, nc, NC, inv_temp, S, s));
    } else {
      GemmArgs g{};
      g.A[0] = x; g.B[0] = y + (int64_t)c0 * D; g.C[0] = S;
      g.a_rs = D; g.a_cs = 1; g.b_rs = D; g.b_cs = 1; g.ldc = NC;
      g.M = Bl; g.N = nc; g.K = D; g.nbatch = 1; g.splitk = 1; g.alpha = inv_temp; g.beta = 0.f;
      TT_TRY(gemm_f32(g, s));
    }
    tt::note_launch(); ce_rowstats_kernel<<<(unsigned)ceil_div(Bl, 8), 256, 0, s>>>(S, NC, Bl, c0, nc, pos_offset, margin, rmax, rsum, plog, c0 == 0);
  }
  tt::note_launch(); ce_finish_kernel<<<(unsigned)ceil_div(Bl, 256), 256, 0, s>>>(rmax, rsum, plog, row_loss, row_lse, Bl);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_score_ce_bwd(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                    float inv_temp, float margin, const float* row_lse, float gscale, float* dx, float* dy,
                    void* workspace, uint64_t workspace_bytes, void* stream) {
  return tt_score_ce_bwd_dev(x, Bl, y, Bg, D, pos_offset, inv_temp, margin, row_lse, gscale, nullptr, dx, dy, workspace,
                             workspace_bytes, stream);
}

int tt_score_ce_bwd_dev(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                        float inv_temp, float margin, const float* row_lse, float gscale, const float* gscale_dev,
                        float* dx, float* dy, void* workspace, uint64_t workspace_bytes, void* stream) {
  TT_CHECK_ARG(x && y && row_lse && dx && dy && Bl > 0 && Bg > 0 && D > 0, "score_ce_bwd: bad args");
  cudaStream_t s = (cudaStream_t)stream;
  const int NC = chunk_cols(Bl, Bg);
  Workspace w(workspace, workspace_bytes);
  float* S = w.take<float>((int64_t)Bl * NC);
  for (int i = 0; i < 3; ++i) w.take<float>(Bl);              // (same carve order as the forward / the workspace query)
  const bool tc = score_tc();
  ScorePlanes sp{};
  if (tc) carve_score_planes(w, Bl, Bg, D, NC, true, &sp);
  if (!w.ok()) { set_error("score_ce_bwd: workspace too small"); return TT_ERR_WORKSPACE; }
  if (tc) TT_TRY(split_xy(x, Bl, y, Bg, D, sp, s));
  // power of two lifting |G| <= |gscale| * inv_temp into [4, 8): keeps small probabilities out of fp16 subnormals
  float lift = 1.f;
  if (tc) {
    const float gmax = fabsf(gscale * inv_temp);
    int ex = 0;
    if (gmax > 0.f && isfinite(gmax)) { frexpf(gmax, &ex); lift = ldexpf(1.f, 3 - ex); }
  }
  for (int c0 = 0; c0 < Bg; c0 += NC) {
    const int nc = (Bg - c0) < NC ? (Bg - c0) : NC;
    if (tc) {
      TT_TRY(score_chunk_tc(sp, Bl, Bg, D, c0, nc, NC, inv_temp, S, s));
      int64_t tot = (int64_t)Bl * sp.ncp, nb = ceil_div(tot, 256);
      if (nb > 148 * 16) nb = 148 * 16;
      tt::note_launch();
      ce_grad_planes_kernel<<<(unsigned)nb, 256, 0, s>>>(S, NC, Bl, c0, nc, pos_offset, margin, row_lse, gscale * inv_temp * lift,
                                                         sp.gh, sp.gl, (int)sp.ncp);
      // dx[Bl,D] (+)= G[Bl,nc] Y[c0:c0+nc, D]      (A K-major over the chunk columns, B = Y rows as an MN-major operand)
      UmmaGemm gx{};
      gx.A[0] = Planes{(const bf16*)sp.gh, (const bf16*)sp.gl, sp.ncp};
      gx.B[0] = Planes{(const bf16*)(sp.yh + (int64_t)c0 * sp.dp), (const bf16*)(sp.yl + (int64_t)c0 * sp.dp), sp.dp};
      gx.C[0] = dx; gx.ldc = D; gx.M = Bl; gx.N = D; gx.K = nc; gx.nbatch = 1; gx.a_mn = false; gx.b_mn = true;
      gx.alpha = 1.f / lift; gx.alpha_ptr = gscale_dev; gx.beta = c0 == 0 ? 0.f : 1.f; gx.splitk = 0; gx.f16 = 1; gx.npass = 3;
      TT_TRY(umma_gemm(gx, s));
      // dy[c0:c0+nc, D] = G^T X                     (both operands MN-major: reduction over the Bl rows)
      UmmaGemm gy{};
      gy.A[0] = Planes{(const bf16*)sp.gh, (const bf16*)sp.gl, sp.ncp};
      gy.B[0] = Planes{(const bf16*)sp.xh, (const bf16*)sp.xl, sp.dp};
      gy.C[0] = dy + (int64_t)c0 * D; gy.ldc = D; gy.M = nc; gy.N = D; gy.K = Bl; gy.nbatch = 1; gy.a_mn = gy.b_mn = true;
      gy.alpha = 1.f / lift; gy.alpha_ptr = gscale_dev; gy.beta = 0.f; gy.splitk = 0; gy.f16 = 1; gy.npass = 3;
      TT_TRY(umma_gemm(gy, s));
      continue;
    }
    GemmArgs g{};
    g.A[0] = x; g.B[0] = y + (int64_t)c0 * D; g.C[0] = S;
    g.a_rs = D; g.a_cs = 1; g.b_rs = D; g.b_cs = 1; g.ldc = NC;
    g.M = Bl; g.N = nc; g.K = D; g.nbatch = 1; g.splitk = 1; g.alpha = inv_temp; g.beta = 0.f;
    TT_TRY(gemm_f32(g, s));
    int64_t tot = (int64_t)Bl * nc, nb = ceil_div(tot, 256);
    if (nb > 148 * 16) nb = 148 * 16;
    tt::note_launch(); ce_grad_kernel<<<(unsigned)nb, 256, 0, s>>>(S, NC, Bl, c0, nc, pos_offset, margin, row_lse, gscale * inv_temp, gscale_dev);
    // dx[Bl,D] (+)= G[Bl,nc] Y[c0:c0+nc, D]
    GemmArgs gx{};
    gx.A[0] = S; gx.B[0] = y + (int64_t)c0 * D; gx.C[0] = dx;
    gx.a_rs = NC; gx.a_cs = 1; gx.b_rs = 1; gx.b_cs = D; gx.ldc = D;
    gx.M = Bl; gx.N = D; gx.K = nc; gx.nbatch = 1; gx.splitk = 0; gx.alpha = 1.f; gx.beta = c0 == 0 ? 0.f : 1.f;
    TT_TRY(gemm_f32(gx, s));
    // dy[c0:c0+nc, D] = G^T X
    GemmArgs gy{};
    gy.A[0] = S; gy.B[0] = x; gy.C[0] = dy + (int64_t)c0 * D;
    gy.a_rs = 1; gy.a_cs = NC; gy.b_rs = 1; gy.b_cs = D; gy.ldc = D;
    gy.M = nc; gy.N = D; gy.K = Bl; gy.nbatch = 1; gy.splitk = 0; gy.alpha = 1.f; gy.beta = 0.f;
    TT_TRY(gemm_f32(gy, s));
  }
  TT_LAUNCH_CHECK();
  return TT_OK;
}

int tt_hinge_fwd(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D, float margin,
                 float* row_loss, void* stream) {
  TT_CHECK_ARG(q && pos && neg && row_loss && B > 0 && k > 0 && D > 0, "hinge_fwd: bad args");
  tt::note_launch(); hinge_kernel<false><<<(unsigned)ceil_div(B, 8), 256, 0, (cudaStream_t)stream>>>(q, pos, neg, B, k, D, margin, 0.f, nullptr,
                                                                                row_loss, nullptr, nullptr, nullptr);
  TT_LAUNCH_CHECK();
  return TT_OK;
}
int tt_hinge_bwd(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D, float margin,
                 float gscale, float* dq, float* dpos, float* dneg, void* stream) {
  return tt_hinge_bwd_dev(q, pos, neg, B, k, D, margin, gscale, nullptr, dq, dpos, dneg, stream);
}
int tt_hinge_bwd_dev(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D, float margin,
                     float gscale, const float* gscale_dev, float* dq, float* dpos, float* dneg, void* stream) {
  TT_CHECK_ARG(q && pos && neg && dq && dpos && dneg && B > 0 && k > 0 && D > 0, "hinge_bwd: bad args");
  tt::note_launch(); hinge_kernel<true><<<(unsigned)ceil_div(B, 8), 256, 0, (cudaStream_t)stream>>>(q, pos, neg, B, k, D, margin, gscale,
                                                                               gscale_dev, nullptr, dq, dpos, dneg);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

}  // extern "C"
