// GRU layer forward / BPTT (K-A gather, K-B input projection, K-C recurrence, K-C' backward).
// Engine: exact fp32 (SIMT GEMM + fused cell kernels).  See DESIGN.md for the data layout.
#include "common.cuh"
#include "gru_umma.cuh"

namespace tt {

// ------------------------------------------------------------------------------------ K-A gather
// one warp per token row; float4 when E % 4 == 0 (300 -> 75 float4), coalesced 16-B lanes.
template <typename IdT>
__global__ void gather_rows_kernel(const IdT* __restrict__ ids, int64_t n_tokens, const float* __restrict__ table,
                                   int64_t vocab, int E, int64_t pad_id, float* __restrict__ out, float p,
                                   float inv_keep, uint64_t seed, uint64_t offset) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t tok = warp; tok < n_tokens; tok += nwarps) {
    const int64_t id = (int64_t)ids[tok];
    const bool zero = id < 0 || id >= vocab || id == pad_id;
    const float* src = table + (zero ? 0 : id) * E;
    float* dst = out + tok * E;
    if ((E & 3) == 0) {
      for (int c = lane; c < (E >> 2); c += 32) {
        float4 v = zero ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(reinterpret_cast<const float4*>(src) + c);
        if (p > 0.f) {
          uint64_t e = (uint64_t)tok * E + 4 * c;
          v.x *= dropout_scale(seed, offset, e, p, inv_keep);
          v.y *= dropout_scale(seed, offset, e + 1, p, inv_keep);
          v.z *= dropout_scale(seed, offset, e + 2, p, inv_keep);
          v.w *= dropout_scale(seed, offset, e + 3, p, inv_keep);
        }
        reinterpret_cast<float4*>(dst)[c] = v;
      }
    } else {
      for (int c = lane; c < E; c += 32) {
        float v = zero ? 0.f : __ldg(src + c);
        if (p > 0.f) v *= dropout_scale(seed, offset, (uint64_t)tok * E + c, p, inv_keep);
        dst[c] = v;
      }
    }
  }
}

static int launch_gather(const void* ids, int ids_are_i64, int64_t n_tokens, const float* table, int64_t vocab,
                         int E, int64_t pad_id, float* out, float p, uint64_t seed, uint64_t offset,
                         cudaStream_t s) {
  if (n_tokens <= 0) return TT_OK;
  int64_t blocks = ceil_div(n_tokens, 8);
  if (blocks > 148 * 16) blocks = 148 * 16;
  float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  note_launch();
  if (ids_are_i64)
    gather_rows_kernel<int64_t><<<(unsigned)blocks, 256, 0, s>>>((const int64_t*)ids, n_tokens, table, vocab, E,
                                                               pad_id, out, p, inv_keep, seed, offset);
  else
    gather_rows_kernel<int32_t><<<(unsigned)blocks, 256, 0, s>>>((const int32_t*)ids, n_tokens, table, vocab, E,
                                                               pad_id, out, p, inv_keep, seed, offset);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

__global__ void dropout_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n, float p,
                               float inv_keep, uint64_t seed, uint64_t offset) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = x[i] * dropout_scale(seed, offset, (uint64_t)i, p, inv_keep);
}
static int launch_dropout(const float* x, float* y, int64_t n, float p, uint64_t seed, uint64_t offset,
                          cudaStream_t s) {
  if (n <= 0) return TT_OK;
  int64_t blocks = ceil_div(n, 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  tt::note_launch(); dropout_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, y, n, p, 1.f / (1.f - p), seed, offset);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

// ------------------------------------------------------------------------------------ K-C cell
// gi: [B*T, ndir*3H] (bias b_ih already added), gh: [ndir][B][3H] (bias b_hh already added)
// hs: [ndir][B][H] running state (in place), out: [B,T,ndir*H], gates: [ndir][B][T][4H]
__global__ void gru_cell_fwd_kernel(const float* __restrict__ gi, const float* __restrict__ gh, float* __restrict__ hs,
                                    float* __restrict__ out, float* __restrict__ gates,
                                    const int32_t* __restrict__ lengths, int B, int T, int H, int ndir, int step) {
  const int64_t total = (int64_t)ndir * B * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % H);
    const int b = (int)((i / H) % B);
    const int dir = (int)(i / ((int64_t)H * B));
    const int t = dir == 0 ? step : T - 1 - step;
    const float hp = hs[i];
    float* o = out + ((int64_t)b * T + t) * (ndir * H) + dir * H + j;
    float* gsave = gates ? gates + (((int64_t)dir * B + b) * T + t) * (4 * H) : nullptr;
    if (lengths && t >= lengths[b]) {
      *o = 0.f;
      if (gsave) { gsave[j] = 0.f; gsave[H + j] = 1.f; gsave[2 * H + j] = 0.f; gsave[3 * H + j] = 0.f; }
      continue;
    }
    const float* gi_ = gi + ((int64_t)b * T + t) * (ndir * 3 * H) + dir * 3 * H;
    const float* gh_ = gh + ((int64_t)dir * B + b) * (3 * H);
    const float r = sigmoidf_(gi_[j] + gh_[j]);
    const float z = sigmoidf_(gi_[H + j] + gh_[H + j]);
    const float hn = gh_[2 * H + j];
    const float n = tanhf(gi_[2 * H + j] + r * hn);
    const float h = (1.f - z) * n + z * hp;
    hs[i] = h;
    *o = h;
    if (gsave) { gsave[j] = r; gsave[H + j] = z; gsave[2 * H + j] = n; gsave[3 * H + j] = hn; }
  }
}

// dhs: [ndir][B][H] running dL/dh_t (in place -> becomes the elementwise part of dL/dh_{t-1})
// dG:  [ndir][B][T][4H] = (dr_pre, dz_pre, dn_pre*r, dn_pre)   (first 3H = d gh, {0,1,3} = d gi)
__global__ void gru_cell_bwd_kernel(float* __restrict__ dhs, const float* __restrict__ d_out,
                                    const float* __restrict__ gates, const float* __restrict__ out,
                                    float* __restrict__ dG, const int32_t* __restrict__ lengths, int B, int T, int H,
                                    int ndir, int step) {
  const int64_t total = (int64_t)ndir * B * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = (int)(i % H);
    const int b = (int)((i / H) % B);
    const int dir = (int)(i / ((int64_t)H * B));
    const int t = dir == 0 ? step : T - 1 - step;
    float* dg = dG + (((int64_t)dir * B + b) * T + t) * (4 * H);
    if (lengths && t >= lengths[b]) {
      dg[j] = 0.f; dg[H + j] = 0.f; dg[2 * H + j] = 0.f; dg[3 * H + j] = 0.f;
      continue;   // dhs passes through unchanged, d_out[t] does not flow (out[t] == 0)
    }
    float dh = dhs[i];
    if (d_out) dh += d_out[((int64_t)b * T + t) * (ndir * H) + dir * H + j];
    const float* g = gates + (((int64_t)dir * B + b) * T + t) * (4 * H);
    const float r = g[j], z = g[H + j], n = g[2 * H + j], hn = g[3 * H + j];
    const int tp = dir == 0 ? t - 1 : t + 1;
    const float hp = (tp >= 0 && tp < T) ? out[((int64_t)b * T + tp) * (ndir * H) + dir * H + j] : 0.f;
    const float dn_pre = dh * (1.f - z) * (1.f - n * n);
    const float dz_pre = dh * (hp - n) * z * (1.f - z);
    const float dr_pre = dn_pre * hn * r * (1.f - r);
    dg[j] = dr_pre; dg[H + j] = dz_pre; dg[2 * H + j] = dn_pre * r; dg[3 * H + j] = dn_pre;
    dhs[i] = dh * z;
  }
}

static inline unsigned ew_blocks(int64_t n) {
  int64_t b = ceil_div(n, 256);
  return (unsigned)(b > 148 * 8 ? 148 * 8 : (b < 1 ? 1 : b));
}

static int check_layer(const tt_gru_layer* a) {
  TT_CHECK_ARG(a != nullptr, "gru: null args");
  TT_CHECK_ARG(a->B > 0 && a->T > 0 && a->In > 0 && a->H > 0, "gru: bad dims B=%d T=%d In=%d H=%d", a->B, a->T, a->In, a->H);
  TT_CHECK_ARG(a->ndir == 1 || a->ndir == 2, "gru: ndir must be 1 or 2");
  TT_CHECK_ARG((a->x != nullptr) != (a->ids != nullptr), "gru: give exactly one of x / ids");
  TT_CHECK_ARG(a->ids == nullptr || a->table != nullptr, "gru: ids without table");
  TT_CHECK_ARG(a->dropout_p >= 0.f && a->dropout_p < 1.f, "gru: dropout_p out of range");
  for (int d = 0; d < a->ndir; ++d)
    TT_CHECK_ARG(a->w_ih[d] && a->w_hh[d] && a->b_ih[d] && a->b_hh[d], "gru: null weight (dir %d)", d);
  TT_CHECK_ARG(a->out && a->h_n, "gru: null out/h_n");
  return TT_OK;
}

static bool use_umma(const tt_gru_layer* a) { return engine_mode() >= 1 && (a->H % 8) == 0; }
static bool needs_xeff(const tt_gru_layer* a) { return a->ids != nullptr || a->dropout_p > 0.f; }

// materialise the effective layer input (gathered and/or dropped-out) when it is not `x` itself
static int make_xeff(const tt_gru_layer* a, float* xeff, cudaStream_t s) {
  const int64_t ntok = (int64_t)a->B * a->T;
  if (a->ids)
    return launch_gather(a->ids, a->ids_are_i64, ntok, a->table, a->vocab, a->In, a->pad_id, xeff, a->dropout_p,
                         a->seed, a->offset, s);
  return launch_dropout(a->x, xeff, ntok * a->In, a->dropout_p, a->seed, a->offset, s);
}

}  // namespace tt

using namespace tt;

extern "C" {

int tt_embed_gather(const void* ids, int ids_are_i64, int64_t n_tokens, const float* table, int64_t vocab, int32_t E,
                    int64_t pad_id, float* out, void* stream) {
  TT_CHECK_ARG(ids && table && out && E > 0 && vocab > 0, "embed_gather: bad args");
  return launch_gather(ids, ids_are_i64, n_tokens, table, vocab, E, pad_id, out, 0.f, 0, 0, (cudaStream_t)stream);
}

int tt_dropout(const float* x, float* y, int64_t n, float p, uint64_t seed, uint64_t offset, void* stream) {
  TT_CHECK_ARG(x && y && p >= 0.f && p < 1.f, "dropout: bad args");
  return launch_dropout(x, y, n, p, seed, offset, (cudaStream_t)stream);
}

uint64_t tt_gru_layer_fwd_workspace(const tt_gru_layer* a) {
  if (!a) return 0;
  if (use_umma(a)) return gru_fwd_workspace_umma(a);
  Workspace w(nullptr, 0);
  const int64_t ntok = (int64_t)a->B * a->T;
  if (needs_xeff(a)) w.take<float>(ntok * a->In);
  w.take<float>(ntok * a->ndir * 3 * a->H);          // gi
  w.take<float>((int64_t)a->ndir * a->B * 3 * a->H);  // gh
  return w.used + 256;
}

int tt_gru_layer_fwd(const tt_gru_layer* a, void* stream) {
  TT_TRY(check_layer(a));
  cudaStream_t s = (cudaStream_t)stream;
  TT_CHECK_ARG(!a->save_for_backward || a->gates, "gru fwd: save_for_backward without gates buffer");
  if (use_umma(a)) return gru_layer_fwd_umma(a, s);
  const int B = a->B, T = a->T, In = a->In, H = a->H, nd = a->ndir;
  const int64_t ntok = (int64_t)B * T;
  Workspace w(a->workspace, a->workspace_bytes);
  float* xeff = needs_xeff(a) ? w.take<float>(ntok * In) : nullptr;
  float* gi = w.take<float>(ntok * nd * 3 * H);
  float* gh = w.take<float>((int64_t)nd * B * 3 * H);
  if (!w.ok()) { set_error("gru fwd: workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
  const float* x = a->x;
  if (xeff) { TT_TRY(make_xeff(a, xeff, s)); x = xeff; }

  // K-B: gi[:, dir*3H:(dir+1)*3H] = x W_ih[dir]^T + b_ih[dir]   (all timesteps, both directions)
  GemmArgs g{};
  for (int d = 0; d < nd; ++d) { g.A[d] = x; g.B[d] = a->w_ih[d]; g.C[d] = gi + (int64_t)d * 3 * H; g.bias[d] = a->b_ih[d]; }
  g.a_rs = In; g.a_cs = 1; g.b_rs = In; g.b_cs = 1; g.ldc = (int64_t)nd * 3 * H;
  g.M = (int)ntok; g.N = 3 * H; g.K = In; g.nbatch = nd; g.splitk = 1; g.alpha = 1.f; g.beta = 0.f;
  TT_TRY(gemm_f32(g, s));

  TT_CUDA(cudaMemsetAsync(a->h_n, 0, sizeof(float) * nd * B * H, s));
  // K-C: per step  gh = h W_hh^T + b_hh  then the fused cell
  GemmArgs r{};
  for (int d = 0; d < nd; ++d) { r.A[d] = a->h_n + (int64_t)d * B * H; r.B[d] = a->w_hh[d]; r.C[d] = gh + (int64_t)d * B * 3 * H; r.bias[d] = a->b_hh[d]; }
  r.a_rs = H; r.a_cs = 1; r.b_rs = H; r.b_cs = 1; r.ldc = 3 * H;
  r.M = B; r.N = 3 * H; r.K = H; r.nbatch = nd; r.splitk = 1; r.alpha = 1.f; r.beta = 0.f;
  const unsigned cb = ew_blocks((int64_t)nd * B * H);
  for (int step = 0; step < T; ++step) {
    TT_TRY(gemm_f32(r, s));
    tt::note_launch(); gru_cell_fwd_kernel<<<cb, 256, 0, s>>>(gi, gh, a->h_n, a->out, a->save_for_backward ? a->gates : nullptr,
                                           a->lengths, B, T, H, nd, step);
  }
  TT_LAUNCH_CHECK();
  return TT_OK;
}

uint64_t tt_gru_layer_bwd_workspace(const tt_gru_layer_grad* a) {
  if (!a) return 0;
  const tt_gru_layer* f = &a->fwd;
  if (use_umma(f)) return gru_bwd_workspace_umma(a);
  Workspace w(nullptr, 0);
  const int64_t ntok = (int64_t)f->B * f->T;
  if (needs_xeff(f)) w.take<float>(ntok * f->In);
  w.take<float>((int64_t)f->ndir * ntok * 4 * f->H);      // dG
  w.take<float>((int64_t)f->ndir * f->B * f->H);          // dhs
  w.take<float>((int64_t)f->ndir * 4 * f->H);             // column sums
  return w.used + 256;
}

int tt_gru_layer_bwd(const tt_gru_layer_grad* a, void* stream) {
  TT_CHECK_ARG(a != nullptr, "gru bwd: null args");
  const tt_gru_layer* f = &a->fwd;
  TT_TRY(check_layer(f));
  TT_CHECK_ARG(f->gates != nullptr, "gru bwd: forward did not save gates");
  cudaStream_t s = (cudaStream_t)stream;
  const int B = f->B, T = f->T, In = f->In, H = f->H, nd = f->ndir;
  const int64_t ntok = (int64_t)B * T;
  for (int d = 0; d < nd; ++d)
    TT_CHECK_ARG(a->d_w_ih[d] && a->d_w_hh[d] && a->d_b_ih[d] && a->d_b_hh[d], "gru bwd: null grad buffer");
  if (use_umma(f)) {
    TT_TRY(gru_layer_bwd_umma(a, s));
    if (a->d_x && f->dropout_p > 0.f && !f->ids)
      TT_TRY(launch_dropout(a->d_x, a->d_x, ntok * In, f->dropout_p, f->seed, f->offset, s));
    return TT_OK;
  }
  Workspace w(a->workspace, a->workspace_bytes);
  float* xeff = needs_xeff(f) ? w.take<float>(ntok * In) : nullptr;
  float* dG = w.take<float>((int64_t)nd * ntok * 4 * H);
  float* dhs = w.take<float>((int64_t)nd * B * H);
  float* csum = w.take<float>((int64_t)nd * 4 * H);
  if (!w.ok()) { set_error("gru bwd: workspace too small (%llu < %llu)", (unsigned long long)a->workspace_bytes, (unsigned long long)w.used); return TT_ERR_WORKSPACE; }
  const float* x = f->x;
  if (xeff) { TT_TRY(make_xeff(f, xeff, s)); x = xeff; }

  if (a->d_h_n) TT_CUDA(cudaMemcpyAsync(dhs, a->d_h_n, sizeof(float) * nd * B * H, cudaMemcpyDeviceToDevice, s));
  else TT_CUDA(cudaMemsetAsync(dhs, 0, sizeof(float) * nd * B * H, s));

  // per step (reverse order): gate grads, then dhs += dGH_t W_hh
  GemmArgs r{};
  r.a_rs = (int64_t)T * 4 * H; r.a_cs = 1; r.b_rs = 1; r.b_cs = H; r.ldc = H;
  r.M = B; r.N = H; r.K = 3 * H; r.nbatch = nd; r.splitk = 1; r.alpha = 1.f; r.beta = 1.f;
  const unsigned cb = ew_blocks((int64_t)nd * B * H);
  for (int step = T - 1; step >= 0; --step) {
    tt::note_launch(); gru_cell_bwd_kernel<<<cb, 256, 0, s>>>(dhs, a->d_out, f->gates, f->out, dG, f->lengths, B, T, H, nd, step);
    if (step == 0) break;
    for (int d = 0; d < nd; ++d) {
      const int t = d == 0 ? step : T - 1 - step;
      r.A[d] = dG + ((int64_t)d * ntok + t) * 4 * H;
      r.B[d] = f->w_hh[d]; r.C[d] = dhs + (int64_t)d * B * H; r.bias[d] = nullptr;
    }
    TT_TRY(gemm_f32(r, s));
  }
  TT_LAUNCH_CHECK();

  for (int d = 0; d < nd; ++d) {
    const float* dGd = dG + (int64_t)d * ntok * 4 * H;
    // biases: column sums of dG -> (r, z, nr, n)
    float* cs = csum + (int64_t)d * 4 * H;
    TT_TRY(colsum_f32(dGd, 4 * H, ntok, 4 * H, cs, s));
    TT_CUDA(cudaMemcpyAsync(a->d_b_hh[d], cs, sizeof(float) * 3 * H, cudaMemcpyDeviceToDevice, s));
    TT_CUDA(cudaMemcpyAsync(a->d_b_ih[d], cs, sizeof(float) * 2 * H, cudaMemcpyDeviceToDevice, s));
    TT_CUDA(cudaMemcpyAsync(a->d_b_ih[d] + 2 * H, cs + 3 * H, sizeof(float) * H, cudaMemcpyDeviceToDevice, s));

    // dW_hh[3H,H] = sum_{b,t} dGH[b,t,:]^T h_{prev}[b,t,:]  over the (T-1) steps that have a predecessor
    if (T > 1) {
      GemmArgs g{};
      g.A[0] = dGd; g.B[0] = f->out + (int64_t)d * H; g.C[0] = a->d_w_hh[d]; g.bias[0] = nullptr;
      g.a_rs = 1; g.a_cs = 4 * H; g.b_rs = 1; g.b_cs = (int64_t)nd * H; g.ldc = H;
      g.M = 3 * H; g.N = H; g.K = B * (T - 1); g.nbatch = 1; g.splitk = 0; g.alpha = 1.f; g.beta = 0.f;
      g.seg_len = T - 1; g.seg_stride = T;
      g.a_seg_off = d == 0 ? 1 : 0; g.b_seg_off = d == 0 ? 0 : 1;
      TT_TRY(gemm_f32(g, s));
    } else {
      TT_CUDA(cudaMemsetAsync(a->d_w_hh[d], 0, sizeof(float) * 3 * H * H, s));
    }
    // dW_ih rows [0,2H) from dG cols [0,2H); rows [2H,3H) from dG cols [3H,4H)
    for (int piece = 0; piece < 2; ++piece) {
      GemmArgs g{};
      g.A[0] = dGd + (piece ? 3 * H : 0); g.B[0] = x; g.C[0] = a->d_w_ih[d] + (int64_t)(piece ? 2 * H : 0) * In;
      g.a_rs = 1; g.a_cs = 4 * H; g.b_rs = 1; g.b_cs = In; g.ldc = In;
      g.M = piece ? H : 2 * H; g.N = In; g.K = (int)ntok; g.nbatch = 1; g.splitk = 0; g.alpha = 1.f; g.beta = 0.f;
      TT_TRY(gemm_f32(g, s));
    }
    // dX += dGI W_ih
    if (a->d_x) {
      for (int piece = 0; piece < 2; ++piece) {
        GemmArgs g{};
        g.A[0] = dGd + (piece ? 3 * H : 0); g.B[0] = f->w_ih[d] + (int64_t)(piece ? 2 * H : 0) * In; g.C[0] = a->d_x;
        g.a_rs = 4 * H; g.a_cs = 1; g.b_rs = 1; g.b_cs = In; g.ldc = In;
        g.M = (int)ntok; g.N = In; g.K = piece ? H : 2 * H; g.nbatch = 1; g.splitk = 1; g.alpha = 1.f;
        g.beta = (d == 0 && piece == 0) ? 0.f : 1.f;
        TT_TRY(gemm_f32(g, s));
      }
    }
  }
  if (a->d_x && f->dropout_p > 0.f && !f->ids)
    TT_TRY(launch_dropout(a->d_x, a->d_x, ntok * In, f->dropout_p, f->seed, f->offset, s));
  return TT_OK;
}

}  // extern "C"
