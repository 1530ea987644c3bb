// tt_gru_stack_*: the whole multi-layer GRU behind one ABI call.  Shapes the v2 pipeline covers (gru_v2.cu) run
// there; everything else (and engines 0 / 1) runs layer by layer through tt_gru_layer_* with fp32 batch-first
// activations, the per-layer `out` / `gates` living in the caller's `saved` buffer.
#include "common.cuh"
#include "gru_v2.cuh"

namespace tt {

__global__ void dropout_inplace_mask_kernel(float* x, int64_t n, float p, float inv_keep, uint64_t seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] *= dropout_scale(seed, 0, (uint64_t)i, p, inv_keep);
}

struct LayerBufs { float* out[8]; float* gates[8]; };
static uint64_t carve_layers(const tt_gru_stack* a, void* base, uint64_t size, LayerBufs* lb, bool gates) {
  Workspace w(base, size);
  for (int l = 0; l < a->L; ++l) {
    float* o = w.take<float>((int64_t)a->B * a->T * a->ndir * a->H);
    float* g = gates ? w.take<float>((int64_t)a->ndir * a->B * a->T * 4 * a->H) : nullptr;
    if (lb) { lb->out[l] = o; lb->gates[l] = g; }
  }
  return w.used + 256;
}
static void fill_layer(const tt_gru_stack* a, int l, const float* x_in, float* out, float* gates, tt_gru_layer* q) {
  memset(q, 0, sizeof(*q));
  q->B = a->B; q->T = a->T; q->H = a->H; q->ndir = a->ndir;
  q->In = l == 0 ? a->In : a->ndir * a->H;
  q->save_for_backward = gates ? 1 : 0;
  if (l == 0 && a->ids) {
    q->ids = a->ids; q->ids_are_i64 = a->ids_are_i64; q->table = a->table; q->vocab = a->vocab; q->pad_id = a->pad_id;
  } else {
    q->x = x_in;
  }
  q->dropout_p = l > 0 ? a->dropout_p : 0.f;
  q->seed = a->seed + (uint64_t)l * 0x9E3779B97F4A7C15ull;
  q->offset = 0;
  for (int d = 0; d < a->ndir; ++d) {
    q->w_ih[d] = a->weights[(l * a->ndir + d) * 4 + 0]; q->w_hh[d] = a->weights[(l * a->ndir + d) * 4 + 1];
    q->b_ih[d] = a->weights[(l * a->ndir + d) * 4 + 2]; q->b_hh[d] = a->weights[(l * a->ndir + d) * 4 + 3];
  }
  q->lengths = a->lengths;
  q->out = out; q->h_n = a->h_n + (int64_t)l * a->ndir * a->B * a->H; q->gates = gates;
}

static int check_stack(const tt_gru_stack* a) {
  TT_CHECK_ARG(a != nullptr, "gru stack: null args");
  TT_CHECK_ARG(a->B > 0 && a->T > 0 && a->In > 0 && a->H > 0 && a->L >= 1 && a->L <= 8, "gru stack: bad dims");
  TT_CHECK_ARG(a->ndir == 1 || a->ndir == 2, "gru stack: ndir must be 1 or 2");
  TT_CHECK_ARG((a->x != nullptr) != (a->ids != nullptr), "gru stack: give exactly one of x / ids");
  TT_CHECK_ARG(a->ids == nullptr || a->table != nullptr, "gru stack: ids without table");
  TT_CHECK_ARG(a->dropout_p >= 0.f && a->dropout_p < 1.f, "gru stack: dropout_p out of range");
  TT_CHECK_ARG(a->weights != nullptr && a->h_n != nullptr, "gru stack: null weights / h_n");
  for (int i = 0; i < a->L * a->ndir * 4; ++i) TT_CHECK_ARG(a->weights[i] != nullptr, "gru stack: null weight %d", i);
  return TT_OK;
}

}  // namespace tt
using namespace tt;

extern "C" {

uint64_t tt_gru_stack_saved_bytes(const tt_gru_stack* a) {
  if (!a) return 0;
  if (v2_supported(a)) return gru_stack_saved_bytes_v2(a);
  return carve_layers(a, nullptr, 0, nullptr, true);
}

uint64_t tt_gru_stack_fwd_workspace(const tt_gru_stack* a) {
  if (!a) return 0;
  if (v2_supported(a)) return gru_stack_fwd_workspace_v2(a);
  uint64_t mx = 0;
  for (int l = 0; l < a->L; ++l) {
    tt_gru_layer q;
    fill_layer(a, l, (const float*)16, (float*)16, nullptr, &q);
    const uint64_t n = tt_gru_layer_fwd_workspace(&q);
    mx = n > mx ? n : mx;
  }
  mx = align_up(mx, 256) + 256;
  if (!(a->save_for_backward && a->saved)) mx += carve_layers(a, nullptr, 0, nullptr, false);
  return mx;
}

int tt_gru_stack_fwd(const tt_gru_stack* a, void* stream) {
  TT_TRY(check_stack(a));
  cudaStream_t s = (cudaStream_t)stream;
  if (v2_supported(a)) return gru_stack_fwd_v2(a, s);
  TT_CHECK_ARG(a->seed_dev == nullptr || a->dropout_p == 0.f, "gru stack: seed_dev (graph-replay dropout) needs engine 2 and H %% 64 == 0");
  const bool save = a->save_for_backward && a->saved;
  LayerBufs lb;
  uint64_t layer_ws_off = 0;
  if (save) {
    if (carve_layers(a, a->saved, a->saved_bytes, &lb, true) > a->saved_bytes) { set_error("gru stack fwd: saved buffer too small"); return TT_ERR_WORKSPACE; }
  } else {
    layer_ws_off = align_up(carve_layers(a, a->workspace, a->workspace_bytes, &lb, false), 256);
    if (layer_ws_off > a->workspace_bytes) { set_error("gru stack fwd: workspace too small"); return TT_ERR_WORKSPACE; }
  }
  for (int l = 0; l < a->L; ++l) {
    tt_gru_layer q;
    fill_layer(a, l, l == 0 ? a->x : lb.out[l - 1], lb.out[l], save ? lb.gates[l] : nullptr, &q);
    q.workspace = (char*)a->workspace + layer_ws_off;
    q.workspace_bytes = a->workspace_bytes - layer_ws_off;
    TT_TRY(tt_gru_layer_fwd(&q, stream));
  }
  if (a->out)
    TT_CUDA(cudaMemcpyAsync(a->out, lb.out[a->L - 1], sizeof(float) * a->B * a->T * a->ndir * a->H, cudaMemcpyDeviceToDevice, s));
  return TT_OK;
}

static void fill_layer_grad(const tt_gru_stack_grad* a, int l, const LayerBufs& lb, const float* d_out, const float* d_h_n,
                            float* d_x, tt_gru_layer_grad* g) {
  const tt_gru_stack* f = &a->fwd;
  memset(g, 0, sizeof(*g));
  fill_layer(f, l, l == 0 ? f->x : lb.out[l - 1], lb.out[l], lb.gates[l], &g->fwd);
  g->d_out = d_out; g->d_h_n = d_h_n; g->d_x = d_x;
  for (int d = 0; d < f->ndir; ++d) {
    g->d_w_ih[d] = a->d_weights[(l * f->ndir + d) * 4 + 0]; g->d_w_hh[d] = a->d_weights[(l * f->ndir + d) * 4 + 1];
    g->d_b_ih[d] = a->d_weights[(l * f->ndir + d) * 4 + 2]; g->d_b_hh[d] = a->d_weights[(l * f->ndir + d) * 4 + 3];
  }
}

uint64_t tt_gru_stack_bwd_workspace(const tt_gru_stack_grad* a) {
  if (!a) return 0;
  const tt_gru_stack* f = &a->fwd;
  if (v2_supported(f)) return gru_stack_bwd_workspace_v2(a);
  LayerBufs lb{};
  uint64_t mx = 0;
  for (int l = 0; l < f->L; ++l) {
    tt_gru_layer_grad g;
    for (int i = 0; i < 8; ++i) { lb.out[i] = (float*)16; lb.gates[i] = (float*)16; }
    fill_layer_grad(a, l, lb, nullptr, nullptr, (float*)16, &g);
    const uint64_t n = tt_gru_layer_bwd_workspace(&g);
    mx = n > mx ? n : mx;
  }
  // two ping-pong buffers for the gradient wrt the inter-layer activations
  return align_up(mx, 256) + 2 * align_up(sizeof(float) * (uint64_t)f->B * f->T * f->ndir * f->H, 256) + 512;
}

int tt_gru_stack_bwd(const tt_gru_stack_grad* a, void* stream) {
  TT_CHECK_ARG(a != nullptr, "gru stack bwd: null args");
  const tt_gru_stack* f = &a->fwd;
  TT_TRY(check_stack(f));
  TT_CHECK_ARG(f->saved != nullptr && a->d_weights != nullptr, "gru stack bwd: forward state / grad buffers missing");
  cudaStream_t s = (cudaStream_t)stream;
  if (v2_supported(f)) return gru_stack_bwd_v2(a, s);
  LayerBufs lb;
  if (carve_layers(f, f->saved, f->saved_bytes, &lb, true) > f->saved_bytes) { set_error("gru stack bwd: saved buffer too small"); return TT_ERR_WORKSPACE; }
  const uint64_t act = align_up(sizeof(float) * (uint64_t)f->B * f->T * f->ndir * f->H, 256);
  if (2 * act + 256 > a->workspace_bytes) { set_error("gru stack bwd: workspace too small"); return TT_ERR_WORKSPACE; }
  float* dact[2] = {(float*)a->workspace, (float*)((char*)a->workspace + act)};
  const float* d_above = a->d_out;
  for (int l = f->L - 1; l >= 0; --l) {
    tt_gru_layer_grad g;
    float* d_x = l > 0 ? dact[l & 1] : a->d_x;
    fill_layer_grad(a, l, lb, d_above, a->d_h_n ? a->d_h_n + (int64_t)l * f->ndir * f->B * f->H : nullptr, d_x, &g);
    g.workspace = (char*)a->workspace + 2 * act;
    g.workspace_bytes = a->workspace_bytes - 2 * act;
    TT_TRY(tt_gru_layer_bwd(&g, stream));
    d_above = d_x;
  }
  return TT_OK;
}

}  // extern "C"
