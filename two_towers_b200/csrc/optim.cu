// N4: clip_grad_norm_ + Adam / AdamW over a whole parameter list in three launches, no host sync.
// Replaces torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0) followed by optimizer.step()
// (simple_two_tower.py:193,239 AdamW(lr 1e-4, wd 0.01); train_margin.py:64-67 Adam + clip; train_enhanced.py:43 Adam).
// HBM-bound elementwise work: 16 B read + 12 B written per parameter (+4 B when the clipped gradient is stored back,
// as clip_grad_norm_ does in place).  One CTA per (tensor, 64 Ki-element chunk); the chunk list is built once by the
// host for a parameter set, so the 44 tensors of a two-tower model are one grid instead of 44 x 3 launches.
#include <mutex>
#include "common.cuh"

namespace tt {

__device__ __forceinline__ float block_sum_256(float v, float* sh) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = lane < 8 ? sh[lane] : 0.f;
    t = warp_sum(t);
  }
  return t;   // valid in warp 0
}

// partial[c] = sum of squares of chunk c of the gradients
__global__ void __launch_bounds__(256) opt_sqnorm_kernel(const tt_opt_tensor* __restrict__ tensors,
                                                         const tt_opt_chunk* __restrict__ chunks,
                                                         float* __restrict__ partial) {
  __shared__ float sh[8];
  const tt_opt_chunk ck = chunks[blockIdx.x];
  const tt_opt_tensor tn = tensors[ck.tensor];
  const int64_t n = min((int64_t)TT_OPT_CHUNK, tn.numel - ck.offset);
  const float* g = tn.grad + ck.offset;
  float acc = 0.f;
  if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    const float4* g4 = reinterpret_cast<const float4*>(g);
    const int64_t n4 = n >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += 256) {
      const float4 v = __ldg(g4 + i);
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (int64_t i = (n4 << 2) + threadIdx.x; i < n; i += 256) acc += g[i] * g[i];
  } else {
    for (int64_t i = threadIdx.x; i < n; i += 256) acc += g[i] * g[i];
  }
  const float t = block_sum_256(acc, sh);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// norm_out[0] = sqrt(sum partial) ; norm_out[1] = min(1, max_norm / (norm + 1e-6))     (clip_grad_norm_'s coefficient)
__global__ void __launch_bounds__(256) opt_finish_norm_kernel(const float* __restrict__ partial, int n, float max_norm,
                                                              float* __restrict__ norm_out) {
  __shared__ double sh[256];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) acc += (double)partial[i];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float norm = (float)sqrt(sh[0]);
    norm_out[0] = norm;
    norm_out[1] = max_norm > 0.f ? fminf(1.f, max_norm / (norm + 1e-6f)) : 1.f;
    const bool bad = !(fabsf(norm) <= 3.0e38f);       // inf or NaN
    norm_out[3] = bad ? 1.f : 0.f;
    if (bad) norm_out[2] += 1.f;
  }
}

// capturable mode: step count and bias corrections on the device (one thread), skipped when the gradients are non-finite
__global__ void opt_prepare_kernel(float* __restrict__ step_dev, float* __restrict__ consts, const float* __restrict__ skip,
                                   double lr, double beta1, double beta2) {
  if (skip && skip[3] != 0.f) return;
  const float t = step_dev[0] + 1.f;
  step_dev[0] = t;
  consts[0] = (float)(lr / (1.0 - pow(beta1, (double)t)));
  consts[1] = (float)sqrt(1.0 - pow(beta2, (double)t));
}

struct AdamConsts {       // every derived constant is formed in double on the host and rounded once, as torch does
  float decay, wd, omb1, beta2, omb2, eps, step_size, bc2_sqrt;
  int decoupled, store_grad;
};

__device__ __forceinline__ void adam_one(float& p, float& g, float& m, float& v, float coef, const AdamConsts& c) {
  g *= coef;
  float gg = g;
  if (c.decoupled) p *= c.decay;                     // AdamW: param.mul_(1 - lr * wd)
  else if (c.wd != 0.f) gg = fmaf(c.wd, p, gg);      // Adam: grad.add(param, alpha=wd)
  m = m + (gg - m) * c.omb1;                         // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(c.omb2, gg * gg, v * c.beta2);            // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = sqrtf(v) / c.bc2_sqrt + c.eps;
  p -= c.step_size * (m / denom);
}

__global__ void __launch_bounds__(256) opt_adam_kernel(const tt_opt_tensor* __restrict__ tensors,
                                                       const tt_opt_chunk* __restrict__ chunks,
                                                       const float* __restrict__ norm, const float* __restrict__ skip,
                                                       const float* __restrict__ consts_dev, AdamConsts c) {
  if (skip && skip[3] != 0.f) return;                  // non-finite gradients: leave everything as it is
  if (consts_dev) { c.step_size = consts_dev[0]; c.bc2_sqrt = consts_dev[1]; }
  const tt_opt_chunk ck = chunks[blockIdx.x];
  const tt_opt_tensor tn = tensors[ck.tensor];
  const int64_t n = min((int64_t)TT_OPT_CHUNK, tn.numel - ck.offset);
  float* p = tn.param + ck.offset;
  float* g = tn.grad + ck.offset;
  float* m = tn.exp_avg + ck.offset;
  float* v = tn.exp_avg_sq + ck.offset;
  const float coef = norm ? norm[1] : 1.f;
  const bool store = c.store_grad && norm;
  const uintptr_t al = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                       reinterpret_cast<uintptr_t>(v);
  int64_t done = 0;
  if ((al & 15) == 0) {
    const int64_t n4 = n >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += 256) {
      float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<const float4*>(g)[i];
      float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
      adam_one(pp.x, gg.x, mm.x, vv.x, coef, c);
      adam_one(pp.y, gg.y, mm.y, vv.y, coef, c);
      adam_one(pp.z, gg.z, mm.z, vv.z, coef, c);
      adam_one(pp.w, gg.w, mm.w, vv.w, coef, c);
      reinterpret_cast<float4*>(p)[i] = pp;
      reinterpret_cast<float4*>(m)[i] = mm;
      reinterpret_cast<float4*>(v)[i] = vv;
      if (store) reinterpret_cast<float4*>(g)[i] = gg;
    }
    done = n4 << 2;
  }
  for (int64_t i = done + threadIdx.x; i < n; i += 256) {
    float pp = p[i], gg = g[i], mm = m[i], vv = v[i];
    adam_one(pp, gg, mm, vv, coef, c);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (store) g[i] = gg;
  }
}

}  // namespace tt
using namespace tt;

extern "C" {

uint64_t tt_opt_step_workspace(int32_t nchunks) { return (uint64_t)(nchunks > 0 ? nchunks : 0) * sizeof(float) + 256; }

int tt_opt_step(const tt_opt_args* a, void* stream) {
  TT_CHECK_ARG(a != nullptr, "opt_step: null args");
  TT_CHECK_ARG(a->ntensors >= 0 && a->nchunks >= 0, "opt_step: negative counts");
  if (a->nchunks == 0) return TT_OK;
  TT_CHECK_ARG(a->tensors && a->chunks, "opt_step: null tensor / chunk table");
  TT_CHECK_ARG(a->lr >= 0.0 && a->eps >= 0.0 && a->beta1 >= 0.0 && a->beta1 < 1.0 && a->beta2 >= 0.0 && a->beta2 < 1.0,
               "opt_step: lr/eps/betas out of range");
  TT_CHECK_ARG(a->step_dev || (a->bias_correction1 > 0.0 && a->bias_correction2_sqrt > 0.0), "opt_step: bias corrections must be > 0 (step >= 1)");
  TT_CHECK_ARG(!a->step_dev || a->consts_dev, "opt_step: step_dev needs consts_dev");
  const bool do_norm = (a->flags & TT_OPT_NORM) != 0, do_step = (a->flags & TT_OPT_STEP) != 0;
  const bool clip = (a->flags & TT_OPT_CLIP) != 0, skip = (a->flags & TT_OPT_SKIP_NONFINITE) != 0;
  TT_CHECK_ARG(!(do_norm || clip || skip) || a->norm_out, "opt_step: TT_OPT_NORM / TT_OPT_CLIP / TT_OPT_SKIP_NONFINITE need norm_out");
  cudaStream_t s = (cudaStream_t)stream;
  if (do_norm) {
    TT_CHECK_ARG(a->workspace && a->workspace_bytes >= tt_opt_step_workspace(a->nchunks), "opt_step: workspace too small");
    float* partial = reinterpret_cast<float*>(a->workspace);
    tt::note_launch();
    opt_sqnorm_kernel<<<(unsigned)a->nchunks, 256, 0, s>>>(a->tensors, a->chunks, partial);
    TT_LAUNCH_CHECK();
    tt::note_launch();
    opt_finish_norm_kernel<<<1, 256, 0, s>>>(partial, a->nchunks, (float)a->max_norm, a->norm_out);
    TT_LAUNCH_CHECK();
  }
  if (!do_step) return TT_OK;
  AdamConsts c;
  c.decay = (float)(1.0 - a->lr * a->weight_decay); c.wd = (float)a->weight_decay;
  c.omb1 = (float)(1.0 - a->beta1); c.beta2 = (float)a->beta2; c.omb2 = (float)(1.0 - a->beta2); c.eps = (float)a->eps;
  c.step_size = a->step_dev ? 0.f : (float)(a->lr / a->bias_correction1);
  c.bc2_sqrt = a->step_dev ? 1.f : (float)a->bias_correction2_sqrt;
  if (a->step_dev) {
    tt::note_launch();
    opt_prepare_kernel<<<1, 1, 0, s>>>(a->step_dev, a->consts_dev, skip ? a->norm_out : nullptr, a->lr, a->beta1, a->beta2);
    TT_LAUNCH_CHECK();
  }
  c.decoupled = a->decoupled_wd != 0 && a->weight_decay != 0.0;
  c.store_grad = clip ? 1 : 0;
  tt::note_launch();
  opt_adam_kernel<<<(unsigned)a->nchunks, 256, 0, s>>>(a->tensors, a->chunks, clip ? a->norm_out : nullptr, skip ? a->norm_out : nullptr,
                                                       a->step_dev ? a->consts_dev : nullptr, c);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

// ------------------------------------------------------------------------------------------ allocation-free variant
// tt_opt_step_noalloc: tt_opt_step with the per-chunk partial sums of the gradient norm (its only scratch memory) held in a
// file-scope device array instead of a caller-provided workspace - for callers that cannot allocate (code under CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// TT_OPT_NOALLOC_MAX_CHUNKS floats (64 KiB) per GPU that loads the library, held until
// the process exits.  Every call on a device shares the array: nothing in it has to start at zero (opt_sqnorm_kernel
// writes partial[c] for every chunk c < nchunks before opt_finish_norm_kernel reads exactly those), but two calls that
// overlap on the GPU (different streams) race on it.
static __device__ __align__(256) float g_opt_noalloc_scratch[TT_OPT_NOALLOC_MAX_CHUNKS];

int tt_opt_step_noalloc(const tt_opt_args* a, void* stream) {
  TT_CHECK_ARG(a != nullptr, "opt_step_noalloc: null args");
  TT_CHECK_ARG(a->nchunks <= TT_OPT_NOALLOC_MAX_CHUNKS,
               "opt_step_noalloc: %d chunks exceed the static scratch (TT_OPT_NOALLOC_MAX_CHUNKS = %d chunks of %d elements); "
               "use tt_opt_step with a caller workspace", a->nchunks, TT_OPT_NOALLOC_MAX_CHUNKS, TT_OPT_CHUNK);
  // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that
  // device current (cudaGetSymbolAddress allocates nothing; it returns the address inside the already loaded module).
  static float* scratch_ptr[64];
  static std::mutex mu;
  int dev = 0;
  TT_CUDA(cudaGetDevice(&dev));
  TT_CHECK_ARG(dev >= 0 && dev < 64, "opt_step_noalloc: device ordinal %d out of range", dev);
  float* ptr = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!scratch_ptr[dev]) {
      float* q = nullptr;
      TT_CUDA(cudaGetSymbolAddress((void**)&q, g_opt_noalloc_scratch));
      scratch_ptr[dev] = q;
    }
    ptr = scratch_ptr[dev];
  }
  tt_opt_args b = *a;
  b.workspace = ptr;
  b.workspace_bytes = sizeof(float) * (uint64_t)TT_OPT_NOALLOC_MAX_CHUNKS;
  return tt_opt_step(&b, stream);
}

}  // extern "C"
