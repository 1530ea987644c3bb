// fp32 SIMT GEMM with generic operand strides, 2-way pointer batching, split-K and segmented
// (batch x time) reduction views.  This is the exact-fp32 engine: it serves the small GEMMs (head,
// per-step recurrence at tiny batch) and is the on-device cross-check for the tcgen05 bf16x3 engine.
// Tile 128x128x16, 256 threads, 8x8 register micro-tile split in 4+4 so shared-memory float4 reads are
// conflict free.  Each C element is ONE fp32 FMA chain over the reduction index in ascending order.
#include <stdarg.h>
#include <mutex>
#include "common.cuh"
#include "simt_tile.cuh"
#include "gru_umma.cuh"

namespace tt {

static thread_local char g_err[512] = "ok";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }
static unsigned long long g_launches = 0;
int g_engine = 2;
int engine_mode() { return g_engine; }
void note_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }
unsigned long long launches() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// ------------------------------------------------------------------------------ per-device one-time state
static std::mutex g_state_mu;
static int cur_dev() { int d = 0; if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d > 63) d = 0; return d; }
bool DeviceOnce::need() { std::lock_guard<std::mutex> l(g_state_mu); return !((mask >> cur_dev()) & 1ull); }
void DeviceOnce::done() { std::lock_guard<std::mutex> l(g_state_mu); mask |= 1ull << cur_dev(); }
int device_sm_count() {
  static int sms[64];
  const int d = cur_dev();
  std::lock_guard<std::mutex> l(g_state_mu);
  if (!sms[d] && (cudaDeviceGetAttribute(&sms[d], cudaDevAttrMultiProcessorCount, d) != cudaSuccess || sms[d] <= 0)) sms[d] = 148;
  return sms[d];
}
int device_cooperative() {
  static int coop[64];            // 0 unknown, 1 no, 2 yes
  const int d = cur_dev();
  std::lock_guard<std::mutex> l(g_state_mu);
  if (!coop[d]) { int v = 0; cudaDeviceGetAttribute(&v, cudaDevAttrCooperativeLaunch, d); coop[d] = v ? 2 : 1; }
  return coop[d] == 2;
}

// ------------------------------------------------------------------------------ per-family event timing
struct ProfSlot {
  static constexpr int CAP = 256;
  cudaEvent_t e0[CAP], e1[CAP];
  int tag[CAP];
  int created = 0, n = 0;
  double ms = 0.0;
  unsigned long long count = 0;
};
struct GemmRow { int M, N, K, nbatch, npass, flags; double ms; unsigned long long count; };
static GemmRow g_gemm_rows[64];
static int g_gemm_nrows = 0;
static thread_local int g_gemm_tag = -1;
static bool g_prof_on = false;
static ProfSlot g_prof[PROF_FAMILIES];
static void prof_drain(ProfSlot& p) {
  for (int i = 0; i < p.n; ++i) {
    float ms = 0.f;
    if (cudaEventSynchronize(p.e1[i]) == cudaSuccess && cudaEventElapsedTime(&ms, p.e0[i], p.e1[i]) == cudaSuccess) {
      p.ms += ms;
      ++p.count;
      if (&p == &g_prof[PROF_GEMM_TC] && p.tag[i] >= 0 && p.tag[i] < g_gemm_nrows) {
        g_gemm_rows[p.tag[i]].ms += ms;
        ++g_gemm_rows[p.tag[i]].count;
      }
    }
  }
  p.n = 0;
}
void prof_begin(int family, cudaStream_t s) {
  if (!g_prof_on) return;
  ProfSlot& p = g_prof[family];
  if (p.n == ProfSlot::CAP) prof_drain(p);
  if (p.n == p.created) {
    cudaEventCreate(&p.e0[p.created]);
    cudaEventCreate(&p.e1[p.created]);
    ++p.created;
  }
  p.tag[p.n] = family == PROF_GEMM_TC ? g_gemm_tag : -1;
  g_gemm_tag = -1;
  cudaEventRecord(p.e0[p.n], s);
}
void prof_begin_gemm(cudaStream_t s, int M, int N, int K, int nbatch, int npass, int flags) {
  if (!g_prof_on) return;
  int r = 0;
  for (; r < g_gemm_nrows; ++r) {
    const GemmRow& g = g_gemm_rows[r];
    if (g.M == M && g.N == N && g.K == K && g.nbatch == nbatch && g.npass == npass && g.flags == flags) break;
  }
  if (r == g_gemm_nrows) {
    if (g_gemm_nrows == 64) r = -1;
    else g_gemm_rows[g_gemm_nrows++] = GemmRow{M, N, K, nbatch, npass, flags, 0.0, 0ull};
  }
  g_gemm_tag = r;                      // consumed by the prof_begin of the launch that follows on this thread
}
void prof_end(int family, cudaStream_t s) {
  if (!g_prof_on) return;
  ProfSlot& p = g_prof[family];
  cudaEventRecord(p.e1[p.n], s);
  ++p.n;
}

template <bool A_RC, bool B_RC>
__global__ void __launch_bounds__(NT) gemm_f32_kernel(GemmArgs g) {
  __shared__ __align__(16) float As[BK][LDS];
  __shared__ __align__(16) float Bs[BK][LDS];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int batch = blockIdx.z / g.splitk, split = blockIdx.z % g.splitk;
  const float* __restrict__ A = g.A[batch];
  const float* __restrict__ B = g.B[batch];
  float* __restrict__ C = g.C[batch];
  const float* __restrict__ bias = g.bias[batch];
  const int bm = blockIdx.y * BM, bn = blockIdx.x * BN;
  int kchunk = (g.K + g.splitk - 1) / g.splitk;
  kchunk = (kchunk + BK - 1) / BK * BK;
  const int kbeg = split * kchunk;
  const int kend = min(g.K, kbeg + kchunk);

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float ra[8], rb[8];
  if (kbeg < kend) {
    load_frag<A_RC>(ra, A, g.a_rs, g.a_cs, bm, g.M, kbeg, kend, g.seg_len, g.seg_stride, g.a_seg_off, t);
    load_frag<B_RC>(rb, B, g.b_rs, g.b_cs, bn, g.N, kbeg, kend, g.seg_len, g.seg_stride, g.b_seg_off, t);
  }
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
    __syncthreads();
    store_frag<A_RC>(ra, As, t);
    store_frag<B_RC>(rb, Bs, t);
    __syncthreads();
    if (k0 + BK < kend) {
      load_frag<A_RC>(ra, A, g.a_rs, g.a_cs, bm, g.M, k0 + BK, kend, g.seg_len, g.seg_stride, g.a_seg_off, t);
      load_frag<B_RC>(rb, B, g.b_rs, g.b_cs, bn, g.N, k0 + BK, kend, g.seg_len, g.seg_stride, g.b_seg_off, t);
    }
    tile_fma(acc, As, Bs, tx, ty);
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = bm + tile_row(ty, i);
    if (m >= g.M) continue;
    float* crow = C + (int64_t)m * g.ldc;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = bn + tile_col(tx, j);
      if (n >= g.N) continue;
      float v = g.alpha * acc[i][j];
      if (g.splitk > 1 || g.shared_c) {
        if (split == 0 && bias) v += bias[n];
        atomicAdd(crow + n, v);
      } else {
        if (g.beta != 0.f) v += g.beta * crow[n];
        if (bias) v += bias[n];
        crow[n] = v;
      }
    }
  }
}

__global__ void scale_rows_kernel(float* C, int64_t ldc, int M, int N, float beta) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)M * N) return;
  float* p = C + (i / N) * ldc + (i % N);
  *p = beta == 0.f ? 0.f : beta * *p;
}

int gemm_f32(const GemmArgs& gin, cudaStream_t s) {
  GemmArgs g = gin;
  if (g.M <= 0 || g.N <= 0) return TT_OK;
  if (g.nbatch < 1) g.nbatch = 1;
  TT_CHECK_ARG(g.nbatch <= 2, "gemm_f32: nbatch %d > 2", g.nbatch);
  const int64_t tiles = ceil_div(g.M, BM) * ceil_div(g.N, BN) * g.nbatch;
  if (g.splitk <= 0) {
    g.splitk = 1;
    if (tiles < 148 && g.K >= 1024) {
      int64_t want = ceil_div(2 * 148, tiles);
      int64_t maxs = g.K / 256;
      g.splitk = (int)(want < maxs ? want : maxs);
      if (g.splitk < 1) g.splitk = 1;
    }
  }
  // two problems summed into one C (e.g. the two halves of a concatenated reduction dimension as ONE launch with twice
  // the CTAs): x + y is commutative, so the result does not depend on which CTA arrives first
  g.shared_c = g.nbatch == 2 && g.C[0] == g.C[1];
  if (g.splitk > 1 || g.shared_c) {
    // atomics accumulate into C: pre-scale it by beta (0 -> clear)
    if (g.beta != 1.f) {
      for (int b = 0; b < (g.shared_c ? 1 : g.nbatch); ++b) {
        int64_t n = (int64_t)g.M * g.N;
        tt::note_launch(); scale_rows_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(g.C[b], g.ldc, g.M, g.N, g.beta);
      }
      TT_LAUNCH_CHECK();
    }
  }
  dim3 grid((unsigned)ceil_div(g.N, BN), (unsigned)ceil_div(g.M, BM), (unsigned)(g.nbatch * g.splitk));
  const bool arc = g.a_cs == 1, brc = g.b_cs == 1;
  note_launch();
  if (arc && brc) gemm_f32_kernel<true, true><<<grid, NT, 0, s>>>(g);
  else if (arc && !brc) gemm_f32_kernel<true, false><<<grid, NT, 0, s>>>(g);
  else if (!arc && brc) gemm_f32_kernel<false, true><<<grid, NT, 0, s>>>(g);
  else gemm_f32_kernel<false, false><<<grid, NT, 0, s>>>(g);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

// ------------------------------------------------------------------------------ column sums
__global__ void colsum_kernel(const float* __restrict__ X, int64_t ld, int64_t M, int N, float* out) {
  __shared__ float red[8][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  float acc = 0.f;
  if (n < N)
    for (int64_t m = (int64_t)blockIdx.y * 8 + threadIdx.y; m < M; m += (int64_t)gridDim.y * 8) acc += X[m * ld + n];
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) v += red[i][threadIdx.x];
    atomicAdd(out + n, v);
  }
}
int colsum_f32(const float* X, int64_t ld, int64_t M, int32_t N, float* out, cudaStream_t s) {
  TT_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * N, s));
  if (M <= 0 || N <= 0) return TT_OK;
  int64_t gy = ceil_div(M, 64);
  if (gy > 128) gy = 128;
  tt::note_launch(); colsum_kernel<<<dim3((unsigned)ceil_div(N, 32), (unsigned)gy), dim3(32, 8), 0, s>>>(X, ld, M, N, out);
  TT_LAUNCH_CHECK();
  return TT_OK;
}

}  // namespace tt

extern "C" {
int tt_abi_version(void) { return 2; }
uint64_t tt_kernel_launches(void) { return tt::launches(); }
const char* tt_engine(void) {
  return tt::engine_mode() == 2 ? "tcgen05-f16x3-fwd/f16-bwd" : tt::engine_mode() == 1 ? "tcgen05-bf16x3" : "simt-fp32";
}
int tt_set_engine(int32_t mode) {
  if (mode < 0 || mode > 2) { tt::set_error("tt_set_engine: mode must be 0 (simt-fp32), 1 (tcgen05-bf16x3) or 2 (tcgen05 v2 stack)"); return TT_ERR_INVALID; }
  tt::g_engine = mode;
  return TT_OK;
}
const char* tt_last_error(void) { return tt::last_error(); }
int tt_prof_enable(int32_t on) {
  for (int f = 0; f < tt::PROF_FAMILIES; ++f) { tt::prof_drain(tt::g_prof[f]); tt::g_prof[f].ms = 0.0; tt::g_prof[f].count = 0; }
  tt::g_gemm_nrows = 0;
  tt::g_prof_on = on != 0;
  return TT_OK;
}
int tt_prof_gemm_rows(tt_prof_gemm_row* out, int32_t cap) {
  tt::prof_drain(tt::g_prof[tt::PROF_GEMM_TC]);
  int n = 0;
  for (; n < tt::g_gemm_nrows && n < cap; ++n) {
    const tt::GemmRow& g = tt::g_gemm_rows[n];
    out[n].M = g.M; out[n].N = g.N; out[n].K = g.K; out[n].nbatch = g.nbatch; out[n].npass = g.npass; out[n].flags = g.flags;
    out[n].total_ms = g.ms; out[n].launches = g.count;
  }
  return n;
}
const char* tt_prof_family(int32_t family) {
  static const char* names[tt::PROF_FAMILIES] = {"gru_seq_kernel (persistent forward recurrence)",
                                                 "gru_bwd_seq_kernel (persistent fused BPTT)",
                                                 "umma_gemm_kernel (tcgen05 GEMM engine)"};
  return family >= 0 && family < tt::PROF_FAMILIES ? names[family] : nullptr;
}
int tt_prof_read(int32_t family, double* total_ms, uint64_t* launches) {
  if (family < 0 || family >= tt::PROF_FAMILIES) { tt::set_error("tt_prof_read: family %d out of range", family); return TT_ERR_INVALID; }
  tt::prof_drain(tt::g_prof[family]);
  if (total_ms) *total_ms = tt::g_prof[family].ms;
  if (launches) *launches = tt::g_prof[family].count;
  return TT_OK;
}

int tt_gemm_nt(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, const float* bias,
               int32_t M, int32_t N, int32_t K, float alpha, float beta, void* stream) {
  tt::GemmArgs g{};
  g.A[0] = A; g.B[0] = B; g.C[0] = C; g.bias[0] = bias;
  g.a_rs = lda; g.a_cs = 1; g.b_rs = ldb; g.b_cs = 1; g.ldc = ldc;
  g.M = M; g.N = N; g.K = K; g.nbatch = 1; g.splitk = 1; g.alpha = alpha; g.beta = beta;
  return tt::gemm_f32(g, (cudaStream_t)stream);
}
}
