// Shared helpers for libtt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/tt_b200.h"
#include "../../include/tt_b200_debug.h"

namespace tt {

void set_error(const char* fmt, ...);
void note_launch();                 // counts kernel launches (bench.py reports them as gpu_launches)
unsigned long long launches();
// CUDA-event timing of a kernel family on its launch stream (off unless tt_prof_enable(1)): bench.py's roofline leg.
enum ProfFamily { PROF_GRU_SEQ_FWD = 0, PROF_GRU_SEQ_BWD = 1, PROF_GEMM_TC = 2, PROF_FAMILIES = 3 };
void prof_begin(int family, cudaStream_t s);
void prof_end(int family, cudaStream_t s);
// per-shape table of the GEMM engine (family PROF_GEMM_TC): launches of one (M, N, K, batches, passes, layout, epilogue)
// share a row; read with tt_prof_gemm_rows
void prof_begin_gemm(cudaStream_t s, int M, int N, int K, int nbatch, int npass, int flags);
struct ProfScope {
  int f; cudaStream_t s;
  ProfScope(int family, cudaStream_t stream) : f(family), s(stream) { prof_begin(f, s); }
  ~ProfScope() { prof_end(f, s); }
};

#define TT_CHECK_ARG(cond, ...)                      \
  do {                                               \
    if (!(cond)) {                                   \
      tt::set_error(__VA_ARGS__);                    \
      return TT_ERR_INVALID;                         \
    }                                                \
  } while (0)

#define TT_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      tt::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return TT_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

#define TT_LAUNCH_CHECK() TT_CUDA(cudaGetLastError())

#define TT_TRY(expr)          \
  do {                        \
    int _r = (expr);          \
    if (_r != TT_OK) return _r; \
  } while (0)

// Per-DEVICE one-time state (cudaFuncSetAttribute applies to the current device's context; SM counts differ per device):
// one process may drive several GPUs, so nothing device-dependent is cached in a plain static.
struct DeviceOnce {
  unsigned long long mask = 0;        // bit d: done on device d (guarded by the library's state mutex)
  bool need();                        // true if the current device has not been marked yet
  void done();
};
int device_sm_count();                // SMs of the current device (cached per device)
int device_cooperative();             // cudaDevAttrCooperativeLaunch of the current device

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline uint64_t align_up(uint64_t a, uint64_t b) { return (a + b - 1) / b * b; }

// fp32 [rows][cols] -> fp16 hi/lo operand planes [rows][pitch] (pad columns zero); defined in score.cu
int split_rows_f16(const float* x, int64_t rows, int cols, int64_t pitch, void* hi, void* lo, cudaStream_t s);
int split_rows_f16_scaled(const float* x, int64_t rows, int cols, int64_t pitch, void* hi, void* lo, const float* row_scale,
                          cudaStream_t s);

// bump allocator over the caller's workspace
struct Workspace {
  char* base;
  uint64_t size, used;
  Workspace(void* p, uint64_t n) : base((char*)p), size(n), used(0) {}
  template <typename T>
  T* take(uint64_t count) {
    uint64_t off = align_up(used, 256);
    used = off + count * sizeof(T);
    return (T*)(base + off);
  }
  bool ok() const { return used <= size && (base != nullptr || used == 0); }
};

// ---- Philox4x32-10 (counter-based; same construction as curand's, own stream)
__host__ __device__ inline void philox_round(uint32_t (&c)[4], uint32_t (&k)[2]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#ifdef __CUDA_ARCH__
  uint32_t hi0 = __umulhi(M0, c[0]), hi1 = __umulhi(M1, c[2]);
#else
  uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c[0]) >> 32), hi1 = (uint32_t)(((uint64_t)M1 * c[2]) >> 32);
#endif
  uint32_t lo0 = M0 * c[0], lo1 = M1 * c[2];
  uint32_t n0 = hi1 ^ c[1] ^ k[0], n1 = lo1, n2 = hi0 ^ c[3] ^ k[1], n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__host__ __device__ inline void philox4x32(uint64_t seed, uint64_t ctr, uint32_t (&out)[4]) {
  uint32_t c[4] = {(uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u};
  uint32_t k[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    philox_round(c, k);
    k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u;
  }
  out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
}
// keep-mask scale for element i: 1/(1-p) or 0.  4 elements share one Philox call.
__device__ inline float dropout_scale(uint64_t seed, uint64_t offset, uint64_t i, float p, float inv_keep) {
  uint64_t e = offset + i;
  uint32_t r[4];
  philox4x32(seed, e >> 2, r);
  float u = (float)(r[e & 3] >> 8) * (1.0f / 16777216.0f);   // [0,1)
  return u >= p ? inv_keep : 0.0f;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// -------------------------------------------------------------------------------- SIMT fp32 GEMM
// C[m,n] = alpha * sum_r A(m,r) * B(n,r) + beta*C[m,n] + bias[n]
// A(m,r) = A[m*a_rs + R(r,a_seg_off)*a_cs], B(n,r) = B[n*b_rs + R(r,b_seg_off)*b_cs]
// R(r,off) = r (seg_len==0) or (r/seg_len)*seg_stride + r%seg_len + off   (strided [batch][time] views)
struct GemmArgs {
  const float* A[2]; const float* B[2]; float* C[2]; const float* bias[2];
  int64_t a_rs, a_cs, b_rs, b_cs, ldc;
  int32_t M, N, K;
  int32_t nbatch, splitk;
  int32_t seg_len, seg_stride, a_seg_off, b_seg_off;
  float alpha, beta;
  int32_t shared_c;   // set by gemm_f32: the two batched problems add into ONE C (C[0] == C[1]) -> atomic accumulation
};
int gemm_f32(const GemmArgs& g, cudaStream_t s);

// column sums of X[M, N] (row stride ld) -> out[N] (written)
int colsum_f32(const float* X, int64_t ld, int64_t M, int32_t N, float* out, cudaStream_t s);

}  // namespace tt
