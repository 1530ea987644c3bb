/* tt_b200_debug.h - test and measurement hooks of libtt_b200.so.
 *
 * NOT part of the drop-in boundary (include/tt_b200.h): nothing here replaces a reference interface.  These entry points
 * expose kernel-level pieces to the parity tests (tests/test_gpu_kernels.py), switch between kernel variants whose results
 * are identical (A/B timing in tools/ and bench.py) and read in-kernel timestamps.  They may change without an ABI version
 * bump. */
#ifndef TT_B200_DEBUG_H
#define TT_B200_DEBUG_H
#include "tt_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* plain fp32 GEMM used by the ops above, exported for kernel-level tests:
 * C[M,N] = alpha * A[M,K] * B[N,K]^T + beta * C + bias[N]   (bias may be NULL) */
int tt_gemm_nt(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
               const float* bias, int32_t M, int32_t N, int32_t K, float alpha, float beta, void* stream);

/* same product through the tcgen05 bf16x3 engine (operands split to bf16 hi/lo planes on the fly); exported for
 * kernel-level tests and roofline measurement.  trans != 0: A is given as [K,M], B as [K,N] (MN-major operands,
 * the wgrad form).  splitk 0 = automatic. */
uint64_t tt_gemm_x3_workspace(int32_t M, int32_t N, int32_t K);
/* operand formats of the v2 pipeline: flags bit0 = fp16 planes (else bf16), bit1 = single pass (else 3), bit2 = A given
 * as [K,M], bit3 = B given as [K,N].  Workspace as tt_gemm_x3_workspace. */
int tt_gemm_test(const float* A, const float* B, float* C, int32_t M, int32_t N, int32_t K, float alpha, float beta,
                 int32_t flags, void* workspace, uint64_t workspace_bytes, void* stream);
int tt_gemm_x3(const float* A, const float* B, float* C, const float* bias, int32_t M, int32_t N, int32_t K,
               float alpha, float beta, int32_t trans, int32_t splitk, void* workspace, uint64_t workspace_bytes,
               void* stream);

/* profiling aid: device buffer of 8 x uint64 %globaltimer marks per CTA of the most recent recurrence-step launch
 * (NULL = off): start, MMA thread ready, first operand tile landed, all MMAs issued, accumulator ready, epilogue
 * math done, TMA store drained, CTA end.  See tools/step_phases.py. */
int tt_debug_step_buffer(void* buf);
/* forward recurrence of the v2 pipeline: 4 (default) / 2 = ONE persistent cooperative kernel for all timesteps of a
 * layer with 4 / 2 epilogue warps per TMEM lane quarter; 0 = one launch per timestep.  Test / measurement aid. */
int tt_debug_seq_mode(int32_t mode);
/* BPTT of the v2 pipeline: 2 (default) = fused persistent kernel (dL/dh resident in tensor memory, hidden size <= 256),
 * run on 2-CTA clusters when the hidden size is 256 and the pairs fit in half the SMs (each CTA owns half the hidden
 * units; staged gate-gradient tiles are exchanged through distributed shared memory: lower latency per timestep, more
 * SM-time); 3 = CTA pairs whenever the hidden size is 256; 1 = always one CTA per (row tile, direction);
 * 0 = gate-gradient kernel + dh GEMM per timestep. */
int tt_debug_bwd_seq_mode(int32_t mode);
/* 1 (default): in a multi-layer backward the weight / bias gradient GEMMs of layer l run on a library-owned helper stream,
 * concurrently with the BPTT kernel of layer l-1, and are joined back into the caller's stream before the call returns
 * (to the caller all work stays ordered on its stream).  0: everything on the caller's stream (per-kernel timing). */
int tt_debug_wgrad_overlap(int32_t on);
/* time segments of the fused BPTT (default 4; 1 = one launch per layer): the sequence runs as n launches of the persistent
 * kernel and the weight-gradient GEMMs over a finished segment's tokens run on the helper stream under the next segment,
 * which takes them off the critical path for every layer.  Needs tt_debug_wgrad_overlap(1); sequences shorter than 16*n
 * steps use fewer segments. */
int tt_debug_bwd_segments(int32_t n);
/* Selection epilogue of the retrieval scan (tt_rows_topk*, D <= 128): 1 (default) keeps the query tile of a work unit in
 * shared memory while its corpus tiles stream through; 0 reloads it with every corpus tile; 5: resident, and 128-column
 * tiles with four accumulators in flight instead of 256-column tiles with two (measured slower).  Results are identical. */
int tt_debug_sel_resident(int32_t on);
/* Tensor passes of the candidate scan: 2 (default where D <= 128: x_hi y_hi + x_lo y_hi, acceptance bound 2^-13 + the
 * measured fp16 rounding residual of the corpus rows) or
 * 3 (adds x_hi y_lo, bound 2^-13).  The final top-k is bit-identical either way (exact re-score + acceptance test).
 * +10: without the per-row lower bound the partitions of a row share; +20: rows that fail the acceptance test go through
 * the tile-shaped exact scan even when they are few (both for measurement / tests only). */
int tt_debug_sel_passes(int32_t passes);
/* Byte offset, inside the workspace of a tt_rows_topk (indexed == 0) / tt_rows_topk_indexed (indexed != 0) call of this
 * shape, of the int32[Q] flags "row failed the acceptance test of the tensor route and went through the exact scan";
 * -1 when the call does not take the tensor route.  Tests use it to check that benign data never falls back. */
int64_t tt_debug_topk_fallback_flags(int32_t Q, int64_t N, int32_t D, int32_t mode, int32_t k, int32_t indexed);
/* GEMM engine: 1 (default) = large K-major problems (the 3-pass input-projection GEMMs) run on 2-CTA clusters with
 * tcgen05.mma.cta_group::2 (256 x 256 tiles, each CTA stages half of B); 0 = always one CTA per 128 x 256 tile.  Same
 * products in the same order per output element: results are bit-identical. */
int tt_debug_gemm_pairs(int32_t on);
/* Same kernel: how many corpus tiles ahead the CTAs of a column range prefetch into L2 (0 = off). */
int tt_debug_sel_prefetch(int32_t tiles);

#ifdef __cplusplus
}
#endif
#endif
