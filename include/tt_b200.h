/*
 * tt_b200.h - C ABI of libtt_b200.so: the two_towers training + retrieval hot path on B200 (sm_100a).
 *
 * The reference (mateomarin/two_towers) has no FFI: its boundary is the Python nn.Module surface and
 * every op below is whatever stock PyTorch dispatches at the cited call site (cuDNN GRU, cuBLAS, ATen).
 * Each entry point names the reference lines it replaces.  The Python host in two_towers_b200/ binds
 * these with ctypes and mirrors the reference classes on top.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to caller-owned memory; row-major, contiguous unless a stride is given
 *   - activations / parameters fp32; token ids int32 or int64 (flag); result indices int64 (torch convention)
 *   - no hidden allocation: scratch comes from the caller, sized by the matching *_workspace() function
 *   - returns 0 on success, negative on error (tt_last_error() gives the text); no exceptions cross the ABI
 *   - stream is a cudaStream_t passed as void*; the caller has made the right device current
 *   - re-entrant per stream; the only global state is one-time cudaFuncSetAttribute and the error string
 */
#ifndef TT_B200_H
#define TT_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TT_OK 0
#define TT_ERR_INVALID (-1)
#define TT_ERR_CUDA (-2)
#define TT_ERR_WORKSPACE (-3)

int tt_abi_version(void);
const char* tt_last_error(void);
/* which engine the GEMM-shaped ops run on: "simt-fp32", "tcgen05-bf16x3" or "tcgen05-f16x3-fwd/f16-bwd" */
const char* tt_engine(void);
/* select the engine for the GEMM-shaped GRU work:
 *   2 (default) = tcgen05 "v2" stack pipeline: time-major fp16 hi/lo planes, 3-pass forward, 1-pass loss-scaled
 *                 backward (hidden size multiple of 64, through tt_gru_stack_*; other shapes use engine 1)
 *   1           = tcgen05 bf16x3, layer by layer (hidden size multiple of 8; other shapes use engine 0)
 *   0           = simt-fp32 (exact fp32).  Test / cross-check aid. */
int tt_set_engine(int32_t mode);
/* number of CUDA kernels this library has launched in this process (monotonic) */
uint64_t tt_kernel_launches(void);
/* CUDA-event timing of the dominant kernel families, recorded on each launch's own stream (off by default; measurement
 * aid for bench.py's roofline object).  Families: 0 persistent forward recurrence, 1 persistent fused BPTT, 2 tcgen05
 * GEMM engine.  tt_prof_enable(on) clears the totals; tt_prof_read waits for the recorded events and returns the summed
 * device time and the number of launches since enable; tt_prof_family names a family (NULL past the last). */
int tt_prof_enable(int32_t on);
int tt_prof_read(int32_t family, double* total_ms, uint64_t* launches);
const char* tt_prof_family(int32_t family);
/* Family 2 broken down by GEMM shape since tt_prof_enable(1): one row per (M, N, K, batched problems, tensor passes,
 * flags); flags bit 0 A MN-major, bit 1 B MN-major, bits 4.. epilogue kind (0 row-major, 1 tiled, 2 top-k selection),
 * bit 8 split-K.  Flops of a row = 2 M N K nbatch (x npass issued).  Returns the number of rows written (<= cap). */
typedef struct { int32_t M, N, K, nbatch, npass, flags; double total_ms; uint64_t launches; } tt_prof_gemm_row;
int tt_prof_gemm_rows(tt_prof_gemm_row* out, int32_t cap);

/* ---------------------------------------------------------------------------------------------
 * K-A  embedding gather.  Replaces the per-word CPU lookup `self.word2vec[word]` + zero padding
 * (simple_two_tower.py:88-110, enhanced_two_tower.py:144-166) and nn.Embedding of the missing token-id
 * model (train_two_tower.py:29).  Row pad_id (if >= 0) reads as all zeros; ids outside [0,vocab) -> zeros.
 * out [n_tokens, E] fp32. */
int tt_embed_gather(const void* ids, int ids_are_i64, int64_t n_tokens, const float* table, int64_t vocab,
                    int32_t E, int64_t pad_id, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K-B + K-C  one GRU layer, all timesteps, 1 or 2 directions (dir 1 = reverse).  Replaces nn.GRU
 * (simple_two_tower.py:17-32, margin_two_tower.py:12-28, enhanced_two_tower.py:17-33): gate order r,z,n,
 *   r = s(gi_r+gh_r)  z = s(gi_z+gh_z)  n = tanh(gi_n + r*gh_n)  h = (1-z)*n + z*h,  zero initial state.
 * The input is either x [B,T,In] fp32, or (ids != NULL) rows of `table` gathered on the fly (In == table E).
 * dropout_p > 0 applies inverted dropout to the INPUT with the library's Philox stream (seed, offset)
 * - this is nn.GRU's inter-layer dropout when used for layer 1 (reference: dropout=0.1).
 * lengths (optional, [B]): steps t >= lengths[b] keep h and emit 0 (pack_padded_sequence semantics);
 * NULL = the reference behaviour (all T steps processed). */
typedef struct tt_gru_layer {
  int32_t B, T, In, H, ndir;
  int32_t save_for_backward;       /* 1: fill `gates` */
  const float* x;                  /* [B,T,In] or NULL when ids given */
  const void* ids;                 /* [B,T] or NULL */
  int32_t ids_are_i64;
  int32_t _pad0;
  const float* table;              /* [vocab,In] */
  int64_t vocab, pad_id;
  float dropout_p;
  int32_t _pad1;
  uint64_t seed, offset;
  const float* w_ih[2];            /* [3H,In] */
  const float* w_hh[2];            /* [3H,H]  */
  const float* b_ih[2];            /* [3H]    */
  const float* b_hh[2];            /* [3H]    */
  const int32_t* lengths;          /* [B] or NULL */
  float* out;                      /* [B,T,ndir*H] */
  float* h_n;                      /* [ndir,B,H]   */
  float* gates;                    /* [ndir,B,T,4H] (r,z,n,gh_n) when save_for_backward */
  void* workspace;
  uint64_t workspace_bytes;
} tt_gru_layer;

uint64_t tt_gru_layer_fwd_workspace(const tt_gru_layer* a);
int tt_gru_layer_fwd(const tt_gru_layer* a, void* stream);

/* K-C' fused BPTT of the same layer (autograd of nn.GRU at loss.backward(), simple_two_tower.py:237).
 * `fwd` carries the forward pointers (x/ids, weights, out, gates).  d_out / d_h_n may be NULL (= zero).
 * Weight grads are WRITTEN (not accumulated).  d_x may be NULL (frozen embedding table below layer 0). */
typedef struct tt_gru_layer_grad {
  tt_gru_layer fwd;
  const float* d_out;              /* [B,T,ndir*H] or NULL */
  const float* d_h_n;              /* [ndir,B,H]   or NULL */
  float* d_x;                      /* [B,T,In]     or NULL */
  float* d_w_ih[2];
  float* d_w_hh[2];
  float* d_b_ih[2];
  float* d_b_hh[2];
  void* workspace;
  uint64_t workspace_bytes;
} tt_gru_layer_grad;

uint64_t tt_gru_layer_bwd_workspace(const tt_gru_layer_grad* a);
int tt_gru_layer_bwd(const tt_gru_layer_grad* a, void* stream);

/* The whole multi-layer (bi)directional GRU in one call == nn.GRU(batch_first=True, num_layers=L) with zero initial
 * state (same reference lines as tt_gru_layer).  This is the entry point the Python host uses: between layers the
 * activations stay in library-private layouts (time-major 16-bit operand planes), the state backward needs lives in
 * the caller-owned opaque buffer `saved`.  weights / d_weights are HOST arrays of L*ndir*4 device pointers in nn.GRU's
 * _flat_weights order (w_ih, w_hh, b_ih, b_hh per layer, per direction).  Inter-layer dropout (p, seed) as above.
 * out may be NULL (the models only use h_n).  Shapes the fast pipeline does not cover run layer by layer. */
typedef struct tt_gru_stack {
  int32_t B, T, In, H, L, ndir;
  int32_t save_for_backward, ids_are_i64;
  const float* x;                  /* [B,T,In] or NULL when ids given */
  const void* ids;                 /* [B,T] or NULL */
  const float* table;              /* [vocab,In] */
  int64_t vocab, pad_id;
  float dropout_p;
  int32_t _pad0;
  uint64_t seed;
  const float* const* weights;     /* host array [L*ndir*4] of device pointers */
  const int32_t* lengths;          /* [B] or NULL */
  float* out;                      /* [B,T,ndir*H] of the last layer, or NULL */
  float* h_n;                      /* [L*ndir,B,H] */
  void* saved; uint64_t saved_bytes;             /* tt_gru_stack_saved_bytes(); needed when save_for_backward */
  void* workspace; uint64_t workspace_bytes;     /* tt_gru_stack_fwd_workspace() */
  const uint64_t* seed_dev;        /* optional DEVICE scalar added to `seed` when the kernels run: a CUDA graph that captured
                                      this call draws a new dropout mask on every replay by bumping it (engine 2 only) */
} tt_gru_stack;
uint64_t tt_gru_stack_saved_bytes(const tt_gru_stack* a);
uint64_t tt_gru_stack_fwd_workspace(const tt_gru_stack* a);
int tt_gru_stack_fwd(const tt_gru_stack* a, void* stream);

typedef struct tt_gru_stack_grad {
  tt_gru_stack fwd;                /* same arguments as the forward call (saved = the buffer it filled) */
  const float* d_out;              /* [B,T,ndir*H] or NULL */
  const float* d_h_n;              /* [L*ndir,B,H] or NULL */
  float* d_x;                      /* [B,T,In] or NULL */
  float* const* d_weights;         /* host array [L*ndir*4] of device pointers, WRITTEN */
  void* workspace; uint64_t workspace_bytes;     /* tt_gru_stack_bwd_workspace() */
} tt_gru_stack_grad;
uint64_t tt_gru_stack_bwd_workspace(const tt_gru_stack_grad* a);
int tt_gru_stack_bwd(const tt_gru_stack_grad* a, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K-D  projection head on the pooled state cat(h_n[-2], h_n[-1]):
 *   y1 = v W1^T + b1 ; y2 = relu(LayerNorm(y1; eps 1e-5)) ; [y3 = y2 W2^T + b2] ; [out = y/max(|y|,1e-12)]
 * Replaces query_proj/doc_proj/projection + F.normalize (simple_two_tower.py:35-61,
 * margin_two_tower.py:30-35,58-62, enhanced_two_tower.py:36-60).  v is given as two halves
 * (v_lo = h_n[-2], v_hi = h_n[-1], each [B,In/2]) so the concat is never materialised.
 * D2 == 0: no second Linear (margin).  dropout_p > 0 applies the reference's nn.Dropout(0.1) that follows the ReLU
 * (simple_two_tower.py:39, margin_two_tower.py:34; train mode only) with the library's Philox stream. */
typedef struct tt_head {
  int32_t B, In, P1, D2, normalize;
  float dropout_p;
  uint64_t seed, offset;
  const float* v_lo; const float* v_hi;
  const float* w1; const float* b1; const float* ln_w; const float* ln_b;
  const float* w2; const float* b2;
  float* y1;        /* [B,P1] pre-LN (saved) */
  float* stats;     /* [B,2] mean, rstd (saved) */
  float* y2;        /* [B,P1] post ReLU (+dropout) (saved; == output when D2==0 and !normalize) */
  float* y3;        /* [B,D2] pre-normalise (saved; == output when !normalize); unused if D2==0 */
  float* out;       /* [B,Dout] normalised output when normalize, else may be NULL */
  float* norm;      /* [B] when normalize */
  const uint64_t* seed_dev;   /* optional DEVICE scalar added to `seed` at run time (CUDA-graph replay), or NULL */
  void* workspace; uint64_t workspace_bytes;   /* tt_head_fwd_workspace(): operand planes of the tensor route; NULL -> the
                                                  exact-fp32 SIMT GEMMs run instead (as under tt_set_engine(0)) */
} tt_head;
uint64_t tt_head_fwd_workspace(const tt_head* a);
int tt_head_fwd(const tt_head* a, void* stream);

typedef struct tt_head_grad {
  tt_head fwd;
  const float* d_out;              /* [B,Dout] */
  float* d_v_lo; float* d_v_hi;    /* [B,In/2] each */
  float* d_w1; float* d_b1; float* d_ln_w; float* d_ln_b; float* d_w2; float* d_b2;   /* written */
  void* workspace; uint64_t workspace_bytes;
} tt_head_grad;
uint64_t tt_head_bwd_workspace(const tt_head_grad* a);
int tt_head_bwd(const tt_head_grad* a, void* stream);

/* F.normalize(p=2, dim=1, eps) alone (margin_two_tower.py:39-40,77-78; enhanced_two_tower.py:74-75) */
int tt_l2norm_fwd(const float* x, int32_t B, int32_t D, float eps, float* y, float* norm, void* stream);
int tt_l2norm_bwd(const float* y, const float* norm, const float* dy, int32_t B, int32_t D, float eps,
                  float* dx, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K-E  in-batch score + cross-entropy, never materialising the [Bl,Bg] matrix.
 *   logit[i][j] = <x_i, y_j> * inv_temp - margin * (j != pos_offset + i),   i < Bl, j < Bg
 *   row_loss[i] = logsumexp_j logit[i][j] - logit[i][pos_offset+i]
 * Replaces matmul/temperature + [eye-margin shift] + F.cross_entropy(sim, arange):
 *   simple_two_tower.py:73-78 (call twice with x,y swapped), margin_two_tower.py:80-85,
 *   enhanced_two_tower.py:77-82, 92-101 (margin > 0).  Under data parallelism x = this rank's rows,
 *   y = the all-gathered matrix and pos_offset = rank * Bl.
 * bwd: dx[i] = gscale*inv_temp*sum_j (p_ij - 1{j=pos}) y_j ;  dy[j] = gscale*inv_temp*sum_i (p_ij - 1{..}) x_i
 *   dx, dy are WRITTEN. */
uint64_t tt_score_ce_workspace(int32_t Bl, int32_t Bg, int32_t D);
int tt_score_ce_fwd(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                    float inv_temp, float margin, float* row_loss, float* row_lse,
                    void* workspace, uint64_t workspace_bytes, void* stream);
int tt_score_ce_bwd(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                    float inv_temp, float margin, const float* row_lse, float gscale,
                    float* dx, float* dy, void* workspace, uint64_t workspace_bytes, void* stream);
/* same, with the upstream gradient left on the device: the effective scale is gscale * (*gscale_dev) (gscale_dev may be
 * NULL).  loss.backward() hands the library a device scalar; reading it on the host would drain the launch queue at the
 * start of every backward pass. */
int tt_score_ce_bwd_dev(const float* x, int32_t Bl, const float* y, int32_t Bg, int32_t D, int32_t pos_offset,
                        float inv_temp, float margin, const float* row_lse, float gscale, const float* gscale_dev,
                        float* dx, float* dy, void* workspace, uint64_t workspace_bytes, void* stream);

/* K-F / K-H  row-wise exact-fp32 top-k of  score(i,j) = <x_i,y_j>  (mode 0)  or cosine (mode 1, norms
 * clamped at 1e-8 like F.cosine_similarity).  Every dot product is one fp32 FMA chain over d = 0..D-1 in
 * ascending order, so equal inputs give bit-equal scores; ties break to the LOWEST index.
 * mask_pos_offset >= 0: score(i, mask_pos_offset+i) := -1 before selection (get_hard_negatives,
 * enhanced_two_tower.py:123-133).  index_base is added to the emitted indices (global ids of a corpus shard:
 * the retrieval scan simple_two_tower.py:333-340 / validate_enhanced.py:73-80 / app.py:94-101).
 * k <= 32.  out_val [Q,k] fp32 descending, out_idx [Q,k] int64.  Rows with N < k are padded (-inf, -1). */
uint64_t tt_rows_topk_workspace(int32_t Q, int64_t N, int32_t D, int32_t mode);
int tt_rows_topk(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                 int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx,
                 void* workspace, uint64_t workspace_bytes, void* stream);

/* K-H with a prebuilt corpus index.  The reference encodes the corpus once and then searches it query by query
 * (simple_two_tower.py:318-340, validate_enhanced.py:61-80, app.py:38-60,94-101); the per-corpus part of tt_rows_topk
 * (row norms and the 16-bit hi/lo planes the tensor route scores against) is likewise built once into a caller-owned
 * device buffer of tt_corpus_index_bytes(N, D) bytes and reused by every tt_rows_topk_indexed call on the same y, N, D
 * and mode (cosine planes hold normalised rows: an index is valid for the mode it was built with only).
 * Results are bit-identical to tt_rows_topk. */
uint64_t tt_corpus_index_bytes(int64_t N, int32_t D);
int tt_corpus_index_build(const float* y, int64_t N, int32_t D, int32_t mode, void* index, uint64_t index_bytes,
                          void* stream);
uint64_t tt_rows_topk_indexed_workspace(int32_t Q, int64_t N, int32_t D, int32_t mode);
int tt_rows_topk_indexed(const float* x, int32_t Q, const float* y, int64_t N, int32_t D, int32_t k, int32_t mode,
                         int64_t mask_pos_offset, int64_t index_base, float* out_val, int64_t* out_idx,
                         const void* index, uint64_t index_bytes, void* workspace, uint64_t workspace_bytes,
                         void* stream);

/* K-I  merge S per-shard top-k lists ([S,Q,k] each, as all-gathered) into the global top-k:
 * order by score descending, then lowest global id.  Entries with idx < 0 are ignored. */
int tt_topk_merge(const float* vals, const int64_t* idx, int32_t S, int32_t Q, int32_t k,
                  float* out_val, int64_t* out_idx, void* stream);

/* MRR@k / recall: rank (1-based) of the first entry of topk_idx[q] equal to relevant[q], 0 if none;
 * rr = 1/rank (simple_two_tower.py:364-373, compare_models.py:55-74). */
int tt_mrr_ranks(const int64_t* topk_idx, const int64_t* relevant, int32_t Q, int32_t k,
                 int32_t* first_rank, float* rr, void* stream);

/* K-G  explicit-negatives hinge (enhanced_two_tower.py:102-121):
 *   row_loss[i] = max(0, margin - cos(q_i,pos_i) + mean_k cos(q_i, neg[i*k+kk])),  cos eps 1e-8
 * bwd writes dq, dpos, dneg scaled by gscale (= upstream_grad / B). */
int tt_hinge_fwd(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D,
                 float margin, float* row_loss, void* stream);
int tt_hinge_bwd(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D,
                 float margin, float gscale, float* dq, float* dpos, float* dneg, void* stream);
int tt_hinge_bwd_dev(const float* q, const float* pos, const float* neg, int32_t B, int32_t k, int32_t D,
                     float margin, float gscale, const float* gscale_dev, float* dq, float* dpos, float* dneg,
                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * N4  fused clip_grad_norm_ + Adam / AdamW step over a parameter list.  Replaces
 * torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1.0) + optimizer.step():
 * simple_two_tower.py:193,239 (AdamW lr 1e-4 wd 0.01), train_margin.py:64-67 (Adam + clip), train_enhanced.py:43.
 * `tensors` and `chunks` are DEVICE arrays the host builds once per parameter set: every tensor is covered by
 * chunks of at most TT_OPT_CHUNK elements (chunk = tensor index + element offset; offsets multiples of 4).
 * Math (per element, torch's single-tensor Adam): g *= coef; AdamW: p *= 1 - lr*wd, Adam: g += wd*p;
 * m += (g-m)(1-b1); v = b2 v + (1-b2) g^2; p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps).
 * flags: TT_OPT_NORM  norm_out[0] = total L2 norm of the gradients of THIS table, norm_out[1] = clip coefficient
 *                     min(1, max_norm / (norm + 1e-6))  (1 when max_norm <= 0); needs the workspace
 *        TT_OPT_STEP  run the update
 *        TT_OPT_CLIP  the update scales every gradient by norm_out[1] (from this or an earlier call - several
 *                     parameter groups share one norm) and stores the scaled gradient back, as clip_grad_norm_ does
 *        TT_OPT_SKIP_NONFINITE  (with or after a TT_OPT_NORM call) a non-finite total norm - an overflow of the fp16
 *                     gradient planes of the single-pass BPTT, DESIGN.md section 3 - sets norm_out[3] = 1, adds 1 to
 *                     norm_out[2] and the update leaves parameters, moments and (capturable) the step count
 *                     untouched: what torch.cuda.amp.GradScaler does with an inf gradient
 * step_dev != NULL ("capturable", for CUDA-graph replay): the step count lives on the device; every TT_OPT_STEP call
 *                     that is not skipped adds 1 to *step_dev and derives both bias corrections from it in double
 *                     on the device (bias_correction1/2_sqrt of the struct are ignored). */
#define TT_OPT_CHUNK 65536
#define TT_OPT_NORM 1
#define TT_OPT_STEP 2
#define TT_OPT_CLIP 4
#define TT_OPT_SKIP_NONFINITE 8
typedef struct { float* param; float* grad; float* exp_avg; float* exp_avg_sq; int64_t numel; } tt_opt_tensor;
typedef struct { int32_t tensor; int32_t reserved; int64_t offset; } tt_opt_chunk;
typedef struct {
  const tt_opt_tensor* tensors; const tt_opt_chunk* chunks;
  int32_t ntensors, nchunks, flags, decoupled_wd;      /* decoupled_wd: 1 = AdamW, 0 = Adam (L2 folded into the gradient) */
  /* hyper-parameters as doubles: torch derives 1-beta, lr/bc1, 1-lr*wd in double and rounds ONCE to fp32 - so does this */
  double lr, beta1, beta2, eps, weight_decay, max_norm;
  double bias_correction1, bias_correction2_sqrt;      /* 1 - beta1^t and sqrt(1 - beta2^t) for this step t >= 1 (host) */
  float* norm_out;                                     /* device [4]: norm, clip coefficient, skipped steps, skip flag */
  void* workspace; uint64_t workspace_bytes;
  float* step_dev;                                     /* device scalar step count or NULL */
  float* consts_dev;                                   /* device [2] scratch for (lr / bc1, sqrt(bc2)); needed with step_dev */
} tt_opt_args;
uint64_t tt_opt_step_workspace(int32_t nchunks);
int tt_opt_step(const tt_opt_args* a, void* stream);
/* Allocation-free variant of tt_opt_step.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments and results as tt_opt_step, except that a->workspace / a->workspace_bytes are IGNORED: the scratch
 * of the gradient-norm pass (one float per chunk) lives in a file-scope `static __device__` array of
 * TT_OPT_NOALLOC_MAX_CHUNKS floats (64 KiB on every GPU that loads the library, held until the process exits), i.e.
 * parameter sets of up to TT_OPT_NOALLOC_MAX_CHUNKS chunks (>= 1.07 G elements when tensors fill their chunks).  More
 * chunks -> TT_ERR_INVALID with a message naming the limit.  The call itself performs no cudaMalloc / cudaFree and calls
 * no allocating routine; the first call on a device resolves the array's device address (cudaGetSymbolAddress) - make
 * that first call outside stream capture.  All calls on a device share the array and it keeps the previous call's
 * contents (none of it needs to start at zero); calls that overlap on the GPU (different streams) race on it. */
#define TT_OPT_NOALLOC_MAX_CHUNKS 16384
int tt_opt_step_noalloc(const tt_opt_args* a, void* stream);

/* ---------------------------------------------------------------------------------------------
 * N1  host input pipeline: text -> token ids.  HOST functions taking HOST pointers - the one exception to
 * the device-pointer convention above.  Replaces the per-word CPU embedding lookup of SimpleDataset.text_to_embedding
 * (simple_two_tower.py:88-110 == enhanced_two_tower.py:144-166): `words = text.lower().split()[:max_len]`, words not
 * in the vocabulary are skipped, the row is padded at the END with pad_id (the table row that reads as zeros).
 * Vocabulary: nwords UTF-8 words, word i = words_blob[word_offsets[i] .. word_offsets[i+1]); its id is ids[i] (ids == NULL:
 * i).  Texts are passed the same way (UTF-8 blob + ntexts+1 offsets); ASCII letters are lower-cased, whitespace is
 * Python's str.isspace() set.  mode 0 = reference order (truncate to max_len words, then drop unknown words),
 * mode 1 = drop unknown words, then truncate.  out_ids [ntexts, max_len] int32 or int64, out_lengths [ntexts] (may be
 * NULL) = number of real tokens per row.  nthreads <= 0: all host threads. */
typedef struct tt_vocab tt_vocab;
tt_vocab* tt_vocab_create(const char* words_blob, const int64_t* word_offsets, int64_t nwords, const int32_t* ids);
void tt_vocab_destroy(tt_vocab* v);
int64_t tt_vocab_size(const tt_vocab* v);
int tt_tokenize_batch(const tt_vocab* v, const char* text_blob, const int64_t* text_offsets, int64_t ntexts,
                      int32_t max_len, int64_t pad_id, int32_t mode, int32_t ids_are_i64, void* out_ids,
                      int32_t* out_lengths, int32_t nthreads);

/* Test / measurement entry points (tt_gemm_*, tt_debug_*) are declared in tt_b200_debug.h: they are exported by the same
 * library but are not part of the interface a host integrates against. */

/* inverted dropout with the library's Philox4x32-10 stream: y = x * keep(seed, offset+i) / (1-p) */
int tt_dropout(const float* x, float* y, int64_t n, float p, uint64_t seed, uint64_t offset, void* stream);

#ifdef __cplusplus
}
#endif
#endif
