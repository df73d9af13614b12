/*
 * m3ae_b200.h — C-ABI of the B200-native M3AE fusion hot path (libm3ae_b200.so).
 *
 * The reference (better62/MM-VQA-Healthcare) is pure Python: it has no FFI/plugin interface, every
 * "kernel" is an implicit PyTorch library call.  This header is therefore the boundary a maintainer
 * would bind with ctypes underneath the reference's nn.Modules; each entry point names the reference
 * call sites (file:line, relative to the reference root) whose arithmetic it replaces.
 *
 * Conventions (all entry points):
 *   - plain pointers and sizes only; every buffer (inputs, outputs, saved-for-backward, workspace) is
 *     caller-allocated DEVICE memory; the library never allocates, frees or retains caller memory;
 *   - returns 0 on success, a negative M3AE_ERR_* code otherwise; m3ae_last_error() (thread-local)
 *     describes the failure; host wrappers raise RuntimeError with that text;
 *   - asynchronous on the cudaStream_t passed as `stream` (void* to keep CUDA types out of the ABI);
 *   - sm_100a only: refuses to run on any other architecture; there is no CPU fallback;
 *   - dtype codes: M3AE_F32 / M3AE_BF16.  bf16 is the product path (tcgen05 tensor cores, fp32
 *     accumulate); f32 is the strict-parity path (1e-4 gate).
 */
#ifndef M3AE_B200_H_
#define M3AE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define M3AE_B200_VERSION 100

#define M3AE_F32 0
#define M3AE_BF16 1

#define M3AE_ACT_NONE 0
#define M3AE_ACT_GELU 1 /* exact erf GELU — HF ACT2FN["gelu"], bert_model.py:421 */
#define M3AE_ACT_RELU 2 /* decoder FFN, m3ae_decoder.py:45 */
#define M3AE_ACT_QUICKGELU 3 /* x * sigmoid(1.702 x): CLIP visual tower MLP, vision_encoders/clip_model.py:36-38 */
#define M3AE_ACT_TANH 4 /* Pooler activation, prediction_heads.py:13,18 */

#define M3AE_MAJOR_K 0  /* operand stored [rows = M or N, cols = K], K contiguous */
#define M3AE_MAJOR_MN 1 /* operand stored [rows = K, cols = M or N], M/N contiguous */

#define M3AE_GEMM_AUTO 0
#define M3AE_GEMM_TCGEN05 1 /* TMA + tcgen05.mma + TMEM (bf16 inputs only) */
#define M3AE_GEMM_SIMT 2    /* fp32 FFMA tiles (f32 or bf16 inputs): strict-parity + debugging path */
#define M3AE_ATTN_MMA_SYNC 3 /* attention only: legacy mma.sync flash tiles instead of the tcgen05 kernels */
#define M3AE_ATTN_NARROW 4   /* attention only: AUTO, but every shape with a side <= 32 tokens goes to the single-pass "narrow" kernels
                                (AUTO itself sends the forward there only where it measured faster) */

#define M3AE_OK 0
#define M3AE_ERR_INVALID (-1)
#define M3AE_ERR_CUDA (-2)
#define M3AE_ERR_ARCH (-3)
#define M3AE_ERR_UNSUPPORTED (-4)

int m3ae_version(void);
const char* m3ae_last_error(void);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int64_t m3ae_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * GEMM with fused epilogue:  out[M,N] (+)= epi( alpha * A[M,K] · B[N,K]^T )
 * Replaces nn.Linear at bert_model.py:263,276-277,285-286 (Q/K/V), :361 (attention out-proj),
 * :426 (FFN up) and :439 (FFN down), m3ae_module.py:235,252 (modality projections),
 * m3ae_decoder.py:60-61,78-79 (in/out projections), :85 (FFN), :136 (vocab projection), and —
 * with MN-major operands — their autograd backward (dgrad: dX = dY·W, wgrad: dW = dY^T·X).
 * Epilogue order = the reference modules': (+bias) -> [aux_out := pre-activation or act'(it)] -> act ->
 * (* act'(aux_in)) -> dropout -> (+residual)   (BertSelfOutput :361-363, BertIntermediate :426-427,
 * BertOutput :439-441).
 * --------------------------------------------------------------------------------------------- */
typedef struct m3ae_gemm_epilogue_t {
  void* out;            /* [M,N] row-major, leading dimension ldo (elements) */
  int32_t out_dtype;    /* M3AE_F32 | M3AE_BF16 */
  int64_t ldo;
  const float* bias;    /* [N] fp32 or NULL */
  int32_t act;          /* M3AE_ACT_* applied after bias */
  void* aux_out;        /* optional: pre-activation copy, dtype out_dtype, ld ldo */
  const void* aux_in;   /* optional: forward pre-activation for dact, ld ld_aux_in */
  int32_t aux_in_dtype;
  int64_t ld_aux_in;
  int32_t dact;         /* M3AE_ACT_*: multiply result by act'(aux_in) (backward through the activation) */
  const void* residual; /* optional [M,N], added last */
  int32_t res_dtype;
  int64_t ldr;
  float drop_p;         /* dropout probability applied before the residual add (0 = off) */
  uint64_t drop_seed;
  uint64_t drop_stream; /* distinguishes dropout sites; element index = row*ldo+col */
  int32_t accumulate;   /* 1: out (fp32) += result (atomic; required when split_k > 1) */
  float alpha;
  int32_t aux_is_grad;  /* 1: aux_out receives act'(pre-activation) instead of the pre-activation, and dact multiplies by
                         * aux_in as stored (the FFN backward then needs one multiply per element instead of a GELU') */
  uint32_t* drop_bits;  /* optional (drop_p > 0, N % 32 == 0, tcgen05 backend): keep bits of the dropout, one uint32 per
                         * (row, 32 columns) at drop_bits[row * (N/32) + col/32], bit j = column col+j kept.  Handed to
                         * m3ae_layernorm_bwd so that backward does not regenerate the Philox stream. */
  int32_t drop_bits_ready; /* 1: drop_bits already holds this site's keep bits (m3ae_dropout_bits): the epilogue READS them
                            * instead of evaluating Philox (same bits, by construction); 0: the epilogue writes them */
  float* amax_val;      /* optional (tcgen05 backend, fp32 out or out == NULL): per-row maximum of every 64-column group of the */
  int32_t* amax_idx;    /* result and its column, amax_val[g * M + row] / amax_idx[g * M + row], g = col / 64, first maximum
                         * wins.  m3ae_argmax_partials reduces them to the row argmax, so the decoder's vocabulary
                         * projection (m3ae_decoder.py:136,151-153) never writes its [B, 30522] logits.  With amax_val set,
                         * out may be NULL (nothing but the partials is written). */
} m3ae_gemm_epilogue_t;

typedef struct m3ae_gemm_problem_t {
  const void* a; /* logical [M,K] */
  const void* b; /* logical [N,K] */
  int64_t lda;   /* elements between consecutive stored rows */
  int64_t ldb;
  int32_t a_major; /* M3AE_MAJOR_* */
  int32_t b_major;
  int32_t M, N, K;
  int32_t split_k; /* >= 1; > 1 requires epi.accumulate = 1 and fp32 out */
  m3ae_gemm_epilogue_t epi;
} m3ae_gemm_problem_t;

/* Runs 1..4 independent problems in ONE persistent launch (the text and image streams of a fusion
 * layer pair are independent, m3ae_module.py:276-277, so their GEMMs are co-scheduled). */
int m3ae_gemm(const m3ae_gemm_problem_t* problems, int32_t n_problems, int32_t in_dtype, int32_t backend, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused scaled-dot-product attention (flash-style: online softmax in registers + warp shuffles,
 * scores never materialised).  Replaces bert_model.py:301,319-322,325,334,340,342-344
 * (BertSelfAttention core, self and cross mode) and the attention inside nn.MultiheadAttention at
 * m3ae_decoder.py:60-61,78-79.
 *   S = scale * Q K^T + mask[b, key] (+ causal -inf);  P = softmax(S);  P' = dropout(P);  O = P' V
 * Row r = b*L + i of q/k/v/out holds token i of batch b; head h occupies columns [h*dh, (h+1)*dh).
 * ld* are row strides in elements, so q/k/v may point into a fused QKV projection buffer.
 * --------------------------------------------------------------------------------------------- */
typedef struct m3ae_attn_params_t {
  const void* q;
  const void* k;
  const void* v;
  int64_t ldq, ldk, ldv;
  const float* mask; /* additive [B, Lk] fp32 (the reference's [B,1,1,Lk] extended mask) or NULL */
  void* out;         /* [B*Lq, H*dh] */
  int64_t ldo;
  float* lse;        /* [B,H,Lq] log-sum-exp of S, saved for backward */
  float* probs;      /* optional [B,H,Lq,Lk] fp32: P' (post-dropout, as the reference returns) or NULL */
  int32_t B, H, Lq, Lk, dh;
  float scale;
  int32_t causal;
  float drop_p;
  uint64_t drop_seed;
  uint64_t drop_stream;
  int32_t dtype;     /* dtype of q,k,v,out (and dq,dk,dv,dout) */
  /* backward only */
  const void* dout;  /* [B*Lq, H*dh], ld ldo */
  void* dq;
  void* dk;
  void* dv;
  int64_t lddq, lddk, lddv;
  float* delta;      /* workspace [B,H,Lq] fp32: rowsum(dO*O) */
  /* optional dropout keep-bit cache [B,H,Lq, 8*ceil(Lk/64)] bytes (bit k%8 of byte k/8 = key k kept): the tensor-core
   * forward writes it, the tensor-core backward reads it instead of re-evaluating Philox.  NULL: bits are recomputed. */
  uint8_t* drop_mask;
  int32_t drop_mask_ready; /* 1: drop_mask was filled by m3ae_dropout_bits: the tensor-core forward reads it too */
} m3ae_attn_params_t;

int m3ae_attn_fwd(const m3ae_attn_params_t* p, void* stream);
/* dq, dk, dv are fully overwritten. */
int m3ae_attn_bwd(const m3ae_attn_params_t* p, void* stream);

/* Attention probabilities only: probs[b,h,i,:] = dropout(exp(scale*q_i.K^T + mask - lse_i)) from the q, k, mask and lse of
 * an earlier m3ae_attn_fwd call (same dropout seed/stream -> same P').  This is what output_attentions=True returns
 * (bert_model.py:346, consumed at m3ae_module.py:282-283); computing it on demand keeps the reference's hard-wired
 * output_attentions=True (m3ae_module.py:276-277) off the hot path. */
int m3ae_attn_probs(const m3ae_attn_params_t* p, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Element-wise pieces of the heads (M = batch rows; HBM-bound).
 * y = act(x): Pooler tanh (prediction_heads.py:13,18), vqa_head GELU (m3ae_module.py:123).  dx = dy * act'(x). */
int m3ae_activation_fwd(const void* x, void* y, int64_t n, int32_t act, int32_t dtype, void* stream);
int m3ae_activation_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t act, int32_t dtype, void* stream);
/* VQA loss (objectives.py:201): loss_out[0] = scale_loss * sum_i bce_with_logits(logits_i, targets_i) over n = B*labels
 * elements (scale_loss = 1/B gives mean * n_labels) and, when dlogits != NULL, dlogits_i = scale_grad * (sigmoid - target). */
int m3ae_bce_with_logits(const float* logits, const float* targets, float* loss_out, float* dlogits, int64_t n,
                         float scale_loss, float scale_grad, void* stream);
/* Decoder fine-tune loss (m3ae_decoder.py:365-368): rows = B*T positions of fp32 logits [M,V]; rows whose target equals
 * ignore_index (or whose weight is 0) contribute nothing.  loss_out[0] = sum_rows w_row * (logsumexp - logit[target]);
 * dlogits (optional) = w_row * (softmax - onehot).  The caller passes w_row = 1 / (number of counted rows). */
int m3ae_cross_entropy(const float* logits, int64_t ld, const int64_t* target, const float* row_weight, float* loss_out,
                       float* dlogits, int64_t ldd, int64_t M, int32_t V, int64_t ignore_index, void* stream);

/* ---------------------------------------------------------------------------------------------
 * LayerNorm over the last dimension (fp32 statistics, two-pass variance).
 * Replaces nn.LayerNorm at bert_model.py:363,441 (eps 1e-12), m3ae_module.py:122 (eps 1e-5),
 * m3ae_decoder.py:46-49 (eps 1e-5).  The residual add and bias are fused into the producing GEMM.
 * --------------------------------------------------------------------------------------------- */
int m3ae_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                       int64_t M, int32_t H, float eps, int32_t dtype, void* stream);
/* dx = LN'(dy); dgamma/dbeta (+)= column reductions (atomic accumulate into caller-zeroed or
 * grad-accumulating fp32 buffers).  When the Linear that produced x applied hidden dropout in its epilogue
 * (drop_p > 0), dx_drop receives dx pushed back through that dropout (same Philox stream, element = row*H+col) —
 * the operand of that Linear's dgrad/wgrad GEMMs — while dx stays the residual-branch gradient.
 * dcolsum (optional, fp32 [H]) += column sums of dx_drop (or dx when drop_p == 0): the Linear's bias gradient.
 * drop_bits (optional): the keep bits the forward GEMM epilogue cached ([M, H/32] uint32); NULL = regenerate. */
int m3ae_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma,
                       void* dx, float* dgamma, float* dbeta, float* dcolsum, int64_t M, int32_t H, int32_t dtype,
                       void* dx_drop, float drop_p, uint64_t drop_seed, uint64_t drop_stream, const uint32_t* drop_bits,
                       void* stream);

/* out[N] (fp32) += column sums of x[M,N] (ld elements).  Bias gradients of the projections. */
int m3ae_colsum(const void* x, float* out, int64_t M, int32_t N, int64_t ld, int32_t dtype, void* stream);

/* dst[i] = (dst_dtype) src[i]  — fp32 master parameters -> bf16 operand cache, activations casts. */
int m3ae_cast(const void* src, int32_t src_dtype, void* dst, int32_t dst_dtype, int64_t n, void* stream);

/* Up to M3AE_MAX_CAST_GROUP casts in ONE launch: a layer's fp32 master parameters (reference state-dict tensors,
 * bert_model.py:214-452) -> its compute-dtype operand cache, refreshed before every training forward. */
#define M3AE_MAX_CAST_GROUP 16
typedef struct m3ae_cast_item {
  const void* src;
  void* dst;
  int64_t n;
  int32_t src_dtype, dst_dtype;
} m3ae_cast_item_t;
int m3ae_cast_group(const m3ae_cast_item_t* items, int32_t n, void* stream);

/* out[r, :] = a[r, :] + b[:]  (row-broadcast add; modality type embedding, m3ae_module.py:260-263). */
int m3ae_add_rowvec(const void* a, const float* b, void* out, int64_t M, int32_t N, int32_t in_dtype, int32_t out_dtype,
                    void* stream);

/* Keep bits of up to M3AE_MAX_DROPBITS dropout sites in ONE streaming launch, ahead of the kernels that apply them: the
 * Philox4x32-7 evaluations (the same counters the kernels would use, so the bits are identical) leave the issue-bound
 * element-wise phases of the attention kernels and the GEMM epilogues (bert_model.py:334 attention-probability dropout,
 * :362,:440 hidden dropouts).
 *   kind 0 (attention): out = drop_mask of m3ae_attn_params_t, rows = B*H*Lq, n = Lk, ld unused
 *   kind 1 (hidden):    out = drop_bits of m3ae_gemm_epilogue_t, rows = M, n = N (N % 32 == 0), ld = the GEMM's ldo */
#define M3AE_MAX_DROPBITS 12
typedef struct m3ae_dropbits_item {
  void* out;
  int64_t rows;
  int64_t ld;
  int32_t n;
  int32_t kind;
  float p;
  uint64_t seed;
  uint64_t site; /* drop_stream of the consuming kernel */
} m3ae_dropbits_item_t;
int m3ae_dropout_bits(const m3ae_dropbits_item_t* items, int32_t n_items, void* stream);

/* In-place dropout-scale of a gradient: x[i] = keep(i) ? x[i]/(1-p) : 0 with the same Philox stream the
 * forward GEMM epilogue used (element index = row*ld+col). */
int m3ae_dropout_apply(const void* x, void* y, int64_t M, int32_t N, int64_t ld, float p, uint64_t seed,
                       uint64_t stream_id, int32_t dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * KV-cached single-token decoder step (greedy generation).  Replaces one iteration of
 * Decoder.search_path (m3ae_decoder.py:149-161), which re-runs Decoder.forward on the whole prefix;
 * only dec_layers[-1] is live (m3ae_decoder.py:131-134 feeds every layer the embedding and keeps
 * the last result).  See mm-vqa-healthcare_b200/decoder.py for the orchestration.
 *   attention for ONE query token per (batch, head) against a cache of `len` keys:
 *   out[b, h*dh:(h+1)*dh] = softmax(scale * q·K^T) V     (no mask: causal prefix / all memory valid)
 * k/v caches are [B, Lmax, H*dh] (row stride ldk/ldv, batch stride = Lmax*ld).
 * --------------------------------------------------------------------------------------------- */
int m3ae_decode_attn(const void* q, int64_t ldq, const void* kcache, const void* vcache, int64_t ldk, int64_t ldv,
                     int64_t batch_stride_k, int64_t batch_stride_v, void* out, int64_t ldo, int32_t B, int32_t H,
                     int32_t len, int32_t dh, float scale, int32_t dtype, void* stream);

/* Row-wise argmax over logits [M,N] fp32/bf16 (first maximal index, as torch.argmax) -> int64 ids.
 * log_softmax (m3ae_decoder.py:151) is monotone and is skipped. */
int m3ae_argmax_rows(const void* logits, int64_t ld, int64_t* out_ids, int64_t M, int32_t N, int32_t dtype, void* stream);

/* Decoder input embedding (m3ae_decoder.py:125-127): out[r,:] = 2*E[ids[r],:] + pe[pos[r],:]. */
int m3ae_decoder_embed(const int64_t* ids, const int32_t* pos, const float* table, const float* pe, void* out, int64_t M,
                       int32_t D, int32_t out_dtype, void* stream);

/* Token hand-over of the greedy decoder in one launch (bf16 activations, D = 768): ids_out[r] = row argmax from the vocabulary
 * projection's partials (as m3ae_argmax_partials; amax_val == NULL: ids are read from ids_in instead), x[r,:] = 2*E[id] + pe[pos]
 * (m3ae_decoder.py:125-127) and, when xn != NULL, xn[r,:] = LayerNorm(x[r,:]) (pre_norm, m3ae_decoder.py:58) — bit-identical to
 * m3ae_argmax_partials + m3ae_decoder_embed + m3ae_layernorm_fwd. */
int m3ae_decoder_next_token(const float* amax_val, const int32_t* amax_idx, int32_t n_groups, int32_t M, const int64_t* ids_in, int64_t* ids_out,
                            int32_t pos, const float* table, const float* pe, const float* ln_gamma, const float* ln_beta, float ln_eps, void* x,
                            void* xn, int32_t D, void* stream);

/* Skinny linear layer of the single-token decoder step (M <= 128 rows = batch; bf16 operands, fp32 accumulate):
 *   out = act( LN?(x)[M,K] . w[N,K]^T + bias ) (+ residual)
 * Replaces, for ONE new token per sequence, the projections of m3ae_decoder.py:54-90 — MHA in-proj / out-proj
 * (nn.MultiheadAttention at :42-43), the FFN (:45, :85-87) — and, when ln_gamma != NULL, the LayerNorm in front of them
 * (pre_norm :58, layernorm1 :66, layernorm2 :83; K must be 768) folded into the activation load.
 * The whole activation panel and the CTA's weight slice are staged in shared memory by one burst of cp.async; every CTA takes
 * 16 or 32 output columns.  Segmented output: column n goes to out[n / seg_cols] + row * ldo[n / seg_cols] + n % seg_cols
 * (the QKV projection writes q and the K / V cache rows in one launch).  ksplit > 1 splits K across CTAs (K / ksplit <= 768
 * per CTA): `ws` ([128, N] fp32) and `counters` (256 int32) must be ZERO on entry and are left zero.  A CTA takes a 2-D tile
 * (128x16 ... 32x32 rows x columns, chosen for the fewest staged bytes per CTA at about one wave of CTAs). */
typedef struct m3ae_skinny_params_t {
  const void* x;        /* [M, K] bf16 */
  int64_t ldx;
  const void* w;        /* [N, K] bf16 (nn.Linear layout) */
  int64_t ldw;
  const float* bias;    /* [N] fp32 or NULL */
  const void* residual; /* [M, N] bf16 or NULL */
  int64_t ldr;
  void* out[3];         /* bf16 */
  int64_t ldo[3];
  int32_t seg_cols;     /* columns per output segment (multiple of 32); = N for a single output */
  int32_t M, N, K;
  int32_t ksplit;       /* 0 / 1: no split */
  int32_t act;          /* M3AE_ACT_NONE or M3AE_ACT_RELU */
  const float* ln_gamma; /* optional LayerNorm over K in front of the projection */
  const float* ln_beta;
  float ln_eps;
  float* ws;
  int32_t* counters;
} m3ae_skinny_params_t;
int m3ae_skinny_linear(const m3ae_skinny_params_t* p, void* stream);
/* m3ae_skinny_linear without any allocation by caller or library for its scratch.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, serving loops, harnesses
 * that forbid it).  Same parameters and results as m3ae_skinny_linear, except that p->ws and p->counters are IGNORED: the
 * split-K workspace and tile counters live in a file-scope static __device__ array inside the library (513 KB per GPU that
 * loads it, held until exit).  Limits: M <= 128 rows, N <= M3AE_SKINNY_STATIC_MAX_N columns; larger sizes return
 * M3AE_ERR_UNSUPPORTED with a message.  The call issues one cudaMemsetAsync of the scratch on `stream` (split-K only) and the
 * kernel; it calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device share the one array: calls that
 * overlap on the GPU (different streams without a dependency) race — serialise them. */
#define M3AE_SKINNY_STATIC_MAX_N 1024
int m3ae_skinny_linear_static(const m3ae_skinny_params_t* p, void* stream);

/* Row argmax from the per-64-column partials of m3ae_gemm (epi.amax_val / amax_idx): ids[row] (int64) = column of the first
 * maximum over groups [0, n_groups).  Replaces argmax(log_softmax(final_linear(x))) of m3ae_decoder.py:151-153. */
int m3ae_argmax_partials(const float* amax_val, const int32_t* amax_idx, int32_t n_groups, int32_t M, int64_t* ids, void* stream);
/* Test hook: 0 = auto (tcgen05/TMEM attention for bf16 where the shape is covered, else mma.sync tiles),
 * 1 = tcgen05 for every covered shape (auto skips the ones where mma.sync is faster), 2 = force the generic fp32-math
 * attention kernels, 3 (M3AE_ATTN_MMA_SYNC) = skip tcgen05, use the mma.sync tiles. */
int m3ae_set_attention_backend(int32_t backend);
/* Test hook: 1 = force the generic per-element GEMM epilogue (cross-checks the TMA-staged one), 0 = auto. */
int m3ae_set_gemm_epilogue(int32_t generic);
/* tcgen05 GEMM tile mode: 0 = auto (clusters of two CTA pairs — tcgen05.mma.cta_group::2 on 256x256 tiles, the operand
 * tile the two pairs share fetched once and TMA-multicast — unless every problem of the launch has M <= 128),
 * 1 = always one CTA per 128x256 tile, 2 = one CTA pair per cluster (no multicast), 3 = always two pairs per cluster.
 * Results are identical.  Process-global test / benchmarking hook: not synchronised with concurrent launches. */
int m3ae_set_gemm_cta_pair(int32_t mode);
/* Number of SMs the persistent tcgen05 GEMM leaves free (0 = none, default) for kernels of other streams — the NCCL gradient
 * all-reduce of data-parallel fine-tuning (ddp.py sets it from M3AE_DDP_RESERVE_SMS when world_size > 1). */
int m3ae_set_gemm_reserved_sms(int32_t n);
/* Single-CTA launches (every problem M <= 128): 0 = auto (128x64 tiles when 128x256 tiles would occupy less than a third of
 * the SMs), 1 = always 128x256, 2 = always 128x64.  Results are identical.  Process-global test hook. */
int m3ae_set_gemm_skinny(int32_t mode);
/* Programmatic dependent launch (every kernel of the library: cudaLaunchAttributeProgrammaticStreamSerialization +
 * griddepcontrol.wait before the first global access): 1 = on, 0 = off (default; M3AE_PDL=1 in the environment turns it
 * on).  Results are identical; measured neutral on the fine-tune step and harmful inside CUDA graphs.  Process-global. */
int m3ae_set_pdl(int32_t on);
/* Test hook: 1 = m3ae_layernorm_bwd loads rows through registers only (cross-checks the cp.async.bulk staging ring), 0 = auto. */
int m3ae_set_layernorm_bwd_staging(int32_t registers_only);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* M3AE_B200_H_ */
