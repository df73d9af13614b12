// extern "C" surface of libm3ae_b200.so (declared in include/m3ae_b200.h): argument validation,
// architecture gate (sm_100 only, no CPU fallback), dispatch to the kernels.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "common.cuh"

namespace m3ae {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static int g_sms[64];
static int g_major[64];
static std::once_flag g_dev_once[64];

static int query_device(int* sms, int* major) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    set_last_error("cudaGetDevice failed: %s", cudaGetErrorString(e));
    return M3AE_ERR_CUDA;
  }
  std::call_once(g_dev_once[dev], [dev] {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess) {
      g_sms[dev] = prop.multiProcessorCount;
      g_major[dev] = prop.major;
    }
  });
  *sms = g_sms[dev];
  *major = g_major[dev];
  return M3AE_OK;
}

int sm_count() {
  int sms = 0, major = 0;
  query_device(&sms, &major);
  return sms > 0 ? sms : 148;
}

int check_device_sm100() {
  int sms = 0, major = 0;
  int rc = query_device(&sms, &major);
  if (rc) return rc;
  if (major != 10) {
    set_last_error("libm3ae_b200 requires an sm_100 (B200) device; current device has compute capability major %d. There is no fallback path.", major);
    return M3AE_ERR_ARCH;
  }
  return M3AE_OK;
}

int gemm_tcgen05_launch(const m3ae_gemm_problem_t* probs, int n, cudaStream_t stream);
int gemm_simt_launch(const m3ae_gemm_problem_t* probs, int n, int in_dtype, cudaStream_t stream);
int attn_fwd(const m3ae_attn_params_t* p, int backend, cudaStream_t st);
int attn_bwd(const m3ae_attn_params_t* p, int backend, cudaStream_t st);
int attn_probs(const m3ae_attn_params_t* p, cudaStream_t st);
int activation_fwd(const void*, void*, int64_t, int, int, cudaStream_t);
int activation_bwd(const void*, const void*, void*, int64_t, int, int, cudaStream_t);
int bce_with_logits(const float*, const float*, float*, float*, int64_t, float, float, cudaStream_t);
int cross_entropy(const float*, int64_t, const int64_t*, const float*, float*, float*, int64_t, int64_t, int, int64_t, cudaStream_t);
int dropout_bits(const m3ae_dropbits_item_t*, int, cudaStream_t);
int layernorm_fwd(const void*, const float*, const float*, void*, float*, float*, int64_t, int, float, int, cudaStream_t);
int layernorm_bwd(const void*, const void*, const float*, const float*, const float*, void*, float*, float*, float*, int64_t, int, int, void*, float, uint64_t, uint64_t, const uint32_t*, cudaStream_t);
int colsum(const void*, float*, int64_t, int, int64_t, int, cudaStream_t);
int cast(const void*, int, void*, int, int64_t, cudaStream_t);
int cast_group(const m3ae_cast_item_t*, int, cudaStream_t);
int add_rowvec(const void*, const float*, void*, int64_t, int, int, int, cudaStream_t);
int dropout_apply(const void*, void*, int64_t, int, int64_t, float, uint64_t, uint64_t, int, cudaStream_t);
int decode_attn(const void*, int64_t, const void*, const void*, int64_t, int64_t, int64_t, int64_t, void*, int64_t, int, int, int, int, float, int, cudaStream_t);
int argmax_rows(const void*, int64_t, int64_t*, int64_t, int, int, cudaStream_t);
int decoder_embed(const int64_t*, const int32_t*, const float*, const float*, void*, int64_t, int, int, cudaStream_t);

extern bool force_generic_epilogue;
extern int gemm_cta_pair_mode;
extern int gemm_skinny_mode;
extern int gemm_reserved_sms;
static int g_pdl_mode = -1;  // -1: not decided yet (M3AE_PDL environment variable), 0 off, 1 on
bool pdl_on() {
  if (g_pdl_mode < 0) {
    const char* e = getenv("M3AE_PDL");
    g_pdl_mode = (e != nullptr && e[0] == '1') ? 1 : 0;  // opt-in: measured neutral on the fusion step, harmful inside CUDA graphs
  }
  return g_pdl_mode == 1;
}
int argmax_partials(const float*, const int32_t*, int, int, int64_t*, cudaStream_t);
int skinny_linear(const m3ae_skinny_params_t*, cudaStream_t);  // skinny_linear.cu
int skinny_linear_static(const m3ae_skinny_params_t*, cudaStream_t);  // skinny_linear.cu
int decoder_next_token(const float*, const int32_t*, int, int, const int64_t*, int64_t*, int, const float*, const float*, const float*, const float*,
                       float, void*, void*, int, cudaStream_t);
extern bool ln_bwd_force_registers;
static int g_attn_backend = M3AE_GEMM_AUTO;

}  // namespace m3ae

using namespace m3ae;

#define ARCH_GATE()                     \
  do {                                  \
    int _rc = check_device_sm100();     \
    if (_rc) return _rc;                \
  } while (0)

extern "C" {

int m3ae_version(void) { return M3AE_B200_VERSION; }
const char* m3ae_last_error(void) { return g_err; }
int64_t m3ae_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int m3ae_gemm(const m3ae_gemm_problem_t* problems, int32_t n, int32_t in_dtype, int32_t backend, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(problems != nullptr && n >= 1 && n <= 4, M3AE_ERR_INVALID, "m3ae_gemm: need 1..4 problems, got %d", n);
  M3AE_CHECK(in_dtype == M3AE_F32 || in_dtype == M3AE_BF16, M3AE_ERR_INVALID, "m3ae_gemm: bad input dtype %d", in_dtype);
  for (int i = 0; i < n; ++i) {
    const m3ae_gemm_problem_t& g = problems[i];
    M3AE_CHECK(g.a && g.b && (g.epi.out || g.epi.amax_val), M3AE_ERR_INVALID, "m3ae_gemm: null operand in problem %d", i);
    M3AE_CHECK((g.epi.amax_val == nullptr) == (g.epi.amax_idx == nullptr), M3AE_ERR_INVALID, "m3ae_gemm: amax_val and amax_idx come together");
    M3AE_CHECK(g.epi.amax_val == nullptr || (in_dtype == M3AE_BF16 && backend != M3AE_GEMM_SIMT && g.split_k <= 1 && !g.epi.accumulate &&
                                              (g.epi.out == nullptr || g.epi.out_dtype == M3AE_F32)),
               M3AE_ERR_UNSUPPORTED, "m3ae_gemm: the row-argmax partials need the tcgen05 backend, no split-K and an fp32 (or no) output");
    M3AE_CHECK(g.epi.drop_p >= 0.f && g.epi.drop_p < 1.f, M3AE_ERR_INVALID, "m3ae_gemm: dropout %f not in [0,1)", g.epi.drop_p);
    M3AE_CHECK(g.epi.dact == M3AE_ACT_NONE || g.epi.aux_in != nullptr, M3AE_ERR_INVALID, "m3ae_gemm: dact needs aux_in");
    M3AE_CHECK(g.epi.drop_bits == nullptr || (g.epi.drop_p > 0.f && g.N % 32 == 0 && in_dtype == M3AE_BF16 && backend != M3AE_GEMM_SIMT),
               M3AE_ERR_UNSUPPORTED, "m3ae_gemm: drop_bits needs drop_p > 0, N %% 32 == 0 and the tcgen05 backend");
    M3AE_CHECK(g.split_k <= 1 || (!g.epi.bias && g.epi.act == M3AE_ACT_NONE && !g.epi.residual && g.epi.dact == M3AE_ACT_NONE && g.epi.drop_p == 0.f && !g.epi.aux_out),
               M3AE_ERR_INVALID, "m3ae_gemm: split_k > 1 supports only the plain accumulate epilogue");
  }
  if (backend == M3AE_GEMM_AUTO) backend = (in_dtype == M3AE_BF16) ? M3AE_GEMM_TCGEN05 : M3AE_GEMM_SIMT;
  if (backend == M3AE_GEMM_TCGEN05) {
    M3AE_CHECK(in_dtype == M3AE_BF16, M3AE_ERR_UNSUPPORTED, "m3ae_gemm: the tcgen05 backend takes bf16 operands (fp32 parity runs use the SIMT backend)");
    return gemm_tcgen05_launch(problems, n, (cudaStream_t)stream);
  }
  M3AE_CHECK(backend == M3AE_GEMM_SIMT, M3AE_ERR_INVALID, "m3ae_gemm: unknown backend %d", backend);
  return gemm_simt_launch(problems, n, in_dtype, (cudaStream_t)stream);
}

/* Test hook: 0 = auto (tensor-core kernels for bf16), 2 = force the generic fp32-math attention kernels. */
int m3ae_set_attention_backend(int32_t backend) { g_attn_backend = backend; return M3AE_OK; }
int m3ae_set_gemm_epilogue(int32_t generic) { force_generic_epilogue = generic != 0; return M3AE_OK; }
/* Test hook: 1 = LayerNorm backward loads rows through registers only (cross-checks the bulk-copy ring), 0 = auto. */
int m3ae_set_layernorm_bwd_staging(int32_t registers_only) { ln_bwd_force_registers = registers_only != 0; return M3AE_OK; }
int m3ae_set_gemm_cta_pair(int32_t mode) {
  M3AE_CHECK(mode >= 0 && mode <= 3, M3AE_ERR_INVALID, "m3ae_set_gemm_cta_pair: mode %d not in {0 auto, 1 single CTA, 2 CTA pair, 3 two pairs with operand multicast}", mode);
  gemm_cta_pair_mode = mode;
  return M3AE_OK;
}

int m3ae_attn_fwd(const m3ae_attn_params_t* p, void* stream) { ARCH_GATE(); return attn_fwd(p, g_attn_backend, (cudaStream_t)stream); }
int m3ae_attn_bwd(const m3ae_attn_params_t* p, void* stream) { ARCH_GATE(); return attn_bwd(p, g_attn_backend, (cudaStream_t)stream); }

int m3ae_attn_probs(const m3ae_attn_params_t* p, void* stream) { ARCH_GATE(); return attn_probs(p, (cudaStream_t)stream); }

int m3ae_activation_fwd(const void* x, void* y, int64_t n, int32_t act, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(x && y, M3AE_ERR_INVALID, "activation_fwd: null pointer");
  return activation_fwd(x, y, n, act, dtype, (cudaStream_t)stream);
}
int m3ae_activation_bwd(const void* dy, const void* x, void* dx, int64_t n, int32_t act, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(dy && x && dx, M3AE_ERR_INVALID, "activation_bwd: null pointer");
  return activation_bwd(dy, x, dx, n, act, dtype, (cudaStream_t)stream);
}
int m3ae_bce_with_logits(const float* logits, const float* targets, float* loss_out, float* dlogits, int64_t n, float scale_loss, float scale_grad, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(logits && targets && loss_out, M3AE_ERR_INVALID, "bce_with_logits: null pointer");
  return bce_with_logits(logits, targets, loss_out, dlogits, n, scale_loss, scale_grad, (cudaStream_t)stream);
}
int m3ae_cross_entropy(const float* logits, int64_t ld, const int64_t* target, const float* row_weight, float* loss_out, float* dlogits, int64_t ldd, int64_t M, int32_t V, int64_t ignore_index, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(logits && target && row_weight && loss_out, M3AE_ERR_INVALID, "cross_entropy: null pointer");
  return cross_entropy(logits, ld, target, row_weight, loss_out, dlogits, ldd, M, V, ignore_index, (cudaStream_t)stream);
}

int m3ae_dropout_bits(const m3ae_dropbits_item_t* items, int32_t n_items, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(items != nullptr && n_items >= 1 && n_items <= M3AE_MAX_DROPBITS, M3AE_ERR_INVALID, "dropout_bits: need 1..%d items, got %d", M3AE_MAX_DROPBITS, n_items);
  return dropout_bits(items, n_items, (cudaStream_t)stream);
}

int m3ae_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd, int64_t M, int32_t H, float eps, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(x && gamma && beta && y && mean && rstd, M3AE_ERR_INVALID, "layernorm_fwd: null pointer");
  return layernorm_fwd(x, gamma, beta, y, mean, rstd, M, H, eps, dtype, (cudaStream_t)stream);
}
int m3ae_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma, void* dx, float* dgamma, float* dbeta, float* dcolsum, int64_t M, int32_t H, int32_t dtype, void* dx_drop, float drop_p, uint64_t drop_seed, uint64_t drop_stream, const uint32_t* drop_bits, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(dy && x && mean && rstd && gamma && dx, M3AE_ERR_INVALID, "layernorm_bwd: null pointer");
  M3AE_CHECK(drop_p == 0.f || dx_drop != nullptr, M3AE_ERR_INVALID, "layernorm_bwd: drop_p > 0 needs dx_drop");
  return layernorm_bwd(dy, x, mean, rstd, gamma, dx, dgamma, dbeta, dcolsum, M, H, dtype, dx_drop, drop_p, drop_seed, drop_stream, drop_bits, (cudaStream_t)stream);
}
int m3ae_colsum(const void* x, float* out, int64_t M, int32_t N, int64_t ld, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(x && out, M3AE_ERR_INVALID, "colsum: null pointer");
  return colsum(x, out, M, N, ld, dtype, (cudaStream_t)stream);
}
int m3ae_cast(const void* src, int32_t sd, void* dst, int32_t dd, int64_t n, void* stream) {
  ARCH_GATE();
  M3AE_CHECK((src && dst) || n == 0, M3AE_ERR_INVALID, "cast: null pointer");
  return cast(src, sd, dst, dd, n, (cudaStream_t)stream);
}
int m3ae_cast_group(const m3ae_cast_item_t* items, int32_t n, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(items != nullptr && n >= 1 && n <= M3AE_MAX_CAST_GROUP, M3AE_ERR_INVALID, "cast_group: need 1..%d items, got %d", M3AE_MAX_CAST_GROUP, n);
  for (int i = 0; i < n; ++i) {
    M3AE_CHECK((items[i].src && items[i].dst) || items[i].n == 0, M3AE_ERR_INVALID, "cast_group: null pointer in item %d", i);
    M3AE_CHECK((items[i].src_dtype == M3AE_F32 || items[i].src_dtype == M3AE_BF16) && (items[i].dst_dtype == M3AE_F32 || items[i].dst_dtype == M3AE_BF16),
               M3AE_ERR_INVALID, "cast_group: bad dtype in item %d", i);
  }
  return cast_group(items, n, (cudaStream_t)stream);
}
int m3ae_add_rowvec(const void* a, const float* b, void* out, int64_t M, int32_t N, int32_t in_dtype, int32_t out_dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(a && b && out, M3AE_ERR_INVALID, "add_rowvec: null pointer");
  return add_rowvec(a, b, out, M, N, in_dtype, out_dtype, (cudaStream_t)stream);
}
int m3ae_dropout_apply(const void* x, void* y, int64_t M, int32_t N, int64_t ld, float p, uint64_t seed, uint64_t stream_id, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(x && y, M3AE_ERR_INVALID, "dropout_apply: null pointer");
  return dropout_apply(x, y, M, N, ld, p, seed, stream_id, dtype, (cudaStream_t)stream);
}
int m3ae_decode_attn(const void* q, int64_t ldq, const void* kcache, const void* vcache, int64_t ldk, int64_t ldv, int64_t bsk, int64_t bsv, void* out, int64_t ldo, int32_t B, int32_t H, int32_t len, int32_t dh, float scale, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(q && kcache && vcache && out, M3AE_ERR_INVALID, "decode_attn: null pointer");
  return decode_attn(q, ldq, kcache, vcache, ldk, ldv, bsk, bsv, out, ldo, B, H, len, dh, scale, dtype, (cudaStream_t)stream);
}
int m3ae_argmax_partials(const float* amax_val, const int32_t* amax_idx, int32_t n_groups, int32_t M, int64_t* ids, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(amax_val && amax_idx && ids && n_groups > 0 && M > 0, M3AE_ERR_INVALID, "argmax_partials: null pointer or empty input");
  return argmax_partials(amax_val, amax_idx, n_groups, M, ids, (cudaStream_t)stream);
}
int m3ae_decoder_next_token(const float* amax_val, const int32_t* amax_idx, int32_t n_groups, int32_t M, const int64_t* ids_in, int64_t* ids_out,
                            int32_t pos, const float* table, const float* pe, const float* ln_gamma, const float* ln_beta, float ln_eps, void* x,
                            void* xn, int32_t D, void* stream) {
  ARCH_GATE();
  return decoder_next_token(amax_val, amax_idx, n_groups, M, ids_in, ids_out, pos, table, pe, ln_gamma, ln_beta, ln_eps, x, xn, D, (cudaStream_t)stream);
}
int m3ae_skinny_linear(const m3ae_skinny_params_t* p, void* stream) {
  ARCH_GATE();
  return skinny_linear(p, (cudaStream_t)stream);
}
int m3ae_skinny_linear_static(const m3ae_skinny_params_t* p, void* stream) {
  ARCH_GATE();
  return skinny_linear_static(p, (cudaStream_t)stream);
}
int m3ae_set_pdl(int32_t on) {
  g_pdl_mode = on ? 1 : 0;
  return M3AE_OK;
}
int m3ae_set_gemm_reserved_sms(int32_t n) {
  M3AE_CHECK(n >= 0 && n <= 64, M3AE_ERR_INVALID, "m3ae_set_gemm_reserved_sms: %d not in [0, 64]", n);
  gemm_reserved_sms = n;
  return M3AE_OK;
}
int m3ae_set_gemm_skinny(int32_t mode) {
  M3AE_CHECK(mode >= 0 && mode <= 2, M3AE_ERR_INVALID, "m3ae_set_gemm_skinny: mode %d not in {0 auto, 1 never, 2 always}", mode);
  gemm_skinny_mode = mode;
  return M3AE_OK;
}
int m3ae_argmax_rows(const void* logits, int64_t ld, int64_t* out_ids, int64_t M, int32_t N, int32_t dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(logits && out_ids, M3AE_ERR_INVALID, "argmax_rows: null pointer");
  return argmax_rows(logits, ld, out_ids, M, N, dtype, (cudaStream_t)stream);
}
int m3ae_decoder_embed(const int64_t* ids, const int32_t* pos, const float* table, const float* pe, void* out, int64_t M, int32_t D, int32_t out_dtype, void* stream) {
  ARCH_GATE();
  M3AE_CHECK(ids && pos && table && pe && out, M3AE_ERR_INVALID, "decoder_embed: null pointer");
  return decoder_embed(ids, pos, table, pe, out, M, D, out_dtype, (cudaStream_t)stream);
}

}  // extern "C"
