// Shared device/host helpers for the M3AE fusion hot path on sm_100a.
// PTX wrappers (mbarrier, TMA, tcgen05/TMEM), Philox dropout RNG, dtype helpers,
// and the GEMM epilogue that is shared by the tcgen05 and the SIMT (fp32 parity) GEMMs.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/m3ae_b200.h"

namespace m3ae {

// ---------------------------------------------------------------- errors
void set_last_error(const char* fmt, ...);
#define M3AE_CHECK(cond, code, ...)                \
  do {                                             \
    if (!(cond)) {                                 \
      ::m3ae::set_last_error(__VA_ARGS__);         \
      return (code);                               \
    }                                              \
  } while (0)

#define M3AE_CUDA(expr)                                                              \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::m3ae::set_last_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                             __FILE__, __LINE__);                                    \
      return M3AE_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

int check_device_sm100();  // 0 when the current device is sm_100, else error code (and sets last error)
// cudaFuncSetAttribute (dynamic shared memory opt-in) is per DEVICE: call sites keep one flag per device ordinal, so a second
// GPU driven from the same process gets its opt-in too
inline int current_device_slot() {
  int dev = 0;
  (void)cudaGetDevice(&dev);
  return dev & 63;
}
int sm_count();

// ---------------------------------------------------------------- programmatic dependent launch
// Opt-in (m3ae_set_pdl(1) or M3AE_PDL=1).  Every kernel of this library can be launched with
// cudaLaunchAttributeProgrammaticStreamSerialization and starts with pdl_sync(): griddepcontrol.launch_dependents lets the
// NEXT kernel of the stream be scheduled as soon as SMs free up (its CTAs then sit in griddepcontrol.wait),
// griddepcontrol.wait blocks until the PREVIOUS kernel has completed and its writes are visible.  Every kernel waits
// before it exits, so completion is transitive; kernels of other libraries (PyTorch) in between are ordinary
// stream-ordered launches.  Without the attribute griddepcontrol.* are no-ops.
// Measured on B200 (profiles/r2_pdl_ab.txt): the ~660 launches of a fine-tune step gain nothing (14.83 / 14.71 ms off vs
// 14.83 / 14.61 ms on: back-to-back launches of one stream already overlap their launch latency), and inside the decoder's
// CUDA graph programmatic edges are slower (157 -> 182 us per token step).  Hence off by default.
bool pdl_on();
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() {
  pdl_launch_dependents();
  pdl_wait();
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_on() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define M3AE_LAUNCH(kernel, grid, block, smem, st, ...) (void)::m3ae::launch_k(kernel, dim3(grid), dim3(block), (size_t)(smem), st, ##__VA_ARGS__)

// ---------------------------------------------------------------- dtype helpers
typedef __nv_bfloat16 bf16;

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float load_as_f32(const void* p, int dtype, size_t idx) {
  return dtype == M3AE_BF16 ? __bfloat162float(((const bf16*)p)[idx]) : ((const float*)p)[idx];
}
__device__ __forceinline__ void store_from_f32(void* p, int dtype, size_t idx, float v) {
  if (dtype == M3AE_BF16) ((bf16*)p)[idx] = __float2bfloat16_rn(v);
  else ((float*)p)[idx] = v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
  __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(t);
}

// ---------------------------------------------------------------- activations
// Exact erf GELU (reference: HF ACT2FN["gelu"], bert_model.py:421,427).
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  const float pdf = 0.39894228040143268f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// Branch-free GELU for the bf16 tensor-core epilogue.  h(x) = 0.5*erfc(|x|/sqrt(2)) ~ 0.5 * t(a1 + t(a2 + a3 t)) exp(-x^2/2),
// t = 1/(1 + p|x|/sqrt(2))  (Abramowitz-Stegun 7.1.25, |abs err| <= 1.3e-5 on h — bf16 resolves 2e-3 relative).  Then
//   gelu(x)  = relu(x) - |x| h          (x Phi(x) with Phi = 1-h for x >= 0, h for x < 0)
//   gelu'(x) = Phi(x) + x pdf(x),  Phi(x) = 0.5 + copysign(0.5 - h, x)
// 11 FMA-pipe instructions + 2 MUFU per element for gelu.  The fp32 parity path keeps erff().
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// returns h = 0.5*erfc(|x|/sqrt(2)) and e = exp(-x^2/2)
__device__ __forceinline__ float fast_half_erfc(float x, float& e) {
  const float t = rcp_approx(fmaf(0.47047f * 0.70710678118654752f, fabsf(x), 1.0f));
  float poly = fmaf(0.5f * 0.7478556f, t, 0.5f * -0.0958798f);
  poly = fmaf(poly, t, 0.5f * 0.3480242f);
  e = ex2_approx((x * -0.72134752044448170f) * x);  // 2^(-x^2 log2(e)/2)
  return (poly * t) * e;
}
__device__ __forceinline__ float fast_gelu(float x) {
  float e;
  const float h = fast_half_erfc(x, e);
  return fmaf(-fabsf(x), h, fmaxf(x, 0.f));
}
__device__ __forceinline__ float fast_gelu_grad(float x) {
  float e;
  const float h = fast_half_erfc(x, e);
  const float cdf = 0.5f + copysignf(0.5f - h, x);
  return fmaf(x * 0.39894228040143268f, e, cdf);
}
// gelu and gelu' sharing one erfc evaluation
__device__ __forceinline__ float fast_gelu_both(float x, float& grad) {
  float e;
  const float h = fast_half_erfc(x, e);
  const float cdf = 0.5f + copysignf(0.5f - h, x);
  grad = fmaf(x * 0.39894228040143268f, e, cdf);
  return fmaf(-fabsf(x), h, fmaxf(x, 0.f));
}

// ---- packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2 on sm_100): two elements per FMA-pipe instruction.  The FMA pipe takes
// one warp instruction every 2 cycles per scheduler, and the GEMM epilogues are bound by exactly that pipe.
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  asm("{\n.reg .b64 ra, rb, rc, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmov.b64 rc, {%6, %7};\n"
      "fma.rn.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c0), "f"(c1));
}
__device__ __forceinline__ void mul2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nmul.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{\n.reg .b64 ra, rb, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %5};\nadd.rn.f32x2 rd, ra, rb;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
// Two elements of fast_gelu_both at once: x0, x1 are replaced by gelu(x), g0, g1 receive gelu'(x).  Same formulas and
// constants as the scalar version (13 packed FMA-pipe instructions + 4 MUFU + 6 ALU per PAIR).
__device__ __forceinline__ void fast_gelu_both2(float& x0, float& x1, float& g0, float& g1) {
  const float n0 = __uint_as_float(__float_as_uint(x0) | 0x80000000u), n1 = __uint_as_float(__float_as_uint(x1) | 0x80000000u);  // -|x|
  constexpr float C = -0.47047f * 0.70710678118654752f;
  float d0, d1;
  fma2(d0, d1, n0, n1, C, C, 1.0f, 1.0f);  // 1 + p|x|/sqrt(2)
  const float t0 = rcp_approx(d0), t1 = rcp_approx(d1);
  float p0, p1;
  fma2(p0, p1, t0, t1, 0.5f * 0.7478556f, 0.5f * 0.7478556f, 0.5f * -0.0958798f, 0.5f * -0.0958798f);
  fma2(p0, p1, p0, p1, t0, t1, 0.5f * 0.3480242f, 0.5f * 0.3480242f);
  float q0, q1;
  mul2(q0, q1, x0, x1, x0, x1);
  mul2(q0, q1, q0, q1, -0.72134752044448170f, -0.72134752044448170f);
  const float e0 = ex2_approx(q0), e1 = ex2_approx(q1);  // exp(-x^2/2)
  mul2(p0, p1, p0, p1, t0, t1);
  float h0, h1;
  mul2(h0, h1, p0, p1, e0, e1);  // 0.5 erfc(|x|/sqrt 2)
  float c0, c1;
  fma2(c0, c1, h0, h1, -1.0f, -1.0f, 0.5f, 0.5f);
  c0 = copysignf(c0, x0);
  c1 = copysignf(c1, x1);
  add2(c0, c1, c0, c1, 0.5f, 0.5f);  // Phi(x)
  float k0, k1;
  mul2(k0, k1, x0, x1, 0.39894228040143268f, 0.39894228040143268f);
  fma2(g0, g1, k0, k1, e0, e1, c0, c1);
  const float r0 = fmaxf(x0, 0.f), r1 = fmaxf(x1, 0.f);
  fma2(x0, x1, n0, n1, h0, h1, r0, r1);
}

// QuickGELU of the CLIP visual tower (reference clip_model.py:36-38): x * sigmoid(1.702 x).
__device__ __forceinline__ float quick_gelu(float x) { return x / (1.0f + __expf(-1.702f * x)); }
__device__ __forceinline__ float quick_gelu_grad(float x) {
  const float s = 1.0f / (1.0f + __expf(-1.702f * x));
  return s * (1.0f + 1.702f * x * (1.0f - s));
}
// bf16 tensor-core epilogue variant: 1 ex2 + 1 rcp (MUFU) + 3 FMA-pipe instructions, gradient shares the sigmoid
__device__ __forceinline__ float fast_quick_gelu_both(float x, float& grad) {
  const float s = rcp_approx(1.0f + ex2_approx(x * (-1.702f * 1.4426950408889634f)));
  const float xs = x * s;
  grad = fmaf(1.702f * xs, 1.0f - s, s);
  return xs;
}
__device__ __forceinline__ float fast_quick_gelu(float x) {
  return x * rcp_approx(1.0f + ex2_approx(x * (-1.702f * 1.4426950408889634f)));
}
// exact-math activation dispatch (generic epilogue, fp32 parity path, standalone activation kernels)
__device__ __forceinline__ float act_exact(int act, float x) {
  return act == M3AE_ACT_GELU ? gelu_erf(x) : act == M3AE_ACT_RELU ? fmaxf(x, 0.0f) : act == M3AE_ACT_QUICKGELU ? quick_gelu(x)
       : act == M3AE_ACT_TANH ? tanhf(x) : x;
}
// derivative with respect to the pre-activation u
__device__ __forceinline__ float act_exact_grad(int act, float u) {
  if (act == M3AE_ACT_GELU) return gelu_erf_grad(u);
  if (act == M3AE_ACT_RELU) return u > 0.0f ? 1.0f : 0.0f;
  if (act == M3AE_ACT_QUICKGELU) return quick_gelu_grad(u);
  if (act == M3AE_ACT_TANH) { const float t = tanhf(u); return 1.0f - t * t; }
  return 1.0f;
}

// ---------------------------------------------------------------- Philox4x32-7 (counter-based RNG for dropout; 7 rounds pass BigCrush)
struct Philox {
  uint32_t key0, key1;
  __device__ __forceinline__ Philox(uint64_t seed) : key0((uint32_t)seed), key1((uint32_t)(seed >> 32)) {}
  __device__ __forceinline__ uint4 operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t k0 = key0, k1 = key1;
#pragma unroll
    for (int r = 0; r < 7; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
// Attention-probability dropout.  One Philox evaluation with counter (row_lin = (b*H+h)*Lq+i, site<<32 | key/8) covers the
// 8 consecutive keys 8g..8g+7 of a row: key 8g+2i takes the low 16-bit lane of word i, key 8g+2i+1 the high lane (the
// order bf16x2 packing uses).  Only 15 bits of a lane are used: adding (0x8000 - thr15) to a 15-bit value carries into the
// lane's top bit exactly when value >= thr15, so two keep decisions cost one AND and one ADD and never cross lanes.
// The drop threshold is p quantised to 1/32768 and the keep scale uses the quantised rate, so every kernel (tcgen05
// tiles, mma.sync tiles that cache 64x64 bit tiles, generic per-element kernels) sees exactly the same mask and scale.
__device__ __forceinline__ uint32_t drop_thr15(float p) { return (uint32_t)(p * 32768.0f + 0.5f); }
__device__ __forceinline__ float drop_keep_scale(float p) {
  return p > 0.f ? 32768.0f / (32768.0f - (float)drop_thr15(p)) : 1.0f;
}
// per-lane addend for attn_keep_flags
__device__ __forceinline__ uint32_t drop_lane_add(uint32_t thr15) { return (0x8000u - thr15) * 0x00010001u; }
// f[i] bit 15 = key 8g+2i kept, bit 31 = key 8g+2i+1 kept (all other bits zero)
__device__ __forceinline__ void attn_keep_flags(uint64_t seed, uint64_t stream_id, uint64_t row_lin, int key_group, uint32_t lane_add,
                                                uint32_t (&f)[4]) {
  Philox ph(seed);
  const uint4 r = ph(row_lin, (stream_id << 32) | (uint64_t)(uint32_t)key_group);
  f[0] = ((r.x & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[1] = ((r.y & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[2] = ((r.z & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[3] = ((r.w & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
}
// flags -> byte with bit j = key 8g+j kept.  umulhi(f, 0x20004) = (f >> 15) + (f >> 30): bit 0 = low lane, bit 1 = high lane
// (bit 16 repeats the high lane and ends up above bit 7).
__device__ __forceinline__ uint32_t attn_flags_to_byte(const uint32_t (&f)[4]) {
  const uint32_t t0 = __umulhi(f[0], 0x20004u), t1 = __umulhi(f[1], 0x20004u), t2 = __umulhi(f[2], 0x20004u), t3 = __umulhi(f[3], 0x20004u);
  return ((t3 * 4u + t2) * 16u + (t1 * 4u + t0)) & 0xFFu;
}
// keep bits for keys 8*key_group .. 8*key_group+7 of one attention row
__device__ __forceinline__ uint32_t attn_keep_bits8(uint64_t seed, uint64_t stream_id, uint64_t row_lin, int key_group, uint32_t thr15) {
  uint32_t f[4];
  attn_keep_flags(seed, stream_id, row_lin, key_group, drop_lane_add(thr15), f);
  return attn_flags_to_byte(f);
}
__device__ __forceinline__ bool attn_dropout_keep(uint64_t seed, uint64_t stream_id, uint64_t row_lin, int key, float p) {
  return (attn_keep_bits8(seed, stream_id, row_lin, key >> 3, drop_thr15(p)) >> (key & 7)) & 1u;
}

// Hidden (activation) dropout of the GEMM epilogues.  Same 15-bit-lane scheme: one Philox evaluation with counter
// (element index / 8, stream) decides the 8 consecutive elements 8g..8g+7 (element index = row*ld + col), so the forward
// epilogue, the generic per-element paths and the LayerNorm backward all see the same mask.  The drop rate is p quantised
// to 1/32768 and the keep scale is drop_keep_scale(p) (the exact inverse of the quantised keep rate).
__device__ __forceinline__ uint32_t dropout_keep8(uint64_t seed, uint64_t stream_id, uint64_t idx8, uint32_t thr15) {
  Philox ph(seed);
  const uint4 r = ph(idx8 >> 3, stream_id);
  const uint32_t lane_add = drop_lane_add(thr15);
  uint32_t f[4];
  f[0] = ((r.x & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[1] = ((r.y & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[2] = ((r.z & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  f[3] = ((r.w & 0x7FFF7FFFu) + lane_add) & 0x80008000u;
  return attn_flags_to_byte(f);
}
__device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t stream_id, uint64_t idx, float p) {
  return (dropout_keep8(seed, stream_id, idx & ~(uint64_t)7, drop_thr15(p)) >> (uint32_t)(idx & 7)) & 1u;
}
// 4 consecutive elements starting at idx (idx % 4 == 0): returns a 4-bit keep mask.
__device__ __forceinline__ uint32_t dropout_keep4(uint64_t seed, uint64_t stream_id, uint64_t idx, float p) {
  return (dropout_keep8(seed, stream_id, idx & ~(uint64_t)7, drop_thr15(p)) >> (uint32_t)(idx & 4)) & 0xFu;
}

// ---------------------------------------------------------------- warp helpers
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

// ---------------------------------------------------------------- GEMM epilogue (shared)
// Device-side view of m3ae_gemm_epilogue_t (include/m3ae_b200.h).
struct EpiDev {
  void* out; int out_dtype; int64_t ldo;
  const float* bias;
  int act;                 // M3AE_ACT_*
  void* aux_out;           // pre-activation copy (dtype = out_dtype, ld = ldo) or null
  const void* aux_in;      // pre-activation of the forward op (dtype = aux_in_dtype, ld = ld_aux_in) for dact
  int aux_in_dtype; int64_t ld_aux_in;
  int dact;                // M3AE_ACT_* whose derivative multiplies the result (uses aux_in)
  const void* residual; int res_dtype; int64_t ldr;
  float drop_p; uint64_t drop_seed; uint64_t drop_stream;
  int accumulate;          // 1: atomicAdd into fp32 out (split-K partials / grad accumulation)
  float alpha;
  int aux_is_grad;         // aux_out := act'(pre-activation); dact multiplies by aux_in as stored
  uint32_t* drop_bits;     // optional keep-bit cache [rows][N/32] (tcgen05 epilogues, 32-column chunks)
  int64_t drop_bits_ld;    // words per row = N / 32 (set by the launcher)
  int drop_bits_ready;     // 1: read the keep bits from drop_bits instead of evaluating Philox
  float* amax_val;         // per-(64-column group, row) maximum / its column (tcgen05 generic epilogue), or null
  int32_t* amax_idx;
  int32_t amax_ld;         // = M (set by the launcher)
};

// Applies the epilogue to `n` consecutive columns [col, col+n) of row `row` whose accumulators are acc[0..n).
// Order matches the reference modules: dense(+bias) -> act -> dropout -> +residual
// (BertSelfOutput bert_model.py:361-363, BertIntermediate :426-427, BertOutput :439-441).
template <int NV>
__device__ __forceinline__ void epilogue_apply(const EpiDev& e, int64_t row, int64_t col, int n_valid, float (&acc)[NV]) {
  const float alpha = e.alpha;
#pragma unroll
  for (int j = 0; j < NV; ++j) acc[j] *= alpha;
  if (e.bias) {
#pragma unroll
    for (int j = 0; j < NV; ++j) if (j < n_valid) acc[j] += __ldg(e.bias + col + j);
  }
  if (e.aux_out) {
#pragma unroll
    for (int j = 0; j < NV; ++j) if (j < n_valid) {
      float a = acc[j];
      if (e.aux_is_grad) a = act_exact_grad(e.act, a);
      store_from_f32(e.aux_out, e.out_dtype, (size_t)(row * e.ldo + col + j), a);
    }
  }
  if (e.act != M3AE_ACT_NONE) {
#pragma unroll
    for (int j = 0; j < NV; ++j) acc[j] = act_exact(e.act, acc[j]);
  }
  if (e.dact != M3AE_ACT_NONE) {
#pragma unroll
    for (int j = 0; j < NV; ++j) if (j < n_valid) {
      const float u = load_as_f32(e.aux_in, e.aux_in_dtype, (size_t)(row * e.ld_aux_in + col + j));
      acc[j] *= e.aux_is_grad ? u : act_exact_grad(e.dact, u);
    }
  }
  if (e.drop_p > 0.0f) {
    const float inv = drop_keep_scale(e.drop_p);
    uint32_t bits = 0;
    // keep-bit cache: one 32-bit word per row and 32 columns; a 16-column caller owns one 16-bit half of a word
    const bool ready = (NV == 32 || NV == 16) && e.drop_bits != nullptr && e.drop_bits_ready;
    uint16_t* const half_bits = reinterpret_cast<uint16_t*>(e.drop_bits) + 2 * row * e.drop_bits_ld + (col >> 4);
    const uint32_t cached = !ready ? 0u : NV == 32 ? e.drop_bits[row * (e.drop_bits_ld) + (col >> 5)] : (uint32_t)*half_bits;
#pragma unroll
    for (int j = 0; j < NV; ++j) if (j < n_valid) {
      const bool keep = ready ? ((cached >> (j & 31)) & 1u) != 0u
                              : dropout_keep(e.drop_seed, e.drop_stream, (uint64_t)(row * e.ldo + col + j), e.drop_p);
      acc[j] = keep ? acc[j] * inv : 0.0f;
      if (NV == 32 || NV == 16) bits |= (keep ? 1u : 0u) << (j & 31);
    }
    if (NV == 32 && e.drop_bits != nullptr && !ready) e.drop_bits[row * (e.drop_bits_ld) + (col >> 5)] = bits;
    if (NV == 16 && e.drop_bits != nullptr && !ready) *half_bits = (uint16_t)bits;
  }
  if (e.residual) {
#pragma unroll
    for (int j = 0; j < NV; ++j) if (j < n_valid) acc[j] += load_as_f32(e.residual, e.res_dtype, (size_t)(row * e.ldr + col + j));
  }
}

// ---------------------------------------------------------------- PTX: mbarrier / TMA / tcgen05
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// A barrier that never completes is a protocol bug: report it and trap instead of hanging the GPU (fail loudly).
#ifndef M3AE_WATCHDOG_CYCLES
#define M3AE_WATCHDOG_CYCLES 20000000000ll  // ~10 s at 2 GHz
#endif
static __device__ __noinline__ void mbar_timeout(uint32_t bar_addr, uint32_t parity) {
  printf("m3ae: mbarrier wait timed out: block (%d,%d) thread %d barrier@smem+0x%x parity %u\n", (int)blockIdx.x, (int)blockIdx.y,
         (int)threadIdx.x, bar_addr, parity);
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > M3AE_WATCHDOG_CYCLES) mbar_timeout(smem_u32(bar), parity);
  }
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)m) : "memory");
}
// 2-D tiled TMA load global -> shared, completion signalled on `bar` (complete_tx::bytes).
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"((uint64_t)m), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
      : "memory");
}

// 2-D tiled TMA store shared -> global (bulk async-group completion); out-of-bounds rows/columns are clipped.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int c_inner, int c_outer) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"((uint64_t)m),
               "r"(smem_u32(smem_src)), "r"(c_inner), "r"(c_outer)
               : "memory");
}
// 2-D tiled TMA reduction shared -> global: global[tile] += smem[tile] (element-wise add done at L2; clipped like a store).
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int c_inner, int c_outer) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"((uint64_t)m),
               "r"(smem_u32(smem_src)), "r"(c_inner), "r"(c_outer)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], single-CTA, kind::f16 (bf16/fp16 in, fp32 accumulate)
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on `bar` once every previously issued tcgen05.mma of this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ---------------------------------------------------------------- CTA-pair (cta_group::2) variants
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Both CTAs of a pair load their share of a stage; the transaction bytes of BOTH land on the leader's (rank 0) barrier:
// clearing bit 24 of the shared::cluster barrier address selects the even CTA of the pair.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"((uint64_t)m), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c_inner), "r"(c_outer)
      : "memory");
}
// The same load, multicast into the same shared-memory offset of every CTA in `mask` (16-bit cluster CTA mask); in each
// destination the bytes complete on the barrier of that CTA's pair leader.
__device__ __forceinline__ void tma_load_2d_pair_mc(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c_inner, int c_outer, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(
          smem_u32(smem_dst)),
      "l"((uint64_t)m), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c_inner), "r"(c_outer), "h"(mask)
      : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// D[tmem of both CTAs] (+)= A * B with M = 256 across the pair; issued by ONE thread of the leader CTA.
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on the barrier at this offset in BOTH CTAs once every previously issued MMA of this thread has completed.
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// ... in every CTA of `mask`
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(mask)
               : "memory");
}
// Arrive on the barrier at the same offset in CTA `rank` of the cluster.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, %1;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(rank)
      : "memory");
}

// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread i <-> TMEM lane base+i).
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n"
      ".reg .b32 rx;\n"
      ".reg .pred px;\n"
      "elect.sync rx|px, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, px;\n"
      "}\n"
      : "=r"(pred)::"memory");
  return pred != 0;
}

}  // namespace m3ae
