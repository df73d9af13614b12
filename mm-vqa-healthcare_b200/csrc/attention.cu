// Fused scaled-dot-product attention, forward and backward, flash-style: scores are never written to
// HBM; the running max / sum / output live in registers and are combined with warp shuffles.
// Reference arithmetic: BertSelfAttention core, bert_model.py:301 (QK^T), :319 (/sqrt(dh)), :320-322 (+mask),
// :325 (softmax), :334 (dropout on probs), :340 (PV), :342-344 (merge heads); nn.MultiheadAttention inside
// DecoderLayer, m3ae_decoder.py:60-61,78-79 (causal / no mask, dh = 96).
//
// This file holds the generic fp32-math kernels (T = float for the strict-parity path, T = bf16 storage
// with fp32 math as the always-correct reference for the tensor-core kernels in attention_mma.cu).
// One lane <-> one key inside a 32-key chunk; one warp <-> a few query rows.
#include "common.cuh"

namespace m3ae {
void count_launch();
int attn_fwd_mma(const m3ae_attn_params_t* p, cudaStream_t st);  // attention_mma.cu; returns M3AE_ERR_UNSUPPORTED if shape not covered
int attn_bwd_mma(const m3ae_attn_params_t* p, cudaStream_t st);
int attn_fwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st);   // attention_tc.cu (tcgen05/TMEM); M3AE_ERR_UNSUPPORTED if not covered / not profitable
int attn_bwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st);
int attn_fwd_narrow(const m3ae_attn_params_t* p, bool force, cudaStream_t st);  // attention_narrow.cu: one side <= 32 tokens (text), HBM-bound single-pass kernels
int attn_bwd_narrow(const m3ae_attn_params_t* p, cudaStream_t st);
// AUTO backend only; M3AE_ATTN_NARROW=0 in the environment sends these shapes to the tcgen05 / mma.sync tiles instead (A/B runs),
// M3AE_ATTN_NARROW=2 makes the forward take every covered shape (tests), default 1 = where it measured faster
static int narrow_mode() {
  static const int mode = [] { const char* e = getenv("M3AE_ATTN_NARROW"); return e ? atoi(e) : 1; }();
  return mode;
}
static bool narrow_on() { return narrow_mode() != 0; }

namespace {

constexpr int KC = 32;  // keys per chunk == warp width

template <typename T, int DH, int ROWS_PER_WARP>
__global__ void __launch_bounds__(128) attn_fwd_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int WARPS = 4;
  constexpr int ROWS = WARPS * ROWS_PER_WARP;
  constexpr int ND = DH / 32;
  __shared__ float sK[KC][DH + 1];
  __shared__ float sV[KC][DH];
  __shared__ float sQ[ROWS][DH];
  const int b = blockIdx.z, h = blockIdx.y;
  const int q0 = blockIdx.x * ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* __restrict__ Q = (const T*)P.q + (size_t)b * P.Lq * P.ldq + (size_t)h * DH;
  const T* __restrict__ K = (const T*)P.k + (size_t)b * P.Lk * P.ldk + (size_t)h * DH;
  const T* __restrict__ V = (const T*)P.v + (size_t)b * P.Lk * P.ldv + (size_t)h * DH;
  const float* __restrict__ mask = P.mask ? P.mask + (size_t)b * P.Lk : nullptr;

  for (int idx = threadIdx.x; idx < ROWS * DH; idx += 128) {
    const int r = idx / DH, d = idx % DH;
    sQ[r][d] = (q0 + r < P.Lq) ? to_f32<T>(Q[(size_t)(q0 + r) * P.ldq + d]) : 0.f;
  }
  float m[ROWS_PER_WARP], l[ROWS_PER_WARP], o[ROWS_PER_WARP][ND];
#pragma unroll
  for (int r = 0; r < ROWS_PER_WARP; ++r) {
    m[r] = -INFINITY; l[r] = 0.f;
#pragma unroll
    for (int i = 0; i < ND; ++i) o[r][i] = 0.f;
  }
  const float keep_scale = drop_keep_scale(P.drop_p);

  for (int k0 = 0; k0 < P.Lk; k0 += KC) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < KC * DH; idx += 128) {
      const int r = idx / DH, d = idx % DH;
      const bool ok = k0 + r < P.Lk;
      sK[r][d] = ok ? to_f32<T>(K[(size_t)(k0 + r) * P.ldk + d]) : 0.f;
      sV[r][d] = ok ? to_f32<T>(V[(size_t)(k0 + r) * P.ldv + d]) : 0.f;
    }
    __syncthreads();
    const int key = k0 + lane;
    const float madd = (key < P.Lk) ? (mask ? mask[key] : 0.f) : -INFINITY;
#pragma unroll
    for (int r = 0; r < ROWS_PER_WARP; ++r) {
      const int lr = warp * ROWS_PER_WARP + r;
      const int qi = q0 + lr;
      if (qi >= P.Lq) break;  // warp-uniform
      float s = 0.f;
#pragma unroll 16
      for (int d = 0; d < DH; ++d) s = fmaf(sQ[lr][d], sK[lane][d], s);
      s = s * P.scale + madd;
      if (P.causal && key > qi) s = -INFINITY;
      const float m_new = fmaxf(m[r], warp_max(s));
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float p = expf(s - m_use);
      const float corr = expf(m[r] - m_use);
      l[r] = l[r] * corr + warp_sum(p);
      m[r] = m_new;
      float pd = p;
      if (P.drop_p > 0.f && key < P.Lk) {
        const uint64_t rl = ((uint64_t)b * P.H + h) * P.Lq + qi;
        pd = attn_dropout_keep(P.drop_seed, P.drop_stream, rl, key, P.drop_p) ? p * keep_scale : 0.f;
      }
#pragma unroll
      for (int i = 0; i < ND; ++i) o[r][i] *= corr;
#pragma unroll 8
      for (int j = 0; j < KC; ++j) {
        const float pj = __shfl_sync(0xffffffffu, pd, j);
#pragma unroll
        for (int i = 0; i < ND; ++i) o[r][i] = fmaf(pj, sV[j][i * 32 + lane], o[r][i]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < ROWS_PER_WARP; ++r) {
    const int qi = q0 + warp * ROWS_PER_WARP + r;
    if (qi >= P.Lq) break;
    const float inv = l[r] > 0.f ? 1.0f / l[r] : 0.f;
    T* out = (T*)P.out + (size_t)(b * P.Lq + qi) * P.ldo + (size_t)h * DH;
#pragma unroll
    for (int i = 0; i < ND; ++i) out[i * 32 + lane] = from_f32<T>(o[r][i] * inv);
    if (lane == 0 && P.lse) P.lse[((size_t)b * P.H + h) * P.Lq + qi] = m[r] + logf(l[r]);
  }
}

// probs[b,h,i,j] = dropout(exp(scale*q_i.k_j + mask_j - lse_i))  -- only for output_attentions=True
template <typename T, int DH>
__global__ void __launch_bounds__(128) attn_probs_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int b = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (qi >= P.Lq) return;
  const T* __restrict__ Q = (const T*)P.q + (size_t)(b * P.Lq + qi) * P.ldq + (size_t)h * DH;
  const T* __restrict__ K = (const T*)P.k + (size_t)b * P.Lk * P.ldk + (size_t)h * DH;
  const float lse = P.lse[((size_t)b * P.H + h) * P.Lq + qi];
  const float keep_scale = drop_keep_scale(P.drop_p);
  float* out = P.probs + (((size_t)b * P.H + h) * P.Lq + qi) * (size_t)P.Lk;
  for (int key = lane; key < P.Lk; key += 32) {
    float s = 0.f;
    for (int d = 0; d < DH; ++d) s = fmaf(to_f32<T>(Q[d]), to_f32<T>(K[(size_t)key * P.ldk + d]), s);
    s = s * P.scale + (P.mask ? P.mask[(size_t)b * P.Lk + key] : 0.f);
    float p = (P.causal && key > qi) ? 0.f : expf(s - lse);
    if (P.drop_p > 0.f) {
      const uint64_t rl = ((uint64_t)b * P.H + h) * P.Lq + qi;
      p = attn_dropout_keep(P.drop_seed, P.drop_stream, rl, key, P.drop_p) ? p * keep_scale : 0.f;
    }
    out[key] = p;
  }
}

// delta[b,h,i] = sum_d dO[i,d] * O[i,d]
template <typename T, int DH>
__global__ void __launch_bounds__(128) attn_delta_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int b = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (qi >= P.Lq) return;
  const size_t off = (size_t)(b * P.Lq + qi) * P.ldo + (size_t)h * DH;
  const T* __restrict__ O = (const T*)P.out + off;
  const T* __restrict__ dO = (const T*)P.dout + off;
  float s = 0.f;
  for (int d = lane; d < DH; d += 32) s += to_f32<T>(O[d]) * to_f32<T>(dO[d]);
  s = warp_sum(s);
  if (lane == 0) P.delta[((size_t)b * P.H + h) * P.Lq + qi] = s;
}

// bf16, dh = 64, heads packed side by side in a row (the fusion stack's ctx layout): one warp per token row, 16-byte loads,
// 8 lanes share a head.  Streams O and dO once at full sector efficiency (the generic kernel above issues 2-byte loads).
__global__ void __launch_bounds__(256) attn_delta_rows_bf16_64_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);  // token row b*Lq + qi
  if (r >= (int64_t)P.B * P.Lq) return;
  const int b = (int)(r / P.Lq), qi = (int)(r % P.Lq);
  const bf16* __restrict__ O = (const bf16*)P.out + r * P.ldo;
  const bf16* __restrict__ dO = (const bf16*)P.dout + r * P.ldo;
  const int nvec = P.H * 8;  // 16-byte vectors per row
  for (int v = lane; v < ((nvec + 31) & ~31); v += 32) {
    float s = 0.f;
    if (v < nvec) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(O) + v), g = __ldg(reinterpret_cast<const uint4*>(dO) + v);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 x = unpack_bf16x2(aw[j]), y = unpack_bf16x2(gw[j]);
        s = fmaf(x.x, y.x, s);
        s = fmaf(x.y, y.y, s);
      }
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if ((lane & 7) == 0 && v < nvec) P.delta[((size_t)b * P.H + (v >> 3)) * P.Lq + qi] = s;
  }
}

// dQ_i = scale * sum_j dS_ij K_j,  dS_ij = P_ij (dP_ij - delta_i),  dP_ij = keep_ij/(1-p) * dO_i.V_j
template <typename T, int DH>
__global__ void __launch_bounds__(128) attn_bwd_dq_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int RPW = 4, ROWS = 16, ND = DH / 32;
  __shared__ float sK[KC][DH + 1];
  __shared__ float sV[KC][DH + 1];
  __shared__ float sQ[ROWS][DH];
  __shared__ float sdO[ROWS][DH];
  const int b = blockIdx.z, h = blockIdx.y;
  const int q0 = blockIdx.x * ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* __restrict__ Q = (const T*)P.q + (size_t)b * P.Lq * P.ldq + (size_t)h * DH;
  const T* __restrict__ K = (const T*)P.k + (size_t)b * P.Lk * P.ldk + (size_t)h * DH;
  const T* __restrict__ V = (const T*)P.v + (size_t)b * P.Lk * P.ldv + (size_t)h * DH;
  const T* __restrict__ dO = (const T*)P.dout + (size_t)b * P.Lq * P.ldo + (size_t)h * DH;
  const float* __restrict__ mask = P.mask ? P.mask + (size_t)b * P.Lk : nullptr;
  for (int idx = threadIdx.x; idx < ROWS * DH; idx += 128) {
    const int r = idx / DH, d = idx % DH;
    const bool ok = q0 + r < P.Lq;
    sQ[r][d] = ok ? to_f32<T>(Q[(size_t)(q0 + r) * P.ldq + d]) : 0.f;
    sdO[r][d] = ok ? to_f32<T>(dO[(size_t)(q0 + r) * P.ldo + d]) : 0.f;
  }
  float lse[RPW], dl[RPW], dq[RPW][ND];
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int qi = q0 + warp * RPW + r;
    const size_t si = ((size_t)b * P.H + h) * P.Lq + (qi < P.Lq ? qi : 0);
    lse[r] = P.lse[si]; dl[r] = P.delta[si];
#pragma unroll
    for (int i = 0; i < ND; ++i) dq[r][i] = 0.f;
  }
  const float keep_scale = drop_keep_scale(P.drop_p);
  for (int k0 = 0; k0 < P.Lk; k0 += KC) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < KC * DH; idx += 128) {
      const int r = idx / DH, d = idx % DH;
      const bool ok = k0 + r < P.Lk;
      sK[r][d] = ok ? to_f32<T>(K[(size_t)(k0 + r) * P.ldk + d]) : 0.f;
      sV[r][d] = ok ? to_f32<T>(V[(size_t)(k0 + r) * P.ldv + d]) : 0.f;
    }
    __syncthreads();
    const int key = k0 + lane;
    const float madd = (key < P.Lk) ? (mask ? mask[key] : 0.f) : -INFINITY;
#pragma unroll
    for (int r = 0; r < RPW; ++r) {
      const int lr = warp * RPW + r;
      const int qi = q0 + lr;
      if (qi >= P.Lq) break;
      float s = 0.f, dp = 0.f;
#pragma unroll 16
      for (int d = 0; d < DH; ++d) {
        s = fmaf(sQ[lr][d], sK[lane][d], s);
        dp = fmaf(sdO[lr][d], sV[lane][d], dp);
      }
      s = s * P.scale + madd;
      if (P.causal && key > qi) s = -INFINITY;
      const float p = expf(s - lse[r]);
      if (P.drop_p > 0.f && key < P.Lk) {
        const uint64_t rl = ((uint64_t)b * P.H + h) * P.Lq + qi;
        dp = attn_dropout_keep(P.drop_seed, P.drop_stream, rl, key, P.drop_p) ? dp * keep_scale : 0.f;
      }
      const float ds = p * (dp - dl[r]) * P.scale;
#pragma unroll 8
      for (int j = 0; j < KC; ++j) {
        const float dsj = __shfl_sync(0xffffffffu, ds, j);
#pragma unroll
        for (int i = 0; i < ND; ++i) dq[r][i] = fmaf(dsj, sK[j][i * 32 + lane], dq[r][i]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int qi = q0 + warp * RPW + r;
    if (qi >= P.Lq) break;
    T* out = (T*)P.dq + (size_t)(b * P.Lq + qi) * P.lddq + (size_t)h * DH;
#pragma unroll
    for (int i = 0; i < ND; ++i) out[i * 32 + lane] = from_f32<T>(dq[r][i]);
  }
}

// dV_j = sum_i P'_ij dO_i,  dK_j = scale * sum_i dS_ij Q_i  (one lane <-> one query inside a 32-query chunk)
template <typename T, int DH>
__global__ void __launch_bounds__(128) attn_bwd_dkv_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int RPW = 4, ROWS = 16, ND = DH / 32;
  __shared__ float sQ[KC][DH + 1];
  __shared__ float sdO[KC][DH + 1];
  __shared__ float sK[ROWS][DH];
  __shared__ float sV[ROWS][DH];
  __shared__ float sLse[KC], sDl[KC];
  const int b = blockIdx.z, h = blockIdx.y;
  const int kbase = blockIdx.x * ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* __restrict__ Q = (const T*)P.q + (size_t)b * P.Lq * P.ldq + (size_t)h * DH;
  const T* __restrict__ K = (const T*)P.k + (size_t)b * P.Lk * P.ldk + (size_t)h * DH;
  const T* __restrict__ V = (const T*)P.v + (size_t)b * P.Lk * P.ldv + (size_t)h * DH;
  const T* __restrict__ dO = (const T*)P.dout + (size_t)b * P.Lq * P.ldo + (size_t)h * DH;
  const float* __restrict__ mask = P.mask ? P.mask + (size_t)b * P.Lk : nullptr;
  for (int idx = threadIdx.x; idx < ROWS * DH; idx += 128) {
    const int r = idx / DH, d = idx % DH;
    const bool ok = kbase + r < P.Lk;
    sK[r][d] = ok ? to_f32<T>(K[(size_t)(kbase + r) * P.ldk + d]) : 0.f;
    sV[r][d] = ok ? to_f32<T>(V[(size_t)(kbase + r) * P.ldv + d]) : 0.f;
  }
  float dk[RPW][ND], dv[RPW][ND], madd[RPW];
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int key = kbase + warp * RPW + r;
    madd[r] = (mask && key < P.Lk) ? mask[key] : 0.f;
#pragma unroll
    for (int i = 0; i < ND; ++i) { dk[r][i] = 0.f; dv[r][i] = 0.f; }
  }
  const float keep_scale = drop_keep_scale(P.drop_p);
  for (int i0 = 0; i0 < P.Lq; i0 += KC) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < KC * DH; idx += 128) {
      const int r = idx / DH, d = idx % DH;
      const bool ok = i0 + r < P.Lq;
      sQ[r][d] = ok ? to_f32<T>(Q[(size_t)(i0 + r) * P.ldq + d]) : 0.f;
      sdO[r][d] = ok ? to_f32<T>(dO[(size_t)(i0 + r) * P.ldo + d]) : 0.f;
    }
    if (threadIdx.x < KC) {
      const int qi = i0 + threadIdx.x;
      const size_t si = ((size_t)b * P.H + h) * P.Lq + (qi < P.Lq ? qi : 0);
      sLse[threadIdx.x] = P.lse[si];
      sDl[threadIdx.x] = P.delta[si];
    }
    __syncthreads();
    const int qi = i0 + lane;
    const bool q_ok = qi < P.Lq;
#pragma unroll
    for (int r = 0; r < RPW; ++r) {
      const int lr = warp * RPW + r;
      const int key = kbase + lr;
      if (key >= P.Lk) break;
      float s = 0.f, dp = 0.f;
#pragma unroll 16
      for (int d = 0; d < DH; ++d) {
        s = fmaf(sK[lr][d], sQ[lane][d], s);
        dp = fmaf(sV[lr][d], sdO[lane][d], dp);
      }
      s = s * P.scale + madd[r];
      float p = (q_ok && !(P.causal && key > qi)) ? expf(s - sLse[lane]) : 0.f;
      float pd = p;
      if (P.drop_p > 0.f && q_ok) {
        const uint64_t rl = ((uint64_t)b * P.H + h) * P.Lq + qi;
        const bool keep = attn_dropout_keep(P.drop_seed, P.drop_stream, rl, key, P.drop_p);
        pd = keep ? p * keep_scale : 0.f;
        dp = keep ? dp * keep_scale : 0.f;
      }
      const float ds = p * (dp - sDl[lane]) * P.scale;
#pragma unroll 8
      for (int j = 0; j < KC; ++j) {
        const float pj = __shfl_sync(0xffffffffu, pd, j);
        const float dsj = __shfl_sync(0xffffffffu, ds, j);
#pragma unroll
        for (int i = 0; i < ND; ++i) {
          dv[r][i] = fmaf(pj, sdO[j][i * 32 + lane], dv[r][i]);
          dk[r][i] = fmaf(dsj, sQ[j][i * 32 + lane], dk[r][i]);
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    const int key = kbase + warp * RPW + r;
    if (key >= P.Lk) break;
    T* ok_ = (T*)P.dk + (size_t)(b * P.Lk + key) * P.lddk + (size_t)h * DH;
    T* ov_ = (T*)P.dv + (size_t)(b * P.Lk + key) * P.lddv + (size_t)h * DH;
#pragma unroll
    for (int i = 0; i < ND; ++i) {
      ok_[i * 32 + lane] = from_f32<T>(dk[r][i]);
      ov_[i * 32 + lane] = from_f32<T>(dv[r][i]);
    }
  }
}

// One query token per (b,h) against a KV cache of `len` keys (greedy decoding step).
// HBM-bound: reads len*dh*2*sizeof(T) bytes per (b,h); grid = B*H CTAs of 4 warps splitting the keys.
template <typename T, int DH>
__global__ void __launch_bounds__(128) decode_attn_kernel(const T* __restrict__ q, int64_t ldq, const T* __restrict__ kc,
                                                         const T* __restrict__ vc, int64_t ldk, int64_t ldv, int64_t bsk,
                                                         int64_t bsv, T* __restrict__ out, int64_t ldo, int H, int len,
                                                         float scale) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int ND = DH / 32;
  __shared__ float sm[4], sl[4], so[4][DH];
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* Q = q + (size_t)b * ldq + (size_t)h * DH;
  const T* K = kc + (size_t)b * bsk + (size_t)h * DH;
  const T* V = vc + (size_t)b * bsv + (size_t)h * DH;
  float qv[ND];
#pragma unroll
  for (int i = 0; i < ND; ++i) qv[i] = to_f32<T>(Q[i * 32 + lane]) * scale;
  float m = -INFINITY, l = 0.f, o[ND];
#pragma unroll
  for (int i = 0; i < ND; ++i) o[i] = 0.f;
  // each warp takes keys warp, warp+4, ...; lanes split the head dimension (coalesced 64/192-byte rows)
  for (int j = warp; j < len; j += 4) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ND; ++i) s = fmaf(qv[i], to_f32<T>(K[(size_t)j * ldk + i * 32 + lane]), s);
    s = warp_sum(s);
    const float m_new = fmaxf(m, s);
    const float corr = expf(m - m_new), p = expf(s - m_new);
    l = l * corr + p;
#pragma unroll
    for (int i = 0; i < ND; ++i) o[i] = o[i] * corr + p * to_f32<T>(V[(size_t)j * ldv + i * 32 + lane]);
    m = m_new;
  }
  if (lane == 0) { sm[warp] = m; sl[warp] = l; }
#pragma unroll
  for (int i = 0; i < ND; ++i) so[warp][i * 32 + lane] = o[i];
  __syncthreads();
  if (warp == 0) {
    float M = fmaxf(fmaxf(sm[0], sm[1]), fmaxf(sm[2], sm[3]));
    float L = 0.f;
    float acc[ND];
#pragma unroll
    for (int i = 0; i < ND; ++i) acc[i] = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float c = (sm[w] == -INFINITY) ? 0.f : expf(sm[w] - M);
      L += sl[w] * c;
#pragma unroll
      for (int i = 0; i < ND; ++i) acc[i] += so[w][i * 32 + lane] * c;
    }
    T* O = out + (size_t)b * ldo + (size_t)h * DH;
#pragma unroll
    for (int i = 0; i < ND; ++i) O[i * 32 + lane] = from_f32<T>(acc[i] / L);
  }
}

// bf16 decode attention built for bandwidth (the cross-attention step streams B*Lm*1536*2 B = 141 MB of cached K/V).
// The generic kernel above walks the keys one by one per warp (2-byte loads -> shuffle reduction -> exp -> V loads, a
// dependent chain per key: ~1 us per key, 102 us per call at B=128, Lm=359).  Here every phase issues independent 16-byte
// loads: (1) scores, 4 lanes per key, 32 keys per CTA iteration; (2) block softmax over the scores in shared memory;
// (3) O = P V with one thread per (16-byte slice of the head dimension, key residue), reduced through shared memory.
// `stream_kv`: the cache is larger than what L2 can keep between two token steps (cross-attention memory: 141 MB per step at
// B = 128): its 16-byte loads carry an L2 evict-first policy (and skip L1), so the pass does not push the decoder's ~66 MB of
// projection / vocabulary weights — re-read by every token step — out of the 126 MB L2.
constexpr int DEC_MAX_LEN = 1024;
__device__ __forceinline__ uint4 ld_kv16(const bf16* p, uint64_t policy, bool stream_kv) {
  uint4 v;
  if (stream_kv)
    asm("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.b32 {%0,%1,%2,%3}, [%4], %5;"
        : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(policy));
  else
    v = __ldg(reinterpret_cast<const uint4*>(p));
  return v;
}
template <int DH>
__global__ void __launch_bounds__(128) decode_attn_bf16_kernel(const bf16* __restrict__ q, int64_t ldq, const bf16* __restrict__ kc,
                                                              const bf16* __restrict__ vc, int64_t ldk, int64_t ldv, int64_t bsk,
                                                              int64_t bsv, bf16* __restrict__ out, int64_t ldo, int H, int len,
                                                              float scale, int stream_kv_i) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const bool stream_kv = stream_kv_i != 0;
  uint64_t policy = 0;
  if (stream_kv) asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  constexpr int NCH = DH / 8;        // 16-byte chunks per row (12 for dh 96, 8 for dh 64)
  constexpr int CPL = NCH / 4;       // chunks per lane in the score phase (4 lanes per key)
  constexpr int NG = 128 / NCH;      // key residues in the P V phase (10 / 16)
  __shared__ float sc[DEC_MAX_LEN];
  __shared__ float red[4];
  __shared__ float part[NG][DH];
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bf16* Q = q + (size_t)b * ldq + (size_t)h * DH;
  const bf16* K = kc + (size_t)b * bsk + (size_t)h * DH;
  const bf16* V = vc + (size_t)b * bsv + (size_t)h * DH;
  // ---- (1) scores: lane quad `kq` of the CTA takes keys kq, kq+32, ...; lane `sub` of the quad takes chunks sub*CPL..
  const int kq = tid >> 2, sub = tid & 3;
  float qf[CPL][8];
#pragma unroll
  for (int c = 0; c < CPL; ++c) {
    const uint4 w = *reinterpret_cast<const uint4*>(Q + (sub * CPL + c) * 8);
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(ww[j]);
      qf[c][2 * j] = f.x * scale;
      qf[c][2 * j + 1] = f.y * scale;
    }
  }
  float mloc = -INFINITY;
  for (int j0 = 0; j0 < len; j0 += 64) {  // two keys in flight per thread
    float s2[2] = {0.f, 0.f};
    uint4 kw[2][CPL];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = j0 + u * 32 + kq;
      if (j < len) {
#pragma unroll
        for (int c = 0; c < CPL; ++c) kw[u][c] = ld_kv16(K + (size_t)j * ldk + (sub * CPL + c) * 8, policy, stream_kv);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = j0 + u * 32 + kq;
      if (j < len) {
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
          const uint32_t ww[4] = {kw[u][c].x, kw[u][c].y, kw[u][c].z, kw[u][c].w};
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const float2 f = unpack_bf16x2(ww[t]);
            s2[u] = fmaf(qf[c][2 * t], f.x, s2[u]);
            s2[u] = fmaf(qf[c][2 * t + 1], f.y, s2[u]);
          }
        }
      }
      // the 4 lanes of a quad hold partial dot products of the same key (all 32 lanes take part in the shuffles)
      s2[u] += __shfl_xor_sync(0xffffffffu, s2[u], 1);
      s2[u] += __shfl_xor_sync(0xffffffffu, s2[u], 2);
      if (j < len) {
        if (sub == 0) sc[j] = s2[u];
        mloc = fmaxf(mloc, s2[u]);
      }
    }
  }
  // ---- (2) softmax over sc[0..len)
  mloc = warp_max(mloc);
  if (lane == 0) red[warp] = mloc;
  __syncthreads();
  const float m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float lsum = 0.f;
  for (int j = tid; j < len; j += 128) {
    const float p = __expf(sc[j] - m);
    sc[j] = p;
    lsum += p;
  }
  lsum = warp_sum(lsum);
  if (lane == 0) red[warp] = lsum;
  __syncthreads();
  const float inv_l = 1.0f / ((red[0] + red[1]) + (red[2] + red[3]));
  // ---- (3) O = P V: thread = (chunk ch, key residue g); keys g, g+NG, ...; 4 keys in flight
  const int ch = tid % NCH, g = tid / NCH;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  if (g < NG) {
    for (int j0 = g; j0 < len; j0 += 4 * NG) {
      uint4 vw[4];
      float p[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + u * NG;
        p[u] = j < len ? sc[j] : 0.f;
        vw[u] = j < len ? ld_kv16(V + (size_t)j * ldv + ch * 8, policy, stream_kv) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t ww[4] = {vw[u].x, vw[u].y, vw[u].z, vw[u].w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = unpack_bf16x2(ww[t]);
          acc[2 * t] = fmaf(p[u], f.x, acc[2 * t]);
          acc[2 * t + 1] = fmaf(p[u], f.y, acc[2 * t + 1]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) part[g][ch * 8 + i] = acc[i];
  }
  __syncthreads();
  if (tid < DH) {
    float o = 0.f;
#pragma unroll
    for (int gg = 0; gg < NG; ++gg) o += part[gg][tid];
    out[(size_t)b * ldo + (size_t)h * DH + tid] = __float2bfloat16_rn(o * inv_l);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) argmax_rows_kernel(const T* __restrict__ x, int64_t ld, int64_t* __restrict__ out, int N) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  __shared__ float sv[8];
  __shared__ int si[8];
  const T* row = x + (size_t)blockIdx.x * ld;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int c = threadIdx.x; c < N; c += 256) {
    const float v = to_f32<T>(row[c]);
    if (v > best || (v == best && c < bi)) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    out[blockIdx.x] = (int64_t)bi;
  }
}

template <typename TO>
__global__ void decoder_embed_kernel(const int64_t* __restrict__ ids, const int32_t* __restrict__ pos,
                                     const float* __restrict__ table, const float* __restrict__ pe, TO* __restrict__ out, int D) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  // target_embed += positional_encoding(target_embed)  ==> 2*E[id] + pe[pos]   (m3ae_decoder.py:125-127, :36)
  const int64_t r = blockIdx.x;
  const float* e = table + (size_t)ids[r] * D;
  const float* p = pe + (size_t)pos[r] * D;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float v = e[d];
    out[(size_t)r * D + d] = from_f32<TO>(v + (v + p[d]));
  }
}

template <typename T>
void launch_probs(const m3ae_attn_params_t* p, cudaStream_t st) {
  dim3 g2((p->Lq + 3) / 4, p->H, p->B);
  if (p->dh == 64) M3AE_LAUNCH((attn_probs_kernel<T, 64>), g2, 128, 0, st, *p);
  else M3AE_LAUNCH((attn_probs_kernel<T, 96>), g2, 128, 0, st, *p);
  count_launch();
}

template <typename T>
int launch_generic_fwd(const m3ae_attn_params_t* p, cudaStream_t st) {
  if (p->dh == 64) {
    dim3 grid((p->Lq + 31) / 32, p->H, p->B);
    M3AE_LAUNCH((attn_fwd_kernel<T, 64, 8>), grid, 128, 0, st, *p);
  } else {
    dim3 grid((p->Lq + 15) / 16, p->H, p->B);
    M3AE_LAUNCH((attn_fwd_kernel<T, 96, 4>), grid, 128, 0, st, *p);
  }
  count_launch();
  if (p->probs) launch_probs<T>(p, st);
  return M3AE_OK;
}

template <typename T>
int launch_generic_bwd(const m3ae_attn_params_t* p, cudaStream_t st) {
  dim3 g0((p->Lq + 3) / 4, p->H, p->B);
  dim3 g1((p->Lq + 15) / 16, p->H, p->B);
  dim3 g2((p->Lk + 15) / 16, p->H, p->B);
  if (p->dh == 64) {
    M3AE_LAUNCH((attn_delta_kernel<T, 64>), g0, 128, 0, st, *p);
    M3AE_LAUNCH((attn_bwd_dq_kernel<T, 64>), g1, 128, 0, st, *p);
    M3AE_LAUNCH((attn_bwd_dkv_kernel<T, 64>), g2, 128, 0, st, *p);
  } else {
    M3AE_LAUNCH((attn_delta_kernel<T, 96>), g0, 128, 0, st, *p);
    M3AE_LAUNCH((attn_bwd_dq_kernel<T, 96>), g1, 128, 0, st, *p);
    M3AE_LAUNCH((attn_bwd_dkv_kernel<T, 96>), g2, 128, 0, st, *p);
  }
  count_launch(); count_launch(); count_launch();
  return M3AE_OK;
}

int validate(const m3ae_attn_params_t* p, bool bwd) {
  M3AE_CHECK(p != nullptr, M3AE_ERR_INVALID, "attention: null params");
  M3AE_CHECK(p->B > 0 && p->H > 0 && p->Lq > 0 && p->Lk > 0, M3AE_ERR_INVALID, "attention: empty shape B=%d H=%d Lq=%d Lk=%d", p->B, p->H, p->Lq, p->Lk);
  M3AE_CHECK(p->dh == 64 || p->dh == 96, M3AE_ERR_UNSUPPORTED, "attention: head dim %d unsupported (64 for the fusion stack, 96 for the decoder)", p->dh);
  M3AE_CHECK(p->dtype == M3AE_F32 || p->dtype == M3AE_BF16, M3AE_ERR_INVALID, "attention: bad dtype %d", p->dtype);
  M3AE_CHECK(p->q && p->k && p->v && p->out && p->lse, M3AE_ERR_INVALID, "attention: null tensor pointer");
  M3AE_CHECK(p->drop_p >= 0.f && p->drop_p < 1.f, M3AE_ERR_INVALID, "attention: dropout %f not in [0,1)", p->drop_p);
  if (bwd) M3AE_CHECK(p->dout && p->dq && p->dk && p->dv && p->delta, M3AE_ERR_INVALID, "attention backward: null gradient pointer");
  return M3AE_OK;
}

}  // namespace

void launch_delta_bf16_64(const m3ae_attn_params_t* p, cudaStream_t st) {
  if ((p->ldo % 8) == 0 && (((uintptr_t)p->out | (uintptr_t)p->dout) & 15) == 0) {
    const int64_t rows = (int64_t)p->B * p->Lq;
    M3AE_LAUNCH((attn_delta_rows_bf16_64_kernel), (unsigned)((rows + 7) / 8), 256, 0, st, *p);
  } else {
    dim3 g0((p->Lq + 3) / 4, p->H, p->B);
    M3AE_LAUNCH((attn_delta_kernel<bf16, 64>), g0, 128, 0, st, *p);
  }
  count_launch();
}

int attn_fwd(const m3ae_attn_params_t* p, int backend, cudaStream_t st) {
  int rc = validate(p, false);
  if (rc) return rc;
  if (backend != M3AE_GEMM_SIMT && p->dtype == M3AE_BF16) {
    rc = ((backend == M3AE_GEMM_AUTO && narrow_on()) || backend == M3AE_ATTN_NARROW) ? attn_fwd_narrow(p, narrow_mode() == 2 || backend == M3AE_ATTN_NARROW, st) : M3AE_ERR_UNSUPPORTED;
    if (rc == M3AE_ERR_UNSUPPORTED) rc = backend == M3AE_ATTN_MMA_SYNC ? M3AE_ERR_UNSUPPORTED : attn_fwd_tc(p, backend == M3AE_GEMM_TCGEN05, st);
    if (rc == M3AE_ERR_UNSUPPORTED) rc = attn_fwd_mma(p, st);
    if (rc == M3AE_OK && p->probs) {
      launch_probs<bf16>(p, st);
      M3AE_CUDA(cudaGetLastError());
    }
    if (rc != M3AE_ERR_UNSUPPORTED) return rc;
  }
  if (p->dtype == M3AE_BF16) launch_generic_fwd<bf16>(p, st);
  else launch_generic_fwd<float>(p, st);
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

// Attention probabilities only, from q, k, mask and the log-sum-exp a forward call saved: what output_attentions=True
// returns (bert_model.py:346).  Run on demand so that the hard-wired output_attentions=True of the reference's fusion
// loop (m3ae_module.py:276-277) does not force 324 MB of fp32 probabilities per image layer through HBM.
int attn_probs(const m3ae_attn_params_t* p, cudaStream_t st) {
  M3AE_CHECK(p != nullptr && p->q && p->k && p->lse && p->probs, M3AE_ERR_INVALID, "attn_probs: q, k, lse and probs are required");
  M3AE_CHECK(p->B > 0 && p->H > 0 && p->Lq > 0 && p->Lk > 0, M3AE_ERR_INVALID, "attn_probs: empty shape");
  M3AE_CHECK(p->dh == 64 || p->dh == 96, M3AE_ERR_UNSUPPORTED, "attn_probs: head dim %d unsupported", p->dh);
  if (p->dtype == M3AE_BF16) launch_probs<bf16>(p, st);
  else launch_probs<float>(p, st);
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int attn_bwd(const m3ae_attn_params_t* p, int backend, cudaStream_t st) {
  int rc = validate(p, true);
  if (rc) return rc;
  if (backend != M3AE_GEMM_SIMT && p->dtype == M3AE_BF16) {
    rc = ((backend == M3AE_GEMM_AUTO && narrow_on()) || backend == M3AE_ATTN_NARROW) ? attn_bwd_narrow(p, st) : M3AE_ERR_UNSUPPORTED;
    if (rc == M3AE_ERR_UNSUPPORTED) rc = backend == M3AE_ATTN_MMA_SYNC ? M3AE_ERR_UNSUPPORTED : attn_bwd_tc(p, backend == M3AE_GEMM_TCGEN05, st);
    if (rc == M3AE_ERR_UNSUPPORTED) rc = attn_bwd_mma(p, st);
    if (rc != M3AE_ERR_UNSUPPORTED) return rc;
  }
  if (p->dtype == M3AE_BF16) launch_generic_bwd<bf16>(p, st);
  else launch_generic_bwd<float>(p, st);
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int decode_attn(const void* q, int64_t ldq, const void* kc, const void* vc, int64_t ldk, int64_t ldv, int64_t bsk, int64_t bsv,
                void* out, int64_t ldo, int B, int H, int len, int dh, float scale, int dtype, cudaStream_t st) {
  M3AE_CHECK(B > 0 && H > 0 && len > 0, M3AE_ERR_INVALID, "decode_attn: empty shape");
  M3AE_CHECK(dh == 64 || dh == 96, M3AE_ERR_UNSUPPORTED, "decode_attn: head dim %d unsupported", dh);
  // bf16 with 16-byte aligned rows and a cache that fits the score buffer: the bandwidth kernel
  if (dtype == M3AE_BF16 && len <= DEC_MAX_LEN && ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && bsk % 8 == 0 && bsv % 8 == 0 &&
      ((((uintptr_t)q) | ((uintptr_t)kc) | ((uintptr_t)vc)) & 15) == 0) {
    // K + V bytes of this call; above 32 MB the cache is streamed with an L2 evict-first policy (M3AE_DECODE_EVICT_FIRST=0: off, A/B)
    static const bool evict_env = [] { const char* e = getenv("M3AE_DECODE_EVICT_FIRST"); return !(e && e[0] == '0'); }();
    const int stream_kv = evict_env && (int64_t)B * len * H * dh * 4 > (int64_t)32 << 20 ? 1 : 0;
    if (dh == 64) M3AE_LAUNCH((decode_attn_bf16_kernel<64>), B * H, 128, 0, st, (const bf16*)q, ldq, (const bf16*)kc, (const bf16*)vc, ldk, ldv, bsk, bsv, (bf16*)out, ldo, H, len, scale, stream_kv);
    else M3AE_LAUNCH((decode_attn_bf16_kernel<96>), B * H, 128, 0, st, (const bf16*)q, ldq, (const bf16*)kc, (const bf16*)vc, ldk, ldv, bsk, bsv, (bf16*)out, ldo, H, len, scale, stream_kv);
    count_launch();
    M3AE_CUDA(cudaGetLastError());
    return M3AE_OK;
  }
#define DA(T, DH) M3AE_LAUNCH((decode_attn_kernel<T, DH>), B * H, 128, 0, st, (const T*)q, ldq, (const T*)kc, (const T*)vc, ldk, ldv, bsk, bsv, (T*)out, ldo, H, len, scale)
  if (dtype == M3AE_BF16) { if (dh == 64) DA(bf16, 64); else DA(bf16, 96); }
  else { if (dh == 64) DA(float, 64); else DA(float, 96); }
#undef DA
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

// Row argmax from per-64-column-group partials [G][M] (GEMM epilogue, epi.amax_val / amax_idx): one block per row, every
// thread takes <= ceil(G/256) groups (independent loads: the first version walked 60 groups per warp in a dependent chain on
// 4 blocks, 20 us per token step); larger value wins, then the smaller column.
__global__ void __launch_bounds__(256) argmax_partials_kernel(const float* __restrict__ val, const int32_t* __restrict__ idx, int G, int M,
                                                             int64_t* __restrict__ out) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  __shared__ float sv[8];
  __shared__ int si[8];
  const int row = blockIdx.x;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int g = threadIdx.x; g < G; g += 256) {
    const float v = val[(size_t)g * M + row];
    const int i = idx[(size_t)g * M + row];
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    out[row] = (int64_t)bi;
  }
}
int argmax_partials(const float* val, const int32_t* idx, int G, int M, int64_t* out, cudaStream_t st) {
  M3AE_LAUNCH((argmax_partials_kernel), (unsigned)M, 256, 0, st, val, idx, G, M, out);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

// Token hand-over of the greedy decoder in ONE launch (one warp per sequence): row argmax from the vocabulary projection's
// per-group partials (as argmax_partials_kernel; skipped when val == NULL: ids are read instead) -> ids_out, then
// x = 2 E[id] + pe[pos] (decoder_embed_kernel) and xn = pre_norm(x) (ln_fwd_kernel's arithmetic, summation order and bf16
// rounding points, so the fused path is bit-identical to the three launches it replaces).  D = 768, bf16 activations.
__global__ void __launch_bounds__(32) decoder_next_token_kernel(const float* __restrict__ val, const int32_t* __restrict__ idx, int G, int M,
                                                                const int64_t* __restrict__ ids_in, int64_t* __restrict__ ids_out, int pos,
                                                                const float* __restrict__ table, const float* __restrict__ pe,
                                                                const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                                                bf16* __restrict__ x, bf16* __restrict__ xn) {
  pdl_sync();
  constexpr int D = 768;
  const int row = blockIdx.x, lane = threadIdx.x;
  int64_t id;
  if (val != nullptr) {
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int g = lane; g < G; g += 32) {
      const float v = val[(size_t)g * M + row];
      const int i = idx[(size_t)g * M + row];
      if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    id = bi;
    if (lane == 0) ids_out[row] = id;
  } else {
    id = ids_in[row];
  }
  const float* e = table + (size_t)id * D;
  const float* p = pe + (size_t)pos * D;
  float v[3][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int c = (i * 32 + lane) * 8;
    const float4 e0 = *reinterpret_cast<const float4*>(e + c), e1 = *reinterpret_cast<const float4*>(e + c + 4);
    const float4 p0 = *reinterpret_cast<const float4*>(p + c), p1 = *reinterpret_cast<const float4*>(p + c + 4);
    const float ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w}, pv[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) w[j] = pack_bf16x2(ev[2 * j] + (ev[2 * j] + pv[2 * j]), ev[2 * j + 1] + (ev[2 * j + 1] + pv[2 * j + 1]));
    *reinterpret_cast<uint4*>(x + (size_t)row * D + c) = make_uint4(w[0], w[1], w[2], w[3]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {  // LayerNorm sees the bf16-rounded x, as the separate launch does
      const float2 f = unpack_bf16x2(w[j]);
      v[i][2 * j] = f.x; v[i][2 * j + 1] = f.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[i][j];
  }
  if (xn == nullptr) return;
  const float mean = warp_sum(s) / (float)D;
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
  const float var = warp_sum(ss) / (float)D;
  const float rstd = rsqrtf(var + eps);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int c = (i * 32 + lane) * 8;
    const float4 g0 = *reinterpret_cast<const float4*>(gamma + c), g1 = *reinterpret_cast<const float4*>(gamma + c + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(beta + c), b1 = *reinterpret_cast<const float4*>(beta + c + 4);
    const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w}, bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * gv[j] + bv[j];
    *reinterpret_cast<uint4*>(xn + (size_t)row * D + c) =
        make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
  }
}
int decoder_next_token(const float* val, const int32_t* idx, int G, int M, const int64_t* ids_in, int64_t* ids_out, int pos, const float* table,
                       const float* pe, const float* gamma, const float* beta, float eps, void* x, void* xn, int D, cudaStream_t st) {
  M3AE_CHECK(D == 768, M3AE_ERR_UNSUPPORTED, "decoder_next_token: D = %d (768 only)", D);
  M3AE_CHECK(M > 0 && table && pe && x && (val ? (idx && ids_out && G > 0) : ids_in != nullptr), M3AE_ERR_INVALID, "decoder_next_token: null pointer");
  M3AE_CHECK(xn == nullptr || (gamma && beta), M3AE_ERR_INVALID, "decoder_next_token: LayerNorm output without gamma / beta");
  M3AE_CHECK((((uintptr_t)table | (uintptr_t)pe | (uintptr_t)x | (uintptr_t)xn | (uintptr_t)gamma | (uintptr_t)beta) & 15) == 0, M3AE_ERR_UNSUPPORTED,
             "decoder_next_token: 16-byte aligned pointers required");
  M3AE_LAUNCH((decoder_next_token_kernel), (unsigned)M, 32, 0, st, val, idx, G, M, ids_in, ids_out, pos, table, pe, gamma, beta, eps, (bf16*)x, (bf16*)xn);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int argmax_rows(const void* x, int64_t ld, int64_t* out, int64_t M, int N, int dtype, cudaStream_t st) {
  M3AE_CHECK(M > 0 && N > 0, M3AE_ERR_INVALID, "argmax_rows: empty input");
  if (dtype == M3AE_BF16) M3AE_LAUNCH((argmax_rows_kernel<bf16>), (unsigned)M, 256, 0, st, (const bf16*)x, ld, out, N);
  else M3AE_LAUNCH((argmax_rows_kernel<float>), (unsigned)M, 256, 0, st, (const float*)x, ld, out, N);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int decoder_embed(const int64_t* ids, const int32_t* pos, const float* table, const float* pe, void* out, int64_t M, int D,
                  int out_dtype, cudaStream_t st) {
  M3AE_CHECK(M > 0 && D > 0, M3AE_ERR_INVALID, "decoder_embed: empty input");
  if (out_dtype == M3AE_BF16) M3AE_LAUNCH((decoder_embed_kernel<bf16>), (unsigned)M, 256, 0, st, ids, pos, table, pe, (bf16*)out, D);
  else M3AE_LAUNCH((decoder_embed_kernel<float>), (unsigned)M, 256, 0, st, ids, pos, table, pe, (float*)out, D);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
