// LayerNorm forward/backward, one warp per row, 128-bit vector loads, fp32 two-pass statistics.
// HBM-bound: forward moves 2*M*H*sizeof(T) (+8*M) bytes, backward 3*M*H*sizeof(T) (+8*M).
// Reference: nn.LayerNorm at bert_model.py:363,441 (eps=1e-12), m3ae_module.py:122, m3ae_decoder.py:46-49.
#include <mutex>

#include "common.cuh"

namespace m3ae {
void count_launch();
namespace {

template <typename T> struct Vec;  // 16-byte vector of T
template <> struct Vec<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ static void store(float* p, const float (&v)[4]) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
};
template <> struct Vec<bf16> {
  static constexpr int N = 8;
  __device__ static void load(const bf16* p, float (&v)[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    float2 f;
    f = unpack_bf16x2(t.x); v[0] = f.x; v[1] = f.y;
    f = unpack_bf16x2(t.y); v[2] = f.x; v[3] = f.y;
    f = unpack_bf16x2(t.z); v[4] = f.x; v[5] = f.y;
    f = unpack_bf16x2(t.w); v[6] = f.x; v[7] = f.y;
  }
  __device__ static void store(bf16* p, const float (&v)[8]) {
    uint4 t;
    t.x = pack_bf16x2(v[0], v[1]); t.y = pack_bf16x2(v[2], v[3]);
    t.z = pack_bf16x2(v[4], v[5]); t.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = t;
  }
};

constexpr int LN_WARPS = 4;

// 16-byte vector of T straight from memory (kept packed so that a prefetched row costs few registers)
template <typename T>
__device__ __forceinline__ uint4 ld_raw(const T* p) { return *reinterpret_cast<const uint4*>(p); }
template <typename T> __device__ __forceinline__ void unpack_raw(const uint4& t, float (&v)[Vec<T>::N]);
template <> __device__ __forceinline__ void unpack_raw<float>(const uint4& t, float (&v)[4]) {
  v[0] = __uint_as_float(t.x); v[1] = __uint_as_float(t.y); v[2] = __uint_as_float(t.z); v[3] = __uint_as_float(t.w);
}
template <> __device__ __forceinline__ void unpack_raw<bf16>(const uint4& t, float (&v)[8]) {
  float2 f;
  f = unpack_bf16x2(t.x); v[0] = f.x; v[1] = f.y;
  f = unpack_bf16x2(t.y); v[2] = f.x; v[3] = f.y;
  f = unpack_bf16x2(t.z); v[4] = f.x; v[5] = f.y;
  f = unpack_bf16x2(t.w); v[6] = f.x; v[7] = f.y;
}
// VN consecutive fp32 parameters (gamma / beta) as 16-byte loads
template <int VN>
__device__ __forceinline__ void ld_param(const float* p, float (&v)[VN]) {
#pragma unroll
  for (int q = 0; q < VN / 4; ++q) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p) + q);
    v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
  }
}

// (A persistent variant — 8 CTAs per SM walking rows with the next row prefetched — measured SLOWER: 20-21 us against 17.9 us
// for an image-sized launch under ncu, round 2; one short-lived 4-row CTA per 4 rows is kept.)
// H == HV * 32 * Vec<T>::N exactly (768 or 1536 columns on this path).  One warp per row; gamma / beta come in as
// 16-byte loads (the first version read them as scalars: 48 extra load instructions per row, 4x slower than the HBM bound).
template <typename T, int HV>
__global__ void __launch_bounds__(LN_WARPS * 32) ln_fwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, T* __restrict__ y,
                                                              float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                              int64_t M, int H, float eps) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int VN = Vec<T>::N;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
  if (row >= M) return;
  const T* xr = x + row * H;
  float v[HV][VN];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < HV; ++i) {
    Vec<T>::load(xr + (i * 32 + lane) * VN, v[i]);
#pragma unroll
    for (int j = 0; j < VN; ++j) s += v[i][j];
  }
  const float mean = warp_sum(s) / (float)H;
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < HV; ++i)
#pragma unroll
    for (int j = 0; j < VN; ++j) { const float d = v[i][j] - mean; ss += d * d; }
  const float var = warp_sum(ss) / (float)H;
  const float rstd = rsqrtf(var + eps);
  if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
  T* yr = y + row * H;
#pragma unroll
  for (int i = 0; i < HV; ++i) {
    const int c = (i * 32 + lane) * VN;
    float g[VN], b[VN], o[VN];
    ld_param<VN>(gamma + c, g);
    ld_param<VN>(beta + c, b);
#pragma unroll
    for (int j = 0; j < VN; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
    Vec<T>::store(yr + c, o);
  }
}

// Backward: each warp walks rows row0, row0+stride, ... and keeps per-lane column partials of
// dgamma / dbeta / colsum(dx) in registers; one smem reduction + one atomicAdd per column per CTA.
// The next row's x and dy are already in flight (packed, 2*HV registers) while the current row is reduced.
// RING (bf16): rows are staged through a per-warp shared-memory ring of LN_RING rows (x and dy, 2 x H x 2 B each) filled
// by cp.async.bulk with one mbarrier per slot.  The register version keeps ONE 3 KB row in flight per warp and the kernel
// needs 168 registers (12 warps/SM): 36 KB in flight per SM capped it at ~2.7 TB/s; the ring keeps LN_RING rows in flight
// per warp without spending registers on them.
constexpr int LN_BWD_CTAS_PER_SM = 3;
constexpr int LN_RING = 3;
__device__ __forceinline__ void bulk_load_row(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"((uint64_t)gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
template <typename T, int HV, bool RING>
__global__ void __launch_bounds__(LN_WARPS * 32, LN_BWD_CTAS_PER_SM) ln_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                              const float* __restrict__ mean, const float* __restrict__ rstd,
                                                              const float* __restrict__ gamma, T* __restrict__ dx,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                              float* __restrict__ dcolsum, int64_t M, int H,
                                                              T* __restrict__ dx_drop, float drop_p, uint64_t drop_seed,
                                                              uint64_t drop_stream, const uint8_t* __restrict__ drop_bits) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int VN = Vec<T>::N;
  extern __shared__ __align__(128) float red[];  // [LN_WARPS][H] floats | RING: [LN_WARPS][LN_RING][2][H] T | [LN_WARPS][LN_RING] barriers
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t row_bytes = (uint32_t)(H * sizeof(T));
  uint8_t* ring = reinterpret_cast<uint8_t*>(red + LN_WARPS * H) + (size_t)warp * LN_RING * 2 * row_bytes;
  uint64_t* rbar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(red + LN_WARPS * H) + (size_t)LN_WARPS * LN_RING * 2 * row_bytes) + warp * LN_RING;
  if (RING) {
    if (lane == 0) {
      for (int s = 0; s < LN_RING; ++s) mbar_init(&rbar[s], 1);
      fence_barrier_init();
    }
    __syncwarp();
  }
  float g[HV][VN], pg[HV][VN], pb[HV][VN], pc[HV][VN];
#pragma unroll
  for (int i = 0; i < HV; ++i) {
    ld_param<VN>(gamma + (i * 32 + lane) * VN, g[i]);
#pragma unroll
    for (int j = 0; j < VN; ++j) { pg[i][j] = 0.f; pb[i][j] = 0.f; pc[i][j] = 0.f; }
  }
  const float invH = 1.0f / (float)H;
  const int64_t stride = (int64_t)gridDim.x * LN_WARPS;
  int64_t row = (int64_t)blockIdx.x * LN_WARPS + warp;
  uint4 rx[HV], rdy[HV];
  // keep bits cached by the forward GEMM epilogue: byte (i*32+lane)*VN/8 of the row covers this lane's VN columns
  constexpr int NB = VN == 8 ? HV : 1;
  uint8_t kb[NB];
  const bool cached = VN == 8 && drop_bits != nullptr;
  float mu = 0.f, rs = 0.f;
  // RING: lane 0 starts the copies of row `r` into slot `k % LN_RING`
  auto ring_issue = [&](int k, int64_t r) {
    uint8_t* slot = ring + (size_t)(k % LN_RING) * 2 * row_bytes;
    uint64_t* bar = &rbar[k % LN_RING];
    mbar_expect_tx(bar, 2 * row_bytes);
    bulk_load_row(slot, x + r * H, row_bytes, bar);
    bulk_load_row(slot + row_bytes, dy + r * H, row_bytes, bar);
  };
  int kiter = 0;
  if (RING) {
    if (lane == 0)
      for (int k = 0; k < LN_RING; ++k)
        if (row + k * stride < M) ring_issue(k, row + k * stride);
  }
  if (row < M) {
    if (!RING) {
#pragma unroll
      for (int i = 0; i < HV; ++i) {
        rx[i] = ld_raw(x + row * H + (i * 32 + lane) * VN);
        rdy[i] = ld_raw(dy + row * H + (i * 32 + lane) * VN);
      }
    }
    if (cached) {
#pragma unroll
      for (int i = 0; i < NB; ++i) kb[i] = drop_bits[row * (H / 8) + i * 32 + lane];
    }
    mu = mean[row]; rs = rstd[row];
  }
  for (; row < M; row += stride, ++kiter) {
    float xh[HV][VN], gd[HV][VN];
    float c1 = 0.f, c2 = 0.f;
    if (RING) {
      const int slot_i = kiter % LN_RING;
      mbar_wait(&rbar[slot_i], (uint32_t)((kiter / LN_RING) & 1));
      const uint8_t* slot = ring + (size_t)slot_i * 2 * row_bytes;
#pragma unroll
      for (int i = 0; i < HV; ++i) {
        rx[i] = *reinterpret_cast<const uint4*>(slot + (i * 32 + lane) * 16);
        rdy[i] = *reinterpret_cast<const uint4*>(slot + row_bytes + (i * 32 + lane) * 16);
      }
      __syncwarp();  // every lane holds its share in registers: the slot may be refilled
      const int64_t rr = row + (int64_t)LN_RING * stride;
      if (lane == 0 && rr < M) {
        fence_proxy_async();
        ring_issue(kiter + LN_RING, rr);
      }
    }
    const float mu_c = mu, rs_c = rs;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
      float dv[VN];
      unpack_raw<T>(rx[i], xh[i]);
      unpack_raw<T>(rdy[i], dv);
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        xh[i][j] = (xh[i][j] - mu_c) * rs_c;
        pg[i][j] += dv[j] * xh[i][j];
        pb[i][j] += dv[j];
        gd[i][j] = dv[j] * g[i][j];
        c1 += gd[i][j];
        c2 += gd[i][j] * xh[i][j];
      }
    }
    const int64_t nrow = row + stride;
    uint8_t kc[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) kc[i] = cached ? kb[i] : (uint8_t)0;
    if (nrow < M) {  // prefetch: the loads fly during the reductions and the dropout regeneration below
      if (!RING) {
#pragma unroll
        for (int i = 0; i < HV; ++i) {
          rx[i] = ld_raw(x + nrow * H + (i * 32 + lane) * VN);
          rdy[i] = ld_raw(dy + nrow * H + (i * 32 + lane) * VN);
        }
      }
      if (cached) {
#pragma unroll
        for (int i = 0; i < NB; ++i) kb[i] = drop_bits[nrow * (H / 8) + i * 32 + lane];
      }
      mu = mean[nrow]; rs = rstd[nrow];
    }
    c1 = warp_sum(c1) * invH;
    c2 = warp_sum(c2) * invH;
#pragma unroll
    for (int i = 0; i < HV; ++i) {
      const int c = (i * 32 + lane) * VN;
      float o[VN];
#pragma unroll
      for (int j = 0; j < VN; ++j) o[j] = rs_c * (gd[i][j] - c1 - xh[i][j] * c2);
      Vec<T>::store(dx + row * H + c, o);
      if (dx_drop != nullptr) {
        // gradient through the hidden dropout of the producing GEMM epilogue: same Philox stream, element = row*H+col
        const float inv = drop_keep_scale(drop_p);
        if (cached) {
          const uint32_t m8 = kc[NB == 1 ? 0 : i];
#pragma unroll
          for (int j = 0; j < VN; ++j) o[j] = ((m8 >> (j & 7)) & 1u) ? o[j] * inv : 0.f;
        } else {
          if (VN == 8) {  // c % 8 == 0 and H % 8 == 0: one Philox evaluation covers the lane's 8 columns
            const uint32_t m8 = dropout_keep8(drop_seed, drop_stream, (uint64_t)(row * H + c), drop_thr15(drop_p));
#pragma unroll
            for (int j = 0; j < VN; ++j) o[j] = ((m8 >> (j & 7)) & 1u) ? o[j] * inv : 0.f;
          } else {
#pragma unroll
            for (int q = 0; q < VN / 4; ++q) {
              const uint32_t m4 = dropout_keep4(drop_seed, drop_stream, (uint64_t)(row * H + c + 4 * q), drop_p);
#pragma unroll
              for (int j = 0; j < 4; ++j) o[4 * q + j] = ((m4 >> j) & 1u) ? o[4 * q + j] * inv : 0.f;
            }
          }
        }
        Vec<T>::store(dx_drop + row * H + c, o);
      }
#pragma unroll
      for (int j = 0; j < VN; ++j) pc[i][j] += to_f32<T>(from_f32<T>(o[j]));  // column sum of what the GEMMs read
    }
  }
  // CTA reduction of the three column partials, one at a time through the same smem buffer.
  for (int which = 0; which < 3; ++which) {
    float* dst = which == 0 ? dgamma : which == 1 ? dbeta : dcolsum;
    if (dst == nullptr) continue;  // uniform
#pragma unroll
    for (int i = 0; i < HV; ++i)
#pragma unroll
      for (int j = 0; j < VN; ++j)
        red[warp * H + (i * 32 + lane) * VN + j] = which == 0 ? pg[i][j] : which == 1 ? pb[i][j] : pc[i][j];
    __syncthreads();
    for (int c = threadIdx.x; c < H; c += LN_WARPS * 32) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < LN_WARPS; ++w) s += red[w * H + c];
      atomicAdd(dst + c, s);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void colsum_kernel(const T* __restrict__ x, float* __restrict__ out, int64_t M, int N, int64_t ld, int rows_per_cta) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  // block = (32 columns) x (8 row lanes); each CTA covers rows [blockIdx.y*rows_per_cta, ...)
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = min(M, r0 + rows_per_cta);
  float s = 0.f;
  if (c < N)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) s += to_f32<T>(x[r * ld + c]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(out + c, t);
  }
}

// bf16 column sums with 16-byte loads: block = 32 lanes x 8 row lanes, a warp covers 256 consecutive columns of a row
// (512 B), 4 rows in flight per thread.  Bias gradients of the fused QKV / KV / Q / FFN-up projections.
__global__ void __launch_bounds__(256) colsum_bf16_vec_kernel(const bf16* __restrict__ x, float* __restrict__ out, int64_t M, int N,
                                                              int64_t ld, int rows_per_cta) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  __shared__ float red[8][32][9];
  const int c = (blockIdx.x * 32 + threadIdx.x) * 8;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = min(M, r0 + rows_per_cta);
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (c < N) {
    int64_t r = r0 + threadIdx.y;
    for (; r + 24 < r1; r += 32) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(x + (r + 8 * u) * ld + c));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = unpack_bf16x2(w[j]);
          s[2 * j] += f.x;
          s[2 * j + 1] += f.y;
        }
      }
    }
    for (; r < r1; r += 8) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + r * ld + c));
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        s[2 * j] += f.x;
        s[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = s[j];
  __syncthreads();
  // 256 threads <-> 256 columns of the block
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const int col = blockIdx.x * 256 + tid;
  if (col < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tid >> 3][tid & 7];
    atomicAdd(out + col, t);
  }
}

template <typename TS, typename TD>
__global__ void cast_kernel(const TS* __restrict__ src, TD* __restrict__ dst, int64_t n) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = from_f32<TD>(to_f32<TS>(src[i]));
}
// 8 fp32 -> 8 bf16 per thread-iteration (n % 8 == 0, aligned)
__global__ void cast_f32_bf16_vec_kernel(const float4* __restrict__ src, uint4* __restrict__ dst, int64_t n8) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = src[2 * i], b = src[2 * i + 1];
    uint4 o;
    o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
    o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
    dst[i] = o;
  }
}

// 8 bf16 -> 8 fp32 per thread-iteration (the gradient exchange's bf16 wire format back into the fp32 gradient buffers)
__global__ void cast_bf16_f32_vec_kernel(const uint4* __restrict__ src, float4* __restrict__ dst, int64_t n8) {
  pdl_sync();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const uint4 v = src[i];
    const float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
    dst[2 * i] = make_float4(a.x, a.y, b.x, b.y);
    dst[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
  }
}

// One launch for a whole list of casts (a layer's 9 weights + 5 biases -> its operand cache): blockIdx.y = item.
struct CastGroup {
  const void* src[M3AE_MAX_CAST_GROUP];
  void* dst[M3AE_MAX_CAST_GROUP];
  int64_t n[M3AE_MAX_CAST_GROUP];
  int8_t sd[M3AE_MAX_CAST_GROUP], dd[M3AE_MAX_CAST_GROUP];
};
__global__ void cast_group_kernel(const __grid_constant__ CastGroup G) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int it = blockIdx.y;
  const int64_t n = G.n[it];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int sd = G.sd[it], dd = G.dd[it];
  if (sd == M3AE_F32 && dd == M3AE_BF16) {
    const float* src = (const float*)G.src[it];
    bf16* dst = (bf16*)G.dst[it];
    if (n % 8 == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
      const float4* s4 = (const float4*)src;
      uint4* d4 = (uint4*)dst;
      for (int64_t i = t0; i < n / 8; i += stride) {
        const float4 a = __ldg(s4 + 2 * i), b = __ldg(s4 + 2 * i + 1);
        uint4 o;
        o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
        o.z = pack_bf16x2(b.x, b.y); o.w = pack_bf16x2(b.z, b.w);
        d4[i] = o;
      }
    } else {
      for (int64_t i = t0; i < n; i += stride) dst[i] = __float2bfloat16_rn(src[i]);
    }
  } else if (sd == M3AE_F32 && dd == M3AE_F32) {
    const float* src = (const float*)G.src[it];
    float* dst = (float*)G.dst[it];
    for (int64_t i = t0; i < n; i += stride) dst[i] = src[i];
  } else if (sd == M3AE_BF16 && dd == M3AE_F32) {
    const bf16* src = (const bf16*)G.src[it];
    float* dst = (float*)G.dst[it];
    for (int64_t i = t0; i < n; i += stride) dst[i] = __bfloat162float(src[i]);
  } else {
    const bf16* src = (const bf16*)G.src[it];
    bf16* dst = (bf16*)G.dst[it];
    for (int64_t i = t0; i < n; i += stride) dst[i] = src[i];
  }
}

template <typename TI, typename TO>
__global__ void add_rowvec_kernel(const TI* __restrict__ a, const float* __restrict__ b, TO* __restrict__ out, int64_t M, int N) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t total = M * N;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
    out[i] = from_f32<TO>(to_f32<TI>(a[i]) + __ldg(b + (i % N)));
}

template <typename T>
__global__ void dropout_apply_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t M, int N, int64_t ld, float p,
                                     uint64_t seed, uint64_t stream_id) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t total = M * N;
  const float inv = drop_keep_scale(p);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / N, c = i % N;
    const int64_t e = r * ld + c;
    y[e] = dropout_keep(seed, stream_id, (uint64_t)e, p) ? from_f32<T>(to_f32<T>(x[e]) * inv) : from_f32<T>(0.f);
  }
}

}  // namespace

int layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd, int64_t M, int H,
                  float eps, int dtype, cudaStream_t st) {
  M3AE_CHECK(M > 0, M3AE_ERR_INVALID, "layernorm: empty input");
  const unsigned grid = (unsigned)((M + LN_WARPS - 1) / LN_WARPS);
#define LN_FWD(T, HV) M3AE_LAUNCH((ln_fwd_kernel<T, HV>), grid, LN_WARPS * 32, 0, st, (const T*)x, gamma, beta, (T*)y, mean, rstd, M, H, eps)
  if (dtype == M3AE_BF16 && H == 768) LN_FWD(bf16, 3);
  else if (dtype == M3AE_BF16 && H == 1536) LN_FWD(bf16, 6);
  else if (dtype == M3AE_F32 && H == 768) LN_FWD(float, 6);
  else if (dtype == M3AE_F32 && H == 1536) LN_FWD(float, 12);
  else M3AE_CHECK(false, M3AE_ERR_UNSUPPORTED, "layernorm: unsupported (dtype=%d, H=%d); H must be 768 or 1536", dtype, H);
#undef LN_FWD
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

bool ln_bwd_force_registers = false;  // test hook (m3ae_set_layernorm_bwd_staging): bypass the shared-memory ring

int layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma, void* dx,
                  float* dgamma, float* dbeta, float* dcolsum, int64_t M, int H, int dtype, void* dx_drop, float drop_p,
                  uint64_t drop_seed, uint64_t drop_stream, const uint32_t* drop_bits, cudaStream_t st) {
  M3AE_CHECK(drop_p >= 0.f && drop_p < 1.f, M3AE_ERR_INVALID, "layernorm_bwd: dropout %f not in [0,1)", drop_p);
  if (drop_p == 0.f) dx_drop = nullptr;
  M3AE_CHECK(drop_bits == nullptr || (dtype == M3AE_BF16 && H % 32 == 0), M3AE_ERR_UNSUPPORTED,
             "layernorm_bwd: cached dropout bits need bf16 activations and H %% 32 == 0");
  M3AE_CHECK(M > 0, M3AE_ERR_INVALID, "layernorm: empty input");
  int64_t want = (M + LN_WARPS - 1) / LN_WARPS;
  const int64_t cap = (int64_t)sm_count() * LN_BWD_CTAS_PER_SM;  // persistent: exactly the resident CTAs
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  const size_t smem = (size_t)LN_WARPS * H * sizeof(float);
#define LN_BWD(T, HV, RING, SMEM)                                                                                              \
  M3AE_LAUNCH((ln_bwd_kernel<T, HV, RING>), grid, LN_WARPS * 32, SMEM, st, (const T*)dy, (const T*)x, mean, rstd, gamma, (T*)dx, dgamma, \
                                                                dbeta, dcolsum, M, H, (T*)dx_drop, drop_p, drop_seed, drop_stream, (const uint8_t*)drop_bits)
  // the fusion stack's case (bf16, H = 768): rows staged through the bulk-copy ring (needs 16-byte aligned rows)
  const bool ring_ok = dtype == M3AE_BF16 && H == 768 && !ln_bwd_force_registers && ((((uintptr_t)x) | ((uintptr_t)dy)) & 15) == 0;
  if (ring_ok) {
    const size_t smem_ring = smem + (size_t)LN_WARPS * LN_RING * 2 * H * sizeof(bf16) + (size_t)LN_WARPS * LN_RING * sizeof(uint64_t);
    static std::once_flag once[64];
    static cudaError_t attr_errs[64];
    const int slot = current_device_slot();
    std::call_once(once[slot], [smem_ring, slot] {
      attr_errs[slot] = cudaFuncSetAttribute(ln_bwd_kernel<bf16, 3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ring);
    });
    const cudaError_t attr_err = attr_errs[slot];
    M3AE_CHECK(attr_err == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(ln_bwd ring smem) failed: %s", cudaGetErrorString(attr_err));
    LN_BWD(bf16, 3, true, smem_ring);
  } else if (dtype == M3AE_BF16 && H == 768) LN_BWD(bf16, 3, false, smem);
  else if (dtype == M3AE_BF16 && H == 1536) LN_BWD(bf16, 6, false, smem);
  else if (dtype == M3AE_F32 && H == 768) LN_BWD(float, 6, false, smem);
  else if (dtype == M3AE_F32 && H == 1536) LN_BWD(float, 12, false, smem);
  else M3AE_CHECK(false, M3AE_ERR_UNSUPPORTED, "layernorm: unsupported (dtype=%d, H=%d); H must be 768 or 1536", dtype, H);
#undef LN_BWD
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int colsum(const void* x, float* out, int64_t M, int N, int64_t ld, int dtype, cudaStream_t st) {
  M3AE_CHECK(M > 0 && N > 0, M3AE_ERR_INVALID, "colsum: empty input");
  if (dtype == M3AE_BF16 && N % 8 == 0 && ld % 8 == 0 && ((uintptr_t)x & 15) == 0) {
    const int cb = (N + 255) / 256;
    int rb = (int)((int64_t)sm_count() * 4 / cb);
    if (rb < 1) rb = 1;
    if ((int64_t)rb * 64 > M) rb = (int)((M + 63) / 64);
    const int rows_per = (int)((M + rb - 1) / rb);
    M3AE_LAUNCH((colsum_bf16_vec_kernel), dim3(cb, rb), dim3(32, 8), 0, st, (const bf16*)x, out, M, N, ld, rows_per);
    count_launch();
    M3AE_CUDA(cudaGetLastError());
    return M3AE_OK;
  }
  const int col_blocks = (N + 31) / 32;
  int row_blocks = (int)((int64_t)sm_count() * 8 / col_blocks);
  if (row_blocks < 1) row_blocks = 1;
  if ((int64_t)row_blocks * 64 > M) row_blocks = (int)((M + 63) / 64);
  const int rows_per = (int)((M + row_blocks - 1) / row_blocks);
  dim3 grid(col_blocks, row_blocks), block(32, 8);
  if (dtype == M3AE_BF16) M3AE_LAUNCH((colsum_kernel<bf16>), grid, block, 0, st, (const bf16*)x, out, M, N, ld, rows_per);
  else M3AE_LAUNCH((colsum_kernel<float>), grid, block, 0, st, (const float*)x, out, M, N, ld, rows_per);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int cast(const void* src, int sd, void* dst, int dd, int64_t n, cudaStream_t st) {
  if (n == 0) return M3AE_OK;
  const int threads = 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (sd == M3AE_F32 && dd == M3AE_BF16 && n % 8 == 0 && ((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0) {
    const int64_t n8 = n / 8;
    int64_t b = (n8 + threads - 1) / threads;
    M3AE_LAUNCH((cast_f32_bf16_vec_kernel), (unsigned)(b < cap ? b : cap), threads, 0, st, (const float4*)src, (uint4*)dst, n8);
  } else if (sd == M3AE_BF16 && dd == M3AE_F32 && n % 8 == 0 && ((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0) {
    const int64_t n8 = n / 8;
    int64_t b = (n8 + threads - 1) / threads;
    M3AE_LAUNCH((cast_bf16_f32_vec_kernel), (unsigned)(b < cap ? b : cap), threads, 0, st, (const uint4*)src, (float4*)dst, n8);
  } else {
    int64_t b = (n + threads - 1) / threads;
    const unsigned grid = (unsigned)(b < cap ? b : cap);
    if (sd == M3AE_F32 && dd == M3AE_BF16) M3AE_LAUNCH((cast_kernel<float, bf16>), grid, threads, 0, st, (const float*)src, (bf16*)dst, n);
    else if (sd == M3AE_BF16 && dd == M3AE_F32) M3AE_LAUNCH((cast_kernel<bf16, float>), grid, threads, 0, st, (const bf16*)src, (float*)dst, n);
    else if (sd == M3AE_F32 && dd == M3AE_F32) M3AE_LAUNCH((cast_kernel<float, float>), grid, threads, 0, st, (const float*)src, (float*)dst, n);
    else M3AE_LAUNCH((cast_kernel<bf16, bf16>), grid, threads, 0, st, (const bf16*)src, (bf16*)dst, n);
  }
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int cast_group(const m3ae_cast_item_t* items, int n, cudaStream_t st) {
  CastGroup G;
  memset(&G, 0, sizeof(G));
  int64_t nmax = 0;
  for (int i = 0; i < n; ++i) {
    G.src[i] = items[i].src; G.dst[i] = items[i].dst; G.n[i] = items[i].n;
    G.sd[i] = (int8_t)items[i].src_dtype; G.dd[i] = (int8_t)items[i].dst_dtype;
    if (items[i].n > nmax) nmax = items[i].n;
  }
  if (nmax == 0) return M3AE_OK;
  const int threads = 256;
  int64_t b = (nmax / 8 + threads - 1) / threads;
  const int64_t cap = (int64_t)sm_count() * 2;
  if (b < 1) b = 1;
  dim3 grid((unsigned)(b < cap ? b : cap), (unsigned)n);
  M3AE_LAUNCH((cast_group_kernel), grid, threads, 0, st, G);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int add_rowvec(const void* a, const float* b, void* out, int64_t M, int N, int in_dtype, int out_dtype, cudaStream_t st) {
  M3AE_CHECK(M > 0 && N > 0, M3AE_ERR_INVALID, "add_rowvec: empty input");
  const int threads = 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  int64_t blocks = (M * N + threads - 1) / threads;
  const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
  if (in_dtype == M3AE_F32 && out_dtype == M3AE_F32) M3AE_LAUNCH((add_rowvec_kernel<float, float>), grid, threads, 0, st, (const float*)a, b, (float*)out, M, N);
  else if (in_dtype == M3AE_F32 && out_dtype == M3AE_BF16) M3AE_LAUNCH((add_rowvec_kernel<float, bf16>), grid, threads, 0, st, (const float*)a, b, (bf16*)out, M, N);
  else if (in_dtype == M3AE_BF16 && out_dtype == M3AE_BF16) M3AE_LAUNCH((add_rowvec_kernel<bf16, bf16>), grid, threads, 0, st, (const bf16*)a, b, (bf16*)out, M, N);
  else M3AE_LAUNCH((add_rowvec_kernel<bf16, float>), grid, threads, 0, st, (const bf16*)a, b, (float*)out, M, N);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int dropout_apply(const void* x, void* y, int64_t M, int N, int64_t ld, float p, uint64_t seed, uint64_t stream_id, int dtype,
                  cudaStream_t st) {
  M3AE_CHECK(M > 0 && N > 0, M3AE_ERR_INVALID, "dropout_apply: empty input");
  M3AE_CHECK(p >= 0.f && p < 1.f, M3AE_ERR_INVALID, "dropout probability %f not in [0,1)", p);
  const int threads = 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  int64_t blocks = (M * N + threads - 1) / threads;
  const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
  if (dtype == M3AE_BF16) M3AE_LAUNCH((dropout_apply_kernel<bf16>), grid, threads, 0, st, (const bf16*)x, (bf16*)y, M, N, ld, p, seed, stream_id);
  else M3AE_LAUNCH((dropout_apply_kernel<float>), grid, threads, 0, st, (const float*)x, (float*)y, M, N, ld, p, seed, stream_id);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
