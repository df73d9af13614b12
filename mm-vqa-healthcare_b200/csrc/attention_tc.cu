// tcgen05 / TMEM flash attention for the fusion stack's bf16 path (12 heads x 64, Lq, Lk <= 384: 325 image tokens at
// 288 px, 32 text tokens).  Reference arithmetic: bert_model.py:301-344 (BertSelfAttention core, self and cross mode).
//
// One CTA = one (batch, head).  K and V of the head stay resident in shared memory (TMA, SWIZZLE_128B: a 64-wide bf16
// head row is exactly one 128-byte swizzle row, so the same tile serves as a K-major or an MN-major UMMA operand).
// Warps 0..7 are the element-wise warps (thread = one row of the 128-row tile, TMEM lane quadrant = warp % 4, the two
// warps of a quadrant split the columns), warp 8 allocates TMEM and one of its lanes issues every TMA and tcgen05.mma.
//
//   forward : per 128-query tile   S = Q K^T           (tcgen05, fp32 S for ALL keys of the row in TMEM: <= 384 columns)
//                                  P = softmax(S)      (two passes over TMEM: max, then exp2/sum/dropout; no online
//                                                       rescaling because the whole row is resident), P -> smem (bf16)
//                                  O = P V             (tcgen05, A = P from smem, B = V MN-major), O/l -> global
//   backward: per (96-key tile, 128-query tile)   S = Q K^T, dP = dO V^T                     (tcgen05 -> TMEM)
//                                  P = exp2(S - lse), dS = P (dP*keep - delta) scale          (thread = query row)
//                                  dV += P'^T dO, dK += dS^T Q  (A = P'/dS tile read MN-major), dQ += dS K
//             dQ of all <= 3 query tiles, dK and dV of the key tile, S and dP fill the 512 TMEM columns exactly
//             (192 + 64 + 64 + 96 + 96), so one kernel produces dQ, dK and dV with every P evaluated once.
//
// Algorithmic work per (b,h): forward 4*Lq*Lk*64 flop, backward 10*Lq*Lk*64 flop; HBM traffic = Q,K,V,O (+dO,dQ,dK,dV)
// once: the scores [B,12,Lq,Lk] (324 MB per image layer at B=64 in the reference) never leave the SM.
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace m3ae {
void count_launch();
void launch_delta_bf16_64(const m3ae_attn_params_t* p, cudaStream_t st);  // attention.cu

namespace {

constexpr int D = 64;            // head dim
constexpr int TQ = 128;          // query rows per tile (UMMA M)
constexpr int KT = 96;           // backward key tile
constexpr int BLK = TQ * 128;    // one [128 rows][64 bf16] swizzled block (16 KB)
constexpr int NT = 288;          // 8 element-wise warps + 1 control warp
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct TcAttn {
  CUtensorMap map_q, map_k, map_v, map_do;  // 2-D [B*L, H*64] bf16 views; box = 64 columns x box rows, SWIZZLE_128B
  m3ae_attn_params_t p;
  int lk_pad;          // Lk rounded up to 16 (UMMA N / K granularity)
  int nbox_k, box_k;   // K and V are staged as nbox_k TMA boxes of box_k rows
  int box_q;           // rows per Q / dO tile load (<= 128)
  int n_qt;            // number of 128-row query tiles
  int tmem_cols;       // power of two >= 32
  int off_v, off_q, off_do, off_p, off_ds, off_mask, off_red, off_bar;  // smem offsets (K at 0)
  float scale_l2;      // scale * log2(e)
};

__device__ __forceinline__ uint64_t sdesc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, M = 128.
__device__ __forceinline__ uint32_t idesc(int n, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(TQ >> 4) << 24);
}

__device__ __forceinline__ void tmem_alloc_n(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_n(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void pair_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

// byte offset of the 16-byte chunk holding keys [c, c+8) (c % 8 == 0) of row r inside a [rows][keys] staging area made of
// 64-key blocks of BLK bytes, SWIZZLE_128B (chunk index ^= row % 8) — the layout TMA would have produced.
__device__ __forceinline__ uint32_t stage_off(int r, int c) {
  return (uint32_t)((c >> 6) * BLK + r * 128 + ((((c >> 3) & 7) ^ (r & 7)) << 4));
}

// 16 probabilities -> bf16 -> two 16-byte chunks of row r at key column c (c % 16 == 0)
__device__ __forceinline__ void store16(uint8_t* area, int r, int c, const float (&v)[16]) {
  uint4 w0, w1;
  w0.x = pack_bf16x2(v[0], v[1]); w0.y = pack_bf16x2(v[2], v[3]); w0.z = pack_bf16x2(v[4], v[5]); w0.w = pack_bf16x2(v[6], v[7]);
  w1.x = pack_bf16x2(v[8], v[9]); w1.y = pack_bf16x2(v[10], v[11]); w1.z = pack_bf16x2(v[12], v[13]); w1.w = pack_bf16x2(v[14], v[15]);
  *reinterpret_cast<uint4*>(area + stage_off(r, c)) = w0;
  *reinterpret_cast<uint4*>(area + stage_off(r, c + 8)) = w1;
}

// 32 fp32 accumulator values of one row -> bf16 -> 64 contiguous bytes in global memory
__device__ __forceinline__ void store_row32(bf16* dst, const uint32_t (&r)[32], float mul) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 w;
    w.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]) * mul, __uint_as_float(r[8 * j + 1]) * mul);
    w.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]) * mul, __uint_as_float(r[8 * j + 3]) * mul);
    w.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]) * mul, __uint_as_float(r[8 * j + 5]) * mul);
    w.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]) * mul, __uint_as_float(r[8 * j + 7]) * mul);
    reinterpret_cast<uint4*>(dst)[j] = w;
  }
}

// ------------------------------------------------------------------------------------------------ forward
// TMEM columns: O at [0, 64), S at [64, 64 + lk_pad).
// Barriers: 0 = K,V landed; 1,2 = Q tile (double-buffered) landed; 3 = S ready (MMA -> element-wise warps);
//           4 = P staged in smem (8 warps -> MMA); 5 = O ready (MMA -> warps); 6 = O read out of TMEM (8 warps -> MMA).
template <bool DROP>
__global__ void __launch_bounds__(NT, 1) attn_fwd_tc_kernel(const __grid_constant__ TcAttn A) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const m3ae_attn_params_t& P = A.p;
  uint8_t* sK = smem;
  uint8_t* sV = smem + A.off_v;
  uint8_t* sQ = smem + A.off_q;
  uint8_t* sP = smem + A.off_p;
  float* sMask = (float*)(smem + A.off_mask);
  float* sRed = (float*)(smem + A.off_red);  // [0..1][128] row max per column half, [2..5][128] row sum (double-buffered by tile parity)
  uint64_t* bars = (uint64_t*)(smem + A.off_bar);
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + 8);
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Lq = P.Lq, Lk = P.Lk, lk_pad = A.lk_pad;

  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&A.map_q);
      tma_prefetch_desc(&A.map_k);
      tma_prefetch_desc(&A.map_v);
      mbar_init(&bars[0], 1);
      mbar_init(&bars[1], 1);
      mbar_init(&bars[2], 1);
      mbar_init(&bars[3], 1);
      mbar_init(&bars[4], 8);
      mbar_init(&bars[5], 1);
      mbar_init(&bars[6], 8);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_n(tmem_ptr_smem, (uint32_t)A.tmem_cols);
  } else {
    const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
    for (int i = threadIdx.x; i < lk_pad; i += 256) sMask[i] = i < Lk ? (mrow ? mrow[i] * LOG2E : 0.f) : -INFINITY;
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tO = tmem_base, tS = tmem_base + 64;

  if (warp == 8) {
    if (lane == 0) {
      // ===================== control thread: TMA + MMA issue =====================
      const int col0 = h * D;
      mbar_expect_tx(&bars[0], (uint32_t)(2 * A.nbox_k * A.box_k * 128));
      for (int i = 0; i < A.nbox_k; ++i) {
        tma_load_2d(sK + i * A.box_k * 128, &A.map_k, &bars[0], col0, b * Lk + i * A.box_k);
        tma_load_2d(sV + i * A.box_k * 128, &A.map_v, &bars[0], col0, b * Lk + i * A.box_k);
      }
      mbar_expect_tx(&bars[1], (uint32_t)(A.box_q * 128));
      tma_load_2d(sQ, &A.map_q, &bars[1], col0, b * Lq);
      mbar_wait(&bars[0], 0);
      // S is issued in one or two N chunks (UMMA N <= 256)
      const int n1 = lk_pad <= 256 ? lk_pad : ((lk_pad / 2 + 15) & ~15);
      const int n2 = lk_pad - n1;
      const uint32_t id_s1 = idesc(n1, 0, 0), id_s2 = idesc(n2 > 0 ? n2 : 16, 0, 0), id_o = idesc(D, 0, 1);
      for (int t = 0; t < A.n_qt; ++t) {
        const int buf = t & 1;
        if (t + 1 < A.n_qt) {  // Q buffer buf^1 was last read by S(t-1), complete since barrier 4 of tile t-1 was observed
          mbar_expect_tx(&bars[1 + (buf ^ 1)], (uint32_t)(A.box_q * 128));
          tma_load_2d(sQ + (buf ^ 1) * BLK, &A.map_q, &bars[1 + (buf ^ 1)], col0, b * Lq + (t + 1) * TQ);
        }
        mbar_wait(&bars[1 + buf], (uint32_t)((t >> 1) & 1));
        tcgen05_fence_after();
        const uint32_t aQ = smem_u32(sQ + buf * BLK), aK = smem_u32(sK);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16(tS, sdesc(aQ + k * 32, 16, 1024), sdesc(aK + k * 32, 16, 1024), id_s1, k > 0);
        if (n2 > 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tS + n1, sdesc(aQ + k * 32, 16, 1024), sdesc(aK + n1 * 128 + k * 32, 16, 1024), id_s2, k > 0);
        }
        umma_commit(&bars[3]);
        mbar_wait(&bars[4], (uint32_t)(t & 1));
        tcgen05_fence_after();
        if (t > 0) {
          mbar_wait(&bars[6], (uint32_t)((t - 1) & 1));
          tcgen05_fence_after();
        }
        const uint32_t aP = smem_u32(sP), aV = smem_u32(sV);
        const int ksteps = lk_pad / 16;
        for (int ks = 0; ks < ksteps; ++ks)  // O[q][d] += P[q][key] V[key][d]: A K-major (64-key blocks), B MN-major
          umma_f16(tO, sdesc(aP + (ks >> 2) * BLK + (ks & 3) * 32, 16, 1024), sdesc(aV + ks * 2048, 8192, 1024), id_o, ks > 0);
        umma_commit(&bars[5]);
      }
    }
  } else {
    // ===================== element-wise warps =====================
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int n16 = lk_pad >> 4;
    const int u0 = half ? (n16 + 1) / 2 : 0, u1 = half ? n16 : (n16 + 1) / 2;
    const float keep_scale = drop_keep_scale(P.drop_p);
    const uint32_t thr = drop_thr16(P.drop_p);
    const int64_t mask_ld = 8 * (int64_t)((Lk + 63) / 64);
    for (int t = 0; t < A.n_qt; ++t) {
      const int qrow = t * TQ + row;
      const bool active = t * TQ + quad * 32 < Lq;  // warp-uniform
      mbar_wait(&bars[3], (uint32_t)(t & 1));
      tcgen05_fence_after();
      float mx = -INFINITY;
      if (active) {
#pragma unroll 1
        for (int u = u0; u < u1; ++u) {
          uint32_t r[16];
          tmem_ld16(tS + lane_addr + u * 16, r);
          tmem_ld_wait();
          const float4* mk = reinterpret_cast<const float4*>(sMask + u * 16);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 m4 = mk[j];
            mx = fmaxf(mx, fmaf(__uint_as_float(r[4 * j + 0]), A.scale_l2, m4.x));
            mx = fmaxf(mx, fmaf(__uint_as_float(r[4 * j + 1]), A.scale_l2, m4.y));
            mx = fmaxf(mx, fmaf(__uint_as_float(r[4 * j + 2]), A.scale_l2, m4.z));
            mx = fmaxf(mx, fmaf(__uint_as_float(r[4 * j + 3]), A.scale_l2, m4.w));
          }
        }
      }
      sRed[half * 128 + row] = mx;
      pair_sync(1 + quad);
      mx = fmaxf(mx, sRed[(half ^ 1) * 128 + row]);
      if (mx == -INFINITY) mx = 0.f;
      float sum = 0.f;
      if (active) {
        const uint64_t row_lin = ((uint64_t)b * P.H + h) * Lq + qrow;
#pragma unroll 1
        for (int u = u0; u < u1; ++u) {
          uint32_t r[16];
          tmem_ld16(tS + lane_addr + u * 16, r);
          tmem_ld_wait();
          const float4* mk = reinterpret_cast<const float4*>(sMask + u * 16);
          float pv[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 m4 = mk[j];
            pv[4 * j + 0] = ex2_approx(fmaf(__uint_as_float(r[4 * j + 0]), A.scale_l2, m4.x - mx));
            pv[4 * j + 1] = ex2_approx(fmaf(__uint_as_float(r[4 * j + 1]), A.scale_l2, m4.y - mx));
            pv[4 * j + 2] = ex2_approx(fmaf(__uint_as_float(r[4 * j + 2]), A.scale_l2, m4.z - mx));
            pv[4 * j + 3] = ex2_approx(fmaf(__uint_as_float(r[4 * j + 3]), A.scale_l2, m4.w - mx));
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) sum += pv[j];
          if (DROP) {
            const uint32_t bits = attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, 2 * u, thr) |
                                  (attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, 2 * u + 1, thr) << 8);
#pragma unroll
            for (int j = 0; j < 16; ++j) pv[j] = ((bits >> j) & 1u) ? pv[j] * keep_scale : 0.f;
            if (P.drop_mask != nullptr && qrow < Lq && 2 * u < mask_ld)
              *reinterpret_cast<uint16_t*>(P.drop_mask + (int64_t)row_lin * mask_ld + 2 * u) = (uint16_t)bits;
          }
          store16(sP, row, u * 16, pv);
        }
      }
      sRed[(2 + 2 * (t & 1) + half) * 128 + row] = sum;
      fence_proxy_async();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[4]);
      mbar_wait(&bars[5], (uint32_t)(t & 1));
      tcgen05_fence_after();
      const float l = sum + sRed[(2 + 2 * (t & 1) + (half ^ 1)) * 128 + row];
      uint32_t o[32];
      if (active) {
        tmem_ld_32x32b_x32(tO + lane_addr + half * 32, o);
        tmem_ld_wait();
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[6]);
      if (active && qrow < Lq) {
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        bf16* dst = (bf16*)P.out + ((int64_t)b * Lq + qrow) * P.ldo + h * D + half * 32;
        store_row32(dst, o, inv);
        if (half == 0) P.lse[((int64_t)b * P.H + h) * Lq + qrow] = (mx + log2f(l)) * LN2;
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 8) {
    tcgen05_fence_after();
    tmem_dealloc_n(tmem_base, (uint32_t)A.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------ backward
// TMEM columns: dQ tiles at [0, 64*n_qt), dK at 64*n_qt, dV at +64, S at +128, dP at +128+kw  (kw = key tile width <= 96).
// Barriers: 0 = K,V landed; 1,2 = Q,dO tile (double-buffered) landed; 3 = S,dP ready; 4 = P',dS staged (8 warps);
//           5 = dV,dK,dQ MMAs of an iteration complete; 6 = dK,dV read out of TMEM (8 warps).
// Loop: key tiles (outer) x query tiles (inner); iteration it = kt * n_qt + qt.
template <bool DROP>
__global__ void __launch_bounds__(NT, 1) attn_bwd_tc_kernel(const __grid_constant__ TcAttn A) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const m3ae_attn_params_t& P = A.p;
  uint8_t* sK = smem;
  uint8_t* sV = smem + A.off_v;
  uint8_t* sQ = smem + A.off_q;
  uint8_t* sdO = smem + A.off_do;
  uint8_t* sP = smem + A.off_p;
  uint8_t* sdS = smem + A.off_ds;
  float* sMask = (float*)(smem + A.off_mask);
  uint64_t* bars = (uint64_t*)(smem + A.off_bar);
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + 8);
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Lq = P.Lq, Lk = P.Lk, lk_pad = A.lk_pad, n_qt = A.n_qt;
  const int kw = lk_pad < KT ? lk_pad : KT;
  const int n_kt = (lk_pad + KT - 1) / KT;
  const int total = n_kt * n_qt;

  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&A.map_q);
      tma_prefetch_desc(&A.map_k);
      tma_prefetch_desc(&A.map_v);
      tma_prefetch_desc(&A.map_do);
      mbar_init(&bars[0], 1);
      mbar_init(&bars[1], 1);
      mbar_init(&bars[2], 1);
      mbar_init(&bars[3], 1);
      mbar_init(&bars[4], 8);
      mbar_init(&bars[5], 1);
      mbar_init(&bars[6], 8);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_n(tmem_ptr_smem, (uint32_t)A.tmem_cols);
  } else {
    const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
    for (int i = threadIdx.x; i < n_kt * KT; i += 256) sMask[i] = i < Lk ? (mrow ? mrow[i] * LOG2E : 0.f) : -INFINITY;
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tDQ = tmem_base, tDK = tmem_base + 64 * n_qt, tDV = tDK + 64, tS = tDV + 64, tDP = tS + kw;

  if (warp == 8) {
    if (lane == 0) {
      // ===================== control thread: TMA + MMA issue =====================
      const int col0 = h * D;
      mbar_expect_tx(&bars[0], (uint32_t)(2 * A.nbox_k * A.box_k * 128));
      for (int i = 0; i < A.nbox_k; ++i) {
        tma_load_2d(sK + i * A.box_k * 128, &A.map_k, &bars[0], col0, b * Lk + i * A.box_k);
        tma_load_2d(sV + i * A.box_k * 128, &A.map_v, &bars[0], col0, b * Lk + i * A.box_k);
      }
      mbar_expect_tx(&bars[1], (uint32_t)(2 * A.box_q * 128));
      tma_load_2d(sQ, &A.map_q, &bars[1], col0, b * Lq);
      tma_load_2d(sdO, &A.map_do, &bars[1], col0, b * Lq);
      mbar_wait(&bars[0], 0);
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aP = smem_u32(sP), aDS = smem_u32(sdS);
      const uint32_t id_acc = idesc(D, 1, 1);  // dV, dK: A (P'/dS tile) and B (dO/Q tile) both MN-major
      const uint32_t id_dq = idesc(D, 0, 1);   // dQ: A = dS K-major, B = K tile MN-major
      int it = 0;
      for (int kt = 0; kt < n_kt; ++kt) {
        const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;  // keys in this tile (multiple of 16)
        const uint32_t id_s = idesc(w, 0, 0);
        for (int qt = 0; qt < n_qt; ++qt, ++it) {
          const int buf = it & 1;
          mbar_wait(&bars[1 + buf], (uint32_t)((it >> 1) & 1));
          tcgen05_fence_after();
          const uint32_t aQ = smem_u32(sQ + buf * BLK), aDO = smem_u32(sdO + buf * BLK);
          const uint32_t kofs = (uint32_t)(kt * KT * 128);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16(tS, sdesc(aQ + k * 32, 16, 1024), sdesc(aK + kofs + k * 32, 16, 1024), id_s, k > 0);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16(tDP, sdesc(aDO + k * 32, 16, 1024), sdesc(aV + kofs + k * 32, 16, 1024), id_s, k > 0);
          umma_commit(&bars[3]);
          if (it + 1 < total) {
            // the other Q/dO buffer was read by the accumulate MMAs of iteration it-1: wait for them before refilling
            if (it > 0) mbar_wait(&bars[5], (uint32_t)((it - 1) & 1));
            const int nqt = (qt + 1 < n_qt) ? qt + 1 : 0;
            mbar_expect_tx(&bars[1 + (buf ^ 1)], (uint32_t)(2 * A.box_q * 128));
            tma_load_2d(sQ + (buf ^ 1) * BLK, &A.map_q, &bars[1 + (buf ^ 1)], col0, b * Lq + nqt * TQ);
            tma_load_2d(sdO + (buf ^ 1) * BLK, &A.map_do, &bars[1 + (buf ^ 1)], col0, b * Lq + nqt * TQ);
          }
          mbar_wait(&bars[4], (uint32_t)(it & 1));
          tcgen05_fence_after();
          if (qt == 0 && kt > 0) {
            mbar_wait(&bars[6], (uint32_t)((kt - 1) & 1));
            tcgen05_fence_after();
          }
          const int qvalid = (Lq - qt * TQ) < TQ ? (Lq - qt * TQ) : TQ;
          const int qsteps = (qvalid + 15) >> 4;
          for (int ks = 0; ks < qsteps; ++ks) {  // reduction over the tile's query rows (16 per step = 2048 B)
            umma_f16(tDV, sdesc(aP + ks * 2048, BLK, 1024), sdesc(aDO + ks * 2048, 8192, 1024), id_acc, (qt > 0 || ks > 0));
            umma_f16(tDK, sdesc(aDS + ks * 2048, BLK, 1024), sdesc(aQ + ks * 2048, 8192, 1024), id_acc, (qt > 0 || ks > 0));
          }
          const int ksteps = w >> 4;
          for (int ks = 0; ks < ksteps; ++ks)    // reduction over the tile's keys
            umma_f16(tDQ + 64 * qt, sdesc(aDS + (ks >> 2) * BLK + (ks & 3) * 32, 16, 1024), sdesc(aK + kofs + ks * 2048, 8192, 1024),
                     id_dq, (kt > 0 || ks > 0));
          umma_commit(&bars[5]);
        }
      }
    }
  } else {
    // ===================== element-wise warps =====================
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const float keep_scale = drop_keep_scale(P.drop_p);
    const float kss = keep_scale * P.scale;
    const uint32_t thr = drop_thr16(P.drop_p);
    const int64_t mask_ld = 8 * (int64_t)((Lk + 63) / 64);
    const int64_t sbase = ((int64_t)b * P.H + h) * Lq;
    int it = 0;
    for (int kt = 0; kt < n_kt; ++kt) {
      const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;
      const int n16 = w >> 4;
      const int u0 = half ? (n16 + 1) / 2 : 0, u1 = half ? n16 : (n16 + 1) / 2;
      for (int qt = 0; qt < n_qt; ++qt, ++it) {
        const int qrow = qt * TQ + row;
        const bool active = qt * TQ + quad * 32 < Lq;  // warp-uniform
        float lse_l2 = INFINITY, delta_s = 0.f;          // rows >= Lq: P = exp2(-inf) = 0
        if (qrow < Lq) {
          lse_l2 = P.lse[sbase + qrow] * LOG2E;
          delta_s = P.delta[sbase + qrow] * P.scale;
        }
        mbar_wait(&bars[3], (uint32_t)(it & 1));
        tcgen05_fence_after();
        if (active) {
          const uint64_t row_lin = (uint64_t)(sbase + qrow);
#pragma unroll 1
          for (int u = u0; u < u1; ++u) {
            uint32_t rs[16], rd[16];
            tmem_ld16(tS + lane_addr + u * 16, rs);
            tmem_ld16(tDP + lane_addr + u * 16, rd);
            tmem_ld_wait();
            const float4* mk = reinterpret_cast<const float4*>(sMask + kt * KT + u * 16);
            float pv[16], ds[16];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 m4 = mk[j];
              pv[4 * j + 0] = ex2_approx(fmaf(__uint_as_float(rs[4 * j + 0]), A.scale_l2, m4.x - lse_l2));
              pv[4 * j + 1] = ex2_approx(fmaf(__uint_as_float(rs[4 * j + 1]), A.scale_l2, m4.y - lse_l2));
              pv[4 * j + 2] = ex2_approx(fmaf(__uint_as_float(rs[4 * j + 2]), A.scale_l2, m4.z - lse_l2));
              pv[4 * j + 3] = ex2_approx(fmaf(__uint_as_float(rs[4 * j + 3]), A.scale_l2, m4.w - lse_l2));
            }
            if (DROP) {
              const int g0 = (kt * KT + u * 16) >> 3;  // 8-key group index of the unit's first key
              uint32_t bits;
              if (P.drop_mask != nullptr && qrow < Lq)
                bits = *reinterpret_cast<const uint16_t*>(P.drop_mask + (int64_t)row_lin * mask_ld + g0);
              else
                bits = attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, g0, thr) |
                       (attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, g0 + 1, thr) << 8);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const bool keep = (bits >> j) & 1u;
                const float t = keep ? __uint_as_float(rd[j]) * kss : 0.f;
                ds[j] = pv[j] * (t - delta_s);
                pv[j] = keep ? pv[j] * keep_scale : 0.f;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) ds[j] = pv[j] * fmaf(__uint_as_float(rd[j]), P.scale, -delta_s);
            }
            store16(sP, row, u * 16, pv);
            store16(sdS, row, u * 16, ds);
          }
        }
        fence_proxy_async();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[4]);
        if (qt == n_qt - 1) {
          // dK, dV of this key tile are complete once the iteration's MMAs have finished: rows of the accumulators = keys
          mbar_wait(&bars[5], (uint32_t)(it & 1));
          tcgen05_fence_after();
          const bool kactive = quad * 32 < w;
          uint32_t rk[32], rv[32];
          if (kactive) {
            tmem_ld_32x32b_x32(tDK + lane_addr + half * 32, rk);
            tmem_ld_32x32b_x32(tDV + lane_addr + half * 32, rv);
            tmem_ld_wait();
          }
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[6]);
          const int key = kt * KT + row;
          if (kactive && row < w && key < Lk) {
            store_row32((bf16*)P.dk + ((int64_t)b * Lk + key) * P.lddk + h * D + half * 32, rk, 1.0f);
            store_row32((bf16*)P.dv + ((int64_t)b * Lk + key) * P.lddv + h * D + half * 32, rv, 1.0f);
          }
        }
      }
    }
    // dQ: every key tile has been accumulated (barrier 5 of the last iteration was observed above)
    for (int qt = 0; qt < n_qt; ++qt) {
      const int qrow = qt * TQ + row;
      if (qt * TQ + quad * 32 < Lq) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tDQ + 64 * qt + lane_addr + half * 32, r);
        tmem_ld_wait();
        if (qrow < Lq) store_row32((bf16*)P.dq + ((int64_t)b * Lq + qrow) * P.lddq + h * D + half * 32, r, 1.0f);
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 8) {
    tcgen05_fence_after();
    tmem_dealloc_n(tmem_base, (uint32_t)A.tmem_cols);
  }
}

// ---------------------------------------------------------------- host side
PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
  });
  return fn;
}

// [rows, 64*H] bf16 view with row pitch ld; box = 64 columns (one head) x box_rows, SWIZZLE_128B
int encode_head_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  auto fn = encode_fn();
  M3AE_CHECK(fn != nullptr, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)(ld * 2)};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  M3AE_CHECK(r == CUDA_SUCCESS, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled (attention) failed with CUresult %d", (int)r);
  return M3AE_OK;
}

// `force`: take every covered shape (test hook, backend M3AE_GEMM_TCGEN05).  Otherwise only where this kernel beats the
// mma.sync tiles (profiles/r1_attn_bench_v3_tcgen05.txt): forward when both lengths exceed one 64-wide tile, backward
// when either does — the 32x32 text self-attention is launch-latency bound either way.
bool tc_supported(const m3ae_attn_params_t* p, bool bwd, bool force) {
  if (p->dh != D || p->dtype != M3AE_BF16 || p->causal) return false;
  if (p->Lq > 3 * TQ || p->Lk > 384) return false;
  if (!force && (bwd ? (p->Lq <= 64 && p->Lk <= 64) : (p->Lq <= 64 || p->Lk <= 64))) return false;
  const int64_t lds[] = {p->ldq, p->ldk, p->ldv, p->ldo};
  for (int64_t l : lds) if (l % 8) return false;
  const void* ptrs[] = {p->q, p->k, p->v, p->out};
  for (const void* q : ptrs) if (((uintptr_t)q) & 15) return false;
  if (bwd) {
    if (p->lddq % 8 || p->lddk % 8 || p->lddv % 8) return false;
    const void* g[] = {p->dout, p->dq, p->dk, p->dv};
    for (const void* q : g) if (((uintptr_t)q) & 15) return false;
  }
  return true;
}

int next_pow2_cols(int c) {
  int n = 32;
  while (n < c) n <<= 1;
  return n;
}

// Fills the geometry shared by both kernels; returns the dynamic shared memory size through *smem_bytes.
int setup(const m3ae_attn_params_t* p, bool bwd, TcAttn* A, int* smem_bytes) {
  A->p = *p;
  A->lk_pad = (p->Lk + 15) & ~15;
  A->nbox_k = (A->lk_pad + 127) / 128;
  A->box_k = (((A->lk_pad + A->nbox_k - 1) / A->nbox_k) + 15) & ~15;
  A->n_qt = (p->Lq + TQ - 1) / TQ;
  A->box_q = p->Lq >= TQ ? TQ : ((p->Lq + 15) & ~15);
  A->scale_l2 = p->scale * LOG2E;
  const int cols = p->H * D;
  int rc;
  if ((rc = encode_head_map(&A->map_q, p->q, (int64_t)p->B * p->Lq, cols, p->ldq, A->box_q))) return rc;
  if ((rc = encode_head_map(&A->map_k, p->k, (int64_t)p->B * p->Lk, cols, p->ldk, A->box_k))) return rc;
  if ((rc = encode_head_map(&A->map_v, p->v, (int64_t)p->B * p->Lk, cols, p->ldv, A->box_k))) return rc;
  if (bwd && (rc = encode_head_map(&A->map_do, p->dout, (int64_t)p->B * p->Lq, cols, p->ldo, A->box_q))) return rc;
  const int kv_bytes = A->nbox_k * A->box_k * 128;  // multiple of 2048
  int off = kv_bytes;
  A->off_v = off; off += kv_bytes;
  A->off_q = off; off += 2 * BLK;
  if (bwd) {
    A->off_do = off; off += 2 * BLK;
    A->off_p = off; off += 2 * BLK;
    A->off_ds = off; off += 2 * BLK;
    A->tmem_cols = next_pow2_cols(64 * A->n_qt + 128 + 2 * (A->lk_pad < KT ? A->lk_pad : KT));
  } else {
    A->off_do = 0;
    A->off_p = off; off += ((A->lk_pad + 63) / 64) * BLK;
    A->off_ds = 0;
    A->tmem_cols = next_pow2_cols(64 + A->lk_pad);
  }
  A->off_mask = off; off += 4 * 384;
  A->off_red = off; off += 6 * 4 * 128;
  A->off_bar = off; off += 8 * 8 + 16;
  *smem_bytes = off + 1024;  // alignment slack
  M3AE_CHECK(A->tmem_cols <= 512, M3AE_ERR_UNSUPPORTED, "attention_tc: %d TMEM columns needed", A->tmem_cols);
  M3AE_CHECK(*smem_bytes <= 232448, M3AE_ERR_UNSUPPORTED, "attention_tc: %d bytes of shared memory needed", *smem_bytes);
  return M3AE_OK;
}

template <typename K>
int set_smem_attr(K kernel) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  M3AE_CHECK(e == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(attention_tc) failed: %s", cudaGetErrorString(e));
  return M3AE_OK;
}

}  // namespace

int attn_fwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st) {
  if (!tc_supported(p, false, force)) return M3AE_ERR_UNSUPPORTED;
  static std::once_flag once;
  static int attr_rc = M3AE_OK;
  std::call_once(once, [] {
    attr_rc = set_smem_attr(attn_fwd_tc_kernel<false>);
    if (!attr_rc) attr_rc = set_smem_attr(attn_fwd_tc_kernel<true>);
  });
  if (attr_rc) return attr_rc;
  TcAttn A;
  memset(&A, 0, sizeof(A));
  int smem = 0;
  int rc = setup(p, false, &A, &smem);
  if (rc) return rc;
  dim3 grid(p->H, p->B);
  if (p->drop_p > 0.f) attn_fwd_tc_kernel<true><<<grid, NT, smem, st>>>(A);
  else attn_fwd_tc_kernel<false><<<grid, NT, smem, st>>>(A);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int attn_bwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st) {
  if (!tc_supported(p, true, force)) return M3AE_ERR_UNSUPPORTED;
  static std::once_flag once;
  static int attr_rc = M3AE_OK;
  std::call_once(once, [] {
    attr_rc = set_smem_attr(attn_bwd_tc_kernel<false>);
    if (!attr_rc) attr_rc = set_smem_attr(attn_bwd_tc_kernel<true>);
  });
  if (attr_rc) return attr_rc;
  TcAttn A;
  memset(&A, 0, sizeof(A));
  int smem = 0;
  int rc = setup(p, true, &A, &smem);
  if (rc) return rc;
  launch_delta_bf16_64(p, st);
  dim3 grid(p->H, p->B);
  if (p->drop_p > 0.f) attn_bwd_tc_kernel<true><<<grid, NT, smem, st>>>(A);
  else attn_bwd_tc_kernel<false><<<grid, NT, smem, st>>>(A);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
