// Tensor-core flash attention for the bf16 product path, head dim 64 (the fusion stack's 12 x 64 heads).
// mma.sync.m16n8k16 (bf16 in, fp32 accumulate); S and P never leave registers: the score accumulator fragments are
// re-packed in place as the A operand of the P·V product, the online softmax runs on the fragments with quad shuffles.
//   forward : one CTA = NWARPS x 16 query rows of one (batch, head); loops over 64-key tiles staged in smem.
//   backward: dQ kernel (same tiling, recomputes P from the saved log-sum-exp) and dK/dV kernel (one CTA = 64 keys,
//             loops over 64-query tiles); no atomics, deterministic.
// Reference arithmetic: bert_model.py:301-344 (see attention.cu).  Algorithmic work per (b,h): 4*Lq*Lk*64 flop forward.
//
// NOTE (round 1): legacy tensor path (HMMA).  Attention is 5.5% of the fusion FLOPs; the GEMMs (94.5%) run on
// tcgen05/TMEM.  A tcgen05 attention is a later-round item.
#include "common.cuh"

namespace m3ae {
void count_launch();
namespace {

constexpr int D = 64;     // head dim
constexpr int LDS = 72;   // padded row length in bf16 (144 B): conflict-free ldmatrix
// Tile sizes are template parameters: NWARPS x 16 rows per CTA, BN (32 / 48 / 64) columns per smem tile.  L = 325
// (ViT-B/16 at 288 px) is 7 x 48 - 11, so 3-warp CTAs and 48-wide tiles waste 3% where 64/64 tiling wasted 28%;
// L = 32 (text) uses 32-wide tiles.

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}

// A-operand fragments (16 rows x 64 k) of rows [r0, r0+16) of a [*][LDS] tile: frag[kk] covers k = 16kk..16kk+15.
__device__ __forceinline__ void load_a_frags(uint32_t (&f)[4][4], const bf16 (*tile)[LDS], int r0, int lane) {
  const int mi = lane >> 3, r = lane & 7;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) ldsm_x4(f[kk], &tile[r0 + (mi & 1) * 8 + r][kk * 16 + (mi >> 1) * 8]);
}
// B fragments for C[m][n] += A[m][k] * T[n][k] (T stored [n][k], k contiguous): n-tiles 2jp, 2jp+1, k-step kk.
__device__ __forceinline__ void load_b_nk(uint32_t (&b)[4], const bf16 (*tile)[LDS], int jp, int kk, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4(b, &tile[(2 * jp + (mi >> 1)) * 8 + r][kk * 16 + (mi & 1) * 8]);
}
// B fragments for C[m][n] += A[m][k] * T[k][n] (T stored [k][n], n contiguous): n-tiles 2jp, 2jp+1, k-step kk.
__device__ __forceinline__ void load_b_kn(uint32_t (&b)[4], const bf16 (*tile)[LDS], int jp, int kk, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4_t(b, &tile[kk * 16 + (mi & 1) * 8 + r][(2 * jp + (mi >> 1)) * 8]);
}

// rows [row0, row0+NROWS) x 64 bf16 of a global matrix (row stride ld) -> smem tile, zero-filling rows >= nrows_total
template <int NROWS, int NTHREADS>
__device__ __forceinline__ void load_tile(bf16 (*tile)[LDS], const bf16* __restrict__ g, int64_t ld, int row0, int nrows_total, int tid) {
#pragma unroll
  for (int i = tid; i < NROWS * 8; i += NTHREADS) {
    const int r = i >> 3, c = (i & 7) * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row0 + r < nrows_total) v = *reinterpret_cast<const uint4*>(g + (int64_t)(row0 + r) * ld + c);
    *reinterpret_cast<uint4*>(&tile[r][c]) = v;
  }
}

__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// ---- dropout keep-bit tiles: sBits[row] holds 64 bits = keys k0..k0+63 of attention row (q0+row).
// Forward generates the tile cooperatively (one Philox evaluation per 8 keys) and stores it to the global bit cache;
// backward kernels load it back.  Without a cache (drop_mask == NULL) every kernel regenerates the tile.
__device__ __forceinline__ int64_t mask_ld_bytes(const m3ae_attn_params_t& P) { return 8 * (int64_t)((P.Lk + 63) / 64); }
// sBits[row] (8 bytes per row) holds the keep bits of keys c0 .. c0+8*GROUPS-1 of attention row row0+row.
// gen: evaluate Philox (8 keys per evaluation) and, if a global cache is given, store the bytes there as well.
template <int ROWS, int GROUPS, int NT>
__device__ __forceinline__ void gen_keep_tile(uint64_t* sBits, const m3ae_attn_params_t& P, int b, int h, int row0, int c0, int tid, bool store) {
  const uint32_t thr = drop_thr15(P.drop_p);
  uint8_t* sb = reinterpret_cast<uint8_t*>(sBits);
  const int64_t ld = mask_ld_bytes(P);
  for (int i = tid; i < ROWS * GROUPS; i += NT) {
    const int r = i / GROUPS, grp = i % GROUPS;
    const uint64_t row_lin = ((uint64_t)b * P.H + h) * P.Lq + (row0 + r);
    const uint8_t bits = (uint8_t)attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, (c0 >> 3) + grp, thr);
    sb[r * 8 + grp] = bits;
    if (store && row0 + r < P.Lq && (c0 >> 3) + grp < ld) P.drop_mask[(int64_t)row_lin * ld + (c0 >> 3) + grp] = bits;
  }
}
template <int ROWS, int GROUPS, int NT>
__device__ __forceinline__ void load_keep_tile(uint64_t* sBits, const m3ae_attn_params_t& P, int b, int h, int row0, int c0, int tid) {
  uint8_t* sb = reinterpret_cast<uint8_t*>(sBits);
  const int64_t ld = mask_ld_bytes(P);
  const uint8_t* base = P.drop_mask + (((int64_t)b * P.H + h) * P.Lq) * ld;
  for (int i = tid; i < ROWS * GROUPS; i += NT) {
    const int r = i / GROUPS, grp = i % GROUPS;
    sb[r * 8 + grp] = (row0 + r < P.Lq && (c0 >> 3) + grp < ld) ? base[(int64_t)(row0 + r) * ld + (c0 >> 3) + grp] : (uint8_t)0;
  }
}

// ------------------------------------------------------------------------------------------------ forward
template <int NWARPS, int BN>
__global__ void __launch_bounds__(NWARPS * 32) attn_fwd_mma_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int BM = NWARPS * 16, NT = NWARPS * 32, NJ = BN / 8, NK2 = BN / 16;
  __shared__ __align__(16) bf16 sQ[BM][LDS];
  __shared__ __align__(16) bf16 sK[BN][LDS];
  __shared__ __align__(16) bf16 sV[BN][LDS];
  __shared__ float sMask[BN];
  __shared__ uint64_t sBits[BM];
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BM;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const bf16* __restrict__ Q = (const bf16*)P.q + (int64_t)b * P.Lq * P.ldq + h * D;
  const bf16* __restrict__ K = (const bf16*)P.k + (int64_t)b * P.Lk * P.ldk + h * D;
  const bf16* __restrict__ V = (const bf16*)P.v + (int64_t)b * P.Lk * P.ldv + h * D;
  const float* __restrict__ mask = P.mask ? P.mask + (int64_t)b * P.Lk : nullptr;
  const bool drop = P.drop_p > 0.f;

  load_tile<BM, NT>(sQ, Q, P.ldq, q0, P.Lq, tid);
  __syncthreads();
  uint32_t qa[4][4];
  load_a_frags(qa, sQ, warp * 16, lane);

  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const int qi0 = q0 + warp * 16 + g, qi1 = qi0 + 8;
  const float keep_scale = drop_keep_scale(P.drop_p);

  for (int k0 = 0; k0 < P.Lk; k0 += BN) {
    __syncthreads();
    load_tile<BN, NT>(sK, K, P.ldk, k0, P.Lk, tid);
    load_tile<BN, NT>(sV, V, P.ldv, k0, P.Lk, tid);
    for (int i = tid; i < BN; i += NT) sMask[i] = (k0 + i < P.Lk) ? (mask ? mask[k0 + i] : 0.f) : -INFINITY;
    if (drop) {
      if (P.drop_mask != nullptr && P.drop_mask_ready) load_keep_tile<BM, NJ, NT>(sBits, P, b, h, q0, k0, tid);  // pre-generated
      else gen_keep_tile<BM, NJ, NT>(sBits, P, b, h, q0, k0, tid, P.drop_mask != nullptr);
    }
    __syncthreads();

    float s[NJ][4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int jp = 0; jp < NJ / 2; ++jp) {
        uint32_t bb[4];
        load_b_nk(bb, sK, jp, kk, lane);
        mma16816(s[2 * jp], qa[kk], bb[0], bb[1]);
        mma16816(s[2 * jp + 1], qa[kk], bb[2], bb[3]);
      }
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = j * 8 + 2 * t;
      const float ma = sMask[c], mb = sMask[c + 1];
      s[j][0] = s[j][0] * P.scale + ma; s[j][1] = s[j][1] * P.scale + mb;
      s[j][2] = s[j][2] * P.scale + ma; s[j][3] = s[j][3] * P.scale + mb;
      if (P.causal) {
        const int key = k0 + c;
        if (key > qi0) s[j][0] = -INFINITY;
        if (key + 1 > qi0) s[j][1] = -INFINITY;
        if (key > qi1) s[j][2] = -INFINITY;
        if (key + 1 > qi1) s[j][3] = -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
      mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
    }
    const float mn0 = fmaxf(m0, quad_max(mx0)), mn1 = fmaxf(m1, quad_max(mx1));
    const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;
    const float c0 = __expf(m0 - mu0), c1 = __expf(m1 - mu1);
    m0 = mn0; m1 = mn1;
    float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      s[j][0] = __expf(s[j][0] - mu0); s[j][1] = __expf(s[j][1] - mu0);
      s[j][2] = __expf(s[j][2] - mu1); s[j][3] = __expf(s[j][3] - mu1);
      ps0 += s[j][0] + s[j][1];
      ps1 += s[j][2] + s[j][3];
    }
    l0 = l0 * c0 + ps0;
    l1 = l1 * c1 + ps1;
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] *= c0; o[j][1] *= c0; o[j][2] *= c1; o[j][3] *= c1; }
    if (drop) {
      const uint64_t bits0 = sBits[warp * 16 + g] >> (2 * t), bits1 = sBits[warp * 16 + g + 8] >> (2 * t);
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const uint32_t ka = (uint32_t)(bits0 >> (8 * j)) & 3u, kb = (uint32_t)(bits1 >> (8 * j)) & 3u;
        s[j][0] = (ka & 1) ? s[j][0] * keep_scale : 0.f; s[j][1] = (ka & 2) ? s[j][1] * keep_scale : 0.f;
        s[j][2] = (kb & 1) ? s[j][2] * keep_scale : 0.f; s[j][3] = (kb & 2) ? s[j][3] * keep_scale : 0.f;
      }
    }
#pragma unroll
    for (int kk = 0; kk < NK2; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t bb[4];
        load_b_kn(bb, sV, jp, kk, lane);
        mma16816(o[2 * jp], pa, bb[0], bb[1]);
        mma16816(o[2 * jp + 1], pa, bb[2], bb[3]);
      }
    }
  }
  l0 = quad_sum(l0);
  l1 = quad_sum(l1);
  const float i0 = l0 > 0.f ? 1.0f / l0 : 0.f, i1 = l1 > 0.f ? 1.0f / l1 : 0.f;
  // stage the warp's 16x64 output block through its own rows of sQ, then 16-byte coalesced stores
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    *reinterpret_cast<uint32_t*>(&sQ[warp * 16 + g][j * 8 + 2 * t]) = pack_bf16x2(o[j][0] * i0, o[j][1] * i0);
    *reinterpret_cast<uint32_t*>(&sQ[warp * 16 + g + 8][j * 8 + 2 * t]) = pack_bf16x2(o[j][2] * i1, o[j][3] * i1);
  }
  __syncwarp();
  bf16* __restrict__ O = (bf16*)P.out + (int64_t)b * P.Lq * P.ldo + h * D;
#pragma unroll
  for (int i = lane; i < 16 * 8; i += 32) {
    const int r = i >> 3, c = (i & 7) * 8;
    const int qi = q0 + warp * 16 + r;
    if (qi < P.Lq) *reinterpret_cast<uint4*>(O + (int64_t)qi * P.ldo + c) = *reinterpret_cast<const uint4*>(&sQ[warp * 16 + r][c]);
  }
  if (t == 0) {
    float* lse = P.lse + ((int64_t)b * P.H + h) * P.Lq;
    if (qi0 < P.Lq) lse[qi0] = m0 + __logf(l0);
    if (qi1 < P.Lq) lse[qi1] = m1 + __logf(l1);
  }
}

// ------------------------------------------------------------------------------------------------ backward: dQ
template <int NWARPS, int BN>
__global__ void __launch_bounds__(NWARPS * 32) attn_bwd_dq_mma_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int BM = NWARPS * 16, NT = NWARPS * 32, NJ = BN / 8, NK2 = BN / 16;
  __shared__ __align__(16) bf16 sQ[BM][LDS];   // Q, then dQ staging
  __shared__ __align__(16) bf16 sdO[BM][LDS];
  __shared__ __align__(16) bf16 sK[BN][LDS];
  __shared__ __align__(16) bf16 sV[BN][LDS];
  __shared__ float sMask[BN];
  __shared__ uint64_t sBits[BM];
  const bool drop = P.drop_p > 0.f;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * BM;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const bf16* __restrict__ Q = (const bf16*)P.q + (int64_t)b * P.Lq * P.ldq + h * D;
  const bf16* __restrict__ K = (const bf16*)P.k + (int64_t)b * P.Lk * P.ldk + h * D;
  const bf16* __restrict__ V = (const bf16*)P.v + (int64_t)b * P.Lk * P.ldv + h * D;
  const bf16* __restrict__ dO = (const bf16*)P.dout + (int64_t)b * P.Lq * P.ldo + h * D;
  const float* __restrict__ mask = P.mask ? P.mask + (int64_t)b * P.Lk : nullptr;
  load_tile<BM, NT>(sQ, Q, P.ldq, q0, P.Lq, tid);
  load_tile<BM, NT>(sdO, dO, P.ldo, q0, P.Lq, tid);
  __syncthreads();
  uint32_t qa[4][4], da[4][4];
  load_a_frags(qa, sQ, warp * 16, lane);
  load_a_frags(da, sdO, warp * 16, lane);
  const int qi0 = q0 + warp * 16 + g, qi1 = qi0 + 8;
  const int64_t sbase = ((int64_t)b * P.H + h) * P.Lq;
  const float lse0 = P.lse[sbase + min(qi0, P.Lq - 1)], lse1 = P.lse[sbase + min(qi1, P.Lq - 1)];
  const float dl0 = P.delta[sbase + min(qi0, P.Lq - 1)], dl1 = P.delta[sbase + min(qi1, P.Lq - 1)];
  const float keep_scale = drop_keep_scale(P.drop_p);
  float dq[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) dq[j][0] = dq[j][1] = dq[j][2] = dq[j][3] = 0.f;

  for (int k0 = 0; k0 < P.Lk; k0 += BN) {
    __syncthreads();
    load_tile<BN, NT>(sK, K, P.ldk, k0, P.Lk, tid);
    load_tile<BN, NT>(sV, V, P.ldv, k0, P.Lk, tid);
    for (int i = tid; i < BN; i += NT) sMask[i] = (k0 + i < P.Lk) ? (mask ? mask[k0 + i] : 0.f) : -INFINITY;
    if (drop) {
      if (P.drop_mask) load_keep_tile<BM, NJ, NT>(sBits, P, b, h, q0, k0, tid);
      else gen_keep_tile<BM, NJ, NT>(sBits, P, b, h, q0, k0, tid, false);
    }
    __syncthreads();
    const uint64_t bits0 = drop ? sBits[warp * 16 + g] >> (2 * t) : 0ull, bits1 = drop ? sBits[warp * 16 + g + 8] >> (2 * t) : 0ull;
    float s[NJ][4], dp[NJ][4];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int jp = 0; jp < NJ / 2; ++jp) {
        uint32_t bb[4];
        load_b_nk(bb, sK, jp, kk, lane);
        mma16816(s[2 * jp], qa[kk], bb[0], bb[1]);
        mma16816(s[2 * jp + 1], qa[kk], bb[2], bb[3]);
        load_b_nk(bb, sV, jp, kk, lane);
        mma16816(dp[2 * jp], da[kk], bb[0], bb[1]);
        mma16816(dp[2 * jp + 1], da[kk], bb[2], bb[3]);
      }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = j * 8 + 2 * t;
      const int key = k0 + c;
      const float ma = sMask[c], mb = sMask[c + 1];
      float p0 = __expf(s[j][0] * P.scale + ma - lse0), p1 = __expf(s[j][1] * P.scale + mb - lse0);
      float p2 = __expf(s[j][2] * P.scale + ma - lse1), p3 = __expf(s[j][3] * P.scale + mb - lse1);
      if (P.causal) {
        if (key > qi0) p0 = 0.f;
        if (key + 1 > qi0) p1 = 0.f;
        if (key > qi1) p2 = 0.f;
        if (key + 1 > qi1) p3 = 0.f;
      }
      float d0 = dp[j][0], d1 = dp[j][1], d2 = dp[j][2], d3 = dp[j][3];
      if (drop) {
        const uint32_t ka = (uint32_t)(bits0 >> (8 * j)) & 3u, kb = (uint32_t)(bits1 >> (8 * j)) & 3u;
        d0 = (ka & 1) ? d0 * keep_scale : 0.f; d1 = (ka & 2) ? d1 * keep_scale : 0.f;
        d2 = (kb & 1) ? d2 * keep_scale : 0.f; d3 = (kb & 2) ? d3 * keep_scale : 0.f;
      }
      s[j][0] = p0 * (d0 - dl0) * P.scale; s[j][1] = p1 * (d1 - dl0) * P.scale;
      s[j][2] = p2 * (d2 - dl1) * P.scale; s[j][3] = p3 * (d3 - dl1) * P.scale;
    }
#pragma unroll
    for (int kk = 0; kk < NK2; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t bb[4];
        load_b_kn(bb, sK, jp, kk, lane);  // dQ[q][d] += dS[q][key] * K[key][d]
        mma16816(dq[2 * jp], pa, bb[0], bb[1]);
        mma16816(dq[2 * jp + 1], pa, bb[2], bb[3]);
      }
    }
  }
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    *reinterpret_cast<uint32_t*>(&sQ[warp * 16 + g][j * 8 + 2 * t]) = pack_bf16x2(dq[j][0], dq[j][1]);
    *reinterpret_cast<uint32_t*>(&sQ[warp * 16 + g + 8][j * 8 + 2 * t]) = pack_bf16x2(dq[j][2], dq[j][3]);
  }
  __syncwarp();
  bf16* __restrict__ dQ = (bf16*)P.dq + (int64_t)b * P.Lq * P.lddq + h * D;
#pragma unroll
  for (int i = lane; i < 16 * 8; i += 32) {
    const int r = i >> 3, c = (i & 7) * 8;
    const int qi = q0 + warp * 16 + r;
    if (qi < P.Lq) *reinterpret_cast<uint4*>(dQ + (int64_t)qi * P.lddq + c) = *reinterpret_cast<const uint4*>(&sQ[warp * 16 + r][c]);
  }
}

// ------------------------------------------------------------------------------------------------ backward: dK, dV
// Works on the transposed problem: rows = keys (16 per warp), columns = queries (64 per smem tile).
template <int NWARPS, int BN>
__global__ void __launch_bounds__(NWARPS * 32) attn_bwd_dkv_mma_kernel(const m3ae_attn_params_t P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  constexpr int BM = NWARPS * 16, NT = NWARPS * 32, NJ = BN / 8, NK2 = BN / 16;
  __shared__ __align__(16) bf16 sK[BM][LDS];   // K, then dK staging
  __shared__ __align__(16) bf16 sV[BM][LDS];   // V, then dV staging
  __shared__ __align__(16) bf16 sQ[BN][LDS];
  __shared__ __align__(16) bf16 sdO[BN][LDS];
  __shared__ float sLse[BN], sDl[BN];
  __shared__ uint64_t sBits[BN];  // [query in tile] -> keep bits of the 64-key group containing this CTA's keys
  const bool drop = P.drop_p > 0.f;
  const int b = blockIdx.z, h = blockIdx.y, kb0 = blockIdx.x * BM;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const bf16* __restrict__ Q = (const bf16*)P.q + (int64_t)b * P.Lq * P.ldq + h * D;
  const bf16* __restrict__ K = (const bf16*)P.k + (int64_t)b * P.Lk * P.ldk + h * D;
  const bf16* __restrict__ V = (const bf16*)P.v + (int64_t)b * P.Lk * P.ldv + h * D;
  const bf16* __restrict__ dO = (const bf16*)P.dout + (int64_t)b * P.Lq * P.ldo + h * D;
  load_tile<BM, NT>(sK, K, P.ldk, kb0, P.Lk, tid);
  load_tile<BM, NT>(sV, V, P.ldv, kb0, P.Lk, tid);
  __syncthreads();
  uint32_t ka[4][4], va[4][4];
  load_a_frags(ka, sK, warp * 16, lane);
  load_a_frags(va, sV, warp * 16, lane);
  const int key0 = kb0 + warp * 16 + g, key1 = key0 + 8;
  const float* mrow = P.mask ? P.mask + (int64_t)b * P.Lk : nullptr;
  const float ma0 = (mrow && key0 < P.Lk) ? mrow[key0] : 0.f, ma1 = (mrow && key1 < P.Lk) ? mrow[key1] : 0.f;
  const int64_t sbase = ((int64_t)b * P.H + h) * P.Lq;
  const float keep_scale = drop_keep_scale(P.drop_p);
  const int kl0 = warp * 16 + g, kl1 = kl0 + 8;  // bit positions of this thread's two keys inside the CTA's key bytes
  float dk[8][4], dv[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    dk[j][0] = dk[j][1] = dk[j][2] = dk[j][3] = 0.f;
    dv[j][0] = dv[j][1] = dv[j][2] = dv[j][3] = 0.f;
  }
  for (int i0 = 0; i0 < P.Lq; i0 += BN) {
    __syncthreads();
    load_tile<BN, NT>(sQ, Q, P.ldq, i0, P.Lq, tid);
    load_tile<BN, NT>(sdO, dO, P.ldo, i0, P.Lq, tid);
    for (int i = tid; i < BN; i += NT) {
      const bool ok = i0 + i < P.Lq;
      sLse[i] = ok ? P.lse[sbase + i0 + i] : INFINITY;  // exp(s - inf) = 0 for padded queries
      sDl[i] = ok ? P.delta[sbase + i0 + i] : 0.f;
    }
    if (drop) {
      if (P.drop_mask) load_keep_tile<BN, BM / 8, NT>(sBits, P, b, h, i0, kb0, tid);
      else gen_keep_tile<BN, BM / 8, NT>(sBits, P, b, h, i0, kb0, tid, false);
    }
    __syncthreads();
    float s[NJ][4], dp[NJ][4];  // S^T[key][query], dP^T[key][query]
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
#pragma unroll
      for (int jp = 0; jp < NJ / 2; ++jp) {
        uint32_t bb[4];
        load_b_nk(bb, sQ, jp, kk, lane);
        mma16816(s[2 * jp], ka[kk], bb[0], bb[1]);
        mma16816(s[2 * jp + 1], ka[kk], bb[2], bb[3]);
        load_b_nk(bb, sdO, jp, kk, lane);
        mma16816(dp[2 * jp], va[kk], bb[0], bb[1]);
        mma16816(dp[2 * jp + 1], va[kk], bb[2], bb[3]);
      }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = j * 8 + 2 * t;
      const int qa_ = i0 + c, qb_ = qa_ + 1;
      const float la = sLse[c], lb = sLse[c + 1], da_ = sDl[c], db_ = sDl[c + 1];
      float p0 = __expf(s[j][0] * P.scale + ma0 - la), p1 = __expf(s[j][1] * P.scale + ma0 - lb);
      float p2 = __expf(s[j][2] * P.scale + ma1 - la), p3 = __expf(s[j][3] * P.scale + ma1 - lb);
      if (P.causal) {
        if (key0 > qa_) p0 = 0.f;
        if (key0 > qb_) p1 = 0.f;
        if (key1 > qa_) p2 = 0.f;
        if (key1 > qb_) p3 = 0.f;
      }
      float d0 = dp[j][0], d1 = dp[j][1], d2 = dp[j][2], d3 = dp[j][3];
      float e0 = p0, e1 = p1, e2 = p2, e3 = p3;  // P' (dropped) for dV
      if (drop) {
        const uint64_t ba = sBits[c], bb_ = sBits[c + 1];
        const bool k00 = (ba >> kl0) & 1ull, k01 = (bb_ >> kl0) & 1ull, k10 = (ba >> kl1) & 1ull, k11 = (bb_ >> kl1) & 1ull;
        d0 = k00 ? d0 * keep_scale : 0.f; d1 = k01 ? d1 * keep_scale : 0.f;
        d2 = k10 ? d2 * keep_scale : 0.f; d3 = k11 ? d3 * keep_scale : 0.f;
        e0 = k00 ? p0 * keep_scale : 0.f; e1 = k01 ? p1 * keep_scale : 0.f;
        e2 = k10 ? p2 * keep_scale : 0.f; e3 = k11 ? p3 * keep_scale : 0.f;
      }
      dp[j][0] = p0 * (d0 - da_) * P.scale; dp[j][1] = p1 * (d1 - db_) * P.scale;
      dp[j][2] = p2 * (d2 - da_) * P.scale; dp[j][3] = p3 * (d3 - db_) * P.scale;
      s[j][0] = e0; s[j][1] = e1; s[j][2] = e2; s[j][3] = e3;
    }
#pragma unroll
    for (int kk = 0; kk < NK2; ++kk) {
      uint32_t pa[4], sa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
      sa[0] = pack_bf16x2(dp[2 * kk][0], dp[2 * kk][1]);
      sa[1] = pack_bf16x2(dp[2 * kk][2], dp[2 * kk][3]);
      sa[2] = pack_bf16x2(dp[2 * kk + 1][0], dp[2 * kk + 1][1]);
      sa[3] = pack_bf16x2(dp[2 * kk + 1][2], dp[2 * kk + 1][3]);
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t bb[4];
        load_b_kn(bb, sdO, jp, kk, lane);  // dV[key][d] += P'^T[key][q] * dO[q][d]
        mma16816(dv[2 * jp], pa, bb[0], bb[1]);
        mma16816(dv[2 * jp + 1], pa, bb[2], bb[3]);
        load_b_kn(bb, sQ, jp, kk, lane);   // dK[key][d] += dS^T[key][q] * Q[q][d]
        mma16816(dk[2 * jp], sa, bb[0], bb[1]);
        mma16816(dk[2 * jp + 1], sa, bb[2], bb[3]);
      }
    }
  }
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    *reinterpret_cast<uint32_t*>(&sK[warp * 16 + g][j * 8 + 2 * t]) = pack_bf16x2(dk[j][0], dk[j][1]);
    *reinterpret_cast<uint32_t*>(&sK[warp * 16 + g + 8][j * 8 + 2 * t]) = pack_bf16x2(dk[j][2], dk[j][3]);
    *reinterpret_cast<uint32_t*>(&sV[warp * 16 + g][j * 8 + 2 * t]) = pack_bf16x2(dv[j][0], dv[j][1]);
    *reinterpret_cast<uint32_t*>(&sV[warp * 16 + g + 8][j * 8 + 2 * t]) = pack_bf16x2(dv[j][2], dv[j][3]);
  }
  __syncwarp();
  bf16* __restrict__ dK = (bf16*)P.dk + (int64_t)b * P.Lk * P.lddk + h * D;
  bf16* __restrict__ dV = (bf16*)P.dv + (int64_t)b * P.Lk * P.lddv + h * D;
#pragma unroll
  for (int i = lane; i < 16 * 8; i += 32) {
    const int r = i >> 3, c = (i & 7) * 8;
    const int key = kb0 + warp * 16 + r;
    if (key < P.Lk) {
      *reinterpret_cast<uint4*>(dK + (int64_t)key * P.lddk + c) = *reinterpret_cast<const uint4*>(&sK[warp * 16 + r][c]);
      *reinterpret_cast<uint4*>(dV + (int64_t)key * P.lddv + c) = *reinterpret_cast<const uint4*>(&sV[warp * 16 + r][c]);
    }
  }
}

bool supported(const m3ae_attn_params_t* p, bool bwd) {
  if (p->dh != 64 || p->dtype != M3AE_BF16) return false;
  // 16-byte vector loads/stores: row pitches multiple of 8 elements, 16-byte aligned bases
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

}  // namespace

// rows per CTA (16 x warps) and smem tile width for a sequence length: 325 -> 48 (7 x 48 = 336), <= 32 -> 32, else 64
inline int pick_tile(int L) { return L <= 32 ? 32 : ((L + 47) / 48 * 48 <= (L + 63) / 64 * 64 ? 48 : 64); }

#define ATTN_DISPATCH(KERNEL, ROWS_L, COLS_L, ROW_COUNT)                                              \
  do {                                                                                                 \
    const int bm = pick_tile(ROWS_L), bn = pick_tile(COLS_L);                                          \
    dim3 grid((ROW_COUNT + bm - 1) / bm, p->H, p->B);                                                  \
    if (bm == 32 && bn == 32) M3AE_LAUNCH((KERNEL<2, 32>), grid, 64, 0, st, *p);                                  \
    else if (bm == 32 && bn == 48) M3AE_LAUNCH((KERNEL<2, 48>), grid, 64, 0, st, *p);                             \
    else if (bm == 32) M3AE_LAUNCH((KERNEL<2, 64>), grid, 64, 0, st, *p);                                         \
    else if (bm == 48 && bn == 32) M3AE_LAUNCH((KERNEL<3, 32>), grid, 96, 0, st, *p);                             \
    else if (bm == 48 && bn == 48) M3AE_LAUNCH((KERNEL<3, 48>), grid, 96, 0, st, *p);                             \
    else if (bm == 48) M3AE_LAUNCH((KERNEL<3, 64>), grid, 96, 0, st, *p);                                         \
    else if (bn == 32) M3AE_LAUNCH((KERNEL<4, 32>), grid, 128, 0, st, *p);                                        \
    else if (bn == 48) M3AE_LAUNCH((KERNEL<4, 48>), grid, 128, 0, st, *p);                                        \
    else M3AE_LAUNCH((KERNEL<4, 64>), grid, 128, 0, st, *p);                                                      \
    count_launch();                                                                                    \
  } while (0)

int attn_fwd_mma(const m3ae_attn_params_t* p, cudaStream_t st) {
  if (!supported(p, false)) return M3AE_ERR_UNSUPPORTED;
  ATTN_DISPATCH(attn_fwd_mma_kernel, p->Lq, p->Lk, p->Lq);
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

void launch_delta_bf16_64(const m3ae_attn_params_t* p, cudaStream_t st);  // attention.cu

int attn_bwd_mma(const m3ae_attn_params_t* p, cudaStream_t st) {
  if (!supported(p, true)) return M3AE_ERR_UNSUPPORTED;
  launch_delta_bf16_64(p, st);
  ATTN_DISPATCH(attn_bwd_dq_mma_kernel, p->Lq, p->Lk, p->Lq);    // CTA rows = queries, tiles over keys
  ATTN_DISPATCH(attn_bwd_dkv_mma_kernel, p->Lk, p->Lq, p->Lk);   // CTA rows = keys, tiles over queries
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
