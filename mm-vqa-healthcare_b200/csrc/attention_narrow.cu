// "Narrow" attention for the bf16 product path, head dim 64: the shapes of the fusion stack in which ONE side is the
// 32-token text sequence — text self-attention (32 x 32), text->image cross-attention (32 queries x 325 keys) and
// image->text cross-attention (325 queries x 32 keys).  Reference arithmetic: bert_model.py:301-344 (self and cross mode).
//
// These shapes are HBM-bound (10*Lq*Lk*64 flop against ~125 KB of Q/dO/dQ or K/V/dK/dV per (batch, head): ~50 flop/B,
// ridge ~210), and on the tcgen05 kernels of attention_tc.cu they were latency chains: one CTA per SM, ~6k cycles of
// prologue + ~5k of tail around ~6k cycles of work (77 / 97 us backward, 36 / 47 us forward on the mma.sync tiles, against
// ~15 us of HBM time each; profiles/r2_attn_bwd_timeline.txt).  Here one CTA = one (batch, head) with EVERYTHING it reads
// brought into shared memory by one burst of cp.async (no staged loop => one DRAM round trip), two CTAs resident per SM so
// one computes while the other loads, mma.sync.m16n8k16 on register fragments (the work is far below the tensor roofline;
// what matters is bytes in flight), and one pass that produces dQ, dK and dV together:
//
//   forward   rows = 16-query blocks, keys in chunks of 32 with an online softmax; when there are fewer row blocks than
//             warps the keys are split across warps and the partial (max, sum, O) states are merged through shared memory.
//   backward  "long side" L (queries when the keys are the 32-token side, keys otherwise) is cut into 16-row blocks, one
//             warp each; S and dP blocks [16 x 32] stay in registers; the gradient of the long side's operand(s) is written
//             per block, the gradient(s) of the narrow side accumulate in registers across the warp's blocks
//             (transposed fragments through movmatrix) and are reduced across the 4 warps once at the end: no atomics,
//             deterministic.  With all keys in one block (keys narrow) delta = rowsum(dO*O) = rowsum(P'*dP) is computed
//             in place, so that case needs no delta launch and never reads O.
// Dropout: the same Philox keep bits as every other attention kernel (common.cuh), staged per (b,h) in shared memory —
// generated (and written to the keep-bit cache) by the forward, read back from the cache by the backward.
#include "common.cuh"

namespace m3ae {
void count_launch();
void launch_delta_bf16_64(const m3ae_attn_params_t* p, cudaStream_t st);  // attention.cu

namespace {

constexpr int D = 64;      // head dim
constexpr int LDS = 72;    // padded smem row (144 B): conflict-free ldmatrix
constexpr int NW = 4;      // warps per CTA
constexpr int NT = NW * 32;
constexpr int NS = 32;     // the narrow side, padded
constexpr int RLD = 72;    // fp32 row pitch of the cross-warp reduction buffer
constexpr int OLD = 68;    // fp32 row pitch of the forward merge buffer
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

typedef bf16 (*Tile)[LDS];

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
// transpose of an 8x8 bf16 tile held one 32-bit register per lane (lane i: row i/4, columns 2(i%4), 2(i%4)+1)
__device__ __forceinline__ uint32_t movm_t(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}

// A-operand fragments (16 rows x 64 k) of rows [0, 16) of `tile`: f[kk] covers k = 16kk .. 16kk+15
__device__ __forceinline__ void load_a_frags(uint32_t (&f)[4][4], Tile tile, int lane) {
  const int mi = lane >> 3, r = lane & 7;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) ldsm_x4(f[kk], &tile[(mi & 1) * 8 + r][kk * 16 + (mi >> 1) * 8]);
}
// B fragments for C[m][n] += A[m][k] * T[n][k] (T stored [n][k], k contiguous): n-tiles 2jp, 2jp+1, k-step kk
__device__ __forceinline__ void load_b_nk(uint32_t (&b)[4], Tile tile, int jp, int kk, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4(b, &tile[(2 * jp + (mi >> 1)) * 8 + r][kk * 16 + (mi & 1) * 8]);
}
// B fragments for C[m][n] += A[m][k] * T[k][n] (T stored [k][n], n contiguous): n-tiles 2jp, 2jp+1, k-step kk
__device__ __forceinline__ void load_b_kn(uint32_t (&b)[4], Tile tile, int jp, int kk, int lane) {
  const int mi = lane >> 3, r = lane & 7;
  ldsm_x4_t(b, &tile[kk * 16 + (mi & 1) * 8 + r][(2 * jp + (mi >> 1)) * 8]);
}
// rows [0, rows_pad) x 64 bf16 of a global matrix -> smem tile (cp.async); rows >= rows_valid are zero-filled
__device__ __forceinline__ void load_rows_async(Tile tile, const bf16* __restrict__ g, int64_t ld, int rows_valid, int rows_pad, int tid) {
  for (int i = tid; i < rows_pad * 8; i += NT) {
    const int r = i >> 3, c = (i & 7) * 8;
    const bool ok = r < rows_valid;
    cp_async16(&tile[r][c], g + (int64_t)(ok ? r : 0) * ld + c, ok ? 16 : 0);
  }
}
// ---- per-warp progressive loading: a warp brings in the 16-row blocks (or 32-key chunks) only IT consumes, in NG cp.async
// groups, and starts computing when the first group has landed.  Every thread's group order is: [shared operands] [g0] .. [g3].
constexpr int NG = 4;
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_pending(int pending) {  // at most `pending` groups of this thread still in flight
  switch (pending) {
    case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
    case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
    case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
    case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
    default: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
  }
}
// `rows` (16 or 32) rows starting at row0 of a global matrix -> the same rows of an smem tile, by ONE warp
__device__ __forceinline__ void warp_load_rows_async(Tile tile, const bf16* __restrict__ g, int64_t ld, int row0, int rows, int rows_valid, int lane) {
  for (int i = lane; i < rows * 8; i += 32) {
    const int r = row0 + (i >> 3), c = (i & 7) * 8;
    const bool ok = r < rows_valid;
    cp_async16(&tile[r][c], g + (int64_t)(ok ? r : 0) * ld + c, ok ? 16 : 0);
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
// 16 x 64 fp32 accumulator block (8 n-tiles) -> bf16 -> rows [0,16) of a warp-private smem tile -> global rows [row0, ..)
__device__ __forceinline__ void store_block(const float (&acc)[8][4], float mul0, float mul1, Tile stage, bf16* __restrict__ dst, int64_t ld,
                                            int row0, int rows_total, int lane) {
  const int g = lane >> 2, t = lane & 3;
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    *reinterpret_cast<uint32_t*>(&stage[g][j * 8 + 2 * t]) = pack_bf16x2(acc[j][0] * mul0, acc[j][1] * mul0);
    *reinterpret_cast<uint32_t*>(&stage[g + 8][j * 8 + 2 * t]) = pack_bf16x2(acc[j][2] * mul1, acc[j][3] * mul1);
  }
  __syncwarp();
#pragma unroll
  for (int i = lane; i < 16 * 8; i += 32) {
    const int r = i >> 3, c = (i & 7) * 8;
    if (row0 + r < rows_total) *reinterpret_cast<uint4*>(dst + (int64_t)(row0 + r) * ld + c) = *reinterpret_cast<const uint4*>(&stage[r][c]);
  }
}

__host__ __device__ inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ inline int keep_ld(int Lk) { return 8 * ((Lk + 63) / 64); }  // bytes per row of the keep-bit cache

// Keep bits of one (b,h): sBits[Lq][mask_ld] (bit k%8 of byte k/8 = key k kept).  from_cache: copy from the global cache;
// otherwise evaluate Philox (8 keys per evaluation) and, if `store`, write the rows to the global cache.
__device__ __forceinline__ void stage_keep_bits(uint8_t* sBits, const m3ae_attn_params_t& P, int64_t sbase, int mask_ld, bool from_cache,
                                                bool store, int tid) {
  const int words = P.Lq * (mask_ld >> 3);
  uint2* s2 = reinterpret_cast<uint2*>(sBits);
  uint2* g2 = reinterpret_cast<uint2*>(P.drop_mask + sbase * mask_ld);
  if (from_cache) {
    for (int i = tid; i < words; i += NT) s2[i] = g2[i];
    return;
  }
  const uint32_t thr = drop_thr15(P.drop_p);
  const int n_units = (P.Lk + 7) >> 3;
  for (int i = tid; i < P.Lq * mask_ld; i += NT) {
    const int r = i / mask_ld, u = i - r * mask_ld;
    sBits[i] = u < n_units ? (uint8_t)attn_keep_bits8(P.drop_seed, P.drop_stream, (uint64_t)(sbase + r), u, thr) : (uint8_t)0;
  }
  if (store) {
    __syncthreads();
    for (int i = tid; i < words; i += NT) g2[i] = s2[i];
  }
}

// ------------------------------------------------------------------------------------------------ forward
// One online-softmax step of a 16-query row block against the 32-key chunk at rows [k0, k0+32) of sK / sV.
// State: o (unnormalised output), m0/m1 (running row max, log2 domain), l0/l1 (this lane's share of the running row sum).
template <bool DROP>
__device__ __forceinline__ void fwd_step(const uint32_t (&qa)[4][4], Tile sK, Tile sV, const float* sMask, int k0, float c, const uint8_t* bits0,
                                         const uint8_t* bits1, float (&o)[8][4], float& m0, float& m1, float& l0, float& l1, int lane) {
  const int t = lane & 3;
  float s[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk)
#pragma unroll
    for (int jp = 0; jp < 2; ++jp) {
      uint32_t bb[4];
      load_b_nk(bb, sK + k0, jp, kk, lane);
      mma16816(s[2 * jp], qa[kk], bb[0], bb[1]);
      mma16816(s[2 * jp + 1], qa[kk], bb[2], bb[3]);
    }
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 mk = *reinterpret_cast<const float2*>(&sMask[k0 + j * 8 + 2 * t]);
    s[j][0] = fmaf(s[j][0], c, mk.x); s[j][1] = fmaf(s[j][1], c, mk.y);
    s[j][2] = fmaf(s[j][2], c, mk.x); s[j][3] = fmaf(s[j][3], c, mk.y);
    mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
    mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
  }
  const float mn0 = fmaxf(m0, quad_max(mx0)), mn1 = fmaxf(m1, quad_max(mx1));
  const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;
  const float c0 = ex2_approx(m0 - mu0), c1 = ex2_approx(m1 - mu1);
  m0 = mn0; m1 = mn1;
  float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    s[j][0] = ex2_approx(s[j][0] - mu0); s[j][1] = ex2_approx(s[j][1] - mu0);
    s[j][2] = ex2_approx(s[j][2] - mu1); s[j][3] = ex2_approx(s[j][3] - mu1);
    ps0 += s[j][0] + s[j][1];
    ps1 += s[j][2] + s[j][3];
  }
  l0 = l0 * c0 + ps0;
  l1 = l1 * c1 + ps1;
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] *= c0; o[j][1] *= c0; o[j][2] *= c1; o[j][3] *= c1; }
  if (DROP) {
    const uint32_t w0 = *reinterpret_cast<const uint32_t*>(bits0) >> (2 * t), w1 = *reinterpret_cast<const uint32_t*>(bits1) >> (2 * t);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t ka = (w0 >> (8 * j)) & 3u, kb = (w1 >> (8 * j)) & 3u;
      s[j][0] = (ka & 1) ? s[j][0] : 0.f; s[j][1] = (ka & 2) ? s[j][1] : 0.f;
      s[j][2] = (kb & 1) ? s[j][2] : 0.f; s[j][3] = (kb & 2) ? s[j][3] : 0.f;
    }
  }
#pragma unroll
  for (int kk = 0; kk < 2; ++kk) {
    uint32_t pa[4];
    pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
    pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
    pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
    pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
    for (int jp = 0; jp < 4; ++jp) {
      uint32_t bb[4];
      load_b_kn(bb, sV + k0, jp, kk, lane);
      mma16816(o[2 * jp], pa, bb[0], bb[1]);
      mma16816(o[2 * jp + 1], pa, bb[2], bb[3]);
    }
  }
}

// NQ = false (keys narrow, Lk <= 32): warp = 16-query row blocks, loaded by the warp itself in NG groups; K, V shared.
// NQ = true  (queries narrow, Lq <= 32 < Lk): warp = every 4th 32-key chunk (loaded by the warp itself), both row blocks;
//            the partial (max, sum, O) states of the 4 warps are merged through shared memory (over the K/V region).
// smem: Q[lq_pad] | K[lk_pad] | V[lk_pad] (NQ: merge buffer aliases K, V) | mask[lk_pad] fp32 | keep bits [lq_pad][mask_ld]
__host__ __device__ inline int fwd_kv_bytes(int Lq, int Lk) {
  const int kv = 2 * round_up(Lk, 32) * LDS * 2, merge = NW * NS * OLD * 4 + 2 * NW * NS * 4;
  return (Lk > NS && kv < merge) ? merge : kv;
}
__host__ __device__ inline int fwd_smem_bytes(int Lq, int Lk) {
  const int lq_pad = round_up(Lq, 16), lk_pad = round_up(Lk, 32);
  return lq_pad * LDS * 2 + fwd_kv_bytes(Lq, Lk) + lk_pad * 4 + lq_pad * keep_ld(Lk);
}

template <bool NQ, bool DROP>
__global__ void __launch_bounds__(NT, 2) attn_fwd_narrow_kernel(const m3ae_attn_params_t P) {
  pdl_sync();
  extern __shared__ __align__(16) uint8_t smem[];
  const int Lq = P.Lq, Lk = P.Lk;
  const int lq_pad = round_up(Lq, 16), lk_pad = round_up(Lk, 32), mask_ld = keep_ld(Lk);
  Tile sQ = reinterpret_cast<Tile>(smem);
  Tile sK = sQ + lq_pad;
  Tile sV = sK + lk_pad;
  float* sMask = reinterpret_cast<float*>(smem + lq_pad * LDS * 2 + fwd_kv_bytes(Lq, Lk));
  uint8_t* sBits = reinterpret_cast<uint8_t*>(sMask + lk_pad);
  const int h = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int64_t sbase = ((int64_t)b * P.H + h) * Lq;
  const bf16* gQ = (const bf16*)P.q + (int64_t)b * Lq * P.ldq + h * D;
  const bf16* gK = (const bf16*)P.k + (int64_t)b * Lk * P.ldk + h * D;
  const bf16* gV = (const bf16*)P.v + (int64_t)b * Lk * P.ldv + h * D;
  const int nrb = lq_pad >> 4, nchunks = lk_pad >> 5;
  // units this warp owns: row blocks (NQ = false) or key chunks (NQ = true): warp, warp + NW, ...
  const int nunits = NQ ? nchunks : nrb;
  const int nmine = nunits > warp ? (nunits - warp + NW - 1) / NW : 0;
  const int per = nmine > 0 ? (nmine + NG - 1) / NG : 1;

  // shared operands first (group 0 of every thread), then this warp's own units in NG groups
  if (NQ) {
    load_rows_async(sQ, gQ, P.ldq, Lq, lq_pad, tid);
  } else {
    load_rows_async(sK, gK, P.ldk, Lk, lk_pad, tid);
    load_rows_async(sV, gV, P.ldv, Lk, lk_pad, tid);
  }
  cp_async_commit();
  for (int gi = 0; gi < NG; ++gi) {
    for (int i = gi * per; i < (gi + 1) * per && i < nmine; ++i) {
      const int u = warp + NW * i;
      if (NQ) {
        warp_load_rows_async(sK, gK, P.ldk, u * 32, 32, Lk, lane);
        warp_load_rows_async(sV, gV, P.ldv, u * 32, 32, Lk, lane);
      } else {
        warp_load_rows_async(sQ, gQ, P.ldq, u * 16, 16, Lq, lane);
      }
    }
    cp_async_commit();
  }
  const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
  for (int i = tid; i < lk_pad; i += NT) sMask[i] = i < Lk ? (mrow ? mrow[i] * LOG2E : 0.f) : -INFINITY;
  if (DROP) stage_keep_bits(sBits, P, sbase, mask_ld, P.drop_mask != nullptr && P.drop_mask_ready != 0, P.drop_mask != nullptr, tid);
  cp_async_wait_pending(NG);
  __syncthreads();

  const float c = P.scale * LOG2E;
  const float keep_scale = drop_keep_scale(P.drop_p);
  bf16* __restrict__ O = (bf16*)P.out + (int64_t)b * Lq * P.ldo + h * D;
  float* __restrict__ lse = P.lse + sbase;

  if constexpr (!NQ) {
    for (int i = 0; i < nmine; ++i) {
      if (i % per == 0) {
        cp_async_wait_pending(NG - 1 - i / per);
        __syncwarp();
      }
      const int rb = warp + NW * i;
      Tile qt = sQ + rb * 16;
      uint32_t qa[4][4];
      load_a_frags(qa, qt, lane);
      float o[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
      float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
      fwd_step<DROP>(qa, sK, sV, sMask, 0, c, &sBits[(rb * 16 + g) * mask_ld], &sBits[(rb * 16 + g + 8) * mask_ld], o, m0, m1, l0, l1, lane);
      l0 = quad_sum(l0);
      l1 = quad_sum(l1);
      const float i0 = l0 > 0.f ? keep_scale / l0 : 0.f, i1 = l1 > 0.f ? keep_scale / l1 : 0.f;
      store_block(o, i0, i1, qt, O, P.ldo, rb * 16, Lq, lane);  // Q fragments are in registers: the block's rows are free
      if (t == 0) {
        const int q0 = rb * 16 + g, q1 = q0 + 8;
        if (q0 < Lq) lse[q0] = (m0 + log2f(l0)) * LN2;
        if (q1 < Lq) lse[q1] = (m1 + log2f(l1)) * LN2;
      }
    }
  } else {
  // ---- queries narrow: this warp's key chunks against both row blocks
  uint32_t qa[2][4][4];
  float o[2][8][4], mm[2][2], ll[2][2];
#pragma unroll
  for (int rb = 0; rb < 2; ++rb) {
    if (rb < nrb) load_a_frags(qa[rb], sQ + rb * 16, lane);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[rb][j][0] = o[rb][j][1] = o[rb][j][2] = o[rb][j][3] = 0.f;
    mm[rb][0] = mm[rb][1] = -INFINITY;
    ll[rb][0] = ll[rb][1] = 0.f;
  }
  for (int i = 0; i < nmine; ++i) {
    if (i % per == 0) {
      cp_async_wait_pending(NG - 1 - i / per);
      __syncwarp();
    }
    const int ch = warp + NW * i;
#pragma unroll
    for (int rb = 0; rb < 2; ++rb)
      if (rb < nrb)
        fwd_step<DROP>(qa[rb], sK, sV, sMask, ch * 32, c, &sBits[(rb * 16 + g) * mask_ld + ch * 4], &sBits[(rb * 16 + g + 8) * mask_ld + ch * 4],
                       o[rb], mm[rb][0], mm[rb][1], ll[rb][0], ll[rb][1], lane);
  }
  __syncthreads();  // every warp is done with K and V: the merge buffer may overwrite them
  float* sO = reinterpret_cast<float*>(sK);  // [NW][32][OLD]
  float* sM = sO + NW * NS * OLD;            // [NW][32]
  float* sL = sM + NW * NS;
#pragma unroll
  for (int rb = 0; rb < 2; ++rb) {
    if (rb < nrb) {
      float* po = sO + (warp * NS + rb * 16) * OLD;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        *reinterpret_cast<float2*>(&po[g * OLD + j * 8 + 2 * t]) = make_float2(o[rb][j][0], o[rb][j][1]);
        *reinterpret_cast<float2*>(&po[(g + 8) * OLD + j * 8 + 2 * t]) = make_float2(o[rb][j][2], o[rb][j][3]);
      }
      const float l0 = quad_sum(ll[rb][0]), l1 = quad_sum(ll[rb][1]);
      if (t == 0) {
        sM[warp * NS + rb * 16 + g] = mm[rb][0]; sM[warp * NS + rb * 16 + g + 8] = mm[rb][1];
        sL[warp * NS + rb * 16 + g] = l0; sL[warp * NS + rb * 16 + g + 8] = l1;
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < nrb * 16 * 8; i += NT) {
    const int q = i >> 3, d0 = (i & 7) * 8;
    if (q >= Lq) continue;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w) M = fmaxf(M, sM[w * NS + q]);
    const float Mu = M == -INFINITY ? 0.f : M;
    float L = 0.f, acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float wt = ex2_approx(sM[w * NS + q] - Mu);
      L += wt * sL[w * NS + q];
      const float4 x = *reinterpret_cast<const float4*>(&sO[(w * NS + q) * OLD + d0]);
      const float4 y = *reinterpret_cast<const float4*>(&sO[(w * NS + q) * OLD + d0 + 4]);
      acc[0] += wt * x.x; acc[1] += wt * x.y; acc[2] += wt * x.z; acc[3] += wt * x.w;
      acc[4] += wt * y.x; acc[5] += wt * y.y; acc[6] += wt * y.z; acc[7] += wt * y.w;
    }
    const float inv = L > 0.f ? keep_scale / L : 0.f;
    uint4 w4;
    w4.x = pack_bf16x2(acc[0] * inv, acc[1] * inv);
    w4.y = pack_bf16x2(acc[2] * inv, acc[3] * inv);
    w4.z = pack_bf16x2(acc[4] * inv, acc[5] * inv);
    w4.w = pack_bf16x2(acc[6] * inv, acc[7] * inv);
    *reinterpret_cast<uint4*>(O + (int64_t)q * P.ldo + d0) = w4;
    if (d0 == 0) lse[q] = (M + log2f(L)) * LN2;
  }
  }  // NQ
}

// ------------------------------------------------------------------------------------------------ backward
// NQ = false: keys are the narrow side.  Long side X1 = Q, X2 = dO (Lx = Lq); narrow Y1 = K, Y2 = V.
//             S = X1 Y1^T, dP = X2 Y2^T; per block dQ = dS Y1; accumulated dV = P'^T X2, dK = dS^T X1.
// NQ = true : queries are the narrow side; the problem is worked transposed.  X1 = K, X2 = V (Lx = Lk); Y1 = Q, Y2 = dO.
//             S^T = X1 Y1^T, dP^T = X2 Y2^T; per block dV = P'^T Y2, dK = dS^T Y1; accumulated dQ = dS X1.
// smem: X1[lx_pad] | X2[lx_pad] (the cross-warp reduction buffer aliases them) | Y1[32] | Y2[32] | row stat [lx_pad] |
//       col stat A [32] | col stat B [32] | keep bits [Lq][mask_ld].
__host__ __device__ inline int bwd_x_bytes(int lx_pad, bool nq) {
  const int x = 2 * lx_pad * LDS * 2, red = NW * (nq ? 1 : 2) * NS * RLD * 4;
  return x > red ? x : red;
}
__host__ __device__ inline int bwd_smem_bytes(int Lq, int Lk, bool nq) {
  const int lx_pad = round_up(nq ? Lk : Lq, 16);
  return bwd_x_bytes(lx_pad, nq) + 2 * NS * LDS * 2 + lx_pad * 4 + 2 * NS * 4 + round_up(Lq * keep_ld(Lk), 16);
}

template <bool NQ, bool DROP>
__global__ void __launch_bounds__(NT, 2) attn_bwd_narrow_kernel(const m3ae_attn_params_t P) {
  pdl_sync();
  extern __shared__ __align__(16) uint8_t smem[];
  const int Lq = P.Lq, Lk = P.Lk;
  const int Lx = NQ ? Lk : Lq, Ly = NQ ? Lq : Lk;
  const int lx_pad = round_up(Lx, 16), mask_ld = keep_ld(Lk);
  Tile sX1 = reinterpret_cast<Tile>(smem);
  Tile sX2 = sX1 + lx_pad;
  Tile sY1 = reinterpret_cast<Tile>(smem + bwd_x_bytes(lx_pad, NQ));
  Tile sY2 = sY1 + NS;
  float* sRow = reinterpret_cast<float*>(sY2 + NS);  // NQ: key mask * log2e (-inf: padded key); else lse * log2e (+inf: padded query)
  float* sColA = sRow + lx_pad;                      // NQ: lse * log2e (+inf: padded query); else key mask * log2e (-inf: padded key)
  float* sColB = sColA + NS;                         // NQ: delta
  uint8_t* sBits = reinterpret_cast<uint8_t*>(sColB + NS);
  const int h = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int64_t sbase = ((int64_t)b * P.H + h) * Lq;
  const bf16* gQ = (const bf16*)P.q + (int64_t)b * Lq * P.ldq + h * D;
  const bf16* gK = (const bf16*)P.k + (int64_t)b * Lk * P.ldk + h * D;
  const bf16* gV = (const bf16*)P.v + (int64_t)b * Lk * P.ldv + h * D;
  const bf16* gdO = (const bf16*)P.dout + (int64_t)b * Lq * P.ldo + h * D;

  // shared (narrow) operands first (group 0 of every thread), then this warp's own 16-row blocks of the long side in NG groups
  const int nblk = lx_pad >> 4;
  const int nmine = nblk > warp ? (nblk - warp + NW - 1) / NW : 0;
  const int per = nmine > 0 ? (nmine + NG - 1) / NG : 1;
  if (NQ) {
    load_rows_async(sY1, gQ, P.ldq, Lq, NS, tid);
    load_rows_async(sY2, gdO, P.ldo, Lq, NS, tid);
  } else {
    load_rows_async(sY1, gK, P.ldk, Lk, NS, tid);
    load_rows_async(sY2, gV, P.ldv, Lk, NS, tid);
  }
  cp_async_commit();
  for (int gi = 0; gi < NG; ++gi) {
    for (int i = gi * per; i < (gi + 1) * per && i < nmine; ++i) {
      const int r0 = (warp + NW * i) * 16;
      if (NQ) {
        warp_load_rows_async(sX1, gK, P.ldk, r0, 16, Lk, lane);
        warp_load_rows_async(sX2, gV, P.ldv, r0, 16, Lk, lane);
      } else {
        warp_load_rows_async(sX1, gQ, P.ldq, r0, 16, Lq, lane);
        warp_load_rows_async(sX2, gdO, P.ldo, r0, 16, Lq, lane);
      }
    }
    cp_async_commit();
  }
  {
    const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
    const float* lse = P.lse + sbase;
    if (NQ) {
      for (int i = tid; i < lx_pad; i += NT) sRow[i] = i < Lk ? (mrow ? mrow[i] * LOG2E : 0.f) : -INFINITY;
      for (int i = tid; i < NS; i += NT) {
        sColA[i] = i < Lq ? lse[i] * LOG2E : INFINITY;
        sColB[i] = i < Lq ? P.delta[sbase + i] : 0.f;
      }
    } else {
      for (int i = tid; i < lx_pad; i += NT) sRow[i] = i < Lq ? lse[i] * LOG2E : INFINITY;
      for (int i = tid; i < NS; i += NT) sColA[i] = i < Lk ? (mrow ? mrow[i] * LOG2E : 0.f) : -INFINITY;
    }
  }
  if (DROP) stage_keep_bits(sBits, P, sbase, mask_ld, P.drop_mask != nullptr, false, tid);
  cp_async_wait_pending(NG);
  __syncthreads();

  const float c = P.scale * LOG2E, scale = P.scale;
  const float ks = drop_keep_scale(P.drop_p);
  constexpr int NACC = NQ ? 1 : 2;  // narrow-side gradients accumulated across the warp's blocks
  float acc[NACC][2][8][4];
#pragma unroll
  for (int a = 0; a < NACC; ++a)
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[a][i][j][0] = acc[a][i][j][1] = acc[a][i][j][2] = acc[a][i][j][3] = 0.f;
  // column statistics of this thread's 8 columns (col = 8j + 2t + e)
  float colA[4][2], colB[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    colA[j][0] = sColA[8 * j + 2 * t]; colA[j][1] = sColA[8 * j + 2 * t + 1];
    colB[j][0] = NQ ? sColB[8 * j + 2 * t] : 0.f; colB[j][1] = NQ ? sColB[8 * j + 2 * t + 1] : 0.f;
  }

  for (int bi = 0; bi < nmine; ++bi) {
    if (bi % per == 0) {
      cp_async_wait_pending(NG - 1 - bi / per);
      __syncwarp();
    }
    const int blk = warp + NW * bi;
    const int r0 = blk * 16;
    Tile x1 = sX1 + r0, x2 = sX2 + r0;
    float s[4][4], dp[4][4];
    {
      uint32_t xa[4][4], xb[4][4];
      load_a_frags(xa, x1, lane);
      load_a_frags(xb, x2, lane);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
        dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
      }
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
#pragma unroll
        for (int jp = 0; jp < 2; ++jp) {
          uint32_t bb[4];
          load_b_nk(bb, sY1, jp, kk, lane);
          mma16816(s[2 * jp], xa[kk], bb[0], bb[1]);
          mma16816(s[2 * jp + 1], xa[kk], bb[2], bb[3]);
          load_b_nk(bb, sY2, jp, kk, lane);
          mma16816(dp[2 * jp], xb[kk], bb[0], bb[1]);
          mma16816(dp[2 * jp + 1], xb[kk], bb[2], bb[3]);
        }
    }
    const float rs0 = sRow[r0 + g], rs1 = sRow[r0 + g + 8];
    // keep bits: !NQ: one 32-bit word per row (bit = key); NQ: per column q a 16-bit word of this key block (bit = key in block)
    uint32_t w0 = 0xFFFFFFFFu, w1 = 0xFFFFFFFFu;
    if (DROP && !NQ) {
      w0 = r0 + g < Lq ? *reinterpret_cast<const uint32_t*>(&sBits[(r0 + g) * mask_ld]) : 0u;
      w1 = r0 + g + 8 < Lq ? *reinterpret_cast<const uint32_t*>(&sBits[(r0 + g + 8) * mask_ld]) : 0u;
    }
    // P (s), P' dropped and scaled (s), dP masked and scaled (dp)
    float pr[4][4];  // undropped probabilities
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float rs = e < 2 ? rs0 : rs1, ca = colA[j][e & 1];
        const float p = NQ ? ex2_approx(fmaf(s[j][e], c, rs) - ca) : ex2_approx(fmaf(s[j][e], c, ca) - rs);
        bool keep = true;
        if (DROP) {
          if (NQ) {
            const int q = 8 * j + 2 * t + (e & 1);
            const uint32_t wq = q < Lq ? *reinterpret_cast<const uint16_t*>(&sBits[q * mask_ld + 2 * blk]) : 0u;
            keep = (wq >> (g + (e < 2 ? 0 : 8))) & 1u;
          } else {
            keep = ((e < 2 ? w0 : w1) >> (8 * j + 2 * t + (e & 1))) & 1u;
          }
        }
        pr[j][e] = p;
        s[j][e] = keep ? p * ks : 0.f;
        dp[j][e] = keep ? dp[j][e] * ks : 0.f;
      }
    }
    // delta: rowsum(dO * O) = rowsum(P' * dP) when the whole key row is in this block (keys narrow); else from the delta kernel
    float dl0 = 0.f, dl1 = 0.f;
    if (!NQ) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dl0 += pr[j][0] * dp[j][0] + pr[j][1] * dp[j][1];
        dl1 += pr[j][2] * dp[j][2] + pr[j][3] * dp[j][3];
      }
      dl0 = quad_sum(dl0);
      dl1 = quad_sum(dl1);
      if (t == 0) {  // kept for callers that inspect the workspace (same values the delta kernel would produce)
        if (r0 + g < Lq) P.delta[sbase + r0 + g] = dl0;
        if (r0 + g + 8 < Lq) P.delta[sbase + r0 + g + 8] = dl1;
      }
    }
    // dS (into dp)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float dl = NQ ? colB[j][e & 1] : (e < 2 ? dl0 : dl1);
        dp[j][e] = pr[j][e] * (dp[j][e] - dl) * scale;
      }
    // bf16 fragments: as A operand [row][col] (direct) and transposed [col][row] (movmatrix)
    uint32_t pa[2][4], da[2][4];  // k-step kk covers columns 16kk .. 16kk+15
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      pa[kk][0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[kk][1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[kk][2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[kk][3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
      da[kk][0] = pack_bf16x2(dp[2 * kk][0], dp[2 * kk][1]);
      da[kk][1] = pack_bf16x2(dp[2 * kk][2], dp[2 * kk][3]);
      da[kk][2] = pack_bf16x2(dp[2 * kk + 1][0], dp[2 * kk + 1][1]);
      da[kk][3] = pack_bf16x2(dp[2 * kk + 1][2], dp[2 * kk + 1][3]);
    }
    // ---- narrow-side gradients: (this block)^T x (long-side rows of this block), M = 32 narrow rows, K = 16, N = 64
    // A fragment of m-tile i: {T(rows 0-7, n-tile 2i)^T, T(rows 0-7, n-tile 2i+1)^T, T(rows 8-15, 2i)^T, T(rows 8-15, 2i+1)^T}
    // and n-tiles 2i, 2i+1 are exactly k-step i of the direct fragments: {x[i][0], x[i][2], x[i][1], x[i][3]} transposed.
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      uint32_t dt[4] = {movm_t(da[i][0]), movm_t(da[i][2]), movm_t(da[i][1]), movm_t(da[i][3])};
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t bb[4];
        load_b_kn(bb, x1, jp, 0, lane);  // !NQ: dK += dS^T Q; NQ: dQ += dS K
        mma16816(acc[NACC - 1][i][2 * jp], dt, bb[0], bb[1]);
        mma16816(acc[NACC - 1][i][2 * jp + 1], dt, bb[2], bb[3]);
      }
      if (!NQ) {
        uint32_t pt[4] = {movm_t(pa[i][0]), movm_t(pa[i][2]), movm_t(pa[i][1]), movm_t(pa[i][3])};
#pragma unroll
        for (int jp = 0; jp < 4; ++jp) {
          uint32_t bb[4];
          load_b_kn(bb, x2, jp, 0, lane);  // dV += P'^T dO
          mma16816(acc[0][i][2 * jp], pt, bb[0], bb[1]);
          mma16816(acc[0][i][2 * jp + 1], pt, bb[2], bb[3]);
        }
      }
    }
    // ---- long-side gradients of this block: [16 x 32] x narrow operand [32 x 64]
    {
      float r1[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) r1[j][0] = r1[j][1] = r1[j][2] = r1[j][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int jp = 0; jp < 4; ++jp) {
          uint32_t bb[4];
          load_b_kn(bb, sY1, jp, kk, lane);  // !NQ: dQ = dS K; NQ: dK = dS^T Q
          mma16816(r1[2 * jp], da[kk], bb[0], bb[1]);
          mma16816(r1[2 * jp + 1], da[kk], bb[2], bb[3]);
        }
      if (NQ) store_block(r1, 1.f, 1.f, x1, (bf16*)P.dk + (int64_t)b * Lk * P.lddk + h * D, P.lddk, r0, Lk, lane);
      else store_block(r1, 1.f, 1.f, x1, (bf16*)P.dq + (int64_t)b * Lq * P.lddq + h * D, P.lddq, r0, Lq, lane);
    }
    if (NQ) {
      float r2[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) r2[j][0] = r2[j][1] = r2[j][2] = r2[j][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int jp = 0; jp < 4; ++jp) {
          uint32_t bb[4];
          load_b_kn(bb, sY2, jp, kk, lane);  // dV = P'^T dO
          mma16816(r2[2 * jp], pa[kk], bb[0], bb[1]);
          mma16816(r2[2 * jp + 1], pa[kk], bb[2], bb[3]);
        }
      store_block(r2, 1.f, 1.f, x2, (bf16*)P.dv + (int64_t)b * Lk * P.lddv + h * D, P.lddv, r0, Lk, lane);
    }
  }
  // ---- cross-warp reduction of the narrow-side gradients: red[warp][a][32][RLD] fp32 over the X region
  __syncthreads();
  float* red = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    float* pr_ = red + ((warp * NACC + a) * NS) * RLD;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        *reinterpret_cast<float2*>(&pr_[(16 * i + g) * RLD + 8 * j + 2 * t]) = make_float2(acc[a][i][j][0], acc[a][i][j][1]);
        *reinterpret_cast<float2*>(&pr_[(16 * i + g + 8) * RLD + 8 * j + 2 * t]) = make_float2(acc[a][i][j][2], acc[a][i][j][3]);
      }
  }
  __syncthreads();
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    bf16* dst;
    int64_t ld;
    if (NQ) { dst = (bf16*)P.dq + (int64_t)b * Lq * P.lddq + h * D; ld = P.lddq; }
    else if (a == 0) { dst = (bf16*)P.dv + (int64_t)b * Lk * P.lddv + h * D; ld = P.lddv; }
    else { dst = (bf16*)P.dk + (int64_t)b * Lk * P.lddk + h * D; ld = P.lddk; }
    for (int i = tid; i < NS * 8; i += NT) {
      const int r = i >> 3, d0 = (i & 7) * 8;
      if (r >= Ly) continue;
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const float* src = red + ((w * NACC + a) * NS + r) * RLD + d0;
        const float4 x = *reinterpret_cast<const float4*>(src), y = *reinterpret_cast<const float4*>(src + 4);
        v[0] += x.x; v[1] += x.y; v[2] += x.z; v[3] += x.w; v[4] += y.x; v[5] += y.y; v[6] += y.z; v[7] += y.w;
      }
      uint4 w4;
      w4.x = pack_bf16x2(v[0], v[1]); w4.y = pack_bf16x2(v[2], v[3]); w4.z = pack_bf16x2(v[4], v[5]); w4.w = pack_bf16x2(v[6], v[7]);
      *reinterpret_cast<uint4*>(dst + (int64_t)r * ld + d0) = w4;
    }
  }
}

bool aligned(const m3ae_attn_params_t* p, bool bwd) {
  const int64_t lds[] = {p->ldq, p->ldk, p->ldv, p->ldo};
  for (int64_t l : lds) if (l % 8) return false;
  const void* ptrs[] = {p->q, p->k, p->v, p->out};
  for (const void* q : ptrs) if (((uintptr_t)q) & 15) return false;
  if (bwd) {
    if (p->lddq % 8 || p->lddk % 8 || p->lddv % 8) return false;
    const void* g[] = {p->dout, p->dq, p->dk, p->dv};
    for (const void* q : g) if (((uintptr_t)q) & 15) return false;
  }
  if (p->drop_p > 0.f && p->drop_mask && (((uintptr_t)p->drop_mask) & 7)) return false;
  return true;
}

template <typename K>
int opt_in_smem(K kernel, int bytes) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  M3AE_CHECK(e == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(attention_narrow) failed: %s", cudaGetErrorString(e));
  return M3AE_OK;
}
constexpr int SMEM_MAX = 227 * 1024;

}  // namespace

// M3AE_ERR_UNSUPPORTED unless one side is <= 32 tokens (bf16, head dim 64, no causal mask, everything fits in smem)
// `force` (test hook, M3AE_ATTN_NARROW=2): every covered shape.  Otherwise the forward takes only the shapes where it
// measured faster than the mma.sync tiles (profiles/r2_attn_narrow_ab.txt): keys narrow and more than 32 queries (325 x 32:
// 33.5 vs 37.9 us with dropout); 32 x 325 (33.0 vs 27.3 us) and 32 x 32 (7.8 vs 5.0 us) stay on the mma.sync tiles.
int attn_fwd_narrow(const m3ae_attn_params_t* p, bool force, cudaStream_t st) {
  if (p->dh != D || p->dtype != M3AE_BF16 || p->causal) return M3AE_ERR_UNSUPPORTED;
  if (p->Lq > NS && p->Lk > NS) return M3AE_ERR_UNSUPPORTED;
  if (!force && (p->Lk > NS || p->Lq <= NS)) return M3AE_ERR_UNSUPPORTED;
  if (!aligned(p, false)) return M3AE_ERR_UNSUPPORTED;
  const int smem = fwd_smem_bytes(p->Lq, p->Lk);
  if (smem > SMEM_MAX) return M3AE_ERR_UNSUPPORTED;
  int dev = 0;
  M3AE_CUDA(cudaGetDevice(&dev));
  static int done[64];
  const bool drop = p->drop_p > 0.f, nq = p->Lk > NS;
  if (dev < 64 && !done[dev]) {
    int rc = opt_in_smem(attn_fwd_narrow_kernel<false, false>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_fwd_narrow_kernel<false, true>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_fwd_narrow_kernel<true, false>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_fwd_narrow_kernel<true, true>, SMEM_MAX);
    if (rc) return rc;
    done[dev] = 1;
  }
  dim3 grid(p->H, p->B);
  if (nq) {
    if (drop) M3AE_LAUNCH((attn_fwd_narrow_kernel<true, true>), grid, NT, smem, st, *p);
    else M3AE_LAUNCH((attn_fwd_narrow_kernel<true, false>), grid, NT, smem, st, *p);
  } else {
    if (drop) M3AE_LAUNCH((attn_fwd_narrow_kernel<false, true>), grid, NT, smem, st, *p);
    else M3AE_LAUNCH((attn_fwd_narrow_kernel<false, false>), grid, NT, smem, st, *p);
  }
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int attn_bwd_narrow(const m3ae_attn_params_t* p, cudaStream_t st) {
  if (p->dh != D || p->dtype != M3AE_BF16 || p->causal) return M3AE_ERR_UNSUPPORTED;
  if (p->Lq > NS && p->Lk > NS) return M3AE_ERR_UNSUPPORTED;
  if (!aligned(p, true)) return M3AE_ERR_UNSUPPORTED;
  const bool nq = p->Lk > NS;  // queries are the narrow side (keys are when both fit)
  const int smem = bwd_smem_bytes(p->Lq, p->Lk, nq);
  if (smem > SMEM_MAX) return M3AE_ERR_UNSUPPORTED;
  int dev = 0;
  M3AE_CUDA(cudaGetDevice(&dev));
  static int done[64];
  if (dev < 64 && !done[dev]) {
    int rc = opt_in_smem(attn_bwd_narrow_kernel<false, false>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_bwd_narrow_kernel<false, true>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_bwd_narrow_kernel<true, false>, SMEM_MAX);
    if (!rc) rc = opt_in_smem(attn_bwd_narrow_kernel<true, true>, SMEM_MAX);
    if (rc) return rc;
    done[dev] = 1;
  }
  const bool drop = p->drop_p > 0.f;
  dim3 grid(p->H, p->B);
  if (nq) {
    launch_delta_bf16_64(p, st);  // keys span several blocks: delta is needed before the first dS
    if (drop) M3AE_LAUNCH((attn_bwd_narrow_kernel<true, true>), grid, NT, smem, st, *p);
    else M3AE_LAUNCH((attn_bwd_narrow_kernel<true, false>), grid, NT, smem, st, *p);
  } else {
    if (drop) M3AE_LAUNCH((attn_bwd_narrow_kernel<false, true>), grid, NT, smem, st, *p);
    else M3AE_LAUNCH((attn_bwd_narrow_kernel<false, false>), grid, NT, smem, st, *p);
  }
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
