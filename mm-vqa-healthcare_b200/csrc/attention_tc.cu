// tcgen05 / TMEM flash attention for the fusion stack's bf16 path (12 heads x 64, Lq, Lk <= 384: 325 image tokens at
// 288 px, 32 text tokens).  Reference arithmetic: bert_model.py:301-344 (BertSelfAttention core, self and cross mode).
//
// One CTA = one (batch, head).  One control warp issues every TMA and tcgen05.mma; the element-wise (EW) warps own one
// accumulator row per thread.  Both kernels pull the score tile out of TMEM into REGISTERS in one burst and release the
// TMEM columns at once, so the tensor pipe computes the next tile's scores while the EW warps are still working on this
// one (round 1's first version read TMEM piecemeal and serialised MMA -> softmax -> MMA: 28-42 % issue utilisation,
// 40 % of the warp samples parked on the MMA barrier — profiles/r1_ncu_attn_v5.txt).
//
//   forward : 16 EW warps (4 TMEM lane quadrants x 4 column quarters).  Per 128-query tile
//               S = Q K^T (whole key row, <= 384 fp32 columns in TMEM) -> registers (<= 96 per thread), TMEM released;
//               row max / exp2 / row sum / dropout entirely in registers, P (bf16) -> swizzled smem;
//               O = P V into one of two TMEM accumulators; O/l is read out one tile later, behind the next P.
//   backward: 8 EW warps (4 quadrants x 2 column halves), key tiles of 64 (outer) x query tiles of 128 (inner).
//               S = Q K^T, dP = dO V^T -> registers (32 + 32 per thread), TMEM released;
//               P = exp2(S - lse), dS = P (dP*keep - delta) scale -> double-buffered smem tiles;
//               dV += P'^T dO, dK += dS^T Q, dQ += dS K.  dQ of all <= 3 query tiles, dK, dV, S, dP = 448 TMEM columns:
//             one kernel, every P evaluated once, no atomics.  Q and dO stay resident in smem, K/V stream through a ring.
//
// Algorithmic work per (b,h): forward 4*Lq*Lk*64 flop, backward 10*Lq*Lk*64 flop; HBM traffic = Q,K,V,O (+dO,dQ,dK,dV)
// once: the scores [B,12,Lq,Lk] (324 MB per image layer at B=64 in the reference) never leave the SM.
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace m3ae {
void count_launch();
void launch_delta_bf16_64(const m3ae_attn_params_t* p, cudaStream_t st);  // attention.cu

#ifdef M3AE_ATTN_TRACE
__device__ long long g_attn_trace[512];
#define TR(i) do { if (blockIdx.x == 0 && blockIdx.y == 0) g_attn_trace[(i)] = clock64(); } while (0)
#else
#define TR(i) do { } while (0)
#endif

namespace {

constexpr int D = 64;            // head dim
constexpr int TQ = 128;          // query rows per tile (UMMA M)
constexpr int KT = 64;           // backward key tile
constexpr int BLK = TQ * 128;    // one [128 rows][64 bf16] swizzled block (16 KB)
constexpr int KBLK = KT * 128;   // one backward K / V tile (8 KB)
constexpr int FWD_EW = 16;       // forward element-wise warps
constexpr int BWD_EW = 16;       // backward element-wise warps (4 TMEM lane quadrants x 4 groups of 16 key columns)
constexpr int NT_FWD = (FWD_EW + 1) * 32;
constexpr int NT_BWD = (BWD_EW + 1) * 32;
constexpr int HELD = 6;          // forward: 8-key units of a thread's row share that stay in registers (the rest: TMEM, read twice)
constexpr int SMEM_LIMIT = 232448;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct TcAttn {
  CUtensorMap map_q, map_k, map_v, map_do;  // 2-D [B*L, H*64] bf16 views; box = 64 columns x box rows, SWIZZLE_128B
  m3ae_attn_params_t p;
  int lk_pad;          // Lk rounded up to 16 (UMMA N / K granularity)
  int nbox_k, box_k;   // forward: K and V are staged as nbox_k TMA boxes of box_k rows; backward: box_k = 64
  int box_q;           // rows per Q / dO tile load (<= 128)
  int n_qt;            // number of 128-row query tiles
  int n_kt, nkv;       // backward: key tiles of 64, K/V ring depth
  int mask_ld;         // bytes per row of the dropout keep-bit cache
  int tmem_cols;       // power of two >= 32
  int off_k, off_v, off_q, off_do, off_p, off_ds, off_mask, off_red, off_bits, off_bar;  // smem offsets
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
// The control thread is ONE lane: every integer instruction it spends per tcgen05.mma is ~4-5 cycles of issue latency, and
// the 128xNx16 MMAs of this kernel (N = 64 mostly) only run 32-90 cycles.  So descriptors are kept as {lo, hi} words: hi
// (SBO = 1024, version, SWIZZLE_128B) is a constant, lo = address field | LBO field << 16, and stepping the operand by
// `bytes` is one 32-bit add of bytes >> 4 (addresses stay inside the CTA's shared window, so the 14-bit field never carries).
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint32_t dlo(uint32_t saddr, uint32_t lbo_bytes) { return ((saddr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16); }
// D[tmem] = A*B (first) / D += A*B (acc); the accumulate flag is an immediate.  `el` predicates the instruction: the whole
// control WARP runs the (convergent) control flow so that addresses and descriptors live in uniform registers, and only the
// tcgen05.mma itself is restricted to one lane.  (Issuing from inside an `if (lane == 0)` region makes the compiler move
// every operand into uniform registers through an ELECT / R2UR.BROADCAST waterfall loop, ~130 cycles per MMA.)
// value known to be warp-uniform -> tell the compiler (it then keeps it, and everything derived from it, in uniform registers)
__device__ __forceinline__ uint32_t uni(uint32_t x) { return __shfl_sync(0xffffffffu, x, 0); }
template <bool ACC>
__device__ __forceinline__ void umma_lo(uint32_t el, uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc) {
  (void)el;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b64 da, db;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "mov.b64 da, {%1, %5};\n"
      "mov.b64 db, {%2, %5};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(idesc), "n"(ACC ? 1 : 0), "r"(DESC_HI));
}
__device__ __forceinline__ void umma_lo_p(uint32_t el, uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t acc) {
  (void)el;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b64 da, db;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "mov.b64 da, {%1, %5};\n"
      "mov.b64 db, {%2, %5};\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(acc), "r"(DESC_HI));
}
__device__ __forceinline__ void umma_commit_el(uint32_t el, uint64_t* bar) {
  (void)el;
  umma_commit(bar);
}
// kind::f16 instruction descriptor: D fp32, A/B bf16, M = 128 (or 64: the accumulator then sits in lanes 0-15 of each
// 32-lane TMEM sub-partition, row i -> lane (i % 16) + 32 * (i / 16), and a second M = 64 accumulator can share its
// columns at lane offset 16).
__device__ __forceinline__ uint32_t idesc(int n, int a_mn, int b_mn, int m = TQ) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void tmem_alloc_n(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_n(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void named_bar(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// 16 fp32 accumulator values of one row -> bf16 -> 32 contiguous bytes in global memory
__device__ __forceinline__ void store_row16(bf16* dst, const uint32_t (&r)[16], float mul) {
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint4 w;
    w.x = pack_bf16x2(__uint_as_float(r[8 * j + 0]) * mul, __uint_as_float(r[8 * j + 1]) * mul);
    w.y = pack_bf16x2(__uint_as_float(r[8 * j + 2]) * mul, __uint_as_float(r[8 * j + 3]) * mul);
    w.z = pack_bf16x2(__uint_as_float(r[8 * j + 4]) * mul, __uint_as_float(r[8 * j + 5]) * mul);
    w.w = pack_bf16x2(__uint_as_float(r[8 * j + 6]) * mul, __uint_as_float(r[8 * j + 7]) * mul);
    reinterpret_cast<uint4*>(dst)[j] = w;
  }
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

// shared-memory accesses by 32-bit shared-space address (generic pointers cost 64-bit address arithmetic per access)
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void sts8(uint32_t addr, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts32f(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
__device__ __forceinline__ float lds32f(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
// bits 0 / 1 of b -> 0xFFFF in the low / high half of a bf16x2 lane mask
__device__ __forceinline__ uint32_t pair_mask(uint32_t b) { return (b & 1u) * 0xFFFFu + (b & 2u) * 0x7FFF8000u; }
// lane mask of a flag word (bit 15 / bit 31 -> 0xFFFF in the low / high half)
__device__ __forceinline__ uint32_t flag_lanes(uint32_t f) { return (f >> 15) * 0xFFFFu; }

// Row maximum (log2 domain) of one 8-key unit.  plain: no additive mask and no padded key in the unit.
__device__ __forceinline__ float fwd_unit_max(const uint32_t (&v)[8], bool plain, uint32_t mask_addr, float c) {
  if (plain) {
    float m = fmaxf(__uint_as_float(v[0]), __uint_as_float(v[1]));
#pragma unroll
    for (int j = 2; j < 8; ++j) m = fmaxf(m, __uint_as_float(v[j]));
    return m * c;
  }
  const float4 ma = lds128f(mask_addr), mb = lds128f(mask_addr + 16);
  const float mk[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
  float m = fmaf(__uint_as_float(v[0]), c, mk[0]);
#pragma unroll
  for (int j = 1; j < 8; ++j) m = fmaxf(m, fmaf(__uint_as_float(v[j]), c, mk[j]));
  return m;
}
// exp2, dropout, bf16 pack and the swizzled 16-byte store of one 8-key unit; returns the unit's (pre-dropout) row sum.
template <bool DROP>
__device__ __forceinline__ float fwd_unit_exp(const uint32_t (&v)[8], bool plain, uint32_t mask_addr, float c, float nmx, uint32_t p_addr,
                                              uint32_t bits_addr, const m3ae_attn_params_t& P, uint64_t row_lin, int u, uint32_t lane_add,
                                              bool ready) {
  float pv[8];
  if (plain) {
#pragma unroll
    for (int j = 0; j < 8; ++j) pv[j] = ex2_approx(fmaf(__uint_as_float(v[j]), c, nmx));
  } else {
    const float4 ma = lds128f(mask_addr), mb = lds128f(mask_addr + 16);
    const float mk[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) pv[j] = ex2_approx(fmaf(__uint_as_float(v[j]), c, mk[j] + nmx));
  }
  const float sum = ((pv[0] + pv[1]) + (pv[2] + pv[3])) + ((pv[4] + pv[5]) + (pv[6] + pv[7]));
  uint32_t w0 = pack_bf16x2(pv[0], pv[1]), w1 = pack_bf16x2(pv[2], pv[3]), w2 = pack_bf16x2(pv[4], pv[5]), w3 = pack_bf16x2(pv[6], pv[7]);
  if (DROP) {
    if (ready) {
      // keep bits staged from the pre-generated cache (m3ae_dropout_bits): bit j of the byte = key 8u+j kept
      const uint32_t b = lds8(bits_addr);
      w0 &= pair_mask(b); w1 &= pair_mask(b >> 2); w2 &= pair_mask(b >> 4); w3 &= pair_mask(b >> 6);
    } else {
      uint32_t f[4];
      attn_keep_flags(P.drop_seed, P.drop_stream, row_lin, u, lane_add, f);
      w0 &= flag_lanes(f[0]); w1 &= flag_lanes(f[1]); w2 &= flag_lanes(f[2]); w3 &= flag_lanes(f[3]);
      sts8(bits_addr, attn_flags_to_byte(f));
    }
  }
  sts128(p_addr, w0, w1, w2, w3);
  return sum;
}

// ------------------------------------------------------------------------------------------------ forward
// TMEM columns: O[0] at [0, 64), O[1] at [64, 128), S at [128, 128 + lk_pad).
// smem: K | V (resident) | Q x2 | P (ceil(lk_pad/64) blocks) | mask | row reductions | keep-bit tile | barriers.
// F_OE0/1: one per O accumulator, so a waiter can never be lapped by two phases (parity waits alias after two).
enum { F_KV = 0, F_QF0 = 1, F_QF1 = 2, F_SF = 3, F_SE = 4, F_PF = 5, F_OF = 6, F_OE0 = 7, F_OE1 = 8, F_NBAR = 9 };

template <bool DROP>
__global__ void __launch_bounds__(NT_FWD, 1) attn_fwd_tc_kernel(const __grid_constant__ TcAttn A) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const m3ae_attn_params_t& P = A.p;
  uint8_t* sK = smem + A.off_k;
  uint8_t* sV = smem + A.off_v;
  uint8_t* sQ = smem + A.off_q;
  uint8_t* sP = smem + A.off_p;
  float* sMask = (float*)(smem + A.off_mask);
  float* sRedM = (float*)(smem + A.off_red);   // [4 column quarters][128 rows] row max
  float* sRedS = sRedM + 4 * TQ;               // [4][128] row sum
  uint8_t* sBits = smem + A.off_bits;          // [128 rows][mask_ld] keep bits of the current tile
  uint64_t* bars = (uint64_t*)(smem + A.off_bar);
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + F_NBAR);
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int Lq = P.Lq, Lk = P.Lk, lk_pad = A.lk_pad, n_qt = A.n_qt;
  int mask_nz = 0;

  if (warp == FWD_EW) {
    if (lane == 0) {
      tma_prefetch_desc(&A.map_q);
      tma_prefetch_desc(&A.map_k);
      tma_prefetch_desc(&A.map_v);
      mbar_init(&bars[F_KV], 1);
      mbar_init(&bars[F_QF0], 1);
      mbar_init(&bars[F_QF1], 1);
      mbar_init(&bars[F_SF], 1);
      mbar_init(&bars[F_SE], FWD_EW);
      mbar_init(&bars[F_PF], FWD_EW);
      mbar_init(&bars[F_OF], 1);
      mbar_init(&bars[F_OE0], FWD_EW);
      mbar_init(&bars[F_OE1], FWD_EW);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_n(tmem_ptr_smem, (uint32_t)A.tmem_cols);
  } else {
    const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
    for (int i = threadIdx.x; i < 384; i += FWD_EW * 32) {
      const float mv = (i < Lk && mrow) ? mrow[i] : 0.f;
      mask_nz |= mv != 0.f;
      sMask[i] = i < Lk ? mv * LOG2E : -INFINITY;
    }
  }
  tcgen05_fence_before();
  // An additive mask that is all zeros for this batch row (the image stream's all-valid mask, m3ae_module.py:253-256) is
  // treated as no mask: the element-wise warps then skip two shared-memory loads and an add per 8 keys in both passes.
  const bool any_mask = __syncthreads_or(mask_nz) != 0;
  tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr_smem, 0);
  const uint32_t tS = tmem_base + 128;

  if (warp == FWD_EW) {
    {
      // ===================== control warp: TMA + MMA issue (lane 0 issues; all lanes run the control flow) =====================
      const uint32_t el = lane == 0 ? 1u : 0u;
      const int col0 = h * D;
      if (el) {
        mbar_expect_tx(&bars[F_KV], (uint32_t)(2 * A.nbox_k * A.box_k * 128));
        for (int i = 0; i < A.nbox_k; ++i) {
          tma_load_2d(sK + i * A.box_k * 128, &A.map_k, &bars[F_KV], col0, b * Lk + i * A.box_k);
          tma_load_2d(sV + i * A.box_k * 128, &A.map_v, &bars[F_KV], col0, b * Lk + i * A.box_k);
        }
        mbar_expect_tx(&bars[F_QF0], (uint32_t)(A.box_q * 128));
        tma_load_2d(sQ, &A.map_q, &bars[F_QF0], col0, b * Lq);
        if (n_qt > 1) {
          mbar_expect_tx(&bars[F_QF1], (uint32_t)(A.box_q * 128));
          tma_load_2d(sQ + BLK, &A.map_q, &bars[F_QF1], col0, b * Lq + TQ);
        }
      }
      // S is issued in one or two N chunks (UMMA N <= 256)
      const int n1 = lk_pad <= 256 ? lk_pad : ((lk_pad / 2 + 15) & ~15);
      const int n2 = lk_pad - n1;
      const uint32_t id_s1 = idesc(n1, 0, 0), id_s2 = idesc(n2 > 0 ? n2 : 16, 0, 0), id_o = idesc(D, 0, 1);
      const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;  // re-derived: see the backward kernel
      const uint32_t aK = sb + (uint32_t)A.off_k, aV = sb + (uint32_t)A.off_v, aP = sb + (uint32_t)A.off_p;
      const uint32_t k_lo = dlo(aK, 16), k2_lo = dlo(aK + n1 * 128, 16), v_lo = dlo(aV, 8192), p_lo = dlo(aP, 16);
      const uint32_t q_lo0 = dlo(sb + (uint32_t)A.off_q, 16);
      auto issue_s = [&](int t) {
        const uint32_t q_lo = q_lo0 + (uint32_t)(t & 1) * (BLK >> 4);
        if (!elect_one()) return;  // one lane issues the whole batch (operand set-up is then scheduled ahead of the MMAs)
        umma_lo<false>(el, tS, q_lo, k_lo, id_s1);
        umma_lo<true>(el, tS, q_lo + 2, k_lo + 2, id_s1);
        umma_lo<true>(el, tS, q_lo + 4, k_lo + 4, id_s1);
        umma_lo<true>(el, tS, q_lo + 6, k_lo + 6, id_s1);
        if (n2 > 0) {
          umma_lo<false>(el, tS + n1, q_lo, k2_lo, id_s2);
          umma_lo<true>(el, tS + n1, q_lo + 2, k2_lo + 2, id_s2);
          umma_lo<true>(el, tS + n1, q_lo + 4, k2_lo + 4, id_s2);
          umma_lo<true>(el, tS + n1, q_lo + 6, k2_lo + 6, id_s2);
        }
        umma_commit_el(el, &bars[F_SF]);
      };
      TR(0);
      mbar_wait(&bars[F_KV], 0);
      mbar_wait(&bars[F_QF0], 0);
      tcgen05_fence_after();
      TR(1);
      issue_s(0);
      const int ksteps = lk_pad / 16;
      for (int t = 0; t < n_qt; ++t) {
        if (t + 1 < n_qt) {
          mbar_wait(&bars[F_QF0 + ((t + 1) & 1)], (uint32_t)(((t + 1) >> 1) & 1));
          mbar_wait(&bars[F_SE], (uint32_t)(t & 1));  // S(t) is in registers: TMEM S and Q buffer t&1 are free
          tcgen05_fence_after();
          TR(10 + 8 * t);
          issue_s(t + 1);
          if (t + 2 < n_qt && el) {
            mbar_expect_tx(&bars[F_QF0 + (t & 1)], (uint32_t)(A.box_q * 128));
            tma_load_2d(sQ + (t & 1) * BLK, &A.map_q, &bars[F_QF0 + (t & 1)], col0, b * Lq + (t + 2) * TQ);
          }
        }
        mbar_wait(&bars[F_PF], (uint32_t)(t & 1));
        TR(11 + 8 * t);
        if (t >= 2) mbar_wait(&bars[F_OE0 + (t & 1)], (uint32_t)(((t - 2) >> 1) & 1));  // O(t-2) left this accumulator
        tcgen05_fence_after();
        const uint32_t tO = tmem_base + (uint32_t)((t & 1) * D);
        // O[q][d] = P[q][key] V[key][d]: A K-major (64-key blocks of BLK bytes, 32 B per 16-key step), B MN-major (2048 B per step)
#pragma unroll 1
        uint32_t pb = p_lo, vb = v_lo;
        if (elect_one()) {
        for (int blk = 0, left = ksteps; left > 0; ++blk, left -= 4, pb += BLK >> 4, vb += 512) {
          if (blk == 0) umma_lo<false>(el, tO, pb, vb, id_o);
          else umma_lo<true>(el, tO, pb, vb, id_o);
          if (left > 1) umma_lo<true>(el, tO, pb + 2, vb + 128, id_o);
          if (left > 2) umma_lo<true>(el, tO, pb + 4, vb + 256, id_o);
          if (left > 3) umma_lo<true>(el, tO, pb + 6, vb + 384, id_o);
        }
        umma_commit_el(el, &bars[F_OF]);
        }
        __syncwarp();
        TR(12 + 8 * t);
      }
    }
  } else {
    // ===================== element-wise warps =====================
    const int quad = warp & 3, cq = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int n8 = lk_pad >> 3;
    const int ub = (n8 * cq) >> 2, nu = ((n8 * (cq + 1)) >> 2) - ub;  // this warp's 8-key units [ub, ub + nu)
    const bool has_mask = any_mask;
    const float c = A.scale_l2;
    const float keep_scale = drop_keep_scale(P.drop_p);
    const uint32_t thr = drop_thr15(P.drop_p);
    const int mask_ld = A.mask_ld;
    const int64_t sbase = ((int64_t)b * P.H + h) * Lq;
    float l_prev = 1.f;

    // O(j)/l -> global; every EW warp arrives on F_OE[j&1] once per tile whether or not it holds live rows
    auto readout = [&](int j, float l) {
      const int qrow = j * TQ + row;
      const bool act = j * TQ + quad * 32 < Lq;
      uint32_t o[16];
      if (act) {
        tmem_ld16(tmem_base + (uint32_t)((j & 1) * D) + lane_addr + cq * 16, o);
        tmem_ld_wait();
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[F_OE0 + (j & 1)]);
      if (act && qrow < Lq) {
        const float inv = l > 0.f ? keep_scale / l : 0.f;
        store_row16((bf16*)P.out + ((int64_t)b * Lq + qrow) * P.ldo + h * D + cq * 16, o, inv);
      }
    };

    const uint32_t lane_add = drop_lane_add(thr);
    const bool bits_ready = DROP && P.drop_mask != nullptr && P.drop_mask_ready != 0;
    const int per_row = mask_ld >> 3;  // 8-byte words per row of the keep-bit cache
    const uint32_t a_mask = smem_u32(sMask), a_redm = smem_u32(sRedM) + row * 4, a_reds = smem_u32(sRedS) + row * 4;
    const uint32_t a_prow = smem_u32(sP) + row * 128, a_bits = smem_u32(sBits) + row * mask_ld;
    const uint32_t swz = ((uint32_t)row & 7u) << 4;
    for (int t = 0; t < n_qt; ++t) {
      const int qrow = t * TQ + row;
      const bool active = t * TQ + quad * 32 < Lq;  // uniform over the 4 warps of a quadrant
      const uint64_t row_lin = (uint64_t)(sbase + qrow);
      // The first HELD units of the row stay in registers from here on; the rest are read from TMEM twice (max pass,
      // exp pass) and go FIRST in the exp pass, so the S columns are released roughly half-way through the tile.
      uint32_t s[HELD][8];
      // pre-generated keep bits of this tile: the quadrant's 32 rows x mask_ld bytes are contiguous in the global cache;
      // issue the (<= 2 per thread) 8-byte loads now, park them in shared memory before the max-exchange barrier below
      uint2 pre0 = make_uint2(0u, 0u), pre1 = make_uint2(0u, 0u);
      const int bits_e0 = cq * 32 + lane;
      const int bits_lim = (Lq - (t * TQ + quad * 32)) * per_row;  // words of rows < Lq (may exceed 32 * per_row)
      if (bits_ready && active) {
        const uint2* src = reinterpret_cast<const uint2*>(P.drop_mask + (sbase + t * TQ + quad * 32) * mask_ld);
        if (bits_e0 < 32 * per_row && bits_e0 < bits_lim) pre0 = __ldg(src + bits_e0);
        if (bits_e0 + 128 < 32 * per_row && bits_e0 + 128 < bits_lim) pre1 = __ldg(src + bits_e0 + 128);
      }
      mbar_wait(&bars[F_SF], (uint32_t)(t & 1));
      tcgen05_fence_after();
      if (threadIdx.x == 0) TR(40 + 10 * t);
      float l = 1.f, mx = -INFINITY, sum = 0.f;
      if (active) {
#pragma unroll
        for (int i = 0; i < HELD; ++i)
          if (i < nu) tmem_ld8(tS + lane_addr + (ub + i) * 8, s[i]);
        tmem_ld_wait();
        if (threadIdx.x == 0) TR(41 + 10 * t);
        // ---- pass 1: row max
#pragma unroll
        for (int i = 0; i < HELD; ++i) {
          if (i < nu) {
            const int k0 = (ub + i) * 8;
            mx = fmaxf(mx, fwd_unit_max(s[i], !has_mask && k0 + 8 <= Lk, a_mask + k0 * 4, c));
          }
        }
#pragma unroll 1
        for (int i = HELD; i < nu; ++i) {
          uint32_t v[8];
          const int k0 = (ub + i) * 8;
          tmem_ld8(tS + lane_addr + k0, v);
          tmem_ld_wait();
          mx = fmaxf(mx, fwd_unit_max(v, !has_mask && k0 + 8 <= Lk, a_mask + k0 * 4, c));
        }
        sts32f(a_redm + cq * (TQ * 4), mx);
        if (bits_ready) {  // every warp of the quadrant finished the previous tile's exp pass (its sum-exchange barrier)
          uint2* dst = reinterpret_cast<uint2*>(sBits + quad * 32 * mask_ld);
          if (bits_e0 < 32 * per_row) dst[bits_e0] = pre0;
          if (bits_e0 + 128 < 32 * per_row) dst[bits_e0 + 128] = pre1;
        }
        if (threadIdx.x == 0) TR(42 + 10 * t);
        named_bar(1 + quad, 128);
        if (threadIdx.x == 0) TR(43 + 10 * t);
        mx = fmaxf(fmaxf(lds32f(a_redm), lds32f(a_redm + TQ * 4)), fmaxf(lds32f(a_redm + 2 * TQ * 4), lds32f(a_redm + 3 * TQ * 4)));
        if (mx == -INFINITY) mx = 0.f;
      }
      // P(t-1) must have been consumed by O(t-1) = P(t-1) V before the staging buffer is overwritten
      if (t > 0) {
        mbar_wait(&bars[F_OF], (uint32_t)((t - 1) & 1));
        tcgen05_fence_after();
        readout(t - 1, l_prev);  // now, while the tensor pipe is idle: TMEM reads crawl while the next P V MMAs run
      }
      if (threadIdx.x == 0) TR(44 + 10 * t);
      if (active) {
        // ---- pass 2, TMEM-resident units first
#pragma unroll 1
        for (int i = HELD; i < nu; ++i) {
          uint32_t v[8];
          const int u = ub + i, k0 = u * 8;
          tmem_ld8(tS + lane_addr + k0, v);
          tmem_ld_wait();
          sum += fwd_unit_exp<DROP>(v, !has_mask && k0 + 8 <= Lk, a_mask + k0 * 4, c, -mx,
                                    a_prow + (u >> 3) * BLK + ((((uint32_t)u & 7u) << 4) ^ swz), a_bits + u, P, row_lin, u, lane_add, bits_ready);
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[F_SE]);  // S(t) is no longer needed in TMEM
      if (threadIdx.x == 0) TR(45 + 10 * t);
      if (active) {
#pragma unroll
        for (int i = 0; i < HELD; ++i) {
          if (i < nu) {
            const int u = ub + i, k0 = u * 8;
            sum += fwd_unit_exp<DROP>(s[i], !has_mask && k0 + 8 <= Lk, a_mask + k0 * 4, c, -mx,
                                      a_prow + (u >> 3) * BLK + ((((uint32_t)u & 7u) << 4) ^ swz), a_bits + u, P, row_lin, u, lane_add, bits_ready);
          }
        }
        sts32f(a_reds + cq * (TQ * 4), sum);
      }
      fence_proxy_async();
      __syncwarp();
      if (threadIdx.x == 0) TR(46 + 10 * t);
      if (lane == 0) mbar_arrive(&bars[F_PF]);
      if (threadIdx.x == 0) TR(47 + 10 * t);
      if (active) {
        named_bar(1 + quad, 128);
        l = (lds32f(a_reds) + lds32f(a_reds + TQ * 4)) + (lds32f(a_reds + 2 * TQ * 4) + lds32f(a_reds + 3 * TQ * 4));
        if (cq == 0 && qrow < Lq) P.lse[sbase + qrow] = (mx + log2f(l)) * LN2;  // l >= 1: the row maximum contributes exp2(0)
        if (DROP && P.drop_mask != nullptr && !bits_ready) {
          // the quadrant's 32 rows x mask_ld bytes are contiguous in the global keep-bit cache: coalesced 8-byte copies
          const int r0 = t * TQ + quad * 32;
          const uint2* src = reinterpret_cast<const uint2*>(sBits + quad * 32 * mask_ld);
          uint2* dst = reinterpret_cast<uint2*>(P.drop_mask + (sbase + r0) * mask_ld);
          for (int e = cq * 32 + lane; e < 32 * per_row && e < bits_lim; e += 128) dst[e] = src[e];
        }
      }
      if (threadIdx.x == 0) TR(48 + 10 * t);
      l_prev = l;
    }
    mbar_wait(&bars[F_OF], (uint32_t)((n_qt - 1) & 1));
    tcgen05_fence_after();
    if (threadIdx.x == 0) TR(98);
    readout(n_qt - 1, l_prev);
    if (threadIdx.x == 0) TR(99);
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == FWD_EW) {
    tcgen05_fence_after();
    tmem_dealloc_n(tmem_base, (uint32_t)A.tmem_cols);
  }
}

// One 16-key sub-block of a backward row: P and dS -> bf16 -> two swizzled 16-byte chunks each (chunk0, chunk0+1 of the row).
template <bool DROP>
__device__ __forceinline__ void bwd_sub(const uint32_t (&rs)[16], const uint32_t (&rd)[16], bool plain, uint32_t mask_addr, float c, float nl2,
                                        float dl, float dscale, uint32_t bits, uint32_t p_row, uint32_t ds_row, uint32_t chunk0, uint32_t swz) {
#pragma unroll
  for (int q2 = 0; q2 < 2; ++q2) {
    float pv[8], ds[8];
    if (plain) {
#pragma unroll
      for (int j = 0; j < 8; ++j) pv[j] = ex2_approx(fmaf(__uint_as_float(rs[8 * q2 + j]), c, nl2));
    } else {
      const float4 ma = lds128f(mask_addr + q2 * 32), mb = lds128f(mask_addr + q2 * 32 + 16);
      const float mk[8] = {ma.x, ma.y, ma.z, ma.w, mb.x, mb.y, mb.z, mb.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) pv[j] = ex2_approx(fmaf(__uint_as_float(rs[8 * q2 + j]), c, mk[j] + nl2));
    }
    if (DROP) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool keep = (bits >> (8 * q2 + j)) & 1u;
        const float a = keep ? __uint_as_float(rd[8 * q2 + j]) : 0.f;
        ds[j] = pv[j] * fmaf(a, dscale, -dl);
        pv[j] = keep ? pv[j] : 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) ds[j] = pv[j] * fmaf(__uint_as_float(rd[8 * q2 + j]), dscale, -dl);
    }
    const uint32_t off = ((chunk0 + q2) ^ swz) << 4;
    sts128(p_row + off, pack_bf16x2(pv[0], pv[1]), pack_bf16x2(pv[2], pv[3]), pack_bf16x2(pv[4], pv[5]), pack_bf16x2(pv[6], pv[7]));
    sts128(ds_row + off, pack_bf16x2(ds[0], ds[1]), pack_bf16x2(ds[2], ds[3]), pack_bf16x2(ds[4], ds[5]), pack_bf16x2(ds[6], ds[7]));
  }
}

// ------------------------------------------------------------------------------------------------ backward
// TMEM columns (all 512): dQ tiles at [0, 64*n_qt); dV and dK share [192, 256) as two M = 64 accumulators (dV in lanes
// 0-15 of each sub-partition, dK in lanes 16-31); S/dP double-buffered: S(b) at 256 + 128 b, dP(b) at 320 + 128 b.
// smem: Q[n_qt] | dO[n_qt] (resident) | K ring | V ring (64-key tiles) | P x2 | dS x2 | mask | barriers.
// Iteration it = kt * n_qt + qt (key tiles outer, query tiles inner).  S/dP of iteration it + 2 are issued as soon as the
// element-wise warps have pulled S/dP of iteration it into registers, so the tensor pipe computes them while iteration it
// and it + 1 are being processed: the single-buffered version spent half of every iteration in the chain
// S-drained -> issue -> MMA -> S-full -> tcgen05.ld (4400 cycles per iteration against ~1800 of work on either side,
// profiles/r2_attn_bwd_timeline.txt).
enum { B_KV0 = 0 /* ..3 */, B_QF0 = 4 /* ..6 */, B_SF0 = 7, B_SF1 = 8, B_SE0 = 9, B_SE1 = 10, B_PF0 = 11, B_PF1 = 12, B_MD0 = 13, B_MD1 = 14,
       B_DE = 15, B_NBAR = 16 };
constexpr int MAX_NKV = 4;

template <bool DROP>
__global__ void __launch_bounds__(NT_BWD, 1) attn_bwd_tc_kernel(const __grid_constant__ TcAttn A) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const m3ae_attn_params_t& P = A.p;
  uint8_t* sK = smem + A.off_k;
  uint8_t* sV = smem + A.off_v;
  uint8_t* sQ = smem + A.off_q;
  uint8_t* sdO = smem + A.off_do;
  uint8_t* sP = smem + A.off_p;
  uint8_t* sdS = smem + A.off_ds;
  float* sMask = (float*)(smem + A.off_mask);
  uint64_t* bars = (uint64_t*)(smem + A.off_bar);
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + B_NBAR);
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int Lq = P.Lq, Lk = P.Lk, lk_pad = A.lk_pad, n_qt = A.n_qt, n_kt = A.n_kt, nkv = A.nkv;
  const int total = n_kt * n_qt;
  int mask_nz = 0;
  if (threadIdx.x == 0) TR(100);

  if (warp == BWD_EW) {
    if (lane == 0) {
      tma_prefetch_desc(&A.map_q);
      tma_prefetch_desc(&A.map_k);
      tma_prefetch_desc(&A.map_v);
      tma_prefetch_desc(&A.map_do);
      for (int i = 0; i < MAX_NKV; ++i) mbar_init(&bars[B_KV0 + i], 1);
      for (int i = 0; i < 3; ++i) mbar_init(&bars[B_QF0 + i], 1);
      mbar_init(&bars[B_SF0], 1);
      mbar_init(&bars[B_SF1], 1);
      mbar_init(&bars[B_SE0], BWD_EW);
      mbar_init(&bars[B_SE1], BWD_EW);
      mbar_init(&bars[B_PF0], BWD_EW);
      mbar_init(&bars[B_PF1], BWD_EW);
      mbar_init(&bars[B_MD0], 1);
      mbar_init(&bars[B_MD1], 1);
      mbar_init(&bars[B_DE], BWD_EW);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc_n(tmem_ptr_smem, (uint32_t)A.tmem_cols);
  } else {
    const float* mrow = P.mask ? P.mask + (int64_t)b * Lk : nullptr;
    for (int i = threadIdx.x; i < 384; i += BWD_EW * 32) {
      const float mv = (i < Lk && mrow) ? mrow[i] : 0.f;
      mask_nz |= mv != 0.f;
      sMask[i] = i < Lk ? mv * LOG2E : -INFINITY;
    }
  }
  tcgen05_fence_before();
  const bool any_mask = __syncthreads_or(mask_nz) != 0;  // all-zero additive mask == no mask (see the forward kernel)
  tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr_smem, 0);
  const uint32_t tDQ = tmem_base, tDV = tmem_base + 192, tDK = tmem_base + 192 + (16u << 16), tS0 = tmem_base + 256;
  if (threadIdx.x == 0) TR(101);

  if (warp == BWD_EW) {
    {
      // ===================== control warp: TMA + MMA issue (lane 0 issues; all lanes run the control flow) =====================
      const uint32_t el = lane == 0 ? 1u : 0u;
      const int col0 = h * D;
      auto load_kv = [&](int kt, int slot) {
        if (!el) return;
        mbar_expect_tx(&bars[B_KV0 + slot], (uint32_t)(2 * KBLK));
        tma_load_2d(sK + slot * KBLK, &A.map_k, &bars[B_KV0 + slot], col0, b * Lk + kt * KT);
        tma_load_2d(sV + slot * KBLK, &A.map_v, &bars[B_KV0 + slot], col0, b * Lk + kt * KT);
      };
      if (el) {
        mbar_expect_tx(&bars[B_QF0], (uint32_t)(2 * A.box_q * 128));
        tma_load_2d(sQ, &A.map_q, &bars[B_QF0], col0, b * Lq);
        tma_load_2d(sdO, &A.map_do, &bars[B_QF0], col0, b * Lq);
      }
      load_kv(0, 0);
      for (int qt = 1; qt < n_qt && el; ++qt) {
        mbar_expect_tx(&bars[B_QF0 + qt], (uint32_t)(2 * A.box_q * 128));
        tma_load_2d(sQ + qt * BLK, &A.map_q, &bars[B_QF0 + qt], col0, b * Lq + qt * TQ);
        tma_load_2d(sdO + qt * BLK, &A.map_do, &bars[B_QF0 + qt], col0, b * Lq + qt * TQ);
      }
      for (int kt = 1; kt < n_kt && kt < nkv; ++kt) load_kv(kt, kt);
      const uint32_t id_acc = idesc(D, 1, 1, KT);  // dV, dK (M = 64 keys): A (P'/dS tile) and B (dO/Q tile) both MN-major
      const uint32_t id_dq = idesc(D, 0, 1);   // dQ: A = dS K-major, B = K tile MN-major
      // Everything below is kept division-free and derived from kernel parameters and plain loop counters so that the
      // compiler can hold descriptors in uniform registers (a value that passes through a vector register costs an
      // R2UR round trip in front of every tcgen05.mma, which is what paces 128x64x16 MMAs otherwise).
      // (the shared-window base is re-derived here from the raw symbol so that it is not merged with the per-thread
      // address arithmetic of the element-wise warps, which lives in vector registers)
      const uint32_t sb = (smem_u32(smem_raw) + 1023u) & ~1023u;
      const uint32_t uQ = sb + (uint32_t)A.off_q, uDO = sb + (uint32_t)A.off_do, uK = sb + (uint32_t)A.off_k, uV = sb + (uint32_t)A.off_v;
      const uint32_t uP = sb + (uint32_t)A.off_p, uDS = sb + (uint32_t)A.off_ds;
      const uint32_t qk0 = dlo(uQ, 16), dok0 = dlo(uDO, 16), kk0 = dlo(uK, 16), vk0 = dlo(uV, 16);
      const uint32_t qm0 = dlo(uQ, 8192), dom0 = dlo(uDO, 8192), km0 = dlo(uK, 8192);
      const uint32_t pm0 = dlo(uP, BLK), dsm0 = dlo(uDS, BLK), dsk0 = dlo(uDS, 16);
      // S, dP of iteration (kt, qt) into TMEM buffer sb2; slot = kt % nkv, use = kt / nkv
      auto issue_sdp = [&](int kt, int qt, int slot, int use, int sb2) {
        const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;  // keys in this tile (multiple of 16)
        const uint32_t id_s = idesc(w, 0, 0);
        if (kt == 0) mbar_wait(&bars[B_QF0 + qt], 0);
        if (qt == 0) mbar_wait(&bars[B_KV0 + slot], (uint32_t)(use & 1));
        tcgen05_fence_after();
        const uint32_t q_lo = qk0 + (uint32_t)qt * (BLK >> 4), do_lo = dok0 + (uint32_t)qt * (BLK >> 4);
        const uint32_t k_lo = kk0 + (uint32_t)slot * (KBLK >> 4), v_lo = vk0 + (uint32_t)slot * (KBLK >> 4);
        const uint32_t tS = tS0 + (uint32_t)sb2 * 128u, tDP = tS + 64u;
        // the two independent accumulation chains are interleaved; one lane issues the whole batch
        if (!elect_one()) return;
        umma_lo<false>(el, tS, q_lo, k_lo, id_s);
        umma_lo<false>(el, tDP, do_lo, v_lo, id_s);
        umma_lo<true>(el, tS, q_lo + 2, k_lo + 2, id_s);
        umma_lo<true>(el, tDP, do_lo + 2, v_lo + 2, id_s);
        umma_lo<true>(el, tS, q_lo + 4, k_lo + 4, id_s);
        umma_lo<true>(el, tDP, do_lo + 4, v_lo + 4, id_s);
        umma_lo<true>(el, tS, q_lo + 6, k_lo + 6, id_s);
        umma_lo<true>(el, tDP, do_lo + 6, v_lo + 6, id_s);
        umma_commit_el(el, &bars[B_SF0 + sb2]);
      };
      // (a_kt, a_qt): the next S/dP to issue, two iterations ahead of the accumulate MMAs
      int a_kt = 0, a_qt = 0, a_slot = 0, a_use = 0;
      auto issue_next = [&](int sb2) {
        issue_sdp(a_kt, a_qt, a_slot, a_use, sb2);
        if (++a_qt == n_qt) {
          a_qt = 0;
          ++a_kt;
          if (++a_slot == nkv) { a_slot = 0; ++a_use; }
        }
      };
      issue_next(0);
      if (total > 1) issue_next(1);
      int it = 0, slot = 0, prev_slot = 0;
      for (int kt = 0; kt < n_kt; ++kt) {
        const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;
        const int ksteps = w >> 4;               // reduction over the tile's keys (dQ)
        for (int qt = 0; qt < n_qt; ++qt, ++it) {
          const int buf = it & 1;
          if (it + 2 < total) {
            mbar_wait(&bars[B_SE0 + buf], (uint32_t)((it >> 1) & 1));  // S, dP of this iteration are in registers
            issue_next(buf);
            TR(110 + 4 * it);
          }
          mbar_wait(&bars[B_PF0 + buf], (uint32_t)((it >> 1) & 1));
          TR(111 + 4 * it);
          if (qt == 0 && kt > 0) mbar_wait(&bars[B_DE], (uint32_t)((kt - 1) & 1));  // dK, dV of the previous key tile were read out
          tcgen05_fence_after();
          // MN-major operands step 16 rows = 2048 B per k-step; the K-major dS operand of dQ steps 32 B
          const uint32_t qm_lo = qm0 + (uint32_t)qt * (BLK >> 4), dom_lo = dom0 + (uint32_t)qt * (BLK >> 4);
          const uint32_t km_lo = km0 + (uint32_t)slot * (KBLK >> 4);
          const uint32_t pm_lo = pm0 + (uint32_t)buf * (BLK >> 4), dsm_lo = dsm0 + (uint32_t)buf * (BLK >> 4);
          const uint32_t dsk_lo = dsk0 + (uint32_t)buf * (BLK >> 4);
          const int qvalid = (Lq - qt * TQ) < TQ ? (Lq - qt * TQ) : TQ;
          const int qsteps = (qvalid + 15) >> 4;   // reduction over the tile's query rows
          const uint32_t tdq = tDQ + 64 * qt;
          // three independent chains (dV, dK, dQ), interleaved so consecutive MMAs never wait on each other's accumulator
          if (elect_one()) {
          umma_lo_p(el, tDV, pm_lo, dom_lo, id_acc, qt > 0 ? 1u : 0u);
          umma_lo_p(el, tDK, dsm_lo, qm_lo, id_acc, qt > 0 ? 1u : 0u);
          umma_lo_p(el, tdq, dsk_lo, km_lo, id_dq, kt > 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 1; ks < 8; ++ks) {
            if (ks < qsteps) {
              umma_lo<true>(el, tDV, pm_lo + ks * 128, dom_lo + ks * 128, id_acc);
              umma_lo<true>(el, tDK, dsm_lo + ks * 128, qm_lo + ks * 128, id_acc);
            }
            if (ks < 4 && ks < ksteps) umma_lo<true>(el, tdq, dsk_lo + ks * 2, km_lo + ks * 128, id_dq);
          }
          umma_commit_el(el, &bars[B_MD0 + buf]);
          }
          __syncwarp();
          TR(112 + 4 * it);
          if (qt == 0 && kt > 0 && kt - 1 + nkv < n_kt) {
            // the K/V slot of key tile kt - 1 is free once its last accumulate MMAs (iteration it - 1, issued a whole
            // iteration ago) have completed: refill it with tile kt - 1 + nkv
            mbar_wait(&bars[B_MD0 + ((it - 1) & 1)], (uint32_t)(((it - 1) >> 1) & 1));
            load_kv(kt - 1 + nkv, prev_slot);
          }
        }
        prev_slot = slot;
        if (++slot == nkv) slot = 0;
      }
    }
  } else {
    // ===================== element-wise warps =====================
    // 16 warps: quad = TMEM lane quadrant (32 query rows), cg = 16-key column group of the 64-key tile.  (8 warps x 32
    // columns left 2 warps per scheduler with ~650-instruction dependent chains per iteration: 28 % issue utilisation and
    // 45 % of the samples on long-scoreboard waits, profiles/r2_ncu_attn_full_baseline.txt.)
    const int quad = warp & 3, cg = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const float c = A.scale_l2;
    const float keep_scale = drop_keep_scale(P.drop_p);
    const float kss = keep_scale * P.scale;
    const uint32_t thr = drop_thr15(P.drop_p);
    const int mask_ld = A.mask_ld;
    const int64_t sbase = ((int64_t)b * P.H + h) * Lq;
    const bool has_mask = any_mask;
    // per-row constants of the <= 3 query tiles; rows >= Lq: lse = +inf -> P = exp2(-inf) = 0
    float lse2[3], dls[3];
#pragma unroll
    for (int qt = 0; qt < 3; ++qt) {
      const int qrow = qt * TQ + row;
      const bool ok = qt < n_qt && qrow < Lq;
      lse2[qt] = ok ? P.lse[sbase + qrow] * LOG2E : INFINITY;
      dls[qt] = ok ? P.delta[sbase + qrow] * P.scale : 0.f;
    }

    // dK, dV of key tile kt (complete: its last MMAs were observed) -> global; all EW warps arrive on B_DE.
    // M = 64 accumulators: key i of the tile sits in lane (i % 16) of sub-partition i / 16 (dV) and 16 lanes above (dK).
    auto store_dkv = [&](int kt) {
      const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;
      const bool kactive = quad * 16 < w;
      uint32_t r[16];
      if (kactive) {
        tmem_ld16(tDV + lane_addr + cg * 16, r);
        tmem_ld_wait();
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[B_DE]);
      const int key = kt * KT + quad * 16 + (lane & 15);
      if (kactive && key < Lk) {
        const bool is_v = lane < 16;
        bf16* dst = is_v ? (bf16*)P.dv + ((int64_t)b * Lk + key) * P.lddv : (bf16*)P.dk + ((int64_t)b * Lk + key) * P.lddk;
        store_row16(dst + h * D + cg * 16, r, (DROP && is_v) ? keep_scale : 1.0f);
      }
    };

    const uint32_t a_p = smem_u32(sP) + row * 128, a_ds = smem_u32(sdS) + row * 128, a_mask = smem_u32(sMask);
    const uint32_t swz = (uint32_t)row & 7u;
    // keep bits of this thread's 16 keys in iteration (kt2, qt2): read one iteration ahead, so that the global-load latency
    // (~550 cycles per iteration when it sat in front of the S-full wait) is covered by the previous iteration's arithmetic
    auto keep_bits16 = [&](int kt2, int qt2) -> uint32_t {
      const int w2 = (lk_pad - kt2 * KT) < KT ? (lk_pad - kt2 * KT) : KT;
      const int qrow2 = qt2 * TQ + row;
      if (!DROP || !((qt2 * TQ + quad * 32 < Lq) && cg * 16 < w2)) return 0xFFFFFFFFu;
      const uint64_t row_lin = (uint64_t)(sbase + qrow2);
      const int g0 = kt2 * 8 + cg * 2;  // 8-key group of this warp's first key
      if (P.drop_mask != nullptr)
        return qrow2 < Lq ? (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(P.drop_mask + (int64_t)row_lin * mask_ld + g0)) : 0u;
      return attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, g0, thr) |
             (attn_keep_bits8(P.drop_seed, P.drop_stream, row_lin, g0 + 1, thr) << 8);
    };
    uint32_t bits_next = keep_bits16(0, 0);
    int it = 0;
    for (int kt = 0; kt < n_kt; ++kt) {
      const int w = (lk_pad - kt * KT) < KT ? (lk_pad - kt * KT) : KT;
      const bool cols_live = cg * 16 < w;                    // this warp's 16 key columns exist in the tile (w % 16 == 0)
      const bool plain = !has_mask && (kt * KT + KT <= Lk);  // no additive mask, no padded keys in this tile
#pragma unroll 1
      for (int qt = 0; qt < n_qt; ++qt, ++it) {
        const bool active = (qt * TQ + quad * 32 < Lq) && cols_live;  // warp-uniform
        const float l2 = qt == 0 ? lse2[0] : (qt == 1 ? lse2[1] : lse2[2]);
        const float dl = qt == 0 ? dls[0] : (qt == 1 ? dls[1] : dls[2]);
        const uint32_t bits = bits_next;
        uint32_t rs0[16], rd0[16];
        const uint32_t tS = tS0 + (uint32_t)(it & 1) * 128u, tDP = tS + 64u;
        mbar_wait(&bars[B_SF0 + (it & 1)], (uint32_t)((it >> 1) & 1));
        tcgen05_fence_after();
        if (threadIdx.x == 0) TR(200 + 6 * it);
        if (active) {
          tmem_ld16(tS + lane_addr + cg * 16, rs0);
          tmem_ld16(tDP + lane_addr + cg * 16, rd0);
          tmem_ld_wait();
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[B_SE0 + (it & 1)]);
        if (it + 1 < total) bits_next = qt + 1 < n_qt ? keep_bits16(kt, qt + 1) : keep_bits16(kt + 1, 0);
        if (threadIdx.x == 0) TR(201 + 6 * it);
        // the staging buffers of parity it&1 were last read by the accumulate MMAs of iteration it-2
        if (it >= 2) mbar_wait(&bars[B_MD0 + (it & 1)], (uint32_t)(((it - 2) >> 1) & 1));
        if (threadIdx.x == 0) TR(202 + 6 * it);
        if (active) {
          // P'' = keep ? P : 0 (dV is scaled by 1/(1-p) on the way out), dS = P * (keep ? dP/(1-p) : 0 - delta) * scale
          const uint32_t pb = a_p + (it & 1) * BLK, db = a_ds + (it & 1) * BLK;
          const uint32_t mk = a_mask + (kt * KT + cg * 16) * 4;
          bwd_sub<DROP>(rs0, rd0, plain, mk, c, -l2, dl, DROP ? kss : P.scale, bits, pb, db, (uint32_t)(cg * 2), swz);
        }
        if (threadIdx.x == 0) TR(203 + 6 * it);
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[B_PF0 + (it & 1)]);
        if (threadIdx.x == 0) TR(204 + 6 * it);
        if (qt == 0 && kt > 0) {
          // dK, dV of the previous key tile: its last accumulate MMAs are those of iteration it-1
          mbar_wait(&bars[B_MD0 + ((it - 1) & 1)], (uint32_t)(((it - 1) >> 1) & 1));
          tcgen05_fence_after();
          store_dkv(kt - 1);
          if (threadIdx.x == 0) TR(205 + 6 * it);
        }
      }
    }
    mbar_wait(&bars[B_MD0 + ((total - 1) & 1)], (uint32_t)(((total - 1) >> 1) & 1));
    tcgen05_fence_after();
    store_dkv(n_kt - 1);
    for (int qt = 0; qt < n_qt; ++qt) {
      const int qrow = qt * TQ + row;
      if (qt * TQ + quad * 32 < Lq) {
        uint32_t r[16];
        tmem_ld16(tDQ + 64 * qt + lane_addr + cg * 16, r);
        tmem_ld_wait();
        if (qrow < Lq) store_row16((bf16*)P.dq + ((int64_t)b * Lq + qrow) * P.lddq + h * D + cg * 16, r, 1.0f);
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TR(102);
  if (warp == BWD_EW) {
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
// mma.sync tiles (profiles/r1_attn_bench_v6.txt): forward when both lengths exceed one 64-wide tile, backward when either
// does.  One CTA per SM with ~6 k cycles of prologue cannot win on the tiny cross / text shapes.
bool tc_supported(const m3ae_attn_params_t* p, bool bwd, bool force) {
  if (p->dh != D || p->dtype != M3AE_BF16 || p->causal) return false;
  if (p->Lq > 3 * TQ || p->Lk > 384) return false;
  if (!bwd && p->Lk > 336) return false;  // forward keeps K, V and the whole P row block resident: 336 keys fill 227 KB
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

// Fills the geometry shared by both kernels; returns the dynamic shared memory size through *smem_bytes.
int setup(const m3ae_attn_params_t* p, bool bwd, TcAttn* A, int* smem_bytes) {
  A->p = *p;
  A->lk_pad = (p->Lk + 15) & ~15;
  A->n_qt = (p->Lq + TQ - 1) / TQ;
  A->box_q = p->Lq >= TQ ? TQ : ((p->Lq + 15) & ~15);
  A->n_kt = (A->lk_pad + KT - 1) / KT;
  A->mask_ld = 8 * ((p->Lk + 63) / 64);
  A->scale_l2 = p->scale * LOG2E;
  const int cols = p->H * D;
  int off = 0;
  if (bwd) {
    A->nbox_k = 1;
    A->box_k = KT;
    A->off_q = off; off += A->n_qt * BLK;
    A->off_do = off; off += A->n_qt * BLK;
    A->off_p = off; off += 2 * BLK;
    A->off_ds = off; off += 2 * BLK;
    const int fixed = off + 4 * 384 + B_NBAR * 8 + 16 + 1024;
    int nkv = (SMEM_LIMIT - fixed) / (2 * KBLK);
    if (nkv > MAX_NKV) nkv = MAX_NKV;
    if (nkv > A->n_kt) nkv = A->n_kt;
    M3AE_CHECK(nkv >= 1, M3AE_ERR_UNSUPPORTED, "attention_tc: no room for a K/V tile");
    A->nkv = nkv;
    A->off_k = off; off += nkv * KBLK;
    A->off_v = off; off += nkv * KBLK;
    A->tmem_cols = 512;
    A->off_red = 0;
    A->off_bits = 0;
  } else {
    A->nkv = 0;
    A->nbox_k = (A->lk_pad + 127) / 128;
    A->box_k = (((A->lk_pad + A->nbox_k - 1) / A->nbox_k) + 15) & ~15;
    const int kv_bytes = A->nbox_k * A->box_k * 128;  // multiple of 2048
    A->off_k = off; off += kv_bytes;
    A->off_v = off; off += kv_bytes;
    A->off_q = off; off += 2 * BLK;
    A->off_do = 0;
    A->off_p = off; off += ((A->lk_pad + 63) / 64) * BLK;
    A->off_ds = 0;
    A->tmem_cols = 128 + A->lk_pad <= 256 ? 256 : 512;
    A->off_red = off + 4 * 384;
    A->off_bits = A->off_red + 8 * TQ * 4;
  }
  A->off_mask = off; off += 4 * 384;
  if (!bwd) {
    off += 8 * TQ * 4;            // row max / row sum exchange
    off += TQ * A->mask_ld;       // keep-bit tile (multiple of 8 bytes per row)
  }
  off = (off + 15) & ~15;
  A->off_bar = off; off += (bwd ? (int)B_NBAR : (int)F_NBAR) * 8 + 16;
  *smem_bytes = off + 1024;  // alignment slack
  M3AE_CHECK(*smem_bytes <= SMEM_LIMIT, M3AE_ERR_UNSUPPORTED, "attention_tc: %d bytes of shared memory needed", *smem_bytes);
  int rc;
  if ((rc = encode_head_map(&A->map_q, p->q, (int64_t)p->B * p->Lq, cols, p->ldq, A->box_q))) return rc;
  if ((rc = encode_head_map(&A->map_k, p->k, (int64_t)p->B * p->Lk, cols, p->ldk, A->box_k))) return rc;
  if ((rc = encode_head_map(&A->map_v, p->v, (int64_t)p->B * p->Lk, cols, p->ldv, A->box_k))) return rc;
  if (bwd && (rc = encode_head_map(&A->map_do, p->dout, (int64_t)p->B * p->Lq, cols, p->ldo, A->box_q))) return rc;
  return M3AE_OK;
}

template <typename K>
int set_smem_attr(K kernel) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  M3AE_CHECK(e == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(attention_tc) failed: %s", cudaGetErrorString(e));
  return M3AE_OK;
}

}  // namespace

int attn_fwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st) {
  if (!tc_supported(p, false, force)) return M3AE_ERR_UNSUPPORTED;
  static std::once_flag once[64];
  static int attr_rcs[64];
  const int slot = current_device_slot();
  std::call_once(once[slot], [slot] {
    attr_rcs[slot] = set_smem_attr(attn_fwd_tc_kernel<false>);
    if (!attr_rcs[slot]) attr_rcs[slot] = set_smem_attr(attn_fwd_tc_kernel<true>);
  });
  if (attr_rcs[slot]) return attr_rcs[slot];
  TcAttn A;
  memset(&A, 0, sizeof(A));
  int smem = 0;
  int rc = setup(p, false, &A, &smem);
  if (rc) return rc;
  dim3 grid(p->H, p->B);
  if (p->drop_p > 0.f) M3AE_LAUNCH((attn_fwd_tc_kernel<true>), grid, NT_FWD, smem, st, A);
  else M3AE_LAUNCH((attn_fwd_tc_kernel<false>), grid, NT_FWD, smem, st, A);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int attn_bwd_tc(const m3ae_attn_params_t* p, bool force, cudaStream_t st) {
  if (!tc_supported(p, true, force)) return M3AE_ERR_UNSUPPORTED;
  static std::once_flag once[64];
  static int attr_rcs[64];
  const int slot = current_device_slot();
  std::call_once(once[slot], [slot] {
    attr_rcs[slot] = set_smem_attr(attn_bwd_tc_kernel<false>);
    if (!attr_rcs[slot]) attr_rcs[slot] = set_smem_attr(attn_bwd_tc_kernel<true>);
  });
  if (attr_rcs[slot]) return attr_rcs[slot];
  TcAttn A;
  memset(&A, 0, sizeof(A));
  int smem = 0;
  int rc = setup(p, true, &A, &smem);
  if (rc) return rc;
  launch_delta_bf16_64(p, st);
  dim3 grid(p->H, p->B);
  if (p->drop_p > 0.f) M3AE_LAUNCH((attn_bwd_tc_kernel<true>), grid, NT_BWD, smem, st, A);
  else M3AE_LAUNCH((attn_bwd_tc_kernel<false>), grid, NT_BWD, smem, st, A);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae

#ifdef M3AE_ATTN_TRACE
extern "C" __attribute__((visibility("default"))) int m3ae_attn_trace(long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, m3ae::g_attn_trace, sizeof(long long) * (n < 512 ? n : 512));
}
#endif
