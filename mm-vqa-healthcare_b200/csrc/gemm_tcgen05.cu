// Persistent, warp-specialised bf16 GEMM for sm_100a.
//   Default (CTAS = 2): a CTA PAIR (cluster of 2 on one TPC) computes a 256x256 tile with tcgen05.mma.cta_group::2.
//     TMA (cp.async.bulk.tensor.2d.cta_group::2, SWIZZLE_128B) -> 4-stage smem ring, each CTA staging its 128 rows of A and
//     half of B -> MMAs issued by one lane of the leader CTA (fp32 accumulators double-buffered in the TMEM of both CTAs,
//     multicast commits) -> per-CTA tcgen05.ld epilogue (bias / GELU / act' / dropout with cached keep bits / residual).
//   CTAS = 1: one CTA per 128x256 tile (cta_group::1, 3 stages) — launches whose problems all have M <= 128, and the
//     cross-check of the pair kernel (bit-identical results, tests/test_kernels_gpu.py).
//   bf16 outputs leave through swizzled 32x32 smem staging tiles and TMA stores, residual / pre-activation inputs
//   arrive through TMA loads into the same kind of tile (each lane owns one accumulator ROW, so direct global
//   accesses would touch 32 different lines per instruction); split-K fp32 outputs are added by TMA reduce-add tiles.
// One launch runs up to M3AE_MAX_GROUP independent problems (the text and image streams of a fusion
// layer pair, reference m3ae_module.py:276-277), each with K-major or MN-major operands so the same kernel
// serves forward (A K-major, B K-major), dgrad (B MN-major) and wgrad (A and B MN-major, split-K).
// Tiles are rasterised N-fastest when A is the larger operand, so an A panel is read from DRAM once (see the launcher).
//
// Algorithmic traffic per problem: (M*K + N*K) * 2 B in, M*N * sizeof(out) out; 2*M*N*K flop.
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace m3ae {

namespace {

constexpr int BM = 128;
constexpr int BN_WIDE = 256;    // output columns per tile
constexpr int BN_SKINNY = 64;   // ... of the skinny variant (launches whose problems all have M <= 128 and few wide tiles)
constexpr int BK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM * BK * 2;
// Warp roles: 0..15 epilogue, 16 TMA producer, 17 MMA issuer, 18 TMEM allocator, 19 idle.  The control warps carry the
// HIGHEST warp ids on purpose: a scheduler picks the eligible warp with the highest id first, and epilogue warps polling an
// mbarrier are always eligible — with the producer and the MMA issuer at ids 0 and 1 the spinning epilogue warps of their
// schedulers kept taking the issue slot (MMA issuer found its operands ready in 95 % of its waits, yet the tensor pipe
// was only 60 % active: profiles/r2_ncu_gemm_full_16w.txt).
constexpr int NUM_THREADS = 640;
constexpr int WARP_TMA = 16, WARP_MMA = 17, WARP_ALLOC = 18;
constexpr int NUM_EPI_WARPS = 16;
constexpr int TMEM_COLS = 512;  // 2 accumulator stages x 256 fp32 columns
// Epilogue staging: every epilogue warp owns 2 buffers of 32 rows x 32 bf16 (64-byte rows, SWIZZLE_64B), one per
// 32-column chunk of its 64 accumulator columns: the TMA-loaded residual / saved act' tile is overwritten in place by the
// result, which a TMA store then drains.  The split-K path uses the same 4 KB as one 32x32 fp32 tile.
constexpr int EPI_BUF_BYTES = 32 * 64;
constexpr int EPI_WARP_BYTES = 2 * EPI_BUF_BYTES;
constexpr int EPI_BYTES = NUM_EPI_WARPS * EPI_WARP_BYTES;
constexpr int BIAS_BYTES = 2 * 4 * 64 * 4;  // per accumulator parity and 64-column group: the tile's bias values (the four
                                           // warps of a column group write the same values)
// CTAS = 1: one CTA computes a 128x256 tile (tcgen05.mma.cta_group::1).  CTAS = 2: a CTA PAIR (cluster of 2 on one TPC)
// computes a 256x256 tile with tcgen05.mma.cta_group::2 — each CTA stages its own 128 rows of A and HALF of B
// (128 of the 256 columns), so the operand bytes an SM reads from shared memory per MMA drop from 12 KB to 8 KB
// (96 -> 64 B/clk against the 128 B/clk an SM can deliver) and a stage shrinks from 48 KB to 32 KB (4 stages instead of 3).
// CL = 4 (opt-in, m3ae_set_gemm_cta_pair(3)): a cluster of TWO CTA pairs.  Both pairs work on neighbouring tiles that share
// one operand (the B tile when the pairs are stacked along M, the A tile when they sit side by side along N); every CTA
// fetches HALF of its share of that operand and TMA-multicasts it to the CTA of the same rank in the other pair: 96 KB
// of L2 reads per 16.8 MFLOP instead of 64 KB per 8.4 MFLOP (ncu: 774 MB of TMA loads for 37 MB of operands on the
// FFN-up GEMM, 8.4 TB/s).  Measured: no gain — see the launcher.
// BNV = 64 (single CTA only): 128x64 tiles.  The decoder's per-token GEMMs have M = batch <= 128 and N = 768..3072: with
// 256-column tiles 3-12 CTAs of 148 stream the whole weight matrix through one k-loop each (14-35 us per GEMM against 1-7 us
// of HBM time, profiles/r1_launches_decoder_v18_summary.txt); 64-column tiles put 12-48 CTAs on it with a 6-stage ring.
template <int CL, int BNV = BN_WIDE>
struct Cfg {
  static constexpr int BN = BNV;
  static constexpr int CTAS = CL >= 2 ? 2 : 1;   // CTAs per output tile (cta_group)
  static constexpr int NPAIR = CL / CTAS;        // tiles a cluster works on at a time
  static constexpr int TILE_M = BM * CTAS;
  static constexpr int B_ROWS = BN / CTAS;  // B rows (output columns) staged per CTA
  static constexpr int B_STAGE_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
#ifndef M3AE_GEMM_STAGES
#define M3AE_GEMM_STAGES 5
#endif
  // Operand bytes in flight per SM = (STAGES - 1) x 32 KB; at 64 B/clk per SM (tensor pipe saturated) and ~1500-2000
  // cycles of loaded L2 latency, Little's law asks for 96-128 KB.  Measured 3 / 4 / 5 stages (profiles/r2_gemm_stages_ab.txt):
  // K = 3072 forward 1011 / 1123 / 1175 TFLOP/s, split-K wgrad 1164 / 1333 / 1377; the K = 768 shapes do not move.
  static constexpr int STAGES = CTAS == 1 ? (BNV == BN_SKINNY ? 6 : 3) : M3AE_GEMM_STAGES;
  static constexpr int NUM_BARS = 2 * STAGES + 4 + 2 * NUM_EPI_WARPS;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BIAS_BYTES + NUM_BARS * 8 + 16;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};
constexpr int MAX_GROUP = 4;

struct alignas(64) TcProblem {
  CUtensorMap map_a;
  CUtensorMap map_b;
  CUtensorMap map_out;   // fast epilogue only: bf16 [M,N], box 32x32, SWIZZLE_64B
  CUtensorMap map_aux;   // aux_out (same geometry as out)
  CUtensorMap map_in;    // residual or aux_in
  int share;             // CL = 4 only: 1 = the cluster's pairs are stacked along M and share (multicast) the B tile,
                         //              2 = side by side along N, sharing the A tile.  tiles_m / tiles_n then count
                         //              two-tile super tiles along that dimension.
  int epi_fast;          // 1: TMA-staged bf16 epilogue; 2: fp32 accumulate through TMA reduce-add tiles (split-K wgrad);
                         // 0: generic per-element epilogue
  int in_is_residual;    // fast path: which operand map_in describes (0 = none/aux_in)
  int has_in;
  EpiDev epi;
  int M, N, K;
  int a_major, b_major;
  int tiles_m, tiles_n, split_k, kblocks, kb_per_split;
  int n_fastest;         // tile order: 1 = the N tiles of one M tile are consecutive units (A panel read once), 0 = M fastest
  int unit_begin, unit_end;
};
struct TcParams {
  TcProblem p[MAX_GROUP];
  int n_problems;
  int total_units;
};

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

struct Unit {
  int p, m_blk, n_blk, kb0, kb1;
};
__device__ __forceinline__ Unit decode_unit(const TcParams& P, int unit) {
  Unit u;
  u.p = 0;
#pragma unroll
  for (int i = 1; i < MAX_GROUP; ++i)
    if (i < P.n_problems && unit >= P.p[i].unit_begin) u.p = i;
  const TcProblem& q = P.p[u.p];
  const int local = unit - q.unit_begin;
  const int split = local % q.split_k;
  const int t = local / q.split_k;
  if (q.n_fastest) {
    u.n_blk = t % q.tiles_n;
    u.m_blk = t / q.tiles_n;
  } else {
    u.m_blk = t % q.tiles_m;
    u.n_blk = t / q.tiles_m;
  }
  u.kb0 = split * q.kb_per_split;
  u.kb1 = min(q.kblocks, u.kb0 + q.kb_per_split);
  return u;
}

template <int CL, int BNV>
__global__ void __launch_bounds__(NUM_THREADS, 1) gemm_tcgen05_kernel(const __grid_constant__ TcParams P) {
  using C = Cfg<CL, BNV>;
  constexpr int CTAS = C::CTAS;
  constexpr int BN = C::BN;
  constexpr int STAGES = C::STAGES;
  constexpr int STAGE_BYTES = C::STAGE_BYTES;
  constexpr int NUM_BARS = C::NUM_BARS;
  constexpr int TILE_M = C::TILE_M;
  // SWIZZLE_128B operand tiles need 1024-byte alignment.  The kernel has no static shared memory, so the dynamic window
  // starts on a 1 KB boundary (checked below: every byte of the 227 KB is spoken for, there is no slack to round up into).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0u) {
    if (threadIdx.x == 0) printf("m3ae: gemm_tcgen05_kernel: dynamic shared memory is not 1 KB aligned (0x%x)\n", smem_u32(smem_raw));
    __trap();
  }
  // CTA pair: rank 0 (leader) issues the MMAs for both; work is distributed per pair
  const int crank = CL >= 2 ? (int)cluster_ctarank() : 0;
  const int rank = crank & 1;         // rank inside the CTA pair
  const int pair = crank >> 1;        // which pair of the cluster (CL = 4)
  const uint16_t pair_mask = (uint16_t)(3u << (2 * pair));  // the two CTAs of this pair
  const int worker = (int)blockIdx.x / CL;
  const int n_workers = (int)gridDim.x / CL;
  uint8_t* epi_smem = smem + STAGES * STAGE_BYTES;
  float* bias_smem = (float*)(epi_smem + EPI_BYTES);
  uint64_t* bars = (uint64_t*)(epi_smem + EPI_BYTES + BIAS_BYTES);
  uint64_t* full_bar = bars;                  // [STAGES]
  uint64_t* empty_bar = bars + STAGES;        // [STAGES]
  uint64_t* tmem_full = bars + 2 * STAGES;    // [2]
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;  // [2]
  uint64_t* in_bar = bars + 2 * STAGES + 4;   // [NUM_EPI_WARPS][2]
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + NUM_BARS);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform by construction: say so
  const int lane = threadIdx.x & 31;

  if (warp == WARP_TMA && lane == 0) {
    for (int i = 0; i < P.n_problems; ++i) {
      tma_prefetch_desc(&P.p[i].map_a);
      tma_prefetch_desc(&P.p[i].map_b);
      if (P.p[i].epi_fast == 2) tma_prefetch_desc(&P.p[i].map_out);
      if (P.p[i].epi_fast == 1) {
        tma_prefetch_desc(&P.p[i].map_out);
        if (P.p[i].epi.aux_out) tma_prefetch_desc(&P.p[i].map_aux);
        if (P.p[i].has_in) tma_prefetch_desc(&P.p[i].map_in);
      }
    }
  }
  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], C::NPAIR);  // a slot is free once the MMAs of EVERY pair that received data in it are done
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], NUM_EPI_WARPS * CTAS);  // the leader's barrier collects the epilogue warps of both CTAs
    }
    for (int s = 0; s < 2 * NUM_EPI_WARPS; ++s) mbar_init(&in_bar[s], 1);
    fence_barrier_init();
  }
  if (warp == WARP_ALLOC) {
    if (CTAS == 2) tmem_alloc_pair<TMEM_COLS>(tmem_ptr_smem);
    else tmem_alloc<TMEM_COLS>(tmem_ptr_smem);
  }
  tcgen05_fence_before();
  if (CL >= 2) cluster_sync_all();  // the other CTAs' barriers must be initialised before anything remote arrives on them
  else __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) overlapped the previous kernel's
  // tail; nothing below may touch global memory before that kernel has completed
  pdl_sync();

  // Producer and MMA warps run their control flow with ALL lanes (convergent) and elect one lane per batch of
  // asynchronous operations.  Issuing from inside an `if (lane == 0)` region makes the compiler feed every TMA / MMA
  // operand into uniform registers through an ELECT + R2UR.BROADCAST waterfall loop in front of EACH instruction
  // (~130 cycles per tcgen05.mma: as long as the 128-cycle MMA itself, so the issue thread paced the tensor pipe).
  if (warp == WARP_TMA) {
    // ===================== TMA producer =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int unit = worker; unit < P.total_units; unit += n_workers) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      const int share = CL == 4 ? q.share : 0;
      const int m_blk = share == 1 ? 2 * u.m_blk + pair : u.m_blk;
      const int n_blk = share == 2 ? 2 * u.n_blk + pair : u.n_blk;
      const int a_row0 = m_blk * TILE_M + rank * BM;         // this CTA's 128 rows of A
      const int b_row0 = n_blk * BN + rank * C::B_ROWS;      // this CTA's share of B (all 256 columns when alone)
      const uint16_t mc_mask = (uint16_t)(5u << rank);       // the CTAs of this rank in both pairs
      for (int kb = u.kb0; kb < u.kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_STAGE_BYTES;
          auto load = [&](void* dst, const CUtensorMap* m, int c0, int c1) {
            if (CTAS == 2) tma_load_2d_pair(dst, m, &full_bar[stage], c0, c1);
            else tma_load_2d(dst, m, &full_bar[stage], c0, c1);
          };
          // the shared operand: this CTA fetches 64 of its 128 rows (one of the two 8 KB swizzle blocks) and multicasts
          // them into the same place of the same-rank CTA of the other pair
          auto load_mc = [&](void* dst, const CUtensorMap* m, int c0, int c1) {
            tma_load_2d_pair_mc(dst, m, &full_bar[stage], c0, c1, mc_mask);
          };
          if (rank == 0) mbar_expect_tx(&full_bar[stage], STAGE_BYTES * CTAS);  // bytes of both CTAs land on the leader's barrier
          if (share == 2) {
            if (q.a_major == M3AE_MAJOR_K) load_mc(sa + pair * (64 * 128), &q.map_a, kb * BK, a_row0 + pair * 64);
            else load_mc(sa + pair * (BK * 128), &q.map_a, a_row0 + pair * 64, kb * BK);
          } else if (q.a_major == M3AE_MAJOR_K) {
            load(sa, &q.map_a, kb * BK, a_row0);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load(sa + j * (BK * 128), &q.map_a, a_row0 + j * 64, kb * BK);
          }
          if (share == 1) {
            if (q.b_major == M3AE_MAJOR_K) load_mc(sb + pair * (64 * 128), &q.map_b, kb * BK, b_row0 + pair * 64);
            else load_mc(sb + pair * (BK * 128), &q.map_b, b_row0 + pair * 64, kb * BK);
          } else if (q.b_major == M3AE_MAJOR_K) {
            load(sb, &q.map_b, kb * BK, b_row0);
          } else {
#pragma unroll
            for (int j = 0; j < C::B_ROWS / 64; ++j) load(sb + j * (BK * 128), &q.map_b, b_row0 + j * 64, kb * BK);
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == WARP_MMA && rank == 0) {
    // ===================== MMA issuer (one elected lane per k-block; the leader CTA of a pair) =====================
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int unit = worker; unit < P.total_units; unit += n_workers) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      const uint32_t idesc = (1u << 4) /*D=f32*/ | (1u << 7) /*A=bf16*/ | (1u << 10) /*B=bf16*/ |
                             ((uint32_t)q.a_major << 15) | ((uint32_t)q.b_major << 16) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)(TILE_M >> 4) << 24);
      // K-major: 8-row groups 1024 B apart, K advance = 32 B inside the swizzle row.
      // MN-major: 64-element column groups BK*128 B apart (LBO), 8-k groups 1024 B apart (SBO), K advance = 16 rows.
      const uint32_t a_lbo = q.a_major == M3AE_MAJOR_K ? 16u : (uint32_t)(BK * 128);
      const uint32_t b_lbo = q.b_major == M3AE_MAJOR_K ? 16u : (uint32_t)(BK * 128);
      const uint32_t a_kstep = q.a_major == M3AE_MAJOR_K ? (UMMA_K * 2) : (UMMA_K * 128);
      const uint32_t b_kstep = q.b_major == M3AE_MAJOR_K ? (UMMA_K * 2) : (UMMA_K * 128);
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
      for (int kb = u.kb0; kb < u.kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tcgen05_fence_after();
        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
        const uint32_t sb = sa + A_STAGE_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = make_smem_desc(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t db = make_smem_desc(sb + k * b_kstep, b_lbo, 1024);
            if (CTAS == 2) umma_f16_pair(tmem_d, da, db, idesc, (kb > u.kb0 || k > 0) ? 1u : 0u);
            else umma_f16(tmem_d, da, db, idesc, (kb > u.kb0 || k > 0) ? 1u : 0u);
          }
          if (CTAS == 2) {
            // frees the smem slot once these MMAs have read it: in both CTAs of the pair, and (CL = 4) in the other pair
            // too, whose producers multicast into this pair's slots
            umma_commit_mc(&empty_bar[stage], CL == 4 ? (uint16_t)0xF : pair_mask);
            if (kb + 1 == u.kb1) umma_commit_mc(&tmem_full[acc], pair_mask);
          } else {
            umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
            if (kb + 1 == u.kb1) umma_commit(&tmem_full[acc]);  // accumulator complete -> epilogue
          }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp < NUM_EPI_WARPS) {
    // ===================== epilogue warps =====================
    // 16 warps: quad = TMEM lane quadrant (32 accumulator rows), cgrp = 64-column group of the 256-column accumulator.
    // (Round 1 ran 8 warps x 128 columns at 168 registers: two warps per scheduler could not cover their own dependency
    // latency and the scalar erfc polynomial alone filled the FMA pipe for longer than a K = 768 tile's MMAs take —
    // tensor pipe 31-41 % active on the epilogue-heavy stages, profiles/r1_ncu_gemm_full_v19_final.txt.)
    const int e = warp;
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int cgrp = e >> 2;
    const bool cg_live = cgrp * 64 < BN;  // skinny tiles: only the first column group exists (the others just keep the barrier protocol)
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t iph0 = 0, iph1 = 0;  // phases of this warp's two in-buffer barriers
    uint8_t* const wbuf = epi_smem + e * EPI_WARP_BYTES;
    uint64_t* const ibar = in_bar + 2 * e;
    float* const sbias_grp = bias_smem + cgrp * 64;  // + acc * 256: this tile's bias values
    auto release_acc = [&](int a) {  // this warp is done reading accumulator `a`
      if (CTAS == 2 && rank != 0) mbar_arrive_cluster(&tmem_empty[a], (uint32_t)(2 * pair));
      else mbar_arrive(&tmem_empty[a]);
    };
    for (int unit = worker; unit < P.total_units; unit += n_workers) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      const EpiDev& epi = q.epi;
      const int share = CL == 4 ? q.share : 0;
      const int m_blk = share == 1 ? 2 * u.m_blk + pair : u.m_blk;
      const int n_blk = share == 2 ? 2 * u.n_blk + pair : u.n_blk;
      const int tile_row0 = m_blk * TILE_M + rank * BM;  // first output row of this CTA's 128-lane accumulator
      const int row0 = tile_row0 + quad * 32;
      const int col_base = n_blk * BN + cgrp * 64;
      const uint32_t tacc = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + cgrp * 64);
      if (q.epi_fast == 2) {
        // ---- fp32 accumulate (out += alpha * acc): one 32x32 fp32 tile staged in swizzled smem (the warp's whole 4 KB),
        //      added into global memory by a TMA reduction.  One bulk reduction replaces 1024 scattered RED.ADD.F32.
        const int nchunks = (row0 < q.M && cg_live) ? min(2, (q.N - col_base + 31) / 32) : 0;
        const float alpha = epi.alpha;
        mbar_wait(&tmem_full[acc], acc_phase);
        tcgen05_fence_after();
#pragma unroll 1
        for (int c = 0; c < nchunks; ++c) {
          uint32_t r0[16], r1[16];
          tmem_ld_32x32b_x16(tacc + (uint32_t)(c * 32), r0);
          tmem_ld_32x32b_x16(tacc + (uint32_t)(c * 32 + 16), r1);
          tmem_ld_wait();
          if (lane == 0) tma_store_wait_read<0>();  // the previous reduction has finished reading the tile
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) {  // SWIZZLE_128B: 16-byte chunk ^= row % 8
            const float4 w0 = make_float4(__uint_as_float(r0[4 * j]) * alpha, __uint_as_float(r0[4 * j + 1]) * alpha,
                                          __uint_as_float(r0[4 * j + 2]) * alpha, __uint_as_float(r0[4 * j + 3]) * alpha);
            const float4 w1 = make_float4(__uint_as_float(r1[4 * j]) * alpha, __uint_as_float(r1[4 * j + 1]) * alpha,
                                          __uint_as_float(r1[4 * j + 2]) * alpha, __uint_as_float(r1[4 * j + 3]) * alpha);
            *reinterpret_cast<float4*>(wbuf + lane * 128 + (((uint32_t)j ^ (uint32_t)(lane & 7)) << 4)) = w0;
            *reinterpret_cast<float4*>(wbuf + lane * 128 + (((uint32_t)(j + 4) ^ (uint32_t)(lane & 7)) << 4)) = w1;
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_reduce_add_2d(&q.map_out, wbuf, col_base + c * 32, row0);
            tma_store_commit();
          }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) {
          release_acc(acc);
          tma_store_wait_read<0>();  // the next unit may be a bf16 problem that writes these bytes after a wait_read<1>
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        continue;
      }
      if (q.epi_fast) {
        // ---- TMA-staged bf16 epilogue: coalesced in/out through swizzled smem sub-tiles of 32 rows x 32 columns.
        // Chunk c (32 columns) stages its input (residual / saved act') AND its output in buf[c]: a thread reads and
        // writes only its own row, so the result overwrites the input in place.  With a second (aux) output every chunk
        // uses buf[0] for in/out and buf[1] for the aux tile.
        const int nchunks = cg_live ? min(2, (q.N - col_base + 31) / 32) : 0;  // <= 0: this warp's columns lie outside N
        const int64_t row = (int64_t)row0 + lane;
        const uint32_t sw = (uint32_t)((lane >> 1) & 3);           // SWIZZLE_64B: 16-byte chunk ^= (row/2)%4
        const bool aux = epi.aux_out != nullptr;
        const bool aux_grad = aux && epi.aux_is_grad;
        if (q.has_in && nchunks > 0 && lane == 0) {
          tma_store_wait_read<0>();  // the previous unit's stores have finished reading both buffers
          mbar_expect_tx(&ibar[0], EPI_BUF_BYTES);
          tma_load_2d(wbuf, &q.map_in, &ibar[0], col_base, row0);
          if (!aux && nchunks > 1) {
            mbar_expect_tx(&ibar[1], EPI_BUF_BYTES);
            tma_load_2d(wbuf + EPI_BUF_BYTES, &q.map_in, &ibar[1], col_base + 32, row0);
          }
        }
        // bias: loaded before the accumulator wait (the global-load latency hides behind the tile's MMAs), published to
        // shared memory after it — by then every warp has finished the tile that last used this parity's slot
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (epi.bias && lane < 16) {
          const int c4 = col_base + 4 * lane;
          if (c4 < q.N) b4 = __ldg(reinterpret_cast<const float4*>(epi.bias + c4));
        }
        float* const sbias = sbias_grp + acc * 256;
        uint32_t cached_bits[2] = {0u, 0u};
        if (epi.drop_p > 0.f && epi.drop_bits_ready && row < q.M) {
#pragma unroll
          for (int c = 0; c < 2; ++c)
            if (c < nchunks) cached_bits[c] = __ldg(epi.drop_bits + row * epi.drop_bits_ld + ((col_base + c * 32) >> 5));
        }
        mbar_wait(&tmem_full[acc], acc_phase);
        tcgen05_fence_after();
        if (epi.bias) {
          if (lane < 16) *reinterpret_cast<float4*>(sbias + 4 * lane) = b4;
          __syncwarp();
        }
#pragma unroll 1
        for (int c = 0; c < nchunks; ++c) {
          const int col = col_base + c * 32;
          const int bi = aux ? 0 : c;
          uint8_t* const iobuf = wbuf + bi * EPI_BUF_BYTES;
          uint8_t* const abuf = wbuf + EPI_BUF_BYTES;
          if (lane == 0) {
            if (aux) {
              if (c > 0 || !q.has_in) tma_store_wait_read<0>();
              if (q.has_in && c > 0) {  // (input + two outputs: no prefetch, the in/out buffer is shared by both chunks)
                mbar_expect_tx(&ibar[0], EPI_BUF_BYTES);
                tma_load_2d(wbuf, &q.map_in, &ibar[0], col, row0);
              }
            } else if (!q.has_in) {
              tma_store_wait_read<1>();  // the store that last read buf[c] (this chunk of the previous unit) is done
            }
          }
          __syncwarp();
          if (q.has_in) {
            if (bi) { mbar_wait(&ibar[1], iph1); iph1 ^= 1; } else { mbar_wait(&ibar[0], iph0); iph0 ^= 1; }
          }
          uint32_t keep_word = 0;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t r[16];
            tmem_ld_32x32b_x16(tacc + (uint32_t)(c * 32 + hh * 16), r);
            tmem_ld_wait();
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
            if (epi.bias) {
              const float4* b4 = reinterpret_cast<const float4*>(sbias + c * 32 + hh * 16);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 t4 = b4[j];
                v[4 * j] += t4.x; v[4 * j + 1] += t4.y; v[4 * j + 2] += t4.z; v[4 * j + 3] += t4.w;
              }
            }
            const uint32_t o0 = (((uint32_t)(2 * hh) ^ sw) << 4), o1 = (((uint32_t)(2 * hh + 1) ^ sw) << 4);
            auto put8 = [&](uint8_t* buf, uint32_t off, const float* f) {
              uint4 w;
              w.x = pack_bf16x2(f[0], f[1]); w.y = pack_bf16x2(f[2], f[3]);
              w.z = pack_bf16x2(f[4], f[5]); w.w = pack_bf16x2(f[6], f[7]);
              *reinterpret_cast<uint4*>(buf + lane * 64 + off) = w;
            };
            if (aux && !aux_grad) {
              put8(abuf, o0, v);
              put8(abuf, o1, v + 8);
            }
            if (aux_grad && epi.act == M3AE_ACT_GELU) {
              // aux := gelu'(v): the cdf/pdf terms are shared with the activation itself; two elements per FFMA2
              float d[16];
#pragma unroll
              for (int j = 0; j < 16; j += 2) fast_gelu_both2(v[j], v[j + 1], d[j], d[j + 1]);
              put8(abuf, o0, d);
              put8(abuf, o1, d + 8);
            } else if (aux_grad && epi.act == M3AE_ACT_QUICKGELU) {
              float d[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = fast_quick_gelu_both(v[j], d[j]);
              put8(abuf, o0, d);
              put8(abuf, o1, d + 8);
            } else if (aux_grad) {
              const bool relu = epi.act == M3AE_ACT_RELU;
              float d[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                d[j] = (!relu || v[j] > 0.f) ? 1.f : 0.f;
                if (relu) v[j] = fmaxf(v[j], 0.f);
              }
              put8(abuf, o0, d);
              put8(abuf, o1, d + 8);
            } else if (epi.act == M3AE_ACT_GELU) {
#pragma unroll
              for (int j = 0; j < 16; j += 2) {
                float d0, d1;  // (unused: the compiler drops the gradient half)
                fast_gelu_both2(v[j], v[j + 1], d0, d1);
              }
            } else if (epi.act == M3AE_ACT_RELU) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
            } else if (epi.act == M3AE_ACT_QUICKGELU) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = fast_quick_gelu(v[j]);
            } else if (epi.act == M3AE_ACT_TANH) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = tanhf(v[j]);
            }
            float in[16];
            if (q.has_in) {
              const uint4 wa = *reinterpret_cast<const uint4*>(iobuf + lane * 64 + o0);
              const uint4 wb = *reinterpret_cast<const uint4*>(iobuf + lane * 64 + o1);
              float2 f;
              f = unpack_bf16x2(wa.x); in[0] = f.x; in[1] = f.y;
              f = unpack_bf16x2(wa.y); in[2] = f.x; in[3] = f.y;
              f = unpack_bf16x2(wa.z); in[4] = f.x; in[5] = f.y;
              f = unpack_bf16x2(wa.w); in[6] = f.x; in[7] = f.y;
              f = unpack_bf16x2(wb.x); in[8] = f.x; in[9] = f.y;
              f = unpack_bf16x2(wb.y); in[10] = f.x; in[11] = f.y;
              f = unpack_bf16x2(wb.z); in[12] = f.x; in[13] = f.y;
              f = unpack_bf16x2(wb.w); in[14] = f.x; in[15] = f.y;
            }
            if (epi.dact != M3AE_ACT_NONE) {
              if (epi.aux_is_grad) {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] *= in[j];
              } else if (epi.dact == M3AE_ACT_GELU) {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] *= fast_gelu_grad(in[j]);
              } else if (epi.dact != M3AE_ACT_RELU) {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] *= act_exact_grad(epi.dact, in[j]);
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = in[j] > 0.f ? v[j] : 0.f;
              }
            }
            if (epi.drop_p > 0.f) {
              const float inv = drop_keep_scale(epi.drop_p);
              uint32_t bits16;
              if (epi.drop_bits_ready) {
                // keep bits generated ahead of the GEMM (m3ae_dropout_bits), loaded before the accumulator wait
                bits16 = ((c == 0 ? cached_bits[0] : cached_bits[1]) >> (16 * hh)) & 0xFFFFu;
              } else {
                const uint32_t thr15 = drop_thr15(epi.drop_p);
                const uint64_t e0 = (uint64_t)(row * epi.ldo + col + 16 * hh);  // multiple of 8 (ldo % 8 == 0)
                bits16 = dropout_keep8(epi.drop_seed, epi.drop_stream, e0, thr15) |
                         (dropout_keep8(epi.drop_seed, epi.drop_stream, e0 + 8, thr15) << 8);
                keep_word |= bits16 << (16 * hh);
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = ((bits16 >> j) & 1u) ? v[j] * inv : 0.f;
            }
            if (q.in_is_residual) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += in[j];
            }
            put8(iobuf, o0, v);
            put8(iobuf, o1, v + 8);
          }
          if (epi.drop_p > 0.f && !epi.drop_bits_ready && epi.drop_bits != nullptr && row < q.M)
            epi.drop_bits[row * epi.drop_bits_ld + (col >> 5)] = keep_word;
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&q.map_out, iobuf, col, row0);
            if (aux) tma_store_2d(&q.map_aux, abuf, col, row0);
            tma_store_commit();
          }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        continue;
      }
      // ---- generic per-element epilogue (odd N, fp32 / atomically accumulated outputs): 16 columns at a time
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      const int64_t row = (int64_t)row0 + lane;
      const bool row_ok = row < q.M;
      float best = -INFINITY;  // row maximum of this warp's 64 columns and its column (epi.amax_val)
      int best_col = col_base < q.N ? col_base : 0;
      // Arg-max only (the decoder's vocabulary projection: bias, no output tensor): the general path below spends ~25 instructions
      // per element (scalar bias loads, run-time epilogue switches, a 16-deep dependent compare-select chain) — 26k warp
      // instructions per 128x256 tile at 19 % issue rate, 30k of the kernel's 34k cycles (profiles/r2_ncu_vocab_gemm.txt).
      // Here: vector bias loads, a max tree, and an index search only in the chunk that raises the maximum.
      const bool amax_only = epi.amax_val != nullptr && epi.out == nullptr && epi.aux_out == nullptr && epi.act == M3AE_ACT_NONE &&
                             epi.dact == M3AE_ACT_NONE && epi.residual == nullptr && epi.drop_p == 0.f && epi.alpha == 1.0f &&
                             (epi.bias == nullptr || (((uintptr_t)epi.bias) & 15) == 0);
      if (amax_only) {
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          const int col = col_base + c * 16;
          if (col >= q.N || !cg_live) break;  // warp-uniform
          uint32_t r[16];
          tmem_ld_32x32b_x16(tacc + (uint32_t)(c * 16), r);
          tmem_ld_wait();
          const int n_valid = min(16, q.N - col);
          float v[16];
          if (n_valid == 16) {
            if (epi.bias != nullptr) {
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 bq = __ldg(reinterpret_cast<const float4*>(epi.bias + col) + j4);
                v[4 * j4 + 0] = __uint_as_float(r[4 * j4 + 0]) + bq.x; v[4 * j4 + 1] = __uint_as_float(r[4 * j4 + 1]) + bq.y;
                v[4 * j4 + 2] = __uint_as_float(r[4 * j4 + 2]) + bq.z; v[4 * j4 + 3] = __uint_as_float(r[4 * j4 + 3]) + bq.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = j < n_valid ? __uint_as_float(r[j]) + (epi.bias ? __ldg(epi.bias + col + j) : 0.f) : -INFINITY;
          }
          float m8[8], m4[4];
#pragma unroll
          for (int j = 0; j < 8; ++j) m8[j] = fmaxf(v[2 * j], v[2 * j + 1]);
#pragma unroll
          for (int j = 0; j < 4; ++j) m4[j] = fmaxf(m8[2 * j], m8[2 * j + 1]);
          const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
          if (m > best) {  // strict: an equal maximum in a later chunk does not displace the first one
            int jj = 15;
#pragma unroll
            for (int j = 14; j >= 0; --j) jj = v[j] == m ? j : jj;  // first column of the chunk that holds the maximum
            best = m;
            best_col = col + jj;
          }
        }
        if (row_ok && cg_live && col_base < q.N) {
          const int64_t g = (int64_t)(col_base >> 6) * epi.amax_ld + row;
          epi.amax_val[g] = best;
          epi.amax_idx[g] = best_col;
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        continue;
      }
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const int col = col_base + c * 16;
        if (col >= q.N || !cg_live) break;  // warp-uniform
        uint32_t r[16];
        tmem_ld_32x32b_x16(tacc + (uint32_t)(c * 16), r);
        tmem_ld_wait();
        if (row_ok) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          const int n_valid = min(16, q.N - col);
          epilogue_apply<16>(epi, row, col, n_valid, v);
          if (epi.amax_val != nullptr) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < n_valid && v[j] > best) { best = v[j]; best_col = col + j; }  // ascending columns, strict >: first maximum
            if (epi.out == nullptr) continue;
          }
          const size_t o = (size_t)(row * epi.ldo + col);
          if (epi.accumulate) {
            float* out = (float*)epi.out + o;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < n_valid) atomicAdd(out + j, v[j]);
          } else if (epi.out_dtype == M3AE_BF16) {
            bf16* out = (bf16*)epi.out + o;
            if (n_valid == 16 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                uint4 w;
                w.x = pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
                w.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
                w.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
                w.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
                ((uint4*)out)[j] = w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < n_valid) out[j] = __float2bfloat16_rn(v[j]);
            }
          } else {
            float* out = (float*)epi.out + o;
            if (n_valid == 16 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j) ((float4*)out)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < n_valid) out[j] = v[j];
            }
          }
        }
      }
      if (epi.amax_val != nullptr && row_ok && cg_live && col_base < q.N) {
        const int64_t g = (int64_t)(col_base >> 6) * epi.amax_ld + row;
        epi.amax_val[g] = best;
        epi.amax_idx[g] = best_col;
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) release_acc(acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all();  // staged output tiles must be drained before the CTA's smem goes away
  }

  tcgen05_fence_before();
  if (CL >= 2) cluster_sync_all();  // no CTA may leave while another still multicasts onto its barriers / into its smem
  else __syncthreads();
  if (warp == WARP_ALLOC) {
    tcgen05_fence_after();
    if (CTAS == 2) tmem_dealloc_pair<TMEM_COLS>(tmem_base);
    else tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------- host side
PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
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

// Operand stored as [rows, cols] bf16 row-major with leading dimension ld; box = [box_rows, 64 cols], SWIZZLE_128B.
int encode_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  auto fn = get_encode_fn();
  M3AE_CHECK(fn != nullptr, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  M3AE_CHECK(((uintptr_t)base & 15) == 0, M3AE_ERR_INVALID, "tcgen05 GEMM operand base must be 16-byte aligned");
  M3AE_CHECK((ld * 2) % 16 == 0, M3AE_ERR_INVALID, "tcgen05 GEMM operand row pitch must be a multiple of 16 bytes (ld=%lld)", (long long)ld);
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)(ld * 2)};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  M3AE_CHECK(r == CUDA_SUCCESS, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return M3AE_OK;
}

// bf16 [rows, cols] row-major (ld) viewed through 32x32 boxes with SWIZZLE_64B: the epilogue staging geometry.
int encode_epi_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld) {
  auto fn = get_encode_fn();
  M3AE_CHECK(fn != nullptr, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)(ld * 2)};
  cuuint32_t box[2] = {32u, 32u};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  M3AE_CHECK(r == CUDA_SUCCESS, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled (epilogue) failed with CUresult %d", (int)r);
  return M3AE_OK;
}

// fp32 [rows, cols] row-major (ld) viewed through 32x32 boxes (128-byte rows) with SWIZZLE_128B: reduce-add staging.
int encode_f32_tile_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld) {
  auto fn = get_encode_fn();
  M3AE_CHECK(fn != nullptr, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)(ld * 4)};
  cuuint32_t box[2] = {32u, 32u};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  M3AE_CHECK(r == CUDA_SUCCESS, M3AE_ERR_CUDA, "cuTensorMapEncodeTiled (fp32 reduce) failed with CUresult %d", (int)r);
  return M3AE_OK;
}

// Pure fp32 accumulation (the split-K weight gradients): out += alpha * acc with nothing else in the epilogue.
bool epilogue_is_reduce(const m3ae_gemm_problem_t& g) {
  const m3ae_gemm_epilogue_t& e = g.epi;
  if (e.out_dtype != M3AE_F32 || !e.accumulate) return false;
  if (e.bias || e.aux_out || e.aux_in || e.residual || e.act != M3AE_ACT_NONE || e.dact != M3AE_ACT_NONE || e.drop_p > 0.f) return false;
  return e.out != nullptr && (((uintptr_t)e.out) & 15) == 0 && (e.ldo % 4) == 0;
}

bool tma_ok(const void* p, int64_t ld) { return p != nullptr && (((uintptr_t)p) & 15) == 0 && (ld % 8) == 0; }

// The TMA-staged epilogue covers the layer's bf16 cases: bf16 out (no atomics), N % 32 == 0, alpha == 1, at most one
// bf16 input operand (residual or aux_in), optional bf16 aux_out.  Everything else uses the generic per-element path.
bool epilogue_is_fast(const m3ae_gemm_problem_t& g) {
  const m3ae_gemm_epilogue_t& e = g.epi;
  if (e.amax_val != nullptr || e.out == nullptr) return false;
  if (e.out_dtype != M3AE_BF16 || e.accumulate || e.alpha != 1.0f || (g.N % 32) != 0) return false;
  if (!tma_ok(e.out, e.ldo)) return false;
  if (e.aux_out && !tma_ok(e.aux_out, e.ldo)) return false;
  if (e.bias && (((uintptr_t)e.bias) & 15)) return false;
  if (e.residual && e.aux_in) return false;
  if (e.residual && (e.res_dtype != M3AE_BF16 || !tma_ok(e.residual, e.ldr))) return false;
  if (e.aux_in && (e.aux_in_dtype != M3AE_BF16 || !tma_ok(e.aux_in, e.ld_aux_in))) return false;
  if (e.dact != M3AE_ACT_NONE && !e.aux_in) return false;
  if (e.drop_p > 0.f && (e.ldo % 8) != 0) return false;
  return true;
}

}  // namespace

void epi_to_dev(const m3ae_gemm_epilogue_t& e, EpiDev* d);
void count_launch();

bool force_generic_epilogue = false;  // test hook (m3ae_set_gemm_epilogue): run the per-element epilogue everywhere
int gemm_cta_pair_mode = 0;           // m3ae_set_gemm_cta_pair: 0 auto, 1 always single CTA, 2 always CTA pair
int gemm_skinny_mode = 0;             // m3ae_set_gemm_skinny: 0 auto, 1 never, 2 always (single-CTA launches only)
int gemm_reserved_sms = 0;            // m3ae_set_gemm_reserved_sms: SMs the persistent kernel leaves to other streams

namespace {
// Clusters of CL CTAs that can be resident at once (the kernel is persistent: a cluster that has to wait for another one to
// retire would serialise the launch).  Queried once per cluster size.
template <int CL, int BNV>
int max_clusters() {
  static std::once_flag once;
  static int n = 0;
  std::call_once(once, [] {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(sm_count() / CL * CL));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = Cfg<CL, BNV>::SMEM_BYTES;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int v = 0;
    if (cudaOccupancyMaxActiveClusters(&v, gemm_tcgen05_kernel<CL, BNV>, &cfg) == cudaSuccess && v > 0) n = v;
    else { (void)cudaGetLastError(); n = sm_count() / CL; }
    if (n > sm_count() / CL) n = sm_count() / CL;
  });
  return n;
}

template <int CL, int BNV>
int launch_tc(TcParams& P, int units, cudaStream_t stream) {
  static std::once_flag attr_once[64];
  static cudaError_t attr_errs[64];
  const int slot = current_device_slot();
  std::call_once(attr_once[slot], [slot] {
    attr_errs[slot] = cudaFuncSetAttribute(gemm_tcgen05_kernel<CL, BNV>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<CL, BNV>::SMEM_BYTES);
  });
  const cudaError_t attr_err = attr_errs[slot];
  M3AE_CHECK(attr_err == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(smem=%d) failed: %s", Cfg<CL, BNV>::SMEM_BYTES,
             cudaGetErrorString(attr_err));
  int workers_max = CL == 1 ? sm_count() : max_clusters<CL, BNV>();
  // SMs left free for concurrently running kernels of other streams (the NCCL gradient all-reduce of data-parallel fine-tuning:
  // a persistent GEMM on every SM makes the collective's CTAs wait for a GEMM launch to retire) — m3ae_set_gemm_reserved_sms
  if (gemm_reserved_sms > 0) {
    const int keep = (sm_count() - gemm_reserved_sms) / CL;
    if (keep >= 1 && keep < workers_max) workers_max = keep;
  }
  const int workers = units < workers_max ? units : workers_max;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)(workers * CL));
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = Cfg<CL, BNV>::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_on() ? 2 : 1;
  M3AE_CUDA(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<CL, BNV>, P));
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}
}  // namespace

int gemm_tcgen05_launch(const m3ae_gemm_problem_t* probs, int n, cudaStream_t stream) {
  M3AE_CHECK(n >= 1 && n <= MAX_GROUP, M3AE_ERR_INVALID, "gemm group size %d not in [1,%d]", n, MAX_GROUP);
  // CTA pairs (256-row tiles) unless every problem fits one 128-row tile (the decoder's per-token GEMMs, M = batch:
  // there the second CTA would only multiply zero rows and halve the number of SMs streaming the weights).
  int max_m = 0;
  for (int i = 0; i < n; ++i) max_m = probs[i].M > max_m ? probs[i].M : max_m;
  // mode 0 (auto): CTA pairs; 1: single CTA; 2: one pair per cluster; 3: clusters of two pairs with operand multicast.
  // Mode 3 is correct (bit-identical, tests/test_kernels_gpu.py) but measured SLOWER on this part: 4-CTA clusters leave
  // TPCs idle in GPCs with an odd TPC count and the two pairs advance in lock step (profiles/r2_gemm_stages_ab.txt:
  // QKV forward 72.9 vs 67.7 us, one-wave split-K wgrads 112 vs 71 us) — the kernel is latency-, not L2-bandwidth-bound.
  const int cl = gemm_cta_pair_mode == 1 ? 1 : gemm_cta_pair_mode == 3 ? 4 : gemm_cta_pair_mode == 2 ? 2 : (max_m > BM ? 2 : 1);
  const int ctas = cl >= 2 ? 2 : 1;
  const int tile_m = BM * ctas;
  // skinny tiles when one CTA per tile is used anyway and 256-column tiles would occupy less than a third of the SMs
  int wide_tiles = 0;
  for (int i = 0; i < n; ++i) wide_tiles += ((probs[i].M + BM - 1) / BM) * ((probs[i].N + BN_WIDE - 1) / BN_WIDE) * (probs[i].split_k > 1 ? probs[i].split_k : 1);
  const int BN = (cl == 1 && gemm_skinny_mode != 1 && (gemm_skinny_mode == 2 || wide_tiles * 3 < sm_count())) ? BN_SKINNY : BN_WIDE;
  const int b_rows = BN / ctas;
  TcParams P;
  memset(&P, 0, sizeof(P));
  P.n_problems = n;
  int units = 0;
  for (int i = 0; i < n; ++i) {
    const m3ae_gemm_problem_t& g = probs[i];
    TcProblem& q = P.p[i];
    M3AE_CHECK(g.M > 0 && g.N > 0 && g.K > 0, M3AE_ERR_INVALID, "gemm problem %d has empty shape %dx%dx%d", i, g.M, g.N, g.K);
    M3AE_CHECK(g.split_k >= 1, M3AE_ERR_INVALID, "split_k must be >= 1");
    M3AE_CHECK(g.split_k == 1 || (g.epi.accumulate && g.epi.out_dtype == M3AE_F32), M3AE_ERR_INVALID,
               "split_k > 1 requires accumulate=1 and an fp32 output");
    M3AE_CHECK(!g.epi.accumulate || g.epi.out_dtype == M3AE_F32, M3AE_ERR_INVALID, "accumulate requires an fp32 output");
    // CL = 4: which operand the two pairs of a cluster share.  Stack the pairs along the dimension with an even number of
    // tiles (the odd tile of an odd count is computed against zero-filled operands and clipped away); ties go to M.
    const int tm1 = (g.M + tile_m - 1) / tile_m, tn1 = (g.N + BN - 1) / BN;
    q.share = 0;
    if (cl == 4) {
      const double waste_m = (double)(tm1 & 1) / (double)(tm1 + (tm1 & 1)), waste_n = (double)(tn1 & 1) / (double)(tn1 + (tn1 & 1));
      q.share = waste_m <= waste_n ? 1 : 2;
    }
    int rc;
    if (g.a_major == M3AE_MAJOR_K) rc = encode_map(&q.map_a, g.a, g.M, g.K, g.lda, q.share == 2 ? BM / 2 : BM);
    else rc = encode_map(&q.map_a, g.a, g.K, g.M, g.lda, BK);
    if (rc) return rc;
    if (g.b_major == M3AE_MAJOR_K) rc = encode_map(&q.map_b, g.b, g.N, g.K, g.ldb, q.share == 1 ? b_rows / 2 : b_rows);
    else rc = encode_map(&q.map_b, g.b, g.K, g.N, g.ldb, BK);
    if (rc) return rc;
    epi_to_dev(g.epi, &q.epi);
    q.epi.drop_bits_ld = g.N / 32;
    q.epi.amax_ld = g.M;
    q.epi_fast = (!force_generic_epilogue && epilogue_is_fast(g)) ? 1 : 0;
    if (!force_generic_epilogue && epilogue_is_reduce(g)) q.epi_fast = 2;
    q.has_in = 0;
    q.in_is_residual = 0;
    if (q.epi_fast == 2) {
      rc = encode_f32_tile_map(&q.map_out, g.epi.out, g.M, g.N, g.epi.ldo);
      if (rc) return rc;
    } else if (q.epi_fast) {
      rc = encode_epi_map(&q.map_out, g.epi.out, g.M, g.N, g.epi.ldo);
      if (rc) return rc;
      if (g.epi.aux_out) {
        rc = encode_epi_map(&q.map_aux, g.epi.aux_out, g.M, g.N, g.epi.ldo);
        if (rc) return rc;
      }
      if (g.epi.residual) {
        rc = encode_epi_map(&q.map_in, g.epi.residual, g.M, g.N, g.epi.ldr);
        q.has_in = 1; q.in_is_residual = 1;
      } else if (g.epi.aux_in) {
        rc = encode_epi_map(&q.map_in, g.epi.aux_in, g.M, g.N, g.epi.ld_aux_in);
        q.has_in = 1;
      }
      if (rc) return rc;
    }
    q.M = g.M; q.N = g.N; q.K = g.K;
    q.a_major = g.a_major; q.b_major = g.b_major;
    q.tiles_m = q.share == 1 ? (tm1 + 1) / 2 : tm1;
    q.tiles_n = q.share == 2 ? (tn1 + 1) / 2 : tn1;
    q.kblocks = (g.K + BK - 1) / BK;
    int split = g.split_k < q.kblocks ? g.split_k : q.kblocks;
    q.kb_per_split = (q.kblocks + split - 1) / split;
    q.split_k = (q.kblocks + q.kb_per_split - 1) / q.kb_per_split;  // no empty splits
    // Rasterisation.  Consecutive units run at the same time on neighbouring CTA pairs, so whatever they share is read
    // from DRAM once and from L2 afterwards.  With M fastest the N tiles of one M tile are a whole sweep over M apart:
    // for the K = 3072 GEMMs the A operand (140 MB at batch 64) exceeds the 126 MB L2 and was re-read once per N tile
    // (ncu: 441 MB read against 180 MB algorithmic, profiles/r1_ncu_gemm_full_v9_cta_pair.txt).  N fastest whenever A
    // is the larger operand (forward / dgrad: A = activations, B = weights); split-K weight gradients keep M fastest
    // (their units form about one wave that walks K in lockstep).
    q.n_fastest = (q.split_k == 1 && (int64_t)g.M >= (int64_t)g.N) ? 1 : 0;
    q.unit_begin = units;
    units += q.tiles_m * q.tiles_n * q.split_k;
    q.unit_end = units;
  }
  P.total_units = units;
  if (cl == 4) return launch_tc<4, BN_WIDE>(P, units, stream);
  if (cl == 2) return launch_tc<2, BN_WIDE>(P, units, stream);
  return BN == BN_SKINNY ? launch_tc<1, BN_SKINNY>(P, units, stream) : launch_tc<1, BN_WIDE>(P, units, stream);
}

}  // namespace m3ae
