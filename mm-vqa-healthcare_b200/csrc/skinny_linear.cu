// Skinny linear layer for the KV-cached decoder step (m3ae_decoder.py:54-90 evaluated for ONE new token per sequence):
//   out[M, N] = act( LN?(x)[M, K] . W[N, K]^T + bias ) (+ residual),  M <= 128 (= batch), bf16 operands, fp32 accumulate.
//
// Why a separate kernel.  At M = 128 these projections are 1-5 MB of weights each: pure latency.  On the persistent
// tcgen05 kernel every launch pays TMEM allocation, barrier set-up, tensor-map fetches and a 12-step TMA->MMA pipeline per
// tile, and only N/64 CTAs stream the weights: ~16 us per launch, six launches per token
// (profiles/r2_launches_decoder_b_summary.txt: 98 of 156 us per token step).  Here the WHOLE activation panel
// [128 x K-chunk] (<= 194 KB) and the CTA's weight slice [NC x K-chunk] are brought into shared memory by one burst of
// row-sized bulk copies (cp.async.bulk onto one mbarrier: one L2/DRAM round trip), every SM takes NC = 16 (or 32) output
// columns, and the arithmetic is mma.sync.m16n8k16 on register fragments (0.2-0.6 GFLOP per launch: three orders of
// magnitude below the tensor roofline, so tcgen05/TMEM plumbing would only add latency).
//   * 2-D tiles: a CTA takes MTR rows x NC columns (128x16, 128x32, 64x32, 64x48 or 32x32, chosen so that about one wave of
//     CTAs covers the output and the bytes a CTA stages — (MTR + NC) x K-chunk x 2 — are smallest): one SM ingests only
//     ~24 B/clk from L2/HBM through bulk copies (profiles/r2_ncu_skinny.txt: 220 KB per CTA = 5 us of an 8-10 us launch), so
//     the staged bytes per CTA, not the weight bytes, set the launch time.
//   * LayerNorm folded into the A load (pre_norm -> QKV, layernorm1 -> cross Q, layernorm2 -> FFN-up): each warp normalises
//     ITS 16 rows in shared memory (same two-pass fp32 statistics and the same summation order as ln_fwd_kernel, so the bf16
//     operand is bit-identical to what the separate LayerNorm launch wrote) — three launches per token disappear.
//   * Segmented output: the QKV projection writes q and the K / V cache rows [b, t, :] (three pointers, three pitches).
//   * Split-K for K = 3072 (FFN-down): K-chunks of 512 across CTAs, fp32 partials added into a zeroed workspace, the last
//     arriving CTA of a column tile (atomic counter) applies bias + residual, writes bf16 and re-zeroes workspace and counter.
#include "common.cuh"

// Scratch of the allocation-free entry point m3ae_skinny_linear_static (bottom of this file): the split-K workspace
// [128 rows x SKINNY_STATIC_MAX_N columns] fp32 followed by the 256 tile counters.  513 KB of device memory on every GPU that
// loads the library, held until the process exits.  In host code this NAME is only a stand-in for the symbol — the device
// address comes from cudaGetSymbolAddress, once per device.
#define SKINNY_STATIC_MAX_N 1024
#define SKINNY_STATIC_WS_BYTES (128 * SKINNY_STATIC_MAX_N * 4)
#define SKINNY_STATIC_BYTES (SKINNY_STATIC_WS_BYTES + 256 * 4)
static __device__ __align__(256) unsigned char g_skinny_static_scratch[SKINNY_STATIC_BYTES];

namespace m3ae {
void count_launch();

namespace {

constexpr int MT = 128;        // rows per CTA (the whole batch)
constexpr int NTH = 256;       // 8 warps x 16 rows
constexpr int KC_MAX = 768;    // largest K-chunk held in shared memory
constexpr int PAD = 8;         // bf16 elements of row padding (16 B): conflict-free ldmatrix

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t saddr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}

struct SkinnyDev {
  m3ae_skinny_params_t p;
  int kc;        // K-chunk in shared memory
  int ntiles;    // ceil(N / NC)
  int nrt;       // row tiles: ceil(M / MTR)
};

template <int MTR, int NC>
__global__ void __launch_bounds__(NTH, 1) skinny_linear_kernel(const SkinnyDev S) {
  constexpr int RW = MTR / 16;        // row warps
  constexpr int CW = 8 / RW;          // column warps
  constexpr int WC = NC / CW;         // columns per warp (multiple of 8)
  constexpr int NTL = WC / 8;         // n-tiles per warp
  static_assert(MTR % 16 == 0 && 8 % RW == 0 && NC % CW == 0 && WC % 8 == 0, "tile shape");
  pdl_sync();
  extern __shared__ __align__(16) uint8_t smem[];
  const m3ae_skinny_params_t& P = S.p;
  const int KC = S.kc, ksplit = P.ksplit, M = P.M, N = P.N;
  const int pitch = (KC + PAD) * 2;  // bytes per smem row
  const uint32_t sA = smem_u32(smem), sW = sA + MTR * pitch;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int mi = lane >> 3, r8 = lane & 7;
  const bf16* __restrict__ X = (const bf16*)P.x;
  const bf16* __restrict__ W = (const bf16*)P.w;
  int& s_last = *reinterpret_cast<int*>(smem + (MTR + NC) * pitch);  // after the two tiles (the launcher adds 32 bytes)
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + (MTR + NC) * pitch + 8);
  const int rw = warp % RW, cw = warp / RW;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  __syncthreads();
  uint32_t phase = 0;
  const uint32_t row_bytes = (uint32_t)KC * 2;

  int a_chunk = -1;  // (row tile, K-chunk) currently resident in sA
  const int nrt = S.nrt;
  const int nunits = S.ntiles * ksplit * nrt;
  // unit = ((column tile * ksplit) + K-chunk) * nrt + row tile; the grid is a multiple of nrt, so a CTA's units share its row tile
  for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
    const int rt = unit % nrt, rest = unit / nrt;
    const int tile = rest / ksplit, ks = rest - tile * ksplit;
    const int n0 = tile * NC, k0 = ks * KC, m0 = rt * MTR;
    const int m_valid = M - m0 < MTR ? (M - m0 > 0 ? M - m0 : 0) : MTR;
    // ---- loads: one bulk copy (cp.async.bulk, 1-1.5 KB) per row of the CTA's weight slice and — unless it is already
    // resident — of the activation panel, all completing on one mbarrier.  (The first version issued 13.8k 16-byte cp.async
    // per CTA: 16k cycles for 220 KB, ~14 B/clk per SM — profiles/r2_ncu_skinny.txt.)
    const bool load_a = a_chunk != ks * nrt + rt;
    const int w_valid = N - n0 < NC ? N - n0 : NC, a_valid = load_a ? m_valid : 0;
    if (tid == 0) mbar_expect_tx(bar, (uint32_t)(w_valid + a_valid) * row_bytes);
    fence_proxy_async();  // this thread's earlier generic-proxy reads of the tiles are ordered before the async-proxy writes
    for (int r = tid; r < NC + (load_a ? MTR : 0); r += NTH) {
      const bool is_w = r < NC;
      const int rr = is_w ? r : r - NC;
      const uint32_t dst = (is_w ? sW : sA) + rr * pitch;
      if (rr < (is_w ? w_valid : m_valid)) {
        const bf16* src = is_w ? W + (int64_t)(n0 + rr) * P.ldw + k0 : X + (int64_t)(m0 + rr) * P.ldx + k0;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"((uint64_t)src),
                     "r"(row_bytes), "r"(smem_u32(bar))
                     : "memory");
      } else {
        for (uint32_t c = 0; c < row_bytes; c += 16) asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(dst + c), "r"(0u) : "memory");
      }
    }
    if (load_a) a_chunk = ks * nrt + rt;
    mbar_wait(bar, phase);
    phase ^= 1u;
    if (MTR == MT && load_a && P.ln_gamma != nullptr) {  // folded LayerNorm: 128-row tiles only (each warp owns its 16 rows)
      // LayerNorm over the K = KC = 768 columns of this warp's rows, in place (ln_fwd_kernel's arithmetic and order)
      __syncwarp();
      float gam[3][8], bet[3][8];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int c = (i * 32 + lane) * 8;
        const float4 g0 = *reinterpret_cast<const float4*>(P.ln_gamma + c), g1 = *reinterpret_cast<const float4*>(P.ln_gamma + c + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(P.ln_beta + c), b1 = *reinterpret_cast<const float4*>(P.ln_beta + c + 4);
        gam[i][0] = g0.x; gam[i][1] = g0.y; gam[i][2] = g0.z; gam[i][3] = g0.w; gam[i][4] = g1.x; gam[i][5] = g1.y; gam[i][6] = g1.z; gam[i][7] = g1.w;
        bet[i][0] = b0.x; bet[i][1] = b0.y; bet[i][2] = b0.z; bet[i][3] = b0.w; bet[i][4] = b1.x; bet[i][5] = b1.y; bet[i][6] = b1.z; bet[i][7] = b1.w;
      }
      for (int rr = 0; rr < 16; ++rr) {
        const int row = warp * 16 + rr;
        if (row >= M) break;
        const uint32_t base = sA + row * pitch;
        float v[3][8];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          uint32_t w0, w1, w2, w3;
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "r"(base + (i * 32 + lane) * 16));
          const float2 a = unpack_bf16x2(w0), b = unpack_bf16x2(w1), c = unpack_bf16x2(w2), d = unpack_bf16x2(w3);
          v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y; v[i][4] = c.x; v[i][5] = c.y; v[i][6] = d.x; v[i][7] = d.y;
#pragma unroll
          for (int j = 0; j < 8; ++j) s += v[i][j];
        }
        const float mean = warp_sum(s) / (float)KC;
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
        const float var = warp_sum(ss) / (float)KC;
        const float rstd = rsqrtf(var + P.ln_eps);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          float o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * gam[i][j] + bet[i][j];
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(base + (i * 32 + lane) * 16), "r"(pack_bf16x2(o[0], o[1])),
                       "r"(pack_bf16x2(o[2], o[3])), "r"(pack_bf16x2(o[4], o[5])), "r"(pack_bf16x2(o[6], o[7]))
                       : "memory");
        }
      }
    }
    __syncthreads();  // weight slice (all threads' copies) visible; A rows are warp-private but ordered by the same barrier

    // ---- [16 rows x NC] per warp over the K-chunk
    float acc[NTL][4];
#pragma unroll
    for (int j = 0; j < NTL; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    const uint32_t a_lane = sA + (rw * 16 + (mi & 1) * 8 + r8) * pitch + (mi >> 1) * 16;
    // B fragments: one ldmatrix.x4 = one 8-column n-tile x two 16-wide k-steps (k 0-7 | 8-15 | 16-23 | 24-31)
    const uint32_t b_lane = sW + (cw * WC + r8) * pitch + mi * 16;
#pragma unroll 2
    for (int kk2 = 0; kk2 < KC / 32; ++kk2) {
      uint32_t af0[4], af1[4];
      ldsm_x4(af0, a_lane + kk2 * 64);
      ldsm_x4(af1, a_lane + kk2 * 64 + 32);
#pragma unroll
      for (int j = 0; j < NTL; ++j) {
        uint32_t bb[4];
        ldsm_x4(bb, b_lane + j * 8 * pitch + kk2 * 64);
        mma16816(acc[j], af0, bb[0], bb[1]);
        mma16816(acc[j], af1, bb[2], bb[3]);
      }
    }
    if (KC % 32) {  // K-chunk = 16 (mod 32): one more k-step (the second half of the x4 load stays inside the padded row)
      uint32_t af0[4];
      ldsm_x4(af0, a_lane + (KC / 32) * 64);
#pragma unroll
      for (int j = 0; j < NTL; ++j) {
        uint32_t bb[4];
        ldsm_x4(bb, b_lane + j * 8 * pitch + (KC / 32) * 64);
        mma16816(acc[j], af0, bb[0], bb[1]);
      }
    }

    // ---- epilogue
    const int row0 = m0 + rw * 16 + g, row1 = row0 + 8;
    if (ksplit == 1) {
#pragma unroll
      for (int j = 0; j < NTL; ++j) {
        const int n = n0 + cw * WC + j * 8 + 2 * t;
        if (n >= N) continue;  // N is even: a column pair is in or out together
        const float b0 = P.bias ? P.bias[n] : 0.f, b1 = P.bias ? P.bias[n + 1] : 0.f;
        float v00 = acc[j][0] + b0, v01 = acc[j][1] + b1, v10 = acc[j][2] + b0, v11 = acc[j][3] + b1;
        if (P.act == M3AE_ACT_RELU) { v00 = fmaxf(v00, 0.f); v01 = fmaxf(v01, 0.f); v10 = fmaxf(v10, 0.f); v11 = fmaxf(v11, 0.f); }
        const int seg = n / P.seg_cols, cn = n - seg * P.seg_cols;
        bf16* o = (bf16*)P.out[seg];
        const int64_t ldo = P.ldo[seg];
        if (row0 < M) {
          if (P.residual) {
            const float2 rr = unpack_bf16x2(*reinterpret_cast<const uint32_t*>((const bf16*)P.residual + (int64_t)row0 * P.ldr + n));
            v00 += rr.x; v01 += rr.y;
          }
          *reinterpret_cast<uint32_t*>(o + (int64_t)row0 * ldo + cn) = pack_bf16x2(v00, v01);
        }
        if (row1 < M) {
          if (P.residual) {
            const float2 rr = unpack_bf16x2(*reinterpret_cast<const uint32_t*>((const bf16*)P.residual + (int64_t)row1 * P.ldr + n));
            v10 += rr.x; v11 += rr.y;
          }
          *reinterpret_cast<uint32_t*>(o + (int64_t)row1 * ldo + cn) = pack_bf16x2(v10, v11);
        }
      }
    } else {
      // split-K: fp32 partials into the zeroed workspace [MT, N]; the last CTA of the column tile finishes it
#pragma unroll
      for (int j = 0; j < NTL; ++j) {
        const int n = n0 + cw * WC + j * 8 + 2 * t;
        if (n >= N) continue;
        if (row0 < M) { atomicAdd(P.ws + (int64_t)row0 * N + n, acc[j][0]); atomicAdd(P.ws + (int64_t)row0 * N + n + 1, acc[j][1]); }
        if (row1 < M) { atomicAdd(P.ws + (int64_t)row1 * N + n, acc[j][2]); atomicAdd(P.ws + (int64_t)row1 * N + n + 1, acc[j][3]); }
      }
      __threadfence();
      __syncthreads();
      if (tid == 0) s_last = atomicAdd(P.counters + tile * nrt + rt, 1) == ksplit - 1;
      __syncthreads();
      if (s_last) {
        __threadfence();
        for (int i = tid; i < MTR * (NC / 2); i += NTH) {
          const int row = m0 + i / (NC / 2), n = n0 + (i % (NC / 2)) * 2;
          if (row >= M || n >= N) continue;
          float* wsp = P.ws + (int64_t)row * N + n;
          const float2 s2 = __ldcg(reinterpret_cast<const float2*>(wsp));
          float v0 = s2.x + (P.bias ? P.bias[n] : 0.f), v1 = s2.y + (P.bias ? P.bias[n + 1] : 0.f);
          if (P.act == M3AE_ACT_RELU) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
          if (P.residual) {
            const float2 rr = unpack_bf16x2(*reinterpret_cast<const uint32_t*>((const bf16*)P.residual + (int64_t)row * P.ldr + n));
            v0 += rr.x; v1 += rr.y;
          }
          const int seg = n / P.seg_cols, cn = n - seg * P.seg_cols;
          *reinterpret_cast<uint32_t*>((bf16*)P.out[seg] + (int64_t)row * P.ldo[seg] + cn) = pack_bf16x2(v0, v1);
          *reinterpret_cast<float2*>(wsp) = make_float2(0.f, 0.f);  // leave the workspace zeroed for the next launch
        }
        if (tid == 0) P.counters[tile * nrt + rt] = 0;
      }
    }
    __syncthreads();  // before the next unit's copies overwrite the weight slice / activation panel
  }
}

template <int MTR, int NC>
int launch(const SkinnyDev& S, int smem, int grid, cudaStream_t st) {
  static int done[64];
  const int slot = current_device_slot();
  if (!done[slot]) {
    cudaError_t e = cudaFuncSetAttribute(skinny_linear_kernel<MTR, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    M3AE_CHECK(e == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(skinny_linear) failed: %s", cudaGetErrorString(e));
    done[slot] = 1;
  }
  M3AE_LAUNCH((skinny_linear_kernel<MTR, NC>), grid, NTH, smem, st, S);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace

int skinny_linear(const m3ae_skinny_params_t* p, cudaStream_t st) {
  M3AE_CHECK(p != nullptr && p->x && p->w && p->out[0], M3AE_ERR_INVALID, "skinny_linear: null pointer");
  M3AE_CHECK(p->M > 0 && p->M <= MT, M3AE_ERR_UNSUPPORTED, "skinny_linear: M = %d rows (1..%d)", p->M, MT);
  M3AE_CHECK(p->N > 0 && p->N % 2 == 0 && p->K > 0, M3AE_ERR_INVALID, "skinny_linear: N = %d must be even, K = %d", p->N, p->K);
  const int ksplit = p->ksplit > 0 ? p->ksplit : 1;
  M3AE_CHECK(p->K % ksplit == 0, M3AE_ERR_INVALID, "skinny_linear: K = %d not divisible by ksplit = %d", p->K, ksplit);
  const int kc = p->K / ksplit;
  M3AE_CHECK(kc % 16 == 0 && kc <= KC_MAX, M3AE_ERR_UNSUPPORTED, "skinny_linear: K-chunk %d must be a multiple of 16, <= %d", kc, KC_MAX);
  M3AE_CHECK(p->ldx % 8 == 0 && p->ldw % 8 == 0 && ((((uintptr_t)p->x) | ((uintptr_t)p->w)) & 15) == 0, M3AE_ERR_UNSUPPORTED,
             "skinny_linear: x and w need 16-byte aligned rows");
  M3AE_CHECK(p->seg_cols > 0 && p->seg_cols % 32 == 0 && (p->N + p->seg_cols - 1) / p->seg_cols <= 3, M3AE_ERR_INVALID,
             "skinny_linear: seg_cols = %d (multiple of 32, at most 3 segments)", p->seg_cols);
  for (int s = 0; s < (p->N + p->seg_cols - 1) / p->seg_cols; ++s)
    M3AE_CHECK(p->out[s] && p->ldo[s] % 2 == 0 && (((uintptr_t)p->out[s]) & 3) == 0, M3AE_ERR_INVALID, "skinny_linear: output segment %d", s);
  M3AE_CHECK(!p->residual || (p->ldr % 2 == 0 && (((uintptr_t)p->residual) & 3) == 0), M3AE_ERR_INVALID, "skinny_linear: residual alignment");
  M3AE_CHECK(!p->ln_gamma || (ksplit == 1 && p->K == 768 && p->ln_beta && ((((uintptr_t)p->ln_gamma) | ((uintptr_t)p->ln_beta)) & 15) == 0),
             M3AE_ERR_UNSUPPORTED, "skinny_linear: fused LayerNorm needs K = 768, no split-K and 16-byte aligned gamma / beta");
  M3AE_CHECK(ksplit == 1 || (p->ws && p->counters && (((uintptr_t)p->ws) & 7) == 0), M3AE_ERR_INVALID, "skinny_linear: split-K needs a zeroed workspace and counters");
  SkinnyDev S;
  S.p = *p;
  S.p.ksplit = ksplit;
  S.kc = kc;
  // Tile shape: the fewest staged bytes per CTA, (MTR + NC) x kc, among the shapes that cover the output with at least ~2/3 of a
  // wave of CTAs and at most ~2.6 waves; the folded LayerNorm needs the 128-row tile (each warp normalises the rows it owns).
  struct Shape { int mtr, nc; };
  const Shape shapes[] = {{32, 32}, {64, 32}, {64, 48}, {128, 16}, {128, 32}};
  const int sms = sm_count();
  int best = -1;
  long best_cost = 0;
  for (int i = 0; i < 5; ++i) {
    const Shape& sh = shapes[i];
    if (p->ln_gamma && sh.mtr != MT) continue;
    const int smem_i = (sh.mtr + sh.nc) * (kc + PAD) * 2 + 32;
    if (smem_i > 227 * 1024) continue;
    const int nrt_i = (p->M + sh.mtr - 1) / sh.mtr;
    const long units_i = (long)((p->N + sh.nc - 1) / sh.nc) * ksplit * nrt_i;
    const long waves_x100 = units_i * 100 / sms;
    // staged bytes per CTA, counted once per wave the CTA takes part in (the activation panel is re-used, the weights are not)
    const long rounds = (units_i + sms - 1) / sms;
    // + a second dependent load round trip (and, with split-K, a second pass of atomics + tile counters)
    long cost = (long)sh.mtr * kc * 2 + rounds * sh.nc * kc * 2 + (rounds - 1) * (ksplit > 1 ? 60000 : 20000);
    if (waves_x100 < 60) cost += cost * (60 - waves_x100) / 30;  // too few CTAs: the launch's latency chain gets no help
    if (best < 0 || cost < best_cost) { best = i; best_cost = cost; }
  }
  M3AE_CHECK(best >= 0, M3AE_ERR_UNSUPPORTED, "skinny_linear: no tile shape fits (K-chunk %d)", kc);
  const Shape sh = shapes[best];
  S.ntiles = (p->N + sh.nc - 1) / sh.nc;
  S.nrt = (p->M + sh.mtr - 1) / sh.mtr;
  const int smem = (sh.mtr + sh.nc) * (kc + PAD) * 2 + 32;
  const int units = S.ntiles * ksplit * S.nrt;
  int grid = units < sms ? units : sms / S.nrt * S.nrt;  // a multiple of the row-tile count: a CTA keeps its row tile
  if (grid < 1) grid = 1;
  M3AE_CHECK(ksplit == 1 || S.ntiles * S.nrt <= 256, M3AE_ERR_UNSUPPORTED, "skinny_linear: split-K needs %d counters (256 at most)", S.ntiles * S.nrt);
  if (sh.mtr == 32) return launch<32, 32>(S, smem, grid, st);
  if (sh.mtr == 64) return sh.nc == 32 ? launch<64, 32>(S, smem, grid, st) : launch<64, 48>(S, smem, grid, st);
  return sh.nc == 16 ? launch<128, 16>(S, smem, grid, st) : launch<128, 32>(S, smem, grid, st);
}

// Allocation-free variant (m3ae_skinny_linear_static): the same launch with the split-K workspace and tile counters taken from
// g_skinny_static_scratch instead of caller memory (p->ws / p->counters are ignored).  No cudaMalloc / cudaFree / allocating
// library call on this path: callers under CUDA-graph capture or in serving loops that forbid allocation can use it.
// Every call on a device shares the one array: it is cleared on the caller's stream ahead of each split-K launch (the
// kernel leaves it zeroed, but a launch that was aborted would not), and two calls that overlap on the GPU (different
// streams, no dependency between them) race on it.
int skinny_linear_static(const m3ae_skinny_params_t* p, cudaStream_t st) {
  M3AE_CHECK(p != nullptr, M3AE_ERR_INVALID, "skinny_linear_static: null pointer");
  M3AE_CHECK(p->N <= SKINNY_STATIC_MAX_N, M3AE_ERR_UNSUPPORTED,
             "skinny_linear_static: N = %d output columns exceed the static scratch (at most %d columns, %d rows)", p->N, SKINNY_STATIC_MAX_N, MT);
  M3AE_CHECK(p->M <= MT, M3AE_ERR_UNSUPPORTED, "skinny_linear_static: M = %d rows exceed the static scratch (at most %d rows)", p->M, MT);
  static unsigned char* scratch_addr[64];  // device address of the array, per device ordinal
  const int slot = current_device_slot();
  if (scratch_addr[slot] == nullptr) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_skinny_static_scratch);  // for the CURRENT device
    M3AE_CHECK(e == cudaSuccess && ptr != nullptr, M3AE_ERR_CUDA, "cudaGetSymbolAddress(g_skinny_static_scratch) failed: %s", cudaGetErrorString(e));
    scratch_addr[slot] = static_cast<unsigned char*>(ptr);
  }
  m3ae_skinny_params_t q = *p;
  q.ws = reinterpret_cast<float*>(scratch_addr[slot]);
  q.counters = reinterpret_cast<int32_t*>(scratch_addr[slot] + SKINNY_STATIC_WS_BYTES);
  if (q.ksplit > 1) M3AE_CUDA(cudaMemsetAsync(scratch_addr[slot], 0, SKINNY_STATIC_BYTES, st));
  return skinny_linear(&q, st);
}

}  // namespace m3ae
