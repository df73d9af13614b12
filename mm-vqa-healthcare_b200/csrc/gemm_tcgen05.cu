// Persistent, warp-specialised bf16 GEMM for sm_100a:
//   TMA (cp.async.bulk.tensor, SWIZZLE_128B) -> 4-stage smem ring -> tcgen05.mma (cta_group::1, 128x256x16,
//   fp32 accumulators double-buffered in TMEM) -> tcgen05.ld epilogue (bias / GELU / act' / dropout /
//   residual / split-K atomics) -> global.
// One launch runs up to M3AE_MAX_GROUP independent problems (the text and image streams of a fusion
// layer pair, reference m3ae_module.py:276-277), each with K-major or MN-major operands so the same kernel
// serves forward (A K-major, B K-major), dgrad (B MN-major) and wgrad (A and B MN-major, split-K).
//
// Algorithmic traffic per problem: (M*K + N*K) * 2 B in, M*N * sizeof(out) out; 2*M*N*K flop.
#include <cudaTypedefs.h>

#include <mutex>

#include "common.cuh"

namespace m3ae {

namespace {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int STAGES = 4;
constexpr int A_STAGE_BYTES = BM * BK * 2;
constexpr int B_STAGE_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int NUM_THREADS = 384;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4..11 epilogue
constexpr int NUM_EPI_WARPS = 8;
constexpr int TMEM_COLS = 512;  // 2 accumulator stages x 256 fp32 columns
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int MAX_GROUP = 4;

struct alignas(64) TcProblem {
  CUtensorMap map_a;
  CUtensorMap map_b;
  EpiDev epi;
  int M, N, K;
  int a_major, b_major;
  int tiles_m, tiles_n, split_k, kblocks, kb_per_split;
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
  u.m_blk = t % q.tiles_m;
  u.n_blk = t / q.tiles_m;
  u.kb0 = split * q.kb_per_split;
  u.kb1 = min(q.kblocks, u.kb0 + q.kb_per_split);
  return u;
}

__global__ void __launch_bounds__(NUM_THREADS, 1) gemm_tcgen05_kernel(const __grid_constant__ TcParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + STAGES * STAGE_BYTES);
  uint64_t* full_bar = bars;                  // [STAGES]
  uint64_t* empty_bar = bars + STAGES;        // [STAGES]
  uint64_t* tmem_full = bars + 2 * STAGES;    // [2]
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_ptr_smem = (uint32_t*)(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < P.n_problems; ++i) {
      tma_prefetch_desc(&P.p[i].map_a);
      tma_prefetch_desc(&P.p[i].map_b);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], NUM_EPI_WARPS);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_ptr_smem);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int unit = blockIdx.x; unit < P.total_units; unit += gridDim.x) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      for (int kb = u.kb0; kb < u.kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * STAGE_BYTES;
        uint8_t* sb = sa + A_STAGE_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        if (q.a_major == M3AE_MAJOR_K) {
          tma_load_2d(sa, &q.map_a, &full_bar[stage], kb * BK, u.m_blk * BM);
        } else {
#pragma unroll
          for (int j = 0; j < BM / 64; ++j)
            tma_load_2d(sa + j * (BK * 128), &q.map_a, &full_bar[stage], u.m_blk * BM + j * 64, kb * BK);
        }
        if (q.b_major == M3AE_MAJOR_K) {
          tma_load_2d(sb, &q.map_b, &full_bar[stage], kb * BK, u.n_blk * BN);
        } else {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_2d(sb + j * (BK * 128), &q.map_b, &full_bar[stage], u.n_blk * BN + j * 64, kb * BK);
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && lane == 0) {
    // ===================== MMA issuer (single thread) =====================
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int unit = blockIdx.x; unit < P.total_units; unit += gridDim.x) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      const uint32_t idesc = (1u << 4) /*D=f32*/ | (1u << 7) /*A=bf16*/ | (1u << 10) /*B=bf16*/ |
                             ((uint32_t)q.a_major << 15) | ((uint32_t)q.b_major << 16) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);
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
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          const uint64_t da = make_smem_desc(sa + k * a_kstep, a_lbo, 1024);
          const uint64_t db = make_smem_desc(sb + k * b_kstep, b_lbo, 1024);
          umma_f16(tmem_d, da, db, idesc, (kb > u.kb0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(&tmem_full[acc]);  // accumulator complete -> epilogue
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int e = warp - 4;
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int half = e >> 2;    // which 128-column half of the accumulator
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int unit = blockIdx.x; unit < P.total_units; unit += gridDim.x) {
      const Unit u = decode_unit(P, unit);
      const TcProblem& q = P.p[u.p];
      const EpiDev& epi = q.epi;
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      const int64_t row = (int64_t)u.m_blk * BM + quad * 32 + lane;
      const bool row_ok = row < q.M;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const int col = u.n_blk * BN + half * 128 + c * 32;
        if (col >= q.N) break;  // warp-uniform
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + half * 128 + c * 32);
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr, r);
        tmem_ld_wait();
        if (row_ok) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          const int n_valid = min(32, q.N - col);
          epilogue_apply<32>(epi, row, col, n_valid, v);
          const size_t o = (size_t)(row * epi.ldo + col);
          if (epi.accumulate) {
            float* out = (float*)epi.out + o;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < n_valid) atomicAdd(out + j, v[j]);
          } else if (epi.out_dtype == M3AE_BF16) {
            bf16* out = (bf16*)epi.out + o;
            if (n_valid == 32 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 w;
                w.x = pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
                w.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
                w.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
                w.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
                ((uint4*)out)[j] = w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < n_valid) out[j] = __float2bfloat16_rn(v[j]);
            }
          } else {
            float* out = (float*)epi.out + o;
            if (n_valid == 32 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
              for (int j = 0; j < 8; ++j) ((float4*)out)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < n_valid) out[j] = v[j];
            }
          }
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
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

}  // namespace

void epi_to_dev(const m3ae_gemm_epilogue_t& e, EpiDev* d);
void count_launch();

int gemm_tcgen05_launch(const m3ae_gemm_problem_t* probs, int n, cudaStream_t stream) {
  M3AE_CHECK(n >= 1 && n <= MAX_GROUP, M3AE_ERR_INVALID, "gemm group size %d not in [1,%d]", n, MAX_GROUP);
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    attr_err = cudaFuncSetAttribute(gemm_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
  });
  M3AE_CHECK(attr_err == cudaSuccess, M3AE_ERR_CUDA, "cudaFuncSetAttribute(smem=%d) failed: %s", SMEM_BYTES,
             cudaGetErrorString(attr_err));
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
    int rc;
    if (g.a_major == M3AE_MAJOR_K) rc = encode_map(&q.map_a, g.a, g.M, g.K, g.lda, BM);
    else rc = encode_map(&q.map_a, g.a, g.K, g.M, g.lda, BK);
    if (rc) return rc;
    if (g.b_major == M3AE_MAJOR_K) rc = encode_map(&q.map_b, g.b, g.N, g.K, g.ldb, BN);
    else rc = encode_map(&q.map_b, g.b, g.K, g.N, g.ldb, BK);
    if (rc) return rc;
    epi_to_dev(g.epi, &q.epi);
    q.M = g.M; q.N = g.N; q.K = g.K;
    q.a_major = g.a_major; q.b_major = g.b_major;
    q.tiles_m = (g.M + BM - 1) / BM;
    q.tiles_n = (g.N + BN - 1) / BN;
    q.kblocks = (g.K + BK - 1) / BK;
    int split = g.split_k < q.kblocks ? g.split_k : q.kblocks;
    q.kb_per_split = (q.kblocks + split - 1) / split;
    q.split_k = (q.kblocks + q.kb_per_split - 1) / q.kb_per_split;  // no empty splits
    q.unit_begin = units;
    units += q.tiles_m * q.tiles_n * q.split_k;
    q.unit_end = units;
  }
  P.total_units = units;
  const int sms = sm_count();
  const int grid = units < sms ? units : sms;
  gemm_tcgen05_kernel<<<grid, NUM_THREADS, SMEM_BYTES, stream>>>(P);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
