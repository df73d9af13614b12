// fp32-FFMA tiled GEMM (strict-parity path, also the on-device cross-check for the tcgen05 kernel).
// out = epi(alpha * A[M,K] · B[N,K]^T), A/B fp32 or bf16, any major (strided reads), fp32 accumulate.
// This is the path behind the "fp32 logits within 1e-4" gate: tcgen05 has no full-fp32 input kind
// (kind::tf32 keeps 10 mantissa bits), so fp32 parity runs use FFMA.
#include "common.cuh"

namespace m3ae {

void epi_to_dev(const m3ae_gemm_epilogue_t& e, EpiDev* d) {
  d->out = e.out; d->out_dtype = e.out_dtype; d->ldo = e.ldo;
  d->bias = e.bias; d->act = e.act; d->aux_out = e.aux_out;
  d->aux_in = e.aux_in; d->aux_in_dtype = e.aux_in_dtype; d->ld_aux_in = e.ld_aux_in; d->dact = e.dact;
  d->residual = e.residual; d->res_dtype = e.res_dtype; d->ldr = e.ldr;
  d->drop_p = e.drop_p; d->drop_seed = e.drop_seed; d->drop_stream = e.drop_stream;
  d->accumulate = e.accumulate; d->alpha = e.alpha; d->aux_is_grad = e.aux_is_grad;
  d->drop_bits = e.drop_bits; d->drop_bits_ld = 0; d->drop_bits_ready = e.drop_bits_ready;
  d->amax_val = e.amax_val; d->amax_idx = e.amax_idx; d->amax_ld = 0;
}
void count_launch();

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

struct SimtProblem {
  const void* a; const void* b;
  int64_t sam, sak, sbn, sbk;  // element strides: A(m,k) = a[m*sam + k*sak], B(n,k) = b[n*sbn + k*sbk]
  int M, N, K, split_k;
  EpiDev epi;
};

template <typename TIn>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const SimtProblem P) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const TIn* __restrict__ A = (const TIn*)P.a;
  const TIn* __restrict__ B = (const TIn*)P.b;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int k_per = ((P.K + P.split_k - 1) / P.split_k + TK - 1) / TK * TK;
  const int kbeg = blockIdx.z * k_per;
  const int kend = min(P.K, kbeg + k_per);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = kbeg; k0 < kend; k0 += TK) {
    // cooperative load of a TM x TK and TN x TK slab (1024 elements each, 4 per thread)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256;
      int mm, kk;
      if (P.sak == 1) { kk = idx & 15; mm = idx >> 4; } else { mm = idx & 63; kk = idx >> 6; }
      const int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < P.M && gk < kend) ? to_f32<TIn>(A[gm * P.sam + gk * P.sak]) : 0.f;
      int nn, kb;
      if (P.sbk == 1) { kb = idx & 15; nn = idx >> 4; } else { nn = idx & 63; kb = idx >> 6; }
      const int gn = n0 + nn, gkb = k0 + kb;
      Bs[kb][nn] = (gn < P.N && gkb < kend) ? to_f32<TIn>(B[gn * P.sbn + gkb * P.sbk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  if (kbeg >= kend) return;
  const EpiDev& e = P.epi;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t row = m0 + ty * 4 + i;
    if (row >= P.M) continue;
    const int64_t col = n0 + tx * 4;
    if (col >= P.N) continue;
    const int n_valid = min(4, P.N - (int)col);
    float v[4] = {acc[i][0], acc[i][1], acc[i][2], acc[i][3]};
    epilogue_apply<4>(e, row, col, n_valid, v);
    const size_t o = (size_t)(row * e.ldo + col);
    for (int j = 0; j < n_valid; ++j) {
      if (e.accumulate) atomicAdd((float*)e.out + o + j, v[j]);
      else store_from_f32(e.out, e.out_dtype, o + j, v[j]);
    }
  }
}

}  // namespace

int gemm_simt_launch(const m3ae_gemm_problem_t* probs, int n, int in_dtype, cudaStream_t stream) {
  for (int i = 0; i < n; ++i) {
    const m3ae_gemm_problem_t& g = probs[i];
    M3AE_CHECK(g.M > 0 && g.N > 0 && g.K > 0, M3AE_ERR_INVALID, "gemm problem %d has empty shape %dx%dx%d", i, g.M, g.N, g.K);
    M3AE_CHECK(g.split_k >= 1, M3AE_ERR_INVALID, "split_k must be >= 1");
    M3AE_CHECK(g.split_k == 1 || (g.epi.accumulate && g.epi.out_dtype == M3AE_F32), M3AE_ERR_INVALID,
               "split_k > 1 requires accumulate=1 and an fp32 output");
    M3AE_CHECK(!g.epi.accumulate || g.epi.out_dtype == M3AE_F32, M3AE_ERR_INVALID, "accumulate requires an fp32 output");
    SimtProblem P;
    P.a = g.a; P.b = g.b;
    if (g.a_major == M3AE_MAJOR_K) { P.sam = g.lda; P.sak = 1; } else { P.sam = 1; P.sak = g.lda; }
    if (g.b_major == M3AE_MAJOR_K) { P.sbn = g.ldb; P.sbk = 1; } else { P.sbn = 1; P.sbk = g.ldb; }
    P.M = g.M; P.N = g.N; P.K = g.K; P.split_k = g.split_k;
    epi_to_dev(g.epi, &P.epi);
    dim3 grid((g.N + TN - 1) / TN, (g.M + TM - 1) / TM, g.split_k);
    if (in_dtype == M3AE_BF16) M3AE_LAUNCH((gemm_simt_kernel<bf16>), grid, 256, 0, stream, P);
    else M3AE_LAUNCH((gemm_simt_kernel<float>), grid, 256, 0, stream, P);
    count_launch();
    M3AE_CUDA(cudaGetLastError());
  }
  return M3AE_OK;
}

}  // namespace m3ae
