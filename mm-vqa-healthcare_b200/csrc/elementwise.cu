// Small element-wise kernels of the heads on the hot path (HBM-bound, M = batch rows):
//   activation fwd / bwd    Pooler tanh (prediction_heads.py:13,18), vqa_head GELU (m3ae_module.py:123)
//   bce_with_logits         VQA loss of compute_vqa_m3ae (objectives.py:201): mean BCE-with-logits * n_labels, and its
//                           gradient with respect to the logits in the same pass
//   cross_entropy           decoder fine-tune loss (m3ae_decoder.py:365-368: CrossEntropyLoss(ignore_index=pad) on the
//                           transposed logits, * padding_mask, / padding_mask.sum()) and its gradient
#include <string.h>

#include "common.cuh"

namespace m3ae {
void count_launch();

namespace {

template <typename T>
__global__ void act_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t n, int act) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    y[i] = from_f32<T>(act_exact(act, to_f32<T>(x[i])));
}

// dx = dy * act'(x), x = the forward pre-activation
template <typename T>
__global__ void act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t n, int act) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dx[i] = from_f32<T>(to_f32<T>(dy[i]) * act_exact_grad(act, to_f32<T>(x[i])));
}

// loss_out[0] += scale_loss * sum_i bce(x_i, t_i);  dlogits_i = scale_grad * (sigmoid(x_i) - t_i)
// bce(x, t) = max(x, 0) - x t + log1p(exp(-|x|))   (torch.nn.functional.binary_cross_entropy_with_logits)
__global__ void __launch_bounds__(256) bce_logits_kernel(const float* __restrict__ x, const float* __restrict__ t,
                                                          float* __restrict__ loss_out, float* __restrict__ dx, int64_t n,
                                                          float scale_loss, float scale_grad) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float xi = x[i], ti = t[i];
    const float e = expf(-fabsf(xi));
    acc += fmaxf(xi, 0.f) - xi * ti + log1pf(e);
    if (dx) {
      const float sig = xi >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
      dx[i] = scale_grad * (sig - ti);
    }
  }
  acc = warp_sum(acc);
  __shared__ float part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += part[w];
    atomicAdd(loss_out, s * scale_loss);
  }
}

// One CTA per row of logits [M, V] (fp32, leading dimension ld): lse over the row, loss and gradient.
//   target == ignore_index or weight[row] == 0: contributes nothing (gradient row = 0)
//   loss_out[0] += w_row * (lse - x[target]);  dlogits[row, j] = w_row * (softmax_j - [j == target])
// The caller folds 1 / sum(weights) into w_row (m3ae_decoder.py:367-368).
__global__ void __launch_bounds__(256) cross_entropy_kernel(const float* __restrict__ x, int64_t ld, const int64_t* __restrict__ target,
                                                             const float* __restrict__ row_weight, float* __restrict__ loss_out,
                                                             float* __restrict__ dx, int64_t ldd, int V, int64_t ignore_index) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t row = blockIdx.x;
  const float* xr = x + row * ld;
  const int64_t tg = target[row];
  const float w = (tg == ignore_index) ? 0.f : row_weight[row];
  __shared__ float red[8];
  __shared__ float bcast;
  if (w == 0.f) {
    if (dx)
      for (int j = threadIdx.x; j < V; j += blockDim.x) dx[row * ldd + j] = 0.f;
    return;
  }
  float m = -INFINITY;
  for (int j = threadIdx.x; j < V; j += blockDim.x) m = fmaxf(m, xr[j]);
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = red[0];
    for (int k = 1; k < 8; ++k) t = fmaxf(t, red[k]);
    bcast = t;
  }
  __syncthreads();
  m = bcast;
  float s = 0.f;
  for (int j = threadIdx.x; j < V; j += blockDim.x) s += expf(xr[j] - m);
  s = warp_sum(s);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += red[k];
    bcast = t;
  }
  __syncthreads();
  s = bcast;
  const float lse = m + logf(s);
  if (threadIdx.x == 0) atomicAdd(loss_out, w * (lse - xr[tg]));
  if (dx) {
    const float inv = 1.f / s;
    for (int j = threadIdx.x; j < V; j += blockDim.x)
      dx[row * ldd + j] = w * (expf(xr[j] - m) * inv - (j == tg ? 1.f : 0.f));
  }
}

struct DropBitsGroup {
  m3ae_dropbits_item_t it[M3AE_MAX_DROPBITS];
  int64_t word_begin[M3AE_MAX_DROPBITS + 1];  // prefix sums of output words (uint32) per item
  int32_t words_per_row[M3AE_MAX_DROPBITS];
  int n;
};

// One thread = one uint32 of keep bits = 4 Philox evaluations (independent: good ILP).  Streaming, no shared memory.
__global__ void __launch_bounds__(256) dropout_bits_kernel(const __grid_constant__ DropBitsGroup G) {
  pdl_sync();  // programmatic dependent launch: let the next kernel be scheduled, then wait for the previous one's results
  const int64_t total = G.word_begin[G.n];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += stride) {
    int i = 0;
#pragma unroll
    for (int k = 1; k < M3AE_MAX_DROPBITS; ++k)
      if (k < G.n && w >= G.word_begin[k]) i = k;
    const m3ae_dropbits_item_t& it = G.it[i];
    const int64_t local = w - G.word_begin[i];
    const int wpr = G.words_per_row[i];
    const int64_t row = local / wpr;
    const int wc = (int)(local - row * wpr);
    const uint32_t thr = drop_thr15(it.p);
    uint32_t bits = 0;
    if (it.kind == 0) {
      const int units = (it.n + 7) >> 3;  // 8-key units of a row; bytes past the last unit stay 0
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (4 * wc + j < units) bits |= attn_keep_bits8(it.seed, it.site, (uint64_t)row, 4 * wc + j, thr) << (8 * j);
    } else {
      const uint64_t e0 = (uint64_t)(row * it.ld + 32 * wc);
#pragma unroll
      for (int j = 0; j < 4; ++j) bits |= dropout_keep8(it.seed, it.site, e0 + 8 * j, thr) << (8 * j);
    }
    reinterpret_cast<uint32_t*>(it.out)[local] = bits;
  }
}

unsigned grid_for(int64_t n, int threads) {
  const int64_t cap = (int64_t)sm_count() * 8;
  const int64_t blocks = (n + threads - 1) / threads;
  return (unsigned)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace

int activation_fwd(const void* x, void* y, int64_t n, int act, int dtype, cudaStream_t st) {
  M3AE_CHECK(n > 0, M3AE_ERR_INVALID, "activation_fwd: empty input");
  M3AE_CHECK(act >= M3AE_ACT_NONE && act <= M3AE_ACT_TANH, M3AE_ERR_INVALID, "activation_fwd: unknown activation %d", act);
  if (dtype == M3AE_BF16) M3AE_LAUNCH((act_fwd_kernel<bf16>), grid_for(n, 256), 256, 0, st, (const bf16*)x, (bf16*)y, n, act);
  else M3AE_LAUNCH((act_fwd_kernel<float>), grid_for(n, 256), 256, 0, st, (const float*)x, (float*)y, n, act);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int activation_bwd(const void* dy, const void* x, void* dx, int64_t n, int act, int dtype, cudaStream_t st) {
  M3AE_CHECK(n > 0, M3AE_ERR_INVALID, "activation_bwd: empty input");
  M3AE_CHECK(act >= M3AE_ACT_NONE && act <= M3AE_ACT_TANH, M3AE_ERR_INVALID, "activation_bwd: unknown activation %d", act);
  if (dtype == M3AE_BF16) M3AE_LAUNCH((act_bwd_kernel<bf16>), grid_for(n, 256), 256, 0, st, (const bf16*)dy, (const bf16*)x, (bf16*)dx, n, act);
  else M3AE_LAUNCH((act_bwd_kernel<float>), grid_for(n, 256), 256, 0, st, (const float*)dy, (const float*)x, (float*)dx, n, act);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int bce_with_logits(const float* logits, const float* targets, float* loss_out, float* dlogits, int64_t n, float scale_loss,
                    float scale_grad, cudaStream_t st) {
  M3AE_CHECK(n > 0, M3AE_ERR_INVALID, "bce_with_logits: empty input");
  M3AE_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), st));
  M3AE_LAUNCH((bce_logits_kernel), grid_for(n, 256), 256, 0, st, logits, targets, loss_out, dlogits, n, scale_loss, scale_grad);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int dropout_bits(const m3ae_dropbits_item_t* items, int n, cudaStream_t st) {
  DropBitsGroup G;
  memset(&G, 0, sizeof(G));
  G.n = n;
  int64_t words = 0;
  for (int i = 0; i < n; ++i) {
    const m3ae_dropbits_item_t& it = items[i];
    M3AE_CHECK(it.out != nullptr && it.rows > 0 && it.n > 0, M3AE_ERR_INVALID, "dropout_bits: empty item %d", i);
    M3AE_CHECK(it.p > 0.f && it.p < 1.f, M3AE_ERR_INVALID, "dropout_bits: p %f not in (0,1)", it.p);
    M3AE_CHECK(it.kind == 0 || (it.kind == 1 && it.n % 32 == 0 && it.ld % 8 == 0), M3AE_ERR_INVALID,
               "dropout_bits: item %d: kind must be 0 (attention) or 1 (hidden, N %% 32 == 0, ld %% 8 == 0)", i);
    M3AE_CHECK((((uintptr_t)it.out) & 3) == 0, M3AE_ERR_INVALID, "dropout_bits: output of item %d is not 4-byte aligned", i);
    G.it[i] = it;
    G.words_per_row[i] = it.kind == 0 ? 2 * ((it.n + 63) / 64) : it.n / 32;  // attention rows: 8 bytes per 64 keys
    G.word_begin[i] = words;
    words += it.rows * G.words_per_row[i];
  }
  G.word_begin[n] = words;
  for (int i = n + 1; i <= M3AE_MAX_DROPBITS; ++i) G.word_begin[i] = words;
  M3AE_LAUNCH((dropout_bits_kernel), grid_for(words, 256), 256, 0, st, G);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

int cross_entropy(const float* logits, int64_t ld, const int64_t* target, const float* row_weight, float* loss_out, float* dlogits,
                  int64_t ldd, int64_t M, int V, int64_t ignore_index, cudaStream_t st) {
  M3AE_CHECK(M > 0 && V > 0, M3AE_ERR_INVALID, "cross_entropy: empty input");
  M3AE_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), st));
  M3AE_LAUNCH((cross_entropy_kernel), (unsigned)M, 256, 0, st, logits, ld, target, row_weight, loss_out, dlogits, ldd, V, ignore_index);
  count_launch();
  M3AE_CUDA(cudaGetLastError());
  return M3AE_OK;
}

}  // namespace m3ae
