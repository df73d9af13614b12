// Micro-benchmark: cycles per tcgen05.mma (cta_group::1, kind::f16, M=128) as a function of N, the number of
// independent accumulator chains and the operand majors.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
//   -I mm-vqa-healthcare_b200/csrc tools/mma_probe.cu -o tools/mma_probe   (run on a B200: prints a table)
#include <cstdio>
#include "common.cuh"
using namespace m3ae;
constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint32_t dlo(uint32_t saddr, uint32_t lbo) { return ((saddr >> 4) & 0x3FFFu) | (((lbo >> 4) & 0x3FFFu) << 16); }
__device__ __forceinline__ void mma(uint32_t d, uint32_t alo, uint32_t blo, uint32_t idesc, uint32_t acc) {
  if (elect_one()) {
    asm volatile("{\n.reg .pred p;\n.reg .b64 da, db;\nsetp.ne.b32 p, %4, 0;\nmov.b64 da, {%1, %5};\nmov.b64 db, {%2, %5};\n"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n}\n" ::"r"(d), "r"(alo), "r"(blo), "r"(idesc), "r"(acc), "r"(DESC_HI) : "memory");
  }
}
struct Cfg { int n, chains, a_mn, b_mn, count, ctas; };
template <int CH>
__global__ void __launch_bounds__(128, 1) probe(Cfg c, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tptr;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;  // bf16 small values
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc<512>(&tptr);
  fence_proxy_async();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tb = tptr;
  if (warp == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)c.a_mn << 15) | ((uint32_t)c.b_mn << 16) | ((uint32_t)(c.n >> 3) << 17) | (8u << 24);
    const uint32_t a0 = dlo(smem_u32(smem), c.a_mn ? 16384u : 16u), b0 = dlo(smem_u32(smem + 65536), c.b_mn ? 8192u : 16u);
    const uint32_t astep = c.a_mn ? 128u : 2u, bstep = c.b_mn ? 128u : 2u;
    const int ncol = c.n;  // columns per accumulator
    long long t0 = clock64();
    // first round (overwrite), then count/CH-1 accumulate rounds; k-step cycles through 4 slices of the operand tiles
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) mma(tb + ch * ncol, a0 + ch * 512, b0, idesc, 0);
    for (int r = 1; r < c.count / CH; r += 4) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) mma(tb + ch * ncol, a0 + k * astep + ch * 512, b0 + k * bstep, idesc, 1);
    }
    long long t1 = clock64();
    if (elect_one()) umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) { tcgen05_fence_after(); tmem_dealloc<512>(tb); }
}
int main() {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(probe<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  printf("   N chains A  B  ctas | issue cyc/MMA | total cyc/MMA\n");
  const int Ns[] = {64, 128, 256};
  for (int ctas : {148})
  for (int am = 0; am < 2; ++am) for (int bm = 0; bm < 2; ++bm)
  for (int n : Ns) for (int ch : {1, 2, 4}) {
    if (n * ch > 512) continue;
    Cfg c{n, ch, am, bm, 132, ctas};
    if (ch == 1) probe<1><<<ctas, 128, 200 * 1024>>>(c, d);
    else if (ch == 2) probe<2><<<ctas, 128, 200 * 1024>>>(c, d);
    else probe<4><<<ctas, 128, 200 * 1024>>>(c, d);
    long long h[2]; cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    printf("%4d %5d  %s  %s %4d | %8.1f | %8.1f\n", n, ch, am ? "MN" : "K ", bm ? "MN" : "K ", ctas, h[0] / (double)(ch + 4 * ch * ((132 / ch - 1 + 3) / 4)), h[1] / (double)(ch + 4 * ch * ((132 / ch - 1 + 3) / 4)));
  }
  return 0;
}
