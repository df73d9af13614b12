// Tensor-core (mma.sync m16n8k16 bf16) flash attention for the bf16 product path.
// Placeholder dispatch: returns M3AE_ERR_UNSUPPORTED so attention.cu uses the generic fp32-math kernels.
#include "common.cuh"

namespace m3ae {
int attn_fwd_mma(const m3ae_attn_params_t* p, cudaStream_t st) { (void)p; (void)st; return M3AE_ERR_UNSUPPORTED; }
int attn_bwd_mma(const m3ae_attn_params_t* p, cudaStream_t st) { (void)p; (void)st; return M3AE_ERR_UNSUPPORTED; }
}  // namespace m3ae
