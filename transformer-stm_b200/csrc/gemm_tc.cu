// tcgen05 / TMEM contraction path (CVT_PREC_TF32 / CVT_PREC_BF16).  Placeholder until the kernel lands: nothing is
// routed here yet, so every contraction runs on the fp32 FFMA kernel.
#include "model.cuh"

namespace cvt {
bool gemm_tc_supports(const GemmP&) { return false; }
cudaError_t launch_gemm_tc(const GemmP&, cudaStream_t) { set_error("gemm_tc: not built"); return cudaErrorNotSupported; }
bool attention_tc_enabled(const Model*, int) { return false; }
cudaError_t attention_forward_tc(Model*, int, int, cudaStream_t) { set_error("attention_tc: not built"); return cudaErrorNotSupported; }
cudaError_t attention_backward_tc(Model*, int, int, const float*, float*, cudaStream_t) { set_error("attention_tc: not built"); return cudaErrorNotSupported; }
}  // namespace cvt
