// Tensor-core (tcgen05 3xTF32) GEMM path — dispatch stub; falls through to the fp32 FFMA tiles when it declines.
#include "common.cuh"

namespace grappa {
struct GemmParams;
int gemm_tc_dispatch(int layout, const GemmParams& p, cudaStream_t st, bool* handled) {
    (void)layout; (void)p; (void)st;
    *handled = false;
    return 0;
}
}  // namespace grappa
