// mlp_tc.cu -- tcgen05 (5th-gen tensor core) forward GEMM for the actor-critic MLP layers.
// Placeholder dispatch until the kind::tf32 kernel lands: reports UNSUPPORTED (never silently falls back).
#include "common.cuh"

namespace b200lab {
int linear_fwd_tc(const float* X, const float* W, const float* b, float* Y, int64_t M, int K, int Nout, int elu,
                  int precision, cudaStream_t s) {
  (void)X; (void)W; (void)b; (void)Y; (void)M; (void)K; (void)Nout; (void)elu; (void)s;
  set_error("linear_fwd: precision %d (tcgen05) not available in this build", precision);
  return B200LAB_ERR_UNSUPPORTED;
}
}  // namespace b200lab
