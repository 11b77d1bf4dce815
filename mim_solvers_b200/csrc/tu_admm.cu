/* translation unit of the condensed-chain CSQP kernels */
#include <cstdlib>

#include "kernels_admm.cuh"

namespace mimgpu {
AdmmPlan admm_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem) { return admm_plan(sh, stage_fx, max_smem); }
int admm_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const AdmmPlan& pl,
                       double* const scratch[3], cudaStream_t st, long long* timing) {
  return admm_csqp_launch(a, q, with_kkt, pl, scratch, st, timing);
}
}  // namespace mimgpu
