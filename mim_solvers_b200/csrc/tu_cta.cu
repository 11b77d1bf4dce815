/* translation unit of the CTA-per-problem DMMA SQP / FDDP kernels */
#include <cstdlib>

#include "kernels_cta.cuh"

namespace mimgpu {
CtaPlan cta_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem) { return cta_plan(sh, stage_fx, max_smem); }
int cta_sqp_launch_x(const KArgs& a, const SqpKPar& kp, bool fddp, const CtaPlan& pl, cudaStream_t st) {
  return cta_sqp_launch(a, kp, fddp, pl, st);
}
}  // namespace mimgpu
