/* translation unit of the streaming CSQP kernels */
#include <cstdlib>

#include "kernels_cta.cuh"
#include "kernels_rec.cuh"

namespace mimgpu {
StreamPlan stream_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem) { return stream_plan(sh, stage_fx, max_smem); }
int stream_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const StreamPlan& pl,
                         cudaStream_t st) {
  return stream_csqp_launch(a, q, with_kkt, pl, st);
}
RecPlan rec_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem) { return rec_plan(sh, stage_fx, max_smem); }
int rec_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const RecPlan& pl, double* rec,
                      cudaStream_t st) {
  return rec_csqp_launch(a, q, with_kkt, pl, rec, st);
}
}  // namespace mimgpu
