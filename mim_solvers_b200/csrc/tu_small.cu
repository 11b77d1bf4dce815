/* translation unit of the warp-per-problem SQP kernels */
#include <cstdlib>

#include "launchers.h"
#include "kernels_small.cuh"

namespace mimgpu {
int small_kernels_configure_x(const mimgpu_shape& s) { return small_kernels_configure(s); }
int small_sqp_direction_launch_x(const KArgs& a, const SqpKPar& kp, int variant, cudaStream_t st) {
  return small_sqp_direction_launch(a, kp, variant, st);
}
int small_pack_layout_x(const mimgpu_shape& s, size_t* pack_doubles, size_t* spill_doubles) {
  return small_pack_layout(s, pack_doubles, spill_doubles);
}
int small_pack_launch_x(const KArgs& a, double* pack, int b0, int nb, cudaStream_t st) {
  return small_pack_launch(a, pack, b0, nb, st);
}
int small_unpack_launch_x(const KArgs& a, cudaStream_t st) { return small_unpack_launch(a, st); }
int small_unpack_inputs_launch_x(const KArgs& a, const double* pack, cudaStream_t st) {
  return small_unpack_inputs_launch(a, pack, st);
}
int small_record_offsets_x(const mimgpu_shape& s, int out[16]) { return small_record_offsets(s, out); }
}  // namespace mimgpu
