/*
 * kernels_small.cuh -- specialised kernels for the small hot shapes (UR5 class).
 * Placeholder interface: until a specialisation is registered the generic kernels
 * (kernels_generic.cuh) serve every shape.
 */
#pragma once
#include <cuda_runtime.h>

#include "kernels_generic.cuh"

namespace mimgpu {

/* returns 1 when a specialised SQP kernel exists for this shape */
static inline int small_kernels_configure(const mimgpu_shape&) { return 0; }

/* returns 0 when launched, < 0 when no specialisation applies, > 0 on error */
static inline int small_sqp_direction_launch(const KArgs&, const SqpKPar&, int, cudaStream_t) { return -1; }

}  // namespace mimgpu
