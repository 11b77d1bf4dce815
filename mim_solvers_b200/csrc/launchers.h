/*
 * launchers.h -- host-side entry points of the specialised kernel families.  Each family is
 * compiled in its own translation unit (tu_small.cu, tu_stream.cu, tu_cta.cu) so that the
 * library builds in parallel; mimgpu.cu (the C ABI) only sees these declarations.
 */
#pragma once
#include <cuda_runtime.h>

#include "kernels_generic.cuh"

namespace mimgpu {

struct StreamPlan {
  int ok, ns, nt, smem_bytes, generic_total, direct;
};
struct AdmmPlan {
  int ok, nt, smem_bytes, ric_total, nchain, use_mma, stage_g, chain_ns;
};
struct CtaPlan {
  int ok, nt, smem_bytes, nbuf, ns;
};

/* warp-per-problem SQP kernels (kernels_small*.cuh) */
int small_kernels_configure_x(const mimgpu_shape& s);
/* variant: 0 = fast (Gauss-Jordan gains, K/k/Vxx/Vx spilled only), 1 = full spill (also Qu, Quuk, LLT) */
int small_sqp_direction_launch_x(const KArgs& a, const SqpKPar& kp, int variant, cudaStream_t st);
/* knot-packed HBM layout of the fast variant: sizes (doubles per problem), repack of the field-major
 * inputs at upload time, on-demand copy of the spill records into the Vxx_ Vx_ K_ k_ arrays */
int small_pack_layout_x(const mimgpu_shape& s, size_t* pack_doubles, size_t* spill_doubles);
int small_pack_launch_x(const KArgs& a, double* pack, int b0, int nb, cudaStream_t st);
int small_unpack_launch_x(const KArgs& a, cudaStream_t st);
/* knot records -> field-major arrays (a.Fx ... writable), and the block offsets inside a record */
int small_unpack_inputs_launch_x(const KArgs& a, const double* pack, cudaStream_t st);
int small_record_offsets_x(const mimgpu_shape& s, int out[16]);
/* streaming CSQP kernel (kernels_stream.cuh, kernels_cta.cuh) */
StreamPlan stream_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int stream_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const StreamPlan& pl,
                         cudaStream_t st);
/* condensed-chain CSQP kernel (kernels_admm.cuh); scratch = {Acl, r, tmp} */
AdmmPlan admm_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int admm_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const AdmmPlan& pl,
                       double* const scratch[3], cudaStream_t st, long long* timing = nullptr);
/* CTA-per-problem DMMA SQP / FDDP kernel (kernels_ric.cuh, kernels_cta.cuh) */
CtaPlan cta_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int cta_sqp_launch_x(const KArgs& a, const SqpKPar& kp, bool fddp, const CtaPlan& pl, cudaStream_t st);

}  // namespace mimgpu
