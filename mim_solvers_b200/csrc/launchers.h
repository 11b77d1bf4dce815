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
struct RecPlan { /* record-based CSQP kernel of the small shapes (kernels_rec.cuh) */
  int ok, ns, nt, smem_bytes, generic_total, w1;
  size_t rec_doubles; /* scratch per problem */
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
/* record-based CSQP kernel of the small shapes (kernels_rec.cuh); rec = [B][rec_doubles] scratch */
RecPlan rec_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int rec_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const RecPlan& pl, double* rec,
                      cudaStream_t st);
/* condensed-chain CSQP kernel (kernels_admm.cuh); scratch = {Acl, r, tmp} */
AdmmPlan admm_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int admm_csqp_launch_x(const KArgs& a, const mimgpu_csqp_params& q, int with_kkt, const AdmmPlan& pl,
                       double* const scratch[3], cudaStream_t st, long long* timing = nullptr);
/* device-side solve() for analytic models (kernels_solve.cuh) */
constexpr int SOLVE_NALPHA = 10;   /* alphas_ = 2^-n, n = 0..9 (src/sqp.cpp:60-70) */
constexpr int SOLVE_MAXFILTER = 16;

struct PolyModelDev { /* device pointers, C-ABI block layouts (col-major), leading [B][T] */
  const double *A, *Bm, *f, *Q, *S, *R, *q, *r, *QT, *qT, *x0;
  double c3, w4;
};

struct SolveDev {
  double *xs, *us;                 /* [B][T+1][nx], [B][T][nu] */
  double *cost, *steplength;       /* [B] */
  double *cost_list, *gap_list;    /* [B][filter_size]: index filter_size-1 = newest */
  double *preg, *dreg;             /* [B] */
  int32_t *iter, *done, *flag;     /* [B]; done: 0 running, 1 converged, 2 failed; flag: 1 = retry direction */
  int32_t *list_a, *list_b;        /* [B] compacted problem lists */
  int32_t* counts;                 /* [4] */
};

struct SolvePar {
  double reg_min, reg_max, reg_inc, reg_dec, th_stepdec, th_stepinc, termination_tol, mu_dynamic;
  int use_filter, filter_size;
};

int solve_init_launch_x(const SolveDev& s, const PolyModelDev& m, const double* xs0, const double* us0,
                        const mimgpu_shape& sh, double reg0, cudaStream_t st);
int solve_calcdiff_launch_x(const KArgs& a, const PolyModelDev& m, const SolveDev& s, const int32_t* list, int n,
                            cudaStream_t st);
int solve_post_direction_launch_x(const KArgs& a, const SolveDev& s, const SolvePar& sp, const int32_t* list, int n,
                                  int32_t* retry_out, cudaStream_t st);
int solve_linesearch_launch_x(const KArgs& a, const PolyModelDev& m, const SolveDev& s, const SolvePar& sp,
                              const int32_t* list, int n, cudaStream_t st);
int solve_running_launch_x(const SolveDev& s, int B, cudaStream_t st);
/* CTA-per-problem DMMA SQP / FDDP kernel (kernels_ric.cuh, kernels_cta.cuh) */
CtaPlan cta_plan_x(const mimgpu_shape& sh, int stage_fx, int max_smem);
int cta_sqp_launch_x(const KArgs& a, const SqpKPar& kp, bool fddp, const CtaPlan& pl, cudaStream_t st);

}  // namespace mimgpu
