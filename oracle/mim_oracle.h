/*
 * mim_oracle.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the computeDirection hot path of mim_solvers'
 * SolverSQP / SolverCSQP / SolverFDDP, following /root/reference/src/ (*.cpp)
 * statement by statement (citations beside each function in mim_oracle.c).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product path
 * (mim_solvers_b200/, libmimgpu.so) never does.
 *
 * Pinning status: the reference ships no stored golden vectors and its C++
 * cannot be built here (Eigen/Boost/crocoddyl absent), so this oracle is
 * pinned against (a) outputs of the reference's own NumPy solvers
 * (python/sqp.py, python/qp_solvers/stagewise_qp.py, python/csqp.py) executed
 * in this container on top of a small crocoddyl/eigenpy stand-in
 * (oracle/refshim, tests/golden/make_golden.py -> tests/golden/ npz files), and (b)
 * the invariants the reference's own tests assert (tests/test_solvers.cpp).
 *
 * Memory layout = the C ABI of include/mimgpu.h (problem-major SoA, column-
 * major blocks, K row-major).
 */
#ifndef MIM_ORACLE_H_
#define MIM_ORACLE_H_

#include "../include/mimgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* All solver members of a batch of B problems; the arrays ARE the persistent
 * solver state (y, z, rho_vec, rho_sparse, dx, du survive between calls like
 * the members of a SolverCSQP object). Allocate with ora_state_create. */
typedef struct {
  mimgpu_shape shape;
  double *dx, *du, *lag_mul, *dxt, *dut; /* [B][T+1][nx] / [B][T][nu] */
  double *y, *z, *rho_vec;               /* [B][T*ng+ngT] */
  double *Vxx, *Vx;                      /* [B][T+1][nx*nx], [B][T+1][nx] */
  double *K, *kff, *Qu, *Quuk, *L;       /* [B][T][...] */
  double *Vxx_fs;                        /* [B][T][nx] */
  double *z_prev, *z_relaxed, *inv_rho_vec; /* [B][T*ng+ngT] */
  double *scal;                          /* [MIMGPU_S_COUNT][B] */
  int32_t *qp_iters, *status;            /* [B] */
  int32_t *converged;                    /* [B] ADMM loop broke on tolerance */
} ora_state;

ora_state* ora_state_create(const mimgpu_shape* shape);
void ora_state_destroy(ora_state* s);

/* nthreads <= 0: all OpenMP threads. Each returns 0 or MIMGPU_ERR_INVALID. */
int ora_sqp_direction(ora_state* s, const mimgpu_derivs* d,
                      const mimgpu_sqp_params* p, int with_kkt, int nthreads);
int ora_fddp_direction(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_sqp_params* p, int nthreads);
int ora_csqp_calc_norms(ora_state* s, const mimgpu_derivs* d, int nthreads);
int ora_csqp_reset_rho(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_csqp_params* p, int nthreads);
int ora_csqp_direction(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_csqp_params* p, int with_kkt, int nthreads);

/* Same with per-problem regularisation arrays (NULL -> params value). */
int ora_sqp_direction_reg(ora_state* s, const mimgpu_derivs* d,
                          const mimgpu_sqp_params* p, const double* preg,
                          const double* dreg, int with_kkt, int nthreads);

int ora_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
