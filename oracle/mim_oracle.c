/*
 * mim_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 * See mim_oracle.h for scope and pinning status.
 *
 * Every function cites the reference lines it restates (paths relative to
 * /root/reference).  Statement order follows the reference; inner products are
 * accumulated in increasing index order.  Matrices are column-major:
 * M(i,j) = M[i + rows*j]; K is row-major nu x nx (include/mim_solvers/ddp.hpp:397).
 */
#include "mim_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------- helpers */

/* crocoddyl::raiseIfNaN (used at src/ddp.cpp:340-345): NaN, +-inf or >= 1e30 */
static int raise_if_nan(double v) { return isnan(v) || isinf(v) || v >= 1e30; }

/* lpNorm<Infinity> with NaN made sticky (Eigen's maxCoeff NaN behaviour is
 * implementation defined; any NaN entry is treated as a failure here). */
static double inf_norm(const double* v, int n) {
  double m = 0.0;
  for (int i = 0; i < n; ++i) {
    double a = fabs(v[i]);
    if (isnan(a)) return a;
    if (a > m) m = a;
  }
  return m;
}
static double l1_norm(const double* v, int n) {
  double s = 0.0;
  for (int i = 0; i < n; ++i) s += fabs(v[i]);
  return s;
}
/* std::max / std::min with their exact NaN behaviour (src/csqp.cpp:522-523) */
static double std_max(double a, double b) { return (a < b) ? b : a; }
static double std_min(double a, double b) { return (b < a) ? b : a; }

/* y (+)= A^T x ; A is rows x cols column-major, x has `rows`, y has `cols` */
static void gemv_t_acc(double* y, const double* A, int rows, int cols,
                       const double* x) {
  for (int j = 0; j < cols; ++j) {
    double s = 0.0;
    for (int i = 0; i < rows; ++i) s += A[i + rows * j] * x[i];
    y[j] += s;
  }
}
/* y (+)= A x ; A rows x cols column-major */
static void gemv_acc(double* y, const double* A, int rows, int cols,
                     const double* x) {
  for (int i = 0; i < rows; ++i) {
    double s = 0.0;
    for (int j = 0; j < cols; ++j) s += A[i + rows * j] * x[j];
    y[i] += s;
  }
}
/* C += A^T B ; A is m x p, B is m x q, C is p x q (all column-major) */
static void gemm_tn_acc(double* C, const double* A, const double* B, int m,
                        int p, int q) {
  for (int j = 0; j < q; ++j)
    for (int i = 0; i < p; ++i) {
      double s = 0.0;
      for (int k = 0; k < m; ++k) s += A[k + m * i] * B[k + m * j];
      C[i + p * j] += s;
    }
}
/* C += A B ; A is p x m, B is m x q, C is p x q */
static void gemm_nn_acc(double* C, const double* A, const double* B, int p,
                        int m, int q) {
  for (int j = 0; j < q; ++j)
    for (int i = 0; i < p; ++i) {
      double s = 0.0;
      for (int k = 0; k < m; ++k) s += A[i + p * k] * B[k + m * j];
      C[i + p * j] += s;
    }
}

/* Eigen::LLT<MatrixXd, Lower>::compute, unblocked kernel
 * (Eigen/src/Cholesky/LLT.h llt_inplace<Scalar,Lower>::unblocked): reads the
 * lower triangle only, fails iff a pivot is <= 0 (a NaN pivot does not fail).
 * Returns -1 on success, else the failing column.  A is n x n column-major,
 * factor written in place in the lower triangle. */
static int llt_lower(double* A, int n) {
  for (int k = 0; k < n; ++k) {
    double x = A[k + n * k];
    for (int j = 0; j < k; ++j) x -= A[k + n * j] * A[k + n * j];
    if (x <= 0.0) return k;
    x = sqrt(x);
    A[k + n * k] = x;
    for (int i = k + 1; i < n; ++i) {
      double v = A[i + n * k];
      for (int j = 0; j < k; ++j) v -= A[i + n * j] * A[k + n * j];
      A[i + n * k] = v / x;
    }
  }
  return -1;
}
/* LLT::solveInPlace on one right-hand side with stride `inc` */
static void llt_solve(const double* L, int n, double* b, int inc) {
  for (int i = 0; i < n; ++i) {
    double v = b[i * inc];
    for (int j = 0; j < i; ++j) v -= L[i + n * j] * b[j * inc];
    b[i * inc] = v / L[i + n * i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double v = b[i * inc];
    for (int j = i + 1; j < n; ++j) v -= L[j + n * i] * b[j * inc];
    b[i * inc] = v / L[i + n * i];
  }
}

/* ---------------------------------------------------- one-problem views */

typedef struct {
  int T, nx, nu, ng, ngT;
  /* inputs */
  const double *Fx, *Fu, *Lxx, *Lxu, *Luu, *Lx, *Lu, *fs, *LxxT, *LxT;
  const double *Gx, *Gu, *g, *glb, *gub, *GxT, *gT, *glbT, *gubT;
  double cost;
  /* solver members */
  double *dx, *du, *lag, *dxt, *dut, *y, *z, *rho, *irho, *zprev, *zrel;
  double *Vxx, *Vx, *K, *k, *Qu, *Quuk, *L, *Vxx_fs;
  /* scratch (per thread) */
  double *Qxx, *Qxu, *Quu, *Qx, *FxTVxx, *FuTVxx, *tmpVx, *tmpx, *tmpu;
  double *tdual, *tc, *trg; /* tmp_dual_cwise_, tmp_Cdx_Cdu_, tmp_rhoG*_mat_ */
  /* scalars */
  double gap_norm, merit, kkt, xgrad, ugrad, constraint_norm;
  double norm_primal, norm_dual, norm_primal_rel, norm_dual_rel;
  double rho_sparse, rho_estimate, dg, dq;
  int qp_iters, status, converged;
} prob;

typedef struct {
  double *Qxx, *Qxu, *Quu, *Qx, *FxTVxx, *FuTVxx, *tmpVx, *tmpx, *tmpu;
  double *tdual, *tc, *trg;
} scratch;

static int imax(int a, int b) { return a > b ? a : b; }

static scratch scratch_alloc(const mimgpu_shape* sh) {
  scratch w;
  int nx = sh->nx, nu = sh->nu, ngm = imax(imax(sh->ng, sh->ngT), 1);
  w.Qxx = (double*)calloc((size_t)nx * nx, 8);
  w.Qxu = (double*)calloc((size_t)nx * imax(nu, 1), 8);
  w.Quu = (double*)calloc((size_t)imax(nu * nu, 1), 8);
  w.Qx = (double*)calloc((size_t)nx, 8);
  w.FxTVxx = (double*)calloc((size_t)nx * nx, 8);
  w.FuTVxx = (double*)calloc((size_t)imax(nu, 1) * nx, 8);
  w.tmpVx = (double*)calloc((size_t)nx, 8);
  w.tmpx = (double*)calloc((size_t)nx, 8);
  w.tmpu = (double*)calloc((size_t)imax(nu, 1), 8);
  w.tdual = (double*)calloc((size_t)ngm, 8);
  w.tc = (double*)calloc((size_t)ngm, 8);
  w.trg = (double*)calloc((size_t)ngm * imax(nx, nu), 8);
  return w;
}
static void scratch_free(scratch* w) {
  free(w->Qxx); free(w->Qxu); free(w->Quu); free(w->Qx); free(w->FxTVxx);
  free(w->FuTVxx); free(w->tmpVx); free(w->tmpx); free(w->tmpu);
  free(w->tdual); free(w->tc); free(w->trg);
}

static void bind(prob* q, const ora_state* s, const mimgpu_derivs* d,
                 const scratch* w, int b) {
  const mimgpu_shape* sh = &s->shape;
  const size_t T = sh->T, nx = sh->nx, nu = sh->nu, ng = sh->ng, ngT = sh->ngT;
  const size_t nc = T * ng + ngT;
  memset(q, 0, sizeof(*q));
  q->T = (int)T; q->nx = (int)nx; q->nu = (int)nu; q->ng = (int)ng; q->ngT = (int)ngT;
  q->Fx = d->Fx + b * T * nx * nx;
  q->Fu = d->Fu + b * T * nx * nu;
  q->Lxx = d->Lxx + b * T * nx * nx;
  q->Lxu = d->Lxu + b * T * nx * nu;
  q->Luu = d->Luu + b * T * nu * nu;
  q->Lx = d->Lx + b * T * nx;
  q->Lu = d->Lu + b * T * nu;
  q->fs = d->fs + b * (T + 1) * nx;
  q->LxxT = d->LxxT + b * nx * nx;
  q->LxT = d->LxT + b * nx;
  if (ng) {
    q->Gx = d->Gx + b * T * ng * nx;
    q->Gu = d->Gu + b * T * ng * nu;
    q->g = d->g + b * T * ng;
    q->glb = d->g_lb + b * T * ng;
    q->gub = d->g_ub + b * T * ng;
  }
  if (ngT) {
    q->GxT = d->GxT + b * ngT * nx;
    q->gT = d->gT + b * ngT;
    q->glbT = d->g_lbT + b * ngT;
    q->gubT = d->g_ubT + b * ngT;
  }
  q->cost = d->cost ? d->cost[b] : 0.0;
  q->dx = s->dx + b * (T + 1) * nx;
  q->du = s->du + b * T * nu;
  q->lag = s->lag_mul + b * (T + 1) * nx;
  q->dxt = s->dxt + b * (T + 1) * nx;
  q->dut = s->dut + b * T * nu;
  q->y = s->y + b * nc; q->z = s->z + b * nc; q->rho = s->rho_vec + b * nc;
  q->irho = s->inv_rho_vec + b * nc;
  q->zprev = s->z_prev + b * nc; q->zrel = s->z_relaxed + b * nc;
  q->Vxx = s->Vxx + b * (T + 1) * nx * nx;
  q->Vx = s->Vx + b * (T + 1) * nx;
  q->K = s->K + b * T * nu * nx;
  q->k = s->kff + b * T * nu;
  q->Qu = s->Qu + b * T * nu;
  q->Quuk = s->Quuk + b * T * nu;
  q->L = s->L + b * T * nu * nu;
  q->Vxx_fs = s->Vxx_fs + b * T * nx;
  q->Qxx = w->Qxx; q->Qxu = w->Qxu; q->Quu = w->Quu; q->Qx = w->Qx;
  q->FxTVxx = w->FxTVxx; q->FuTVxx = w->FuTVxx; q->tmpVx = w->tmpVx;
  q->tmpx = w->tmpx; q->tmpu = w->tmpu; q->tdual = w->tdual; q->tc = w->tc;
  q->trg = w->trg;
  /* scalars and counters round-trip through the state arrays */
  {
    const size_t B = sh->B;
    const double* sc = s->scal;
    q->kkt = sc[MIMGPU_S_KKT * B + b];
    q->gap_norm = sc[MIMGPU_S_GAP_NORM * B + b];
    q->merit = sc[MIMGPU_S_MERIT * B + b];
    q->xgrad = sc[MIMGPU_S_XGRAD_NORM * B + b];
    q->ugrad = sc[MIMGPU_S_UGRAD_NORM * B + b];
    q->norm_primal = sc[MIMGPU_S_NORM_PRIMAL * B + b];
    q->norm_dual = sc[MIMGPU_S_NORM_DUAL * B + b];
    q->norm_primal_rel = sc[MIMGPU_S_NORM_PRIMAL_REL * B + b];
    q->norm_dual_rel = sc[MIMGPU_S_NORM_DUAL_REL * B + b];
    q->rho_sparse = sc[MIMGPU_S_RHO_SPARSE * B + b];
    q->constraint_norm = sc[MIMGPU_S_CONSTRAINT_NORM * B + b];
    q->dg = sc[MIMGPU_S_DG * B + b];
    q->dq = sc[MIMGPU_S_DQ * B + b];
    q->qp_iters = s->qp_iters[b];
    q->status = s->status[b];
    q->converged = s->converged[b];
  }
}

static void unbind(const prob* q, ora_state* s, int b) {
  const size_t B = s->shape.B;
  double* sc = s->scal;
  sc[MIMGPU_S_KKT * B + b] = q->kkt;
  sc[MIMGPU_S_GAP_NORM * B + b] = q->gap_norm;
  sc[MIMGPU_S_MERIT * B + b] = q->merit;
  sc[MIMGPU_S_XGRAD_NORM * B + b] = q->xgrad;
  sc[MIMGPU_S_UGRAD_NORM * B + b] = q->ugrad;
  sc[MIMGPU_S_NORM_PRIMAL * B + b] = q->norm_primal;
  sc[MIMGPU_S_NORM_DUAL * B + b] = q->norm_dual;
  sc[MIMGPU_S_NORM_PRIMAL_REL * B + b] = q->norm_primal_rel;
  sc[MIMGPU_S_NORM_DUAL_REL * B + b] = q->norm_dual_rel;
  sc[MIMGPU_S_RHO_SPARSE * B + b] = q->rho_sparse;
  sc[MIMGPU_S_CONSTRAINT_NORM * B + b] = q->constraint_norm;
  sc[MIMGPU_S_DG * B + b] = q->dg;
  sc[MIMGPU_S_DQ * B + b] = q->dq;
  s->qp_iters[b] = q->qp_iters;
  s->status[b] = q->status;
  s->converged[b] = q->converged;
}

/* ------------------------------------------------------------ DDP / SQP */

/* SolverDDP::computeGains, src/ddp.cpp:401-429.  Quu is in q->Quu; the factor
 * is stored per knot in q->L (Quu_llt_[t]).  Returns 0 or MIMGPU_STATUS_LLT. */
static int compute_gains(prob* q, int t) {
  const int nx = q->nx, nu = q->nu;
  if (nu > 0) {
    double* L = q->L + (size_t)t * nu * nu;
    memcpy(L, q->Quu, sizeof(double) * nu * nu);   /* Quu_llt_[t].compute(Quu_[t]) */
    if (llt_lower(L, nu) >= 0) return MIMGPU_STATUS_LLT; /* info != Success */
    double* K = q->K + (size_t)t * nu * nx;        /* row-major nu x nx */
    for (int a = 0; a < nu; ++a)                    /* K_[t] = Qxu_[t].transpose() */
      for (int i = 0; i < nx; ++i) K[a * nx + i] = q->Qxu[i + nx * a];
    for (int i = 0; i < nx; ++i) llt_solve(L, nu, K + i, nx); /* solveInPlace(K_) */
    double* k = q->k + (size_t)t * nu;
    memcpy(k, q->Qu + (size_t)t * nu, sizeof(double) * nu);  /* k_[t] = Qu_[t] */
    llt_solve(L, nu, k, 1);
  }
  return 0;
}

/* Value-function update shared by the three full backward passes:
 * Vx_[t] = Qx_ - K^T Qu ; Vxx_[t] = Qxx_ - Qxu_ K ; symmetrise ; + preg
 * (src/ddp.cpp:319-333, src/csqp.cpp:784-799, :881-895). with_quuk: DDP only. */
static void value_update(prob* q, int t, double preg, int with_quuk) {
  const int nx = q->nx, nu = q->nu;
  double* Vx = q->Vx + (size_t)t * nx;
  double* Vxx = q->Vxx + (size_t)t * nx * nx;
  const double* K = q->K + (size_t)t * nu * nx;
  const double* Qu = q->Qu + (size_t)t * nu;
  memcpy(Vx, q->Qx, sizeof(double) * nx);
  memcpy(Vxx, q->Qxx, sizeof(double) * nx * nx);
  if (nu != 0) {
    if (with_quuk) { /* Quuk_[t] = Quu_[t] * k_[t]  (src/ddp.cpp:322) */
      double* Quuk = q->Quuk + (size_t)t * nu;
      const double* k = q->k + (size_t)t * nu;
      for (int a = 0; a < nu; ++a) {
        double s = 0.0;
        for (int c = 0; c < nu; ++c) s += q->Quu[a + nu * c] * k[c];
        Quuk[a] = s;
      }
    }
    for (int i = 0; i < nx; ++i) { /* Vx_[t] -= K_[t]^T Qu_[t] */
      double s = 0.0;
      for (int a = 0; a < nu; ++a) s += K[a * nx + i] * Qu[a];
      Vx[i] -= s;
    }
    for (int j = 0; j < nx; ++j) /* Vxx_[t] -= Qxu_[t] K_[t] */
      for (int i = 0; i < nx; ++i) {
        double s = 0.0;
        for (int a = 0; a < nu; ++a) s += q->Qxu[i + nx * a] * K[a * nx + j];
        Vxx[i + nx * j] -= s;
      }
  }
  /* Vxx_tmp_ = 0.5 (Vxx + Vxx^T); Vxx_[t] = Vxx_tmp_ */
  for (int j = 0; j < nx; ++j)
    for (int i = j + 1; i < nx; ++i) {
      double v = 0.5 * (Vxx[i + nx * j] + Vxx[j + nx * i]);
      Vxx[i + nx * j] = v;
      Vxx[j + nx * i] = v;
    }
  for (int i = 0; i < nx; ++i) Vxx[i + nx * i] = 0.5 * (Vxx[i + nx * i] + Vxx[i + nx * i]);
  if (!isnan(preg))
    for (int i = 0; i < nx; ++i) Vxx[i + nx * i] += preg;
}

/* SolverDDP::backwardPass, src/ddp.cpp:265-348 (used by SQP and FDDP). */
static int ddp_backward(prob* q, double preg, double dreg, int is_feasible) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  double* VxxT = q->Vxx + (size_t)T * nx * nx;
  double* VxT = q->Vx + (size_t)T * nx;
  memcpy(VxxT, q->LxxT, sizeof(double) * nx * nx);      /* :268 */
  memcpy(VxT, q->LxT, sizeof(double) * nx);             /* :269 */
  if (!isnan(preg))                                     /* :271-273 */
    for (int i = 0; i < nx; ++i) VxxT[i + nx * i] += preg;
  if (!is_feasible)                                     /* :275-277 */
    gemv_acc(VxT, VxxT, nx, nx, q->fs + (size_t)T * nx);

  for (int t = T - 1; t >= 0; --t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* Vxx_p = q->Vxx + (size_t)(t + 1) * nx * nx;
    const double* Vx_p = q->Vx + (size_t)(t + 1) * nx;
    double* Qu = q->Qu + (size_t)t * nu;

    memset(q->FxTVxx, 0, sizeof(double) * nx * nx);     /* :289 */
    gemm_tn_acc(q->FxTVxx, Fx, Vxx_p, nx, nx, nx);
    memcpy(q->Qx, q->Lx + (size_t)t * nx, sizeof(double) * nx); /* :291-292 */
    gemv_t_acc(q->Qx, Fx, nx, nx, Vx_p);
    memcpy(q->Qxx, q->Lxx + (size_t)t * nx * nx, sizeof(double) * nx * nx); /* :295-296 */
    gemm_nn_acc(q->Qxx, q->FxTVxx, Fx, nx, nx, nx);
    if (nu != 0) {
      memset(q->FuTVxx, 0, sizeof(double) * nu * nx);   /* :299 */
      gemm_tn_acc(q->FuTVxx, Fu, Vxx_p, nx, nu, nx);
      memcpy(Qu, q->Lu + (size_t)t * nu, sizeof(double) * nu); /* :301-302 */
      gemv_t_acc(Qu, Fu, nx, nu, Vx_p);
      memcpy(q->Quu, q->Luu + (size_t)t * nu * nu, sizeof(double) * nu * nu); /* :305-306 */
      gemm_nn_acc(q->Quu, q->FuTVxx, Fu, nu, nx, nu);
      memcpy(q->Qxu, q->Lxu + (size_t)t * nx * nu, sizeof(double) * nx * nu); /* :309-310 */
      gemm_nn_acc(q->Qxu, q->FxTVxx, Fu, nx, nx, nu);
      if (!isnan(dreg))                                 /* :312-314 */
        for (int a = 0; a < nu; ++a) q->Quu[a + nu * a] += dreg;
    }
    if (compute_gains(q, t)) return MIMGPU_STATUS_LLT | (t << 8); /* :317 */
    value_update(q, t, preg, 1);                        /* :319-333 */
    double* Vx = q->Vx + (size_t)t * nx;
    double* Vxx = q->Vxx + (size_t)t * nx * nx;
    if (!is_feasible)                                   /* :336-338 */
      gemv_acc(Vx, Vxx, nx, nx, q->fs + (size_t)t * nx);
    if (raise_if_nan(inf_norm(Vx, nx)) || raise_if_nan(inf_norm(Vxx, nx * nx)))
      return MIMGPU_STATUS_NONFINITE | (t << 8);        /* :340-345 */
  }
  return 0;
}

/* gap_norm_ of SolverSQP/FDDP::computeDirection, src/sqp.cpp:229-234,
 * src/fddp.cpp:195-200: includes fs_[0]. */
static double gap_norm_all(const prob* q) {
  double s = 0.0;
  for (int t = 0; t < q->T; ++t) s += l1_norm(q->fs + (size_t)t * q->nx, q->nx);
  s += l1_norm(q->fs + (size_t)q->T * q->nx, q->nx);
  return s;
}

/* SolverSQP::forwardPass, src/sqp.cpp:270-305.  dx_[0] is never written. */
static void sqp_forward(prob* q) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  q->xgrad = 0; q->ugrad = 0;
  for (int t = 0; t < T; ++t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* Vxx = q->Vxx + (size_t)t * nx * nx;
    const double* Vx = q->Vx + (size_t)t * nx;
    const double* K = q->K + (size_t)t * nu * nx;
    const double* k = q->k + (size_t)t * nu;
    const double* fs = q->fs + (size_t)t * nx;
    double* dx = q->dx + (size_t)t * nx;
    double* du = q->du + (size_t)t * nu;
    double* dxn = q->dx + (size_t)(t + 1) * nx;
    double* lag = q->lag + (size_t)t * nx;
    for (int i = 0; i < nx; ++i) q->tmpx[i] = dx[i] - fs[i];    /* :282 */
    memset(lag, 0, sizeof(double) * nx);                        /* :283-284 */
    gemv_acc(lag, Vxx, nx, nx, q->tmpx);
    for (int i = 0; i < nx; ++i) lag[i] += Vx[i];
    for (int a = 0; a < nu; ++a) {                              /* :285-286 */
      double s = 0.0;
      for (int i = 0; i < nx; ++i) s += K[a * nx + i] * dx[i];
      du[a] = -s;
      du[a] -= k[a];
    }
    memcpy(dxn, q->fs + (size_t)(t + 1) * nx, sizeof(double) * nx); /* :287 */
    gemv_acc(dxn, Fu, nx, nu, du);                              /* :288 */
    gemv_acc(dxn, Fx, nx, nx, dx);                              /* :289 */
    q->xgrad += l1_norm(dx, nx);                                /* :290-293 */
    q->ugrad += l1_norm(du, nu);
  }
  {
    const double* Vxx = q->Vxx + (size_t)T * nx * nx;
    const double* Vx = q->Vx + (size_t)T * nx;
    const double* dx = q->dx + (size_t)T * nx;
    const double* fs = q->fs + (size_t)T * nx;
    double* lag = q->lag + (size_t)T * nx;
    memcpy(lag, Vx, sizeof(double) * nx);                       /* :296 */
    for (int i = 0; i < nx; ++i) q->tmpx[i] = dx[i] - fs[i];    /* :297 */
    gemv_acc(lag, Vxx, nx, nx, q->tmpx);                        /* :298 */
    q->xgrad += l1_norm(dx, nx);                                /* :299-301 */
  }
  q->xgrad = q->xgrad / (double)(T + 1);                        /* :302 */
  q->ugrad = q->ugrad / (double)T;                              /* :303 */
}

/* SolverSQP::checkKKTConditions, src/sqp.cpp:244-268 (same maths in
 * src/ddp.cpp:177-198 and src/fddp.cpp:206-227). */
static void sqp_kkt(prob* q) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  double kkt = 0.0;
  for (int t = 0; t < T; ++t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* lagn = q->lag + (size_t)(t + 1) * nx;
    const double* lag = q->lag + (size_t)t * nx;
    memcpy(q->tmpx, q->Lx + (size_t)t * nx, sizeof(double) * nx);
    gemv_t_acc(q->tmpx, Fx, nx, nx, lagn);
    for (int i = 0; i < nx; ++i) q->tmpx[i] -= lag[i];
    memcpy(q->tmpu, q->Lu + (size_t)t * nu, sizeof(double) * nu);
    gemv_t_acc(q->tmpu, Fu, nx, nu, lagn);
    kkt = std_max(kkt, inf_norm(q->tmpx, nx));
    kkt = std_max(kkt, inf_norm(q->tmpu, nu));
  }
  for (int i = 0; i < nx; ++i)
    q->tmpx[i] = q->LxT[i] - q->lag[(size_t)T * nx + i];
  kkt = std_max(kkt, inf_norm(q->tmpx, nx));
  kkt = std_max(kkt, inf_norm(q->fs, (T + 1) * nx));  /* fs_flat_ */
  q->kkt = kkt;
}

/* SolverFDDP::updateExpectedImprovement, src/fddp.cpp:256-279. */
static void fddp_update_expected_improvement(prob* q, int is_feasible) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  double dg = 0, dq = 0;
  if (!is_feasible) {
    const double* Vx = q->Vx + (size_t)T * nx;
    const double* Vxx = q->Vxx + (size_t)T * nx * nx;
    const double* fs = q->fs + (size_t)T * nx;
    double s = 0; for (int i = 0; i < nx; ++i) s += Vx[i] * fs[i];
    dg -= s;
    memset(q->tmpx, 0, sizeof(double) * nx);
    gemv_acc(q->tmpx, Vxx, nx, nx, fs);
    s = 0; for (int i = 0; i < nx; ++i) s += fs[i] * q->tmpx[i];
    dq += s;
  }
  for (int t = 0; t < T; ++t) {
    if (nu != 0) {
      const double* Qu = q->Qu + (size_t)t * nu;
      const double* k = q->k + (size_t)t * nu;
      const double* Quuk = q->Quuk + (size_t)t * nu;
      double s = 0; for (int a = 0; a < nu; ++a) s += Qu[a] * k[a];
      dg += s;
      s = 0; for (int a = 0; a < nu; ++a) s += k[a] * Quuk[a];
      dq -= s;
    }
    if (!is_feasible) {
      const double* Vx = q->Vx + (size_t)t * nx;
      const double* Vxx = q->Vxx + (size_t)t * nx * nx;
      const double* fs = q->fs + (size_t)t * nx;
      double s = 0; for (int i = 0; i < nx; ++i) s += Vx[i] * fs[i];
      dg -= s;
      memset(q->tmpx, 0, sizeof(double) * nx);
      gemv_acc(q->tmpx, Vxx, nx, nx, fs);
      s = 0; for (int i = 0; i < nx; ++i) s += fs[i] * q->tmpx[i];
      dq += s;
    }
  }
  q->dg = dg; q->dq = dq;
}

/* ------------------------------------------------------------------ CSQP */

static int nc_at(const prob* q, int t) { return t < q->T ? q->ng : q->ngT; }
static size_t coff(const prob* q, int t) { return (size_t)t * q->ng; }

/* SolverCSQP::apply_rho_update, src/csqp.cpp:539-582. */
static void csqp_apply_rho_update(prob* q, double rho_sparse_tmp, double rho_min) {
  for (int t = 0; t <= q->T; ++t) {
    const int nc = nc_at(q, t);
    const double* lb = t < q->T ? q->glb + coff(q, t) : q->glbT;
    const double* ub = t < q->T ? q->gub + coff(q, t) : q->gubT;
    double* rho = q->rho + coff(q, t);
    double* irho = q->irho + coff(q, t);
    for (int k = 0; k < nc; ++k) {
      if (lb[k] == -INFINITY && ub[k] == INFINITY) {
        rho[k] = rho_min; irho[k] = 1. / rho_min;
      } else if (fabs(lb[k] - ub[k]) <= 1e-6) {
        rho[k] = 1e3 * rho_sparse_tmp; irho[k] = 1. / (1e3 * rho_sparse_tmp);
      } else if (lb[k] < ub[k]) {
        rho[k] = rho_sparse_tmp; irho[k] = 1. / rho_sparse_tmp;
      }
    }
  }
}

/* SolverCSQP::reset_rho_vec, src/csqp.cpp:534-537. */
static void csqp_reset_rho_vec(prob* q, const mimgpu_csqp_params* p) {
  q->rho_sparse = p->rho_sparse_base;
  csqp_apply_rho_update(q, q->rho_sparse, p->rho_min);
}

/* SolverCSQP::reset_params, src/csqp.cpp:170-193. */
static void csqp_reset_params(prob* q, const mimgpu_csqp_params* p) {
  if (p->reset_rho) csqp_reset_rho_vec(q, p);
  const size_t nc = (size_t)q->T * q->ng + q->ngT;
  memset(q->z, 0, sizeof(double) * nc);
  memset(q->zprev, 0, sizeof(double) * nc);
  memset(q->zrel, 0, sizeof(double) * nc);
  if (p->reset_y) memset(q->y, 0, sizeof(double) * nc);
}

/* SolverCSQP::calc (recalc=false part), src/csqp.cpp:417-451. */
static void csqp_calc_norms(prob* q) {
  const int T = q->T, nx = q->nx;
  double gap = 0., con = 0.;
  for (int t = 0; t < T; ++t) {
    gap += l1_norm(q->fs + (size_t)(t + 1) * nx, nx);
    const int nc = q->ng;
    double s1 = 0, s2 = 0;
    for (int k = 0; k < nc; ++k) {
      double a = q->glb[coff(q, t) + k] - q->g[coff(q, t) + k];
      s1 += fabs(a > 0. ? a : 0.);
    }
    con += s1;
    for (int k = 0; k < nc; ++k) {
      double a = q->g[coff(q, t) + k] - q->gub[coff(q, t) + k];
      s2 += fabs(a > 0. ? a : 0.);
    }
    con += s2;
  }
  {
    double s1 = 0, s2 = 0;
    for (int k = 0; k < q->ngT; ++k) {
      double a = q->glbT[k] - q->gT[k];
      s1 += fabs(a > 0. ? a : 0.);
    }
    con += s1;
    for (int k = 0; k < q->ngT; ++k) {
      double a = q->gT[k] - q->gubT[k];
      s2 += fabs(a > 0. ? a : 0.);
    }
    con += s2;
  }
  q->gap_norm = gap; q->constraint_norm = con;
}

/* SolverCSQP::backwardPass_without_constraints, src/csqp.cpp:804-898. */
static int csqp_backward_noconstr(prob* q, double preg, double dreg) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  double* VxxT = q->Vxx + (size_t)T * nx * nx;
  memcpy(VxxT, q->LxxT, sizeof(double) * nx * nx);            /* :827 */
  memcpy(q->Vx + (size_t)T * nx, q->LxT, sizeof(double) * nx); /* :828 */
  if (!isnan(preg)) for (int i = 0; i < nx; ++i) VxxT[i + nx * i] += preg;
  for (int t = T - 1; t >= 0; --t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* Vxx_p = q->Vxx + (size_t)(t + 1) * nx * nx;
    const double* Vx_p = q->Vx + (size_t)(t + 1) * nx;
    double* Qu = q->Qu + (size_t)t * nu;
    memset(q->tmpVx, 0, sizeof(double) * nx);                  /* :844-845 */
    gemv_acc(q->tmpVx, Vxx_p, nx, nx, q->fs + (size_t)(t + 1) * nx);
    for (int i = 0; i < nx; ++i) q->tmpVx[i] += Vx_p[i];
    memset(q->FxTVxx, 0, sizeof(double) * nx * nx);            /* :848 */
    gemm_tn_acc(q->FxTVxx, Fx, Vxx_p, nx, nx, nx);
    memcpy(q->Qx, q->Lx + (size_t)t * nx, sizeof(double) * nx); /* :850-851 */
    gemv_t_acc(q->Qx, Fx, nx, nx, q->tmpVx);
    memcpy(q->Qxx, q->Lxx + (size_t)t * nx * nx, sizeof(double) * nx * nx);
    gemm_nn_acc(q->Qxx, q->FxTVxx, Fx, nx, nx, nx);            /* :854-856 */
    if (nu != 0) {
      memset(q->FuTVxx, 0, sizeof(double) * nu * nx);          /* :859 */
      gemm_tn_acc(q->FuTVxx, Fu, Vxx_p, nx, nu, nx);
      memcpy(Qu, q->Lu + (size_t)t * nu, sizeof(double) * nu); /* :861-862 */
      gemv_t_acc(Qu, Fu, nx, nu, q->tmpVx);
      memcpy(q->Quu, q->Luu + (size_t)t * nu * nu, sizeof(double) * nu * nu);
      gemm_nn_acc(q->Quu, q->FuTVxx, Fu, nu, nx, nu);          /* :866-867 */
      memcpy(q->Qxu, q->Lxu + (size_t)t * nx * nu, sizeof(double) * nx * nu);
      gemm_nn_acc(q->Qxu, q->FxTVxx, Fu, nx, nx, nu);          /* :870-871 */
      if (!isnan(dreg)) for (int a = 0; a < nu; ++a) q->Quu[a + nu * a] += dreg;
    }
    if (compute_gains(q, t)) return MIMGPU_STATUS_LLT | (t << 8); /* :879 */
    value_update(q, t, preg, 0);                                /* :881-895 */
  }
  return 0;
}

/* SolverCSQP::forwardPass_without_constraints, src/csqp.cpp:649-669
 * (tilde = 0), and SolverCSQP::forwardPass, src/csqp.cpp:630-647 (tilde = 1). */
static void csqp_forward(prob* q, int tilde) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  double* X = tilde ? q->dxt : q->dx;
  double* U = tilde ? q->dut : q->du;
  for (int t = 0; t < T; ++t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* K = q->K + (size_t)t * nu * nx;
    const double* k = q->k + (size_t)t * nu;
    double* x = X + (size_t)t * nx;
    double* u = U + (size_t)t * nu;
    double* xn = X + (size_t)(t + 1) * nx;
    for (int a = 0; a < nu; ++a) {
      double s = 0.0;
      for (int i = 0; i < nx; ++i) s += K[a * nx + i] * x[i];
      u[a] = -k[a];
      u[a] -= s;
    }
    memcpy(xn, q->fs + (size_t)(t + 1) * nx, sizeof(double) * nx);
    gemv_acc(xn, Fx, nx, nx, x);
    gemv_acc(xn, Fu, nx, nu, u);
  }
}

/* tmp_dual_cwise_[t] = y_[t] - rho_vec_[t] o z_[t] */
static void dual_term(prob* q, int t) {
  const int nc = nc_at(q, t);
  const size_t o = coff(q, t);
  for (int k = 0; k < nc; ++k) {
    q->tdual[k] = q->y[o + k];
    q->tdual[k] -= q->rho[o + k] * q->z[o + k];
  }
}

/* SolverCSQP::backwardPass (full sweep), src/csqp.cpp:671-802. */
static int csqp_backward_full(prob* q, const mimgpu_csqp_params* p) {
  const int T = q->T, nx = q->nx, nu = q->nu, ng = q->ng, ngT = q->ngT;
  const double sigma = p->sigma, preg = p->preg, dreg = p->dreg;
  double* VxxT = q->Vxx + (size_t)T * nx * nx;
  double* VxT = q->Vx + (size_t)T * nx;
  memcpy(VxxT, q->LxxT, sizeof(double) * nx * nx);             /* :693 */
  for (int i = 0; i < nx; ++i) VxxT[i + nx * i] += sigma;      /* :694 */
  memcpy(VxT, q->LxT, sizeof(double) * nx);                    /* :695 */
  for (int i = 0; i < nx; ++i) VxT[i] -= sigma * q->dx[(size_t)T * nx + i]; /* :696 */
  if (ngT) {                                                    /* :698-704 */
    const double* rho = q->rho + coff(q, T);
    for (int j = 0; j < nx; ++j)
      for (int r = 0; r < ngT; ++r) q->trg[r + ngT * j] = rho[r] * q->GxT[r + ngT * j];
    gemm_tn_acc(VxxT, q->GxT, q->trg, ngT, nx, nx);
    dual_term(q, T);
    gemv_t_acc(VxT, q->GxT, ngT, nx, q->tdual);
  }
  if (!isnan(preg)) for (int i = 0; i < nx; ++i) VxxT[i + nx * i] += preg; /* :705-707 */

  for (int t = T - 1; t >= 0; --t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* Gx = ng ? q->Gx + (size_t)t * ng * nx : 0;
    const double* Gu = ng ? q->Gu + (size_t)t * ng * nu : 0;
    const double* rho = q->rho + coff(q, t);
    const double* Vxx_p = q->Vxx + (size_t)(t + 1) * nx * nx;
    const double* Vx_p = q->Vx + (size_t)(t + 1) * nx;
    double* Vxx_fs = q->Vxx_fs + (size_t)t * nx;
    double* Qu = q->Qu + (size_t)t * nu;
    const int nc = ng;

    memset(Vxx_fs, 0, sizeof(double) * nx);                    /* :720 */
    gemv_acc(Vxx_fs, Vxx_p, nx, nx, q->fs + (size_t)(t + 1) * nx);
    for (int i = 0; i < nx; ++i) q->tmpVx[i] = Vxx_fs[i] + Vx_p[i]; /* :721 */

    memset(q->FxTVxx, 0, sizeof(double) * nx * nx);            /* :725 */
    gemm_tn_acc(q->FxTVxx, Fx, Vxx_p, nx, nx, nx);
    memcpy(q->Qx, q->Lx + (size_t)t * nx, sizeof(double) * nx); /* :727 */
    for (int i = 0; i < nx; ++i) q->Qx[i] -= sigma * q->dx[(size_t)t * nx + i]; /* :728 */
    if (nc != 0) {                                              /* :729-737 */
      if (t > 0 || nu != 0) dual_term(q, t);
      if (t > 0) gemv_t_acc(q->Qx, Gx, nc, nx, q->tdual);
    }
    gemv_t_acc(q->Qx, Fx, nx, nx, q->tmpVx);                   /* :739 */

    memcpy(q->Qxx, q->Lxx + (size_t)t * nx * nx, sizeof(double) * nx * nx); /* :743 */
    for (int i = 0; i < nx; ++i) q->Qxx[i + nx * i] += sigma;  /* :744 */
    if (t > 0 && nc != 0) {                                     /* :745-748 */
      for (int j = 0; j < nx; ++j)
        for (int r = 0; r < nc; ++r) q->trg[r + nc * j] = rho[r] * Gx[r + nc * j];
      gemm_tn_acc(q->Qxx, Gx, q->trg, nc, nx, nx);
    }
    gemm_nn_acc(q->Qxx, q->FxTVxx, Fx, nx, nx, nx);            /* :749 */

    if (nu != 0) {
      memset(q->FuTVxx, 0, sizeof(double) * nu * nx);          /* :754 */
      gemm_tn_acc(q->FuTVxx, Fu, Vxx_p, nx, nu, nx);
      for (int a = 0; a < nu; ++a)                              /* :755 */
        Qu[a] = q->Lu[(size_t)t * nu + a] - sigma * q->du[(size_t)t * nu + a];
      if (nc != 0) gemv_t_acc(Qu, Gu, nc, nu, q->tdual);       /* :756-758 */
      gemv_t_acc(Qu, Fu, nx, nu, q->tmpVx);                    /* :759 */

      memcpy(q->Quu, q->Luu + (size_t)t * nu * nu, sizeof(double) * nu * nu); /* :763 */
      for (int a = 0; a < nu; ++a) q->Quu[a + nu * a] += sigma; /* :764 */
      gemm_nn_acc(q->Quu, q->FuTVxx, Fu, nu, nx, nu);          /* :765 */
      if (nc != 0) {                                            /* :766-769 */
        for (int j = 0; j < nu; ++j)
          for (int r = 0; r < nc; ++r) q->trg[r + nc * j] = rho[r] * Gu[r + nc * j];
        gemm_tn_acc(q->Quu, Gu, q->trg, nc, nu, nu);
      }
      if (!isnan(dreg)) for (int a = 0; a < nu; ++a) q->Quu[a + nu * a] += dreg; /* :770-772 */

      memcpy(q->Qxu, q->Lxu + (size_t)t * nx * nu, sizeof(double) * nx * nu); /* :776 */
      if (t > 0 && nc != 0)                                     /* :777-779, trg = rho o Gu */
        gemm_tn_acc(q->Qxu, Gx, q->trg, nc, nx, nu);
      gemm_nn_acc(q->Qxu, q->FxTVxx, Fu, nx, nx, nu);          /* :780 */
    }
    if (compute_gains(q, t)) return MIMGPU_STATUS_LLT | (t << 8); /* :783 */
    value_update(q, t, preg, 0);                                /* :784-799 */
  }
  return 0;
}

/* SolverCSQP::backwardPass_without_rho_update (light sweep),
 * src/csqp.cpp:1025-1108.  Re-uses Vxx_, K_, Quu_llt_, Vxx_fs_. */
static void csqp_backward_light(prob* q, const mimgpu_csqp_params* p) {
  const int T = q->T, nx = q->nx, nu = q->nu, ng = q->ng, ngT = q->ngT;
  const double sigma = p->sigma;
  double* VxT = q->Vx + (size_t)T * nx;
  memcpy(VxT, q->LxT, sizeof(double) * nx);                    /* :1045 */
  for (int i = 0; i < nx; ++i) VxT[i] -= sigma * q->dx[(size_t)T * nx + i]; /* :1046 */
  if (ngT) {                                                    /* :1048-1052 */
    dual_term(q, T);
    gemv_t_acc(VxT, q->GxT, ngT, nx, q->tdual);
  }
  for (int t = T - 1; t >= 0; --t) {
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* Gx = ng ? q->Gx + (size_t)t * ng * nx : 0;
    const double* Gu = ng ? q->Gu + (size_t)t * ng * nu : 0;
    const double* Vx_p = q->Vx + (size_t)(t + 1) * nx;
    const double* Vxx_fs = q->Vxx_fs + (size_t)t * nx;
    const double* K = q->K + (size_t)t * nu * nx;
    double* Qu = q->Qu + (size_t)t * nu;
    double* k = q->k + (size_t)t * nu;
    double* Vx = q->Vx + (size_t)t * nx;
    const int nc = ng;
    for (int i = 0; i < nx; ++i) q->tmpVx[i] = Vxx_fs[i] + Vx_p[i]; /* :1068 */
    memcpy(q->Qx, q->Lx + (size_t)t * nx, sizeof(double) * nx); /* :1069 */
    for (int i = 0; i < nx; ++i) q->Qx[i] -= sigma * q->dx[(size_t)t * nx + i]; /* :1070 */
    if (nc != 0) {                                              /* :1071-1079 */
      if (t > 0 || nu != 0) dual_term(q, t);
      if (t > 0) gemv_t_acc(q->Qx, Gx, nc, nx, q->tdual);
    }
    gemv_t_acc(q->Qx, Fx, nx, nx, q->tmpVx);                   /* :1080 */
    if (nu != 0) {                                              /* :1084-1093 */
      memcpy(Qu, q->Lu + (size_t)t * nu, sizeof(double) * nu);
      for (int a = 0; a < nu; ++a) Qu[a] -= sigma * q->du[(size_t)t * nu + a];
      if (nc != 0) gemv_t_acc(Qu, Gu, nc, nu, q->tdual);
      gemv_t_acc(Qu, Fu, nx, nu, q->tmpVx);
    }
    memcpy(k, Qu, sizeof(double) * nu);                         /* :1096-1097 */
    llt_solve(q->L + (size_t)t * nu * nu, nu, k, 1);
    memcpy(Vx, q->Qx, sizeof(double) * nx);                     /* :1101-1104 */
    if (nu != 0)
      for (int i = 0; i < nx; ++i) {
        double s = 0.0;
        for (int a = 0; a < nu; ++a) s += K[a * nx + i] * Qu[a];
        Vx[i] -= s;
      }
  }
}

/* SolverCSQP::update_lagrangian_parameters, src/csqp.cpp:1196-1340. */
static void csqp_update_lagrangian(prob* q, const mimgpu_csqp_params* p, int iter) {
  const int T = q->T, nx = q->nx, nu = q->nu;
  const double alpha = p->alpha;
  const int interval = p->rho_update_interval;
  q->norm_primal = -INFINITY; q->norm_dual = -INFINITY;        /* :1201-1204 */
  q->norm_primal_rel = -INFINITY; q->norm_dual_rel = -INFINITY;
  for (int t = 0; t <= T; ++t) {
    const int nc = nc_at(q, t);
    const int last = (t == T);
    double* dx = q->dx + (size_t)t * nx;
    const double* dxt = q->dxt + (size_t)t * nx;
    if (nc == 0) {                                              /* :1222-1226, :1283 */
      memcpy(dx, dxt, sizeof(double) * nx);
      if (!last) memcpy(q->du + (size_t)t * nu, q->dut + (size_t)t * nu, sizeof(double) * nu);
      continue;
    }
    const size_t o = coff(q, t);
    const double* Gx = last ? q->GxT : q->Gx + (size_t)t * nc * nx;
    const double* Gu = last ? 0 : q->Gu + (size_t)t * nc * nu;
    const double* g = last ? q->gT : q->g + o;
    const double* lb = last ? q->glbT : q->glb + o;
    const double* ub = last ? q->gubT : q->gub + o;
    double *z = q->z + o, *zp = q->zprev + o, *zr = q->zrel + o, *y = q->y + o;
    const double *rho = q->rho + o, *irho = q->irho + o;
    memcpy(zp, z, sizeof(double) * nc);                         /* :1228 */
    memset(q->tc, 0, sizeof(double) * nc);                      /* :1229-1230 */
    gemv_acc(q->tc, Gx, nc, nx, dxt);
    if (!last) gemv_acc(q->tc, Gu, nc, nu, q->dut + (size_t)t * nu);
    for (int k = 0; k < nc; ++k) {                              /* :1231-1232 */
      zr[k] = alpha * q->tc[k];
      zr[k] += (1. - alpha) * z[k];
    }
    for (int k = 0; k < nc; ++k) q->tdual[k] = y[k] * irho[k];  /* :1234 */
    for (int k = 0; k < nc; ++k) {                              /* :1236-1237 */
      double v = zr[k] + q->tdual[k];
      double lo = lb[k] - g[k], hi = ub[k] - g[k];
      v = (v < lo) ? lo : v;      /* cwiseMax(lb - g) */
      v = (hi < v) ? hi : v;      /* cwiseMin(ub - g) */
      z[k] = v;
    }
    for (int k = 0; k < nc; ++k) y[k] += rho[k] * (zr[k] - z[k]); /* :1239 */
    memcpy(dx, dxt, sizeof(double) * nx);                       /* :1241-1242 */
    if (!last) memcpy(q->du + (size_t)t * nu, q->dut + (size_t)t * nu, sizeof(double) * nu);

    if (iter % interval == 0) {                                 /* :1244-1280 */
      for (int k = 0; k < nc; ++k) q->tdual[k] = rho[k] * (z[k] - zp[k]);
      if (p->update_rho_with_heuristic) {
        q->norm_dual = std_max(q->norm_dual, inf_norm(q->tdual, nc));
      } else {
        memset(q->tmpx, 0, sizeof(double) * nx);
        gemv_t_acc(q->tmpx, Gx, nc, nx, q->tdual);
        q->norm_dual = std_max(q->norm_dual, inf_norm(q->tmpx, nx));
        if (!last) {
          memset(q->tmpu, 0, sizeof(double) * nu);
          gemv_t_acc(q->tmpu, Gu, nc, nu, q->tdual);
          q->norm_dual = std_max(q->norm_dual, inf_norm(q->tmpu, nu));
        }
      }
      {
        double m = 0.0;
        for (int k = 0; k < nc; ++k) { double a = fabs(q->tc[k] - z[k]); if (a > m) m = a; }
        q->norm_primal = std_max(q->norm_primal, m);
      }
      q->norm_primal_rel = std_max(q->norm_primal_rel, inf_norm(q->tc, nc));
      q->norm_primal_rel = std_max(q->norm_primal_rel, inf_norm(z, nc));
      if (p->update_rho_with_heuristic) {
        q->norm_dual_rel = std_max(q->norm_dual_rel, inf_norm(y, nc));
      } else {
        memset(q->tmpx, 0, sizeof(double) * nx);
        gemv_t_acc(q->tmpx, Gx, nc, nx, y);
        q->norm_dual_rel = std_max(q->norm_dual_rel, inf_norm(q->tmpx, nx));
        if (!last) {
          memset(q->tmpu, 0, sizeof(double) * nu);
          gemv_t_acc(q->tmpu, Gu, nc, nu, y);
          q->norm_dual_rel = std_max(q->norm_dual_rel, inf_norm(q->tmpu, nu));
        }
      }
    }
  }
}

/* SolverCSQP::update_rho_vec, src/csqp.cpp:519-532. */
static void csqp_update_rho_vec(prob* q, const mimgpu_csqp_params* p, int iter) {
  const double scale = sqrt((q->norm_primal * q->norm_dual_rel) /
                            (q->norm_dual * q->norm_primal_rel));
  q->rho_estimate = std_min(std_max(scale * q->rho_sparse, p->rho_min), p->rho_max);
  if (iter % p->rho_update_interval == 0) {
    if (q->rho_estimate > q->rho_sparse * p->adaptive_rho_tolerance ||
        q->rho_estimate < q->rho_sparse / p->adaptive_rho_tolerance) {
      q->rho_sparse = q->rho_estimate;
      csqp_apply_rho_update(q, q->rho_sparse, p->rho_min);
    }
  }
}

/* SolverCSQP::computeDirection, src/csqp.cpp:454-517 (no time budget). */
static int csqp_compute_direction(prob* q, const mimgpu_csqp_params* p) {
  int st;
  q->converged = 0;
  csqp_reset_params(q, p);                                      /* :459 */
  if (p->equality_qp_initial_guess) {                           /* :461-464 */
    if ((st = csqp_backward_noconstr(q, p->preg, p->dreg))) return st;
    csqp_forward(q, 0);
  }
  const int interval = p->rho_update_interval;
  for (q->qp_iters = 1; q->qp_iters < p->max_qp_iters + 1; ++q->qp_iters) { /* :470 */
    if (q->qp_iters % interval == 1 || interval == 1) {         /* :477 */
      if ((st = csqp_backward_full(q, p))) return st;
    } else {
      csqp_backward_light(q, p);
    }
    csqp_forward(q, 1);                                         /* :492 */
    csqp_update_lagrangian(q, p, q->qp_iters);                  /* :493 */
    csqp_update_rho_vec(q, p, q->qp_iters);                     /* :494 */
    if (q->qp_iters % interval == 0) {                          /* :497-512 */
      double tol_p, tol_d;
      if (fabs(p->eps_rel) <= 2.220446049250313e-16) {
        tol_p = p->eps_abs; tol_d = p->eps_abs;
      } else {
        tol_p = p->eps_abs + p->eps_rel * q->norm_primal_rel;
        tol_d = p->eps_abs + p->eps_rel * q->norm_dual_rel;
      }
      if (q->norm_primal <= tol_p && q->norm_dual <= tol_d) { q->converged = 1; break; }
    }
  }
  return 0;
}

/* SolverCSQP::checkKKTConditions, src/csqp.cpp:584-628. */
static void csqp_kkt(prob* q) {
  const int T = q->T, nx = q->nx, nu = q->nu, ng = q->ng, ngT = q->ngT;
  double kkt = 0.0;
  q->xgrad = 0.; q->ugrad = 0.;
  for (int t = 0; t < T + 1; ++t) {                             /* :590-593 */
    double* lag = q->lag + (size_t)t * nx;
    memcpy(lag, q->Vx + (size_t)t * nx, sizeof(double) * nx);
    gemv_acc(lag, q->Vxx + (size_t)t * nx * nx, nx, nx, q->dxt + (size_t)t * nx);
  }
  for (int t = 0; t < T; ++t) {                                 /* :598-614 */
    const double* Fx = q->Fx + (size_t)t * nx * nx;
    const double* Fu = q->Fu + (size_t)t * nx * nu;
    const double* lagn = q->lag + (size_t)(t + 1) * nx;
    const double* lag = q->lag + (size_t)t * nx;
    memcpy(q->tmpx, q->Lx + (size_t)t * nx, sizeof(double) * nx);
    gemv_t_acc(q->tmpx, Fx, nx, nx, lagn);
    for (int i = 0; i < nx; ++i) q->tmpx[i] -= lag[i];
    if (t > 0 && ng) gemv_t_acc(q->tmpx, q->Gx + (size_t)t * ng * nx, ng, nx, q->y + coff(q, t));
    kkt = std_max(kkt, inf_norm(q->tmpx, nx));
    memcpy(q->tmpu, q->Lu + (size_t)t * nu, sizeof(double) * nu);
    gemv_t_acc(q->tmpu, Fu, nx, nu, lagn);
    if (ng) gemv_t_acc(q->tmpu, q->Gu + (size_t)t * ng * nu, ng, nu, q->y + coff(q, t));
    kkt = std_max(kkt, inf_norm(q->tmpu, nu));
    q->xgrad += l1_norm(q->dxt + (size_t)t * nx, nx);
    q->ugrad += l1_norm(q->dut + (size_t)t * nu, nu);
  }
  for (int i = 0; i < nx; ++i) q->tmpx[i] = q->LxT[i] - q->lag[(size_t)T * nx + i]; /* :619-620 */
  if (ngT) gemv_t_acc(q->tmpx, q->GxT, ngT, nx, q->y + coff(q, T));     /* :621 */
  kkt = std_max(kkt, inf_norm(q->tmpx, nx));
  kkt = std_max(kkt, inf_norm(q->fs, (T + 1) * nx));            /* :623 */
  kkt = std_max(kkt, q->constraint_norm);                       /* :624 */
  q->xgrad += l1_norm(q->dxt + (size_t)T * nx, nx);
  q->xgrad = q->xgrad / (double)(T + 1);
  q->ugrad = q->ugrad / (double)T;
  q->kkt = kkt;
}

/* ------------------------------------------------------------ batch API */

static void* zalloc(size_t n, size_t sz) { return calloc(n ? n : 1, sz); }

ora_state* ora_state_create(const mimgpu_shape* sh) {
  if (!sh || sh->B < 1 || sh->T < 1 || sh->nx < 1 || sh->nu < 0 || sh->ng < 0 || sh->ngT < 0)
    return 0;
  ora_state* s = (ora_state*)calloc(1, sizeof(ora_state));
  s->shape = *sh;
  const size_t B = sh->B, T = sh->T, nx = sh->nx, nu = sh->nu;
  const size_t nc = T * sh->ng + sh->ngT;
  s->dx = zalloc(B * (T + 1) * nx, 8); s->du = zalloc(B * T * nu, 8);
  s->lag_mul = zalloc(B * (T + 1) * nx, 8);
  s->dxt = zalloc(B * (T + 1) * nx, 8); s->dut = zalloc(B * T * nu, 8);
  s->y = zalloc(B * nc, 8); s->z = zalloc(B * nc, 8); s->rho_vec = zalloc(B * nc, 8);
  s->Vxx = zalloc(B * (T + 1) * nx * nx, 8); s->Vx = zalloc(B * (T + 1) * nx, 8);
  s->K = zalloc(B * T * nu * nx, 8); s->kff = zalloc(B * T * nu, 8);
  s->Qu = zalloc(B * T * nu, 8); s->Quuk = zalloc(B * T * nu, 8);
  s->L = zalloc(B * T * nu * nu, 8); s->Vxx_fs = zalloc(B * T * nx, 8);
  s->z_prev = zalloc(B * nc, 8); s->z_relaxed = zalloc(B * nc, 8);
  s->inv_rho_vec = zalloc(B * nc, 8);
  s->scal = zalloc((size_t)MIMGPU_S_COUNT * B, 8);
  s->qp_iters = zalloc(B, 4); s->status = zalloc(B, 4); s->converged = zalloc(B, 4);
  /* KKT_ starts at +inf (include/mim_solvers/sqp.hpp:182-183) */
  for (size_t b = 0; b < B; ++b) s->scal[MIMGPU_S_KKT * B + b] = INFINITY;
  /* rho_sparse_ = rho_sparse_base_ default 1e-1 (src/csqp.cpp:50) */
  for (size_t b = 0; b < B; ++b) s->scal[MIMGPU_S_RHO_SPARSE * B + b] = 1e-1;
  return s;
}

void ora_state_destroy(ora_state* s) {
  if (!s) return;
  free(s->dx); free(s->du); free(s->lag_mul); free(s->dxt); free(s->dut);
  free(s->y); free(s->z); free(s->rho_vec); free(s->Vxx); free(s->Vx);
  free(s->K); free(s->kff); free(s->Qu); free(s->Quuk); free(s->L); free(s->Vxx_fs);
  free(s->z_prev); free(s->z_relaxed); free(s->inv_rho_vec); free(s->scal);
  free(s->qp_iters); free(s->status); free(s->converged);
  free(s);
}

int ora_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

static int nthreads_of(int n) {
  /* n <= 0: OpenMP's default team (honours OMP_NUM_THREADS); n > 0: exactly n, capped by the
   * number of processors (torchrun exports OMP_NUM_THREADS=1, the timed CPU arm asks explicitly) */
  int m = ora_max_threads();
  if (n <= 0) return m;
#ifdef _OPENMP
  m = omp_get_num_procs();
#endif
  return n > m ? m : n;
}

#define ORA_PARALLEL_BEGIN(s, nthreads)                                      \
  const int B_ = (s)->shape.B;                                               \
  const int nt_ = nthreads_of(nthreads);                                     \
  _Pragma("omp parallel num_threads(nt_)") {                                 \
    scratch w_ = scratch_alloc(&(s)->shape);                                 \
    _Pragma("omp for schedule(dynamic, 1)") for (int b = 0; b < B_; ++b) {   \
      prob q_;                                                               \
      prob* q = &q_;
#define ORA_PARALLEL_END(s)                                                  \
      unbind(q, (s), b);                                                     \
    }                                                                        \
    scratch_free(&w_);                                                       \
  }

int ora_sqp_direction_reg(ora_state* s, const mimgpu_derivs* d,
                          const mimgpu_sqp_params* p, const double* preg,
                          const double* dreg, int with_kkt, int nthreads) {
  if (!s || !d || !p) return MIMGPU_ERR_INVALID;
  ORA_PARALLEL_BEGIN(s, nthreads)
    bind(q, s, d, &w_, b);
    /* SolverSQP::computeDirection, src/sqp.cpp:224-242 */
    q->gap_norm = gap_norm_all(q);
    q->merit = q->cost + p->mu_dynamic * q->gap_norm;
    q->qp_iters = 0;
    q->status = ddp_backward(q, preg ? preg[b] : p->preg, dreg ? dreg[b] : p->dreg,
                             p->is_feasible);
    if (q->status == 0) {
      sqp_forward(q);
      if (with_kkt) sqp_kkt(q);
    }
  ORA_PARALLEL_END(s)
  return 0;
}

int ora_sqp_direction(ora_state* s, const mimgpu_derivs* d,
                      const mimgpu_sqp_params* p, int with_kkt, int nthreads) {
  return ora_sqp_direction_reg(s, d, p, 0, 0, with_kkt, nthreads);
}

int ora_fddp_direction(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_sqp_params* p, int nthreads) {
  if (!s || !d || !p) return MIMGPU_ERR_INVALID;
  ORA_PARALLEL_BEGIN(s, nthreads)
    bind(q, s, d, &w_, b);
    /* SolverFDDP::computeDirection, src/fddp.cpp:189-204 */
    q->gap_norm = gap_norm_all(q);
    q->status = ddp_backward(q, p->preg, p->dreg, p->is_feasible);
    if (q->status == 0) fddp_update_expected_improvement(q, p->is_feasible);
  ORA_PARALLEL_END(s)
  return 0;
}

int ora_csqp_calc_norms(ora_state* s, const mimgpu_derivs* d, int nthreads) {
  if (!s || !d) return MIMGPU_ERR_INVALID;
  ORA_PARALLEL_BEGIN(s, nthreads)
    bind(q, s, d, &w_, b);
    csqp_calc_norms(q);
  ORA_PARALLEL_END(s)
  return 0;
}

int ora_csqp_reset_rho(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_csqp_params* p, int nthreads) {
  if (!s || !d || !p) return MIMGPU_ERR_INVALID;
  ORA_PARALLEL_BEGIN(s, nthreads)
    bind(q, s, d, &w_, b);
    csqp_reset_rho_vec(q, p);
  ORA_PARALLEL_END(s)
  return 0;
}

int ora_csqp_direction(ora_state* s, const mimgpu_derivs* d,
                       const mimgpu_csqp_params* p, int with_kkt, int nthreads) {
  if (!s || !d || !p || p->rho_update_interval < 1) return MIMGPU_ERR_INVALID;
  ORA_PARALLEL_BEGIN(s, nthreads)
    bind(q, s, d, &w_, b);
    q->status = csqp_compute_direction(q, p);
    if (q->status == 0 && with_kkt) csqp_kkt(q);
  ORA_PARALLEL_END(s)
  return 0;
}
