/*
 * mimgpu.h -- C ABI of the B200-native batched OCP-QP engine.
 *
 * This is the drop-in boundary for the `computeDirection` hot path of
 * mim_solvers' SolverSQP / SolverCSQP / SolverFDDP, for a batch of B independent
 * shooting problems of one shape (T, nx, nu, ng, ngT).  The reference has no
 * FFI layer: its seam is C++ virtual dispatch on crocoddyl::SolverAbstract
 * subclasses.  Each entry point below names the reference member function(s) it
 * replaces (paths relative to the reference tree).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, all reals are IEEE binary64, all
 *     sizes int32; every function returns an int status (MIMGPU_OK == 0) and
 *     never throws.
 *   - Matrices are dense COLUMN-MAJOR blocks, i.e. the memory image of the
 *     Eigen::MatrixXd members of crocoddyl::ActionDataAbstract (d->Fx, d->Fu,
 *     d->Lxx, ...), so an adapter packs them with memcpy.  The only row-major
 *     block is the feedback gain K (nu x nx), as in the reference
 *     (include/mim_solvers/ddp.hpp:69-70,397).
 *   - Batch layout ("problem-major SoA"): one array per field, shaped
 *     [B][T][block] (running knots) or [B][block] (terminal knot), contiguous.
 *     `nx` is the tangent dimension ndx (src/ddp.cpp:467).
 *   - A ctx is bound to one device and one shape bucket; it owns all device
 *     buffers and the persistent per-problem ADMM state (y, z, rho_vec,
 *     rho_sparse, dx, du), mirroring the solver-object members of
 *     include/mim_solvers/csqp.hpp:250-357.  One host thread at a time per ctx.
 *     All device work is enqueued on the caller's stream.
 *   - Numerical failure is reported per problem in `status[b]` (the reference
 *     throws "backward_error": src/ddp.cpp:340-345,415-419), never by a
 *     non-zero return code.
 */
#ifndef MIMGPU_H_
#define MIMGPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIMGPU_ABI_VERSION 3

/* ---- return codes ------------------------------------------------------ */
#define MIMGPU_OK 0
#define MIMGPU_ERR_INVALID 1   /* bad argument / shape            */
#define MIMGPU_ERR_CUDA 2      /* CUDA runtime error (see mimgpu_last_error) */
#define MIMGPU_ERR_NODEVICE 3  /* no CUDA device: there is no CPU fallback */
#define MIMGPU_ERR_STATE 4     /* call order (e.g. direction before upload) */
#define MIMGPU_ERR_UNSUPPORTED 5

/* ---- per-problem status word (status[b]) --------------------------------
 * 0 ok; 1 = Quu not positive definite at some knot (Eigen LLT info != Success,
 * src/ddp.cpp:413-419); 2 = non-finite / >=1e30 Vx or Vxx (raiseIfNaN,
 * src/ddp.cpp:340-345).  Bits 8.. hold the knot index t of the first failure. */
#define MIMGPU_STATUS_OK 0
#define MIMGPU_STATUS_LLT 1
#define MIMGPU_STATUS_NONFINITE 2

typedef struct mimgpu_ctx mimgpu_ctx; /* opaque */

typedef struct {
  int32_t B;   /* problems in this ctx (this rank's shard)            */
  int32_t T;   /* running knots; there are T+1 states                  */
  int32_t nx;  /* tangent-space dimension ndx                          */
  int32_t nu;  /* controls (same at every running knot)                */
  int32_t ng;  /* inequality rows per running knot (0 = none)          */
  int32_t ngT; /* inequality rows at the terminal knot (0 = none)      */
} mimgpu_shape;

/* Per-knot derivatives, as produced by problem_->calcDiff (src/ddp.cpp:258,
 * src/csqp.cpp:413-414).  Host or device pointers (see mimgpu_upload). CSQP
 * pointers may be NULL when ng == 0 / ngT == 0. */
typedef struct {
  const double* Fx;   /* [B][T][nx*nx]  d->Fx  (col-major nx x nx)   */
  const double* Fu;   /* [B][T][nx*nu]  d->Fu  (nx x nu)             */
  const double* Lxx;  /* [B][T][nx*nx]                                 */
  const double* Lxu;  /* [B][T][nx*nu]  (nx x nu)                     */
  const double* Luu;  /* [B][T][nu*nu]                                 */
  const double* Lx;   /* [B][T][nx]                                    */
  const double* Lu;   /* [B][T][nu]                                    */
  const double* fs;   /* [B][T+1][nx]   gaps fs_[0..T]; fs_[t+1] is the
                         defect of knot t (src/csqp.cpp:431)           */
  const double* LxxT; /* [B][nx*nx]     terminal d_T->Lxx              */
  const double* LxT;  /* [B][nx]        terminal d_T->Lx               */
  const double* Gx;   /* [B][T][ng*nx]  d->Gx  (ng x nx)              */
  const double* Gu;   /* [B][T][ng*nu]  d->Gu  (ng x nu)              */
  const double* g;    /* [B][T][ng]     d->g                           */
  const double* g_lb; /* [B][T][ng]     m->get_g_lb()  (+-inf allowed) */
  const double* g_ub; /* [B][T][ng]     m->get_g_ub()                  */
  const double* GxT;  /* [B][ngT*nx]                                   */
  const double* gT;   /* [B][ngT]                                      */
  const double* g_lbT;/* [B][ngT]                                      */
  const double* g_ubT;/* [B][ngT]                                      */
  const double* cost; /* [B]            cost_ (may be NULL -> 0)       */
} mimgpu_derivs;

#define MIMGPU_MEM_HOST 0   /* pointers are host memory: async H2D copy      */
#define MIMGPU_MEM_DEVICE 1 /* pointers are device memory in the layout above:
                               attached zero-copy, caller keeps them alive and
                               may update them IN PLACE between direction
                               calls: every direction call re-reads them (for
                               the small shapes this includes the re-layout
                               into knot records -- use mimgpu_upload_records
                               to skip that pass)                             */

/* Knot-record input layout (the HBM layout the warp-per-problem SQP kernel reads; available for
 * the shapes mimgpu_get_record_layout accepts, ng = ngT = 0).  One problem = T running-knot records
 * followed by one terminal record, problems back to back:
 *   running knot t : [fs_t | Lxx | Lxu | Luu | Lx Lu | Fx Fu]
 *   terminal       : [fs_T | Lxx_T | Lx_T]
 * Lxu, Fx, Fu and the vectors are the dense column-major Eigen image: one memcpy each at
 * base + b * problem_doubles + t * rec_doubles + off_*.  The SYMMETRIC blocks Lxx, Luu, Lxx_T travel as
 * their lower triangle when sym_packed != 0 (16 % fewer bytes over PCIe and from HBM): column j of an
 * n x n block holds rows j .. n-1 -- contiguous in the Eigen image, one memcpy of n - j doubles from
 * &M(j, j) -- at off + j * n - j (j - 1) / 2; the strict upper triangle is not transmitted (the hot path
 * reads Lxx, Luu as symmetric; the reference's own are, up to rounding). */
typedef struct {
  int32_t rec_doubles;     /* doubles of one running-knot record                    */
  int32_t term_doubles;    /* doubles of the terminal record                        */
  int32_t problem_doubles; /* T * rec_doubles + term_doubles                        */
  int32_t off_fs, off_Lxx, off_Lxu, off_Luu, off_Lx, off_Lu, off_Fx, off_Fu; /* in a running record */
  int32_t toff_fs, toff_Lxx, toff_Lx;                                          /* in the terminal record */
  int32_t sym_packed;      /* != 0: Lxx, Luu, Lxx_T are packed lower triangles (see above) */
  int32_t reserved;
} mimgpu_record_layout;

/* SolverSQP / SolverFDDP direction parameters (src/sqp.cpp:91-97,224-242). */
typedef struct {
  double preg;        /* preg_  (NaN = skip, src/ddp.cpp:271)          */
  double dreg;        /* dreg_                                          */
  double mu_dynamic;  /* merit_ = cost_ + mu_dynamic_*gap_norm_         */
  int32_t is_feasible;/* SQP forces 0 (src/sqp.cpp:85)                  */
  int32_t reserved;
} mimgpu_sqp_params;

/* SolverCSQP QP parameters; defaults in include/mim_solvers/csqp.hpp:292-331 */
typedef struct {
  double preg, dreg;              /* 0 when remove_reg_ (src/csqp.cpp:216-219) */
  double sigma;                   /* 1e-6 */
  double alpha;                   /* 1.6  */
  double rho_sparse_base;         /* 1e-1 */
  double rho_min, rho_max;        /* 1e-6, 1e3 */
  double adaptive_rho_tolerance;  /* 5 */
  double eps_abs, eps_rel;        /* 1e-4, 1e-4 */
  int32_t max_qp_iters;           /* 1000 */
  int32_t rho_update_interval;    /* 25 */
  int32_t equality_qp_initial_guess; /* 1 */
  int32_t reset_y;                /* 0 */
  int32_t reset_rho;              /* 0 */
  int32_t update_rho_with_heuristic; /* 0 */
  int32_t warm_start_y;           /* unused, reserved (0) */
  int32_t reserved;
} mimgpu_csqp_params;

/* Downloadable / uploadable per-problem fields (solver members). */
enum mimgpu_field {
  MIMGPU_F_DX = 0,        /* dx_        [B][T+1][nx]  (SQP: direction; CSQP: prox centre) */
  MIMGPU_F_DU = 1,        /* du_        [B][T][nu]    */
  MIMGPU_F_LAG_MUL = 2,   /* lag_mul_   [B][T+1][nx]  */
  MIMGPU_F_DX_TILDE = 3,  /* dxtilde_   [B][T+1][nx]  */
  MIMGPU_F_DU_TILDE = 4,  /* dutilde_   [B][T][nu]    */
  MIMGPU_F_Y = 5,         /* y_         [B][T*ng+ngT] */
  MIMGPU_F_Z = 6,         /* z_         [B][T*ng+ngT] */
  MIMGPU_F_RHO_VEC = 7,   /* rho_vec_   [B][T*ng+ngT] */
  MIMGPU_F_VXX = 8,       /* Vxx_       [B][T+1][nx*nx] */
  MIMGPU_F_VX = 9,        /* Vx_        [B][T+1][nx]  */
  MIMGPU_F_K = 10,        /* K_         [B][T][nu*nx] row-major nu x nx */
  MIMGPU_F_KFF = 11,      /* k_         [B][T][nu]    */
  MIMGPU_F_QU = 12,       /* Qu_        [B][T][nu]    */
  MIMGPU_F_QUUK = 13,     /* Quuk_      [B][T][nu]    (SQP/FDDP only) */
  MIMGPU_F_QUU_LLT = 14,  /* lower Cholesky factor of Quu_ [B][T][nu*nu] col-major */
  MIMGPU_F_XS = 15,       /* xs_        [B][T+1][nx]  (device-side solve(): mimgpu_sqp_solve) */
  MIMGPU_F_US = 16,       /* us_        [B][T][nu]    */
  MIMGPU_F_COUNT = 17
};

/* Per-problem scalars ([B] each). */
enum mimgpu_scalar {
  MIMGPU_S_KKT = 0,
  MIMGPU_S_GAP_NORM = 1,
  MIMGPU_S_MERIT = 2,          /* SQP only: cost_ + mu_dynamic_*gap_norm_ */
  MIMGPU_S_XGRAD_NORM = 3,
  MIMGPU_S_UGRAD_NORM = 4,
  MIMGPU_S_NORM_PRIMAL = 5,
  MIMGPU_S_NORM_DUAL = 6,
  MIMGPU_S_NORM_PRIMAL_REL = 7,
  MIMGPU_S_NORM_DUAL_REL = 8,
  MIMGPU_S_RHO_SPARSE = 9,
  MIMGPU_S_CONSTRAINT_NORM = 10,
  MIMGPU_S_DG = 11,            /* FDDP dg_ (src/fddp.cpp:256-279) */
  MIMGPU_S_DQ = 12,            /* FDDP dq_ */
  MIMGPU_S_COST = 13,          /* cost_       (device-side solve()) */
  MIMGPU_S_STEPLENGTH = 14,    /* steplength_ of the last line search */
  MIMGPU_S_PREG = 15,          /* preg_ = dreg_ */
  MIMGPU_S_COUNT = 16
};

/* Batch-wide summary (the quantities of SURVEY 8(e) all-reduce). */
typedef struct {
  double max_kkt;
  int64_t sum_qp_iters;
  int64_t max_qp_iters;
  int64_t n_converged;  /* CSQP: problems whose ADMM loop broke on tolerance */
  int64_t n_failed;     /* status != 0 */
} mimgpu_summary;

/* ---- lifecycle ----------------------------------------------------------- */
int mimgpu_abi_version(void);
int mimgpu_device_count(int* n);
const char* mimgpu_last_error(void);

/* Allocates all device buffers for `shape` on `device`.  Replaces
 * SolverDDP::allocateData (src/ddp.cpp:447-502) and the SolverCSQP ctor
 * allocations (src/csqp.cpp:24-168): dx, du, dxtilde, dutilde, lag_mul, y, z
 * are zero-initialised exactly as there. */
int mimgpu_create(mimgpu_ctx** out, int device, const mimgpu_shape* shape);
void mimgpu_destroy(mimgpu_ctx* ctx);
int mimgpu_get_shape(const mimgpu_ctx* ctx, mimgpu_shape* shape);

/* ---- data in ------------------------------------------------------------- */
/* Makes the derivatives of all B problems visible to the kernels; the upload
 * point is right after problem_->calcDiff (src/ddp.cpp:253-263,
 * src/csqp.cpp:411-416).  MIMGPU_MEM_HOST: cudaMemcpyAsync on `stream` (pin
 * the host memory for overlap).  MIMGPU_MEM_DEVICE: pointers attached. */
int mimgpu_upload(mimgpu_ctx* ctx, const mimgpu_derivs* derivs, int memkind,
                  void* cuda_stream);

/* Block offsets of the knot-record layout for this ctx's shape; MIMGPU_ERR_UNSUPPORTED when the
 * shape is not served by the record kernel (use mimgpu_upload then). */
int mimgpu_get_record_layout(const mimgpu_ctx* ctx, mimgpu_record_layout* out);

/* Same upload point as mimgpu_upload (right after problem_->calcDiff, src/ddp.cpp:253-263) for a
 * producer that writes knot records: `records` = [B][problem_doubles], `cost` = [B] or NULL.
 * MIMGPU_MEM_DEVICE: attached zero-copy (16-byte aligned), the kernel reads the caller's memory at
 * every direction call and there is NO re-layout pass.  MIMGPU_MEM_HOST: one async H2D copy.
 * Entry points the record kernel does not serve (FDDP, CSQP with ng = 0, downloads of Qu_ / Quuk_ /
 * Quu_llt_) rebuild the field-major arrays on the device first. */
int mimgpu_upload_records(mimgpu_ctx* ctx, const double* records, const double* cost, int memkind,
                          void* cuda_stream);

/* Per-problem regularisation (solver members preg_/dreg_, src/ddp.cpp:431-445)
 * for batches whose problems are at different regularisation levels.  NULL
 * restores the scalar values of the params struct.  Host pointers, [B]. */
int mimgpu_set_reg(mimgpu_ctx* ctx, const double* preg, const double* dreg,
                   void* cuda_stream);

/* Restrict the next direction calls to a subset (the increaseRegularization
 * retry of src/sqp.cpp:103-116 / src/csqp.cpp:256-269, or the still-active
 * problems of a batched solve()).  idx == NULL or n < 0 -> all problems. */
int mimgpu_set_active(mimgpu_ctx* ctx, const int32_t* idx, int32_t n,
                      void* cuda_stream);

/* Write access to solver members that the reference exposes read-write
 * (bindings/csqp.cpp:85-98: y, z, rho_vec, dx, du ...).  Host pointer. */
int mimgpu_set_field(mimgpu_ctx* ctx, int field, const double* src,
                     void* cuda_stream);

/* ---- the hot path ---------------------------------------------------------- */
/* SolverSQP::computeDirection (src/sqp.cpp:224-242) without calcDiff:
 * gap_norm_/merit_, SolverDDP::backwardPass + computeGains (src/ddp.cpp:265-348,
 * 401-429), SolverSQP::forwardPass (src/sqp.cpp:270-305) and
 * SolverSQP::checkKKTConditions (src/sqp.cpp:244-268), one fused launch. */
int mimgpu_sqp_direction(mimgpu_ctx* ctx, const mimgpu_sqp_params* p,
                         void* cuda_stream);

/* The same computeDirection for HOST-resident derivatives, end to end (the call a solver makes
 * right after problem_->calcDiff when the data lives in the ActionData objects): upload, solve and
 * download of dx_ / du_ / lag_mul_ [B][..] and KKT_ [B] (any of them may be NULL) in `n_chunks`
 * slices of the batch on two internal streams, so that the PCIe copies of both directions overlap
 * each other and the kernels.  Synchronous: returns when the host arrays are filled.  Pin the
 * host memory.  Results are those of mimgpu_upload + mimgpu_sqp_direction + mimgpu_download. */
int mimgpu_sqp_direction_host(mimgpu_ctx* ctx, const mimgpu_derivs* host_derivs,
                              const mimgpu_sqp_params* p, double* dx, double* du,
                              double* lag_mul, double* kkt, int32_t n_chunks);

/* The same for host-resident KNOT RECORDS ([B][problem_doubles], cost [B] or NULL): no re-layout
 * kernel at all, one H2D copy per slice. */
int mimgpu_sqp_direction_host_records(mimgpu_ctx* ctx, const double* records, const double* cost,
                                      const mimgpu_sqp_params* p, double* dx, double* du,
                                      double* lag_mul, double* kkt, int32_t n_chunks);

/* SolverFDDP::computeDirection (src/fddp.cpp:189-204) without calcDiff +
 * updateExpectedImprovement (src/fddp.cpp:256-279): backward pass, gap_norm_,
 * dg_, dq_.  Vxx, Vx, K, k, Qu, Quuk stay on the device for download. */
int mimgpu_fddp_direction(mimgpu_ctx* ctx, const mimgpu_sqp_params* p,
                          void* cuda_stream);

/* SolverCSQP::calc norms (src/csqp.cpp:417-451, recalc = false):
 * gap_norm_ = sum_{t=1..T} |fs_t|_1 and constraint_norm_. */
int mimgpu_csqp_calc_norms(mimgpu_ctx* ctx, void* cuda_stream);

/* SolverCSQP::reset_rho_vec (src/csqp.cpp:534-537): rho_sparse_ = base and
 * apply_rho_update (src/csqp.cpp:539-582).  solve() calls it at iter_ == 0
 * (src/csqp.cpp:231-234,248-250). */
int mimgpu_csqp_reset_rho(mimgpu_ctx* ctx, const mimgpu_csqp_params* p,
                          void* cuda_stream);

/* SolverCSQP::computeDirection (src/csqp.cpp:454-517) with everything it
 * drives: reset_params (:170-193), backwardPass_without_constraints /
 * forwardPass_without_constraints (:804-898, :649-669), backwardPass (:671-802),
 * backwardPass_without_rho_update (:1025-1108), forwardPass (:630-647),
 * update_lagrangian_parameters (:1196-1340), update_rho_vec /
 * apply_rho_update (:519-582); followed, when with_kkt != 0, by
 * SolverCSQP::checkKKTConditions (:584-628).  The whole ADMM loop runs on the
 * device; qp_iters[b] follows the reference's for-loop semantics
 * (max_qp_iters+1 on exhaustion).  The first call allocates the kernel's
 * per-ctx scratch (cudaMalloc, freed by mimgpu_destroy): knot records
 * [B][T][nx^2 + 2 nx nu + ng (nx + nu) + 3 nx + nu + 4 ng] for the small shapes,
 * [B][T][nx^2 + 2 nx] for nx >= 32. */
int mimgpu_csqp_direction(mimgpu_ctx* ctx, const mimgpu_csqp_params* p,
                          int with_kkt, void* cuda_stream);

/* The same computeDirection for HOST-resident derivatives, end to end: upload, (optionally) calc norms and
 * rho reset, the ADMM loop and the download of dx~ / du~ / lag_mul / y / z [B][..], KKT [B] and qp_iters [B]
 * (any of them may be NULL) in `n_chunks` slices of the batch on two internal streams, so that the PCIe
 * copies overlap the kernels.  Synchronous; pin the host memory. */
int mimgpu_csqp_direction_host(mimgpu_ctx* ctx, const mimgpu_derivs* host_derivs, const mimgpu_csqp_params* p,
                               int with_kkt, int calc_norms, int reset_rho, double* dx_tilde, double* du_tilde,
                               double* lag_mul, double* y, double* z, double* kkt, int32_t* qp_iters,
                               int32_t n_chunks);

/* ---- solve() on the device (analytic models) -------------------------------- */
/* For problems whose running / terminal models can be evaluated on the GPU the whole
 * SolverSQP::solve (src/sqp.cpp:75-222) runs there: model calc / calcDiff, computeDirection,
 * checkKKTConditions, the merit / filter line search over alphas_ = 2^-n (tryStep,
 * src/sqp.cpp:307-374), setCandidate and the regularisation schedule (src/ddp.cpp:431-445), with
 * one iteration counter, filter memory and regularisation per problem.  No derivative ever
 * crosses PCIe; the host reads two counters per SQP iteration.
 *
 * Model family (LQ plus separable polynomial terms; c3 = w4 = 0 is the ActionModelLQR of the
 * reference's fixtures, tests/python/problems.py, tests/test_solvers.cpp:255-300):
 *     x+  = A_t x + B_t u + f_t + c3 x^3            l_T = 1/2 x^T Q_T x + q_T.x + w4/4 sum x_i^4
 *     l_t = 1/2 [x;u]^T [[Q,S],[S^T,R]] [x;u] + q.x + r.u + w4/4 sum x_i^4
 * Arrays in the block layouts of the derivative struct: A like Fx, Bm like Fu, Q S R like Lxx Lxu Luu. */
typedef struct {
  const double *A, *Bm, *f;        /* [B][T][nx*nx], [B][T][nx*nu], [B][T][nx] */
  const double *Q, *S, *R, *q, *r; /* [B][T][nx*nx], [B][T][nx*nu], [B][T][nu*nu], [B][T][nx], [B][T][nu] */
  const double *QT, *qT;           /* [B][nx*nx], [B][nx] */
  const double* x0;                /* [B][nx] problem_->get_x0() */
  double c3, w4;
} mimgpu_poly_model;

typedef struct {                   /* members of SolverSQP / SolverDDP that steer solve() */
  int32_t maxiter;                 /* 100 */
  int32_t use_filter_line_search;  /* 1   (include/mim_solvers/sqp.hpp:164-181) */
  int32_t filter_size;             /* 1   (<= 16) */
  int32_t extra_iteration_for_last_kkt; /* 0 */
  double reg_init;                 /* NaN -> reg_min (src/sqp.cpp:91-97) */
  double reg_min, reg_max;         /* 1e-9, 1e9 (src/ddp.cpp:27-28) */
  double reg_incfactor, reg_decfactor; /* 10, 10 */
  double th_stepdec, th_stepinc;   /* 0.5, 0.01 */
  double termination_tol;          /* 1e-6 */
  double mu_dynamic;               /* 1.0 */
} mimgpu_sqp_solve_params;

/* Gives the ctx its model (ng = ngT = 0).  MIMGPU_MEM_HOST: copied; MIMGPU_MEM_DEVICE: attached. */
int mimgpu_model_attach(mimgpu_ctx* ctx, const mimgpu_poly_model* model, int memkind, void* cuda_stream);
/* SolverSQP::solve(init_xs, init_us, maxiter, false, reg_init) for all B problems.  xs0 [B][T+1][nx],
 * us0 [B][T][nu] are HOST arrays (NULL = zeros, as crocoddyl's setCandidate).  Synchronous.  Results:
 * MIMGPU_F_XS / MIMGPU_F_US, the members of the last computeDirection, MIMGPU_S_KKT / _COST / _MERIT /
 * _GAP_NORM / _STEPLENGTH / _PREG, mimgpu_solve_info.  *n_sqp_iters (may be NULL) = iterations of the
 * slowest problem. */
int mimgpu_sqp_solve(mimgpu_ctx* ctx, const mimgpu_sqp_solve_params* p, const double* xs0, const double* us0,
                     int32_t* n_sqp_iters, void* cuda_stream);
/* iter_ [B] and the outcome [B] (1 = solve() returned true, 2 = false by regularisation limit, 0 = false by
 * maxiter) of the last mimgpu_sqp_solve; either may be NULL. */
int mimgpu_solve_info(mimgpu_ctx* ctx, int32_t* iters, int32_t* outcome, void* cuda_stream);

/* ---- no-allocation entry point ------------------------------------------------ */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * SolverSQP::computeDirection without a ctx: the same kernel as mimgpu_sqp_direction on knot records
 * (k_sqp_direction_small_gj), on the CURRENT device, enqueued on `cuda_stream`.  It calls no cudaMalloc /
 * cudaFree / allocating routine: the kernel's scratch (the spill records its backward sweep hands to its
 * forward sweep) is a static __device__ array of MIMGPU_NOALLOC_SCRATCH_BYTES that every GPU loading the
 * library holds until exit.  Shapes: those mimgpu_get_record_layout accepts (ng = ngT = 0); sizes:
 * mimgpu_noalloc_scratch_needed(shape) <= MIMGPU_NOALLOC_SCRATCH_BYTES (B <= 177 for UR5 nx=12 nu=6 T=100),
 * otherwise MIMGPU_ERR_INVALID with the numbers in mimgpu_last_error.  All pointers are DEVICE memory:
 *   records [B][problem_doubles] (16-byte aligned), cost [B] or NULL              -- inputs
 *   dx [B][T+1][nx], du [B][T][nu], lag_mul [B][T+1][nx], status [B],
 *   scalars [MIMGPU_S_COUNT][B] (rows KKT, GAP_NORM, MERIT, XGRAD_NORM, UGRAD_NORM are written) -- outputs
 * Every call shares the one scratch array (it keeps the last call's contents; nothing in it needs clearing):
 * calls that overlap on the GPU -- two streams, two host threads -- race.  Asynchronous. */
#define MIMGPU_NOALLOC_SCRATCH_BYTES (32 * 1024 * 1024)
int mimgpu_sqp_direction_records_noalloc(const mimgpu_shape* shape, const double* records, const double* cost,
                                         const mimgpu_sqp_params* p, double* dx, double* du, double* lag_mul,
                                         double* scalars, int32_t* status, void* cuda_stream);
/* Scratch bytes a call with this shape needs (-1: shape not served by the record kernel). */
int64_t mimgpu_noalloc_scratch_needed(const mimgpu_shape* shape);

/* ---- data out -------------------------------------------------------------- */
/* Copies one solver member of all B problems to host memory `dst`
 * (async on `stream`; synchronise the stream before reading). */
int mimgpu_download(mimgpu_ctx* ctx, int field, double* dst, void* cuda_stream);
/* Device pointer of a field (for callers that keep results on the GPU).  Members the fast SQP
 * kernel materialises lazily (Vxx_ Vx_ K_ k_ from its spill records; Qu_ Quuk_ Quu_llt_ by a
 * recomputation) are produced by work enqueued on `cuda_stream`: order later reads after it. */
int mimgpu_field_ptr(mimgpu_ctx* ctx, int field, double** dev_ptr,
                     int64_t* n_doubles, void* cuda_stream);
/* Per-problem scalars: dst [B] doubles. */
int mimgpu_scalar(mimgpu_ctx* ctx, int which, double* dst, void* cuda_stream);
/* qp_iters[B] (CSQP; 0 for SQP) and status[B]; either may be NULL. */
int mimgpu_iters_status(mimgpu_ctx* ctx, int32_t* qp_iters, int32_t* status,
                        void* cuda_stream);
/* converged[B] (CSQP): 1 when the ADMM loop broke on the tolerance test
 * (src/csqp.cpp:508-511), 0 when max_qp_iters was exhausted. */
int mimgpu_converged(mimgpu_ctx* ctx, int32_t* converged, void* cuda_stream);
/* Batch-wide reduction on the device, result copied to host (synchronises
 * `stream`).  With several GPUs the caller all-reduces these five numbers. */
int mimgpu_summary_reduce(mimgpu_ctx* ctx, double termination_tol,
                          mimgpu_summary* out, void* cuda_stream);
/* cudaStreamSynchronize(stream) on the ctx's device. */
int mimgpu_stream_sync(mimgpu_ctx* ctx, void* cuda_stream);
/* Roofline denominators that MEASURED_PEAKS.json does not carry: FP64 FMA (DFMA,
 * CUDA cores) and FP64 tensor (DMMA m8n8k4) throughput in TFLOP/s, and streaming-read
 * HBM bandwidth in GB/s, measured on `device` with register-resident loops. */
int mimgpu_measure_peaks(int device, double* dfma_tflops, double* dmma_tflops, double* hbm_read_gbs);
/* Number of kernel launches this ctx has enqueued since creation. */
int64_t mimgpu_launch_count(const mimgpu_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* MIMGPU_H_ */
