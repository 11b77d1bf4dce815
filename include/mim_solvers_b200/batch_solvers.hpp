/*
 * batch_solvers.hpp -- thin C++ host layer above the C ABI of include/mimgpu.h.
 *
 * BatchSolverSQP / BatchSolverCSQP / BatchSolverFDDP keep the C++ API of the reference
 * solver classes (include/mim_solvers/sqp.hpp:39-191, csqp.hpp:41-357, fddp.hpp:42-166,
 * ddp.hpp:44-420): solve(init_xs, init_us, maxiter, is_feasible, reginit), computeDirection,
 * tryStep, checkKKTConditions, setCandidate and the get_/set_ accessors -- for a batch of B
 * independent shooting problems of one shape.  `computeDirection` is ONE call into
 * libmimgpu.so (CUDA); everything here is the host logic of `solve()` restated from
 * src/sqp.cpp:75-222,307-374, src/csqp.cpp:197-452,1342-1435, src/fddp.cpp:57-187,229-383 and
 * src/ddp.cpp:431-445, with one "still running" flag per problem so that every problem
 * reports the iteration count its own reference solver would.
 *
 * What differs by construction: every per-solver scalar is a vector of B values and the
 * model evaluations (crocoddyl's ShootingProblem::calc / calcDiff) come through the small
 * BatchShootingProblem interface below (a crocoddyl adapter implements it by looping over
 * B ShootingProblems: INTEGRATION.md).  Errors follow the reference: invalid arguments throw
 * std::invalid_argument to the caller (src/sqp.cpp:308-311); numerical failures of a problem
 * ("backward_error", "step_error") never throw, they drive that problem's regularisation
 * schedule / step rejection exactly as the catch blocks of solve() do.  There is no CPU
 * fallback: Engine's constructor throws when libmimgpu.so finds no CUDA device.
 *
 * Header-only, C++17, depends on mimgpu.h and the standard library only.
 */
#ifndef MIM_SOLVERS_B200_BATCH_SOLVERS_HPP_
#define MIM_SOLVERS_B200_BATCH_SOLVERS_HPP_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../mimgpu.h"

namespace mim_solvers_b200 {

using Vec = std::vector<double>;
using Mask = std::vector<char>;

/* crocoddyl::raiseIfNaN: NaN, +-inf or >= 1e30 */
inline bool raiseIfNaN(double v) { return !std::isfinite(v) || v >= 1e30; }
inline bool any(const Mask& m) { return std::find(m.begin(), m.end(), (char)1) != m.end(); }

/* ------------------------------------------------------------------ derivatives (host, SoA) */
/* Pointers to ONE problem's blocks inside the batch arrays of the C ABI (column-major blocks,
 * i.e. the memory image of the Eigen members of crocoddyl::ActionDataAbstract). */
struct KnotDerivs {
  double *Fx, *Fu, *Lxx, *Lxu, *Luu, *Lx, *Lu, *LxxT, *LxT;
  double *Gx, *Gu, *g, *g_lb, *g_ub, *GxT, *gT, *g_lbT, *g_ubT;
};

/* Owning host copy of mimgpu_derivs for one batch (what calcDiff leaves behind). */
struct HostDerivs {
  mimgpu_shape sh{};
  Vec Fx, Fu, Lxx, Lxu, Luu, Lx, Lu, fs, LxxT, LxT, Gx, Gu, g, g_lb, g_ub, GxT, gT, g_lbT, g_ubT, cost;
  HostDerivs() = default;
  explicit HostDerivs(const mimgpu_shape& s) : sh(s) {
    const size_t B = s.B, T = s.T, nx = s.nx, nu = s.nu, ng = s.ng, ngT = s.ngT;
    Fx.assign(B * T * nx * nx, 0.0); Fu.assign(B * T * nx * nu, 0.0);
    Lxx.assign(B * T * nx * nx, 0.0); Lxu.assign(B * T * nx * nu, 0.0); Luu.assign(B * T * nu * nu, 0.0);
    Lx.assign(B * T * nx, 0.0); Lu.assign(B * T * nu, 0.0); fs.assign(B * (T + 1) * nx, 0.0);
    LxxT.assign(B * nx * nx, 0.0); LxT.assign(B * nx, 0.0);
    Gx.assign(B * T * ng * nx, 0.0); Gu.assign(B * T * ng * nu, 0.0);
    g.assign(B * T * ng, 0.0); g_lb.assign(B * T * ng, 0.0); g_ub.assign(B * T * ng, 0.0);
    GxT.assign(B * ngT * nx, 0.0); gT.assign(B * ngT, 0.0); g_lbT.assign(B * ngT, 0.0); g_ubT.assign(B * ngT, 0.0);
    cost.assign(B, 0.0);
  }
  KnotDerivs problem(int b) {
    const size_t T = sh.T, nx = sh.nx, nu = sh.nu, ng = sh.ng, ngT = sh.ngT, s = (size_t)b;
    auto at = [](Vec& v, size_t off) { return v.empty() ? nullptr : v.data() + off; };
    KnotDerivs k;
    k.Fx = at(Fx, s * T * nx * nx); k.Fu = at(Fu, s * T * nx * nu);
    k.Lxx = at(Lxx, s * T * nx * nx); k.Lxu = at(Lxu, s * T * nx * nu); k.Luu = at(Luu, s * T * nu * nu);
    k.Lx = at(Lx, s * T * nx); k.Lu = at(Lu, s * T * nu);
    k.LxxT = at(LxxT, s * nx * nx); k.LxT = at(LxT, s * nx);
    k.Gx = at(Gx, s * T * ng * nx); k.Gu = at(Gu, s * T * ng * nu);
    k.g = at(g, s * T * ng); k.g_lb = at(g_lb, s * T * ng); k.g_ub = at(g_ub, s * T * ng);
    k.GxT = at(GxT, s * ngT * nx); k.gT = at(gT, s * ngT); k.g_lbT = at(g_lbT, s * ngT); k.g_ubT = at(g_ubT, s * ngT);
    return k;
  }
  mimgpu_derivs c_view() const {
    auto p = [](const Vec& v) { return v.empty() ? nullptr : v.data(); };
    mimgpu_derivs d;
    d.Fx = p(Fx); d.Fu = p(Fu); d.Lxx = p(Lxx); d.Lxu = p(Lxu); d.Luu = p(Luu); d.Lx = p(Lx); d.Lu = p(Lu);
    d.fs = p(fs); d.LxxT = p(LxxT); d.LxT = p(LxT); d.Gx = p(Gx); d.Gu = p(Gu); d.g = p(g);
    d.g_lb = p(g_lb); d.g_ub = p(g_ub); d.GxT = p(GxT); d.gT = p(gT); d.g_lbT = p(g_lbT); d.g_ubT = p(g_ubT);
    d.cost = p(cost);
    return d;
  }
};

/* ------------------------------------------------------------------ the problem interface */
/* What the solvers need from B crocoddyl::ShootingProblem objects of one shape.  One call =
 * one problem, as in crocoddyl; the solvers loop (the batch is the outer dimension everywhere).
 * xs: [T+1][nx], us: [T][nu] of that problem; nx is the tangent dimension (ndx). */
class BatchShootingProblem {
 public:
  virtual ~BatchShootingProblem() = default;
  virtual mimgpu_shape shape() const = 0;
  virtual const double* x0(int b) const = 0;
  /* ShootingProblem::calc: returns the total cost; xnext [T][nx], g [T][ng], gT [ngT] */
  virtual double calc(int b, const double* xs, const double* us, double* xnext, double* g, double* gT) = 0;
  /* ShootingProblem::calcDiff: fills every block of `d` (fs and cost are the solver's) */
  virtual void calcDiff(int b, const double* xs, const double* us, const KnotDerivs& d) = 0;
  /* one running / terminal action model (FDDP's nonlinear roll-out, src/fddp.cpp:281-383) */
  virtual double calcKnot(int b, int t, const double* x, const double* u, double* xnext) = 0;
  virtual double calcTerminal(int b, const double* x) = 0;
  /* StateAbstract::integrate / diff (diff(x0, x1) = x1 (-) x0); vector space by default */
  virtual void integrate(const double* x, const double* dx, double* out) const {
    const int n = shape().nx;
    for (int i = 0; i < n; ++i) out[i] = x[i] + dx[i];
  }
  virtual void diff(const double* xa, const double* xb, double* out) const {
    const int n = shape().nx;
    for (int i = 0; i < n; ++i) out[i] = xb[i] - xa[i];
  }
};

/* B discrete-time LQ(-constrained) problems -- the batched analogue of the ActionModelLQR
 * fixtures of the reference tests (tests/python/problems.py, tests/test_solvers.cpp:255-300):
 *   x+ = A_t x + B_t u + f_t,   l_t = 1/2 [x;u]^T [[Q,S],[S^T,R]] [x;u] + q.x + r.u,
 *   l_T = 1/2 x^T Q_T x + q_T.x,   g_lb <= C_t x + D_t u + e_t <= g_ub,  g_lbT <= C_T x + e_T <= g_ubT.
 * All matrices are given in mathematical ROW-major order with leading [B][T]. */
class LQBatchProblem : public BatchShootingProblem {
 public:
  mimgpu_shape sh{};
  Vec X0, A, Bm, f, Q, S, R, q, r, QT, qT, C, D, e, g_lb, g_ub, CT, eT, g_lbT, g_ubT;

  mimgpu_shape shape() const override { return sh; }
  const double* x0(int b) const override { return X0.data() + (size_t)b * sh.nx; }

  double calcKnot(int b, int t, const double* x, const double* u, double* xnext) override {
    const int nx = sh.nx, nu = sh.nu;
    const size_t k = (size_t)b * sh.T + t;
    const double *a = A.data() + k * nx * nx, *bm = Bm.data() + k * nx * nu, *ff = f.data() + k * nx;
    const double *qq = Q.data() + k * nx * nx, *ss = S.data() + k * nx * nu, *rr = R.data() + k * nu * nu;
    for (int i = 0; i < nx; ++i) {
      double v = 0.0, w = 0.0;
      for (int j = 0; j < nx; ++j) v += a[i * nx + j] * x[j];
      for (int j = 0; j < nu; ++j) w += bm[i * nu + j] * u[j];
      xnext[i] = v + w + ff[i];
    }
    double cxx = 0.0, cxu = 0.0, cuu = 0.0, cq = 0.0, cr = 0.0;
    for (int i = 0; i < nx; ++i) {
      double v = 0.0, w = 0.0;
      for (int j = 0; j < nx; ++j) v += qq[i * nx + j] * x[j];
      for (int j = 0; j < nu; ++j) w += ss[i * nu + j] * u[j];
      cxx += x[i] * v; cxu += x[i] * w; cq += q[k * nx + i] * x[i];
    }
    for (int i = 0; i < nu; ++i) {
      double v = 0.0;
      for (int j = 0; j < nu; ++j) v += rr[i * nu + j] * u[j];
      cuu += u[i] * v; cr += r[k * nu + i] * u[i];
    }
    return 0.5 * cxx + cxu + 0.5 * cuu + cq + cr;
  }
  double calcTerminal(int b, const double* x) override {
    const int nx = sh.nx;
    const double* qq = QT.data() + (size_t)b * nx * nx;
    double c = 0.0, l = 0.0;
    for (int i = 0; i < nx; ++i) {
      double v = 0.0;
      for (int j = 0; j < nx; ++j) v += qq[i * nx + j] * x[j];
      c += x[i] * v; l += qT[(size_t)b * nx + i] * x[i];
    }
    return 0.5 * c + l;
  }
  double calc(int b, const double* xs, const double* us, double* xnext, double* g, double* gT) override {
    const int T = sh.T, nx = sh.nx, nu = sh.nu, ng = sh.ng, ngT = sh.ngT;
    double cost = 0.0;
    for (int t = 0; t < T; ++t) {
      cost += calcKnot(b, t, xs + (size_t)t * nx, us + (size_t)t * nu, xnext + (size_t)t * nx);
      if (ng && g) {
        const size_t k = (size_t)b * T + t;
        for (int i = 0; i < ng; ++i) {
          double v = 0.0, w = 0.0;
          for (int j = 0; j < nx; ++j) v += C[(k * ng + i) * nx + j] * xs[(size_t)t * nx + j];
          for (int j = 0; j < nu; ++j) w += D[(k * ng + i) * nu + j] * us[(size_t)t * nu + j];
          g[(size_t)t * ng + i] = v + w + e[k * ng + i];
        }
      }
    }
    cost += calcTerminal(b, xs + (size_t)T * nx);
    if (ngT && gT)
      for (int i = 0; i < ngT; ++i) {
        double v = 0.0;
        for (int j = 0; j < nx; ++j) v += CT[((size_t)b * ngT + i) * nx + j] * xs[(size_t)T * nx + j];
        gT[i] = v + eT[(size_t)b * ngT + i];
      }
    return cost;
  }
  void calcDiff(int b, const double* xs, const double* us, const KnotDerivs& d) override {
    const int T = sh.T, nx = sh.nx, nu = sh.nu, ng = sh.ng, ngT = sh.ngT;
    for (int t = 0; t < T; ++t) {
      const size_t k = (size_t)b * T + t;
      const double *x = xs + (size_t)t * nx, *u = us + (size_t)t * nu;
      const double *a = A.data() + k * nx * nx, *bm = Bm.data() + k * nx * nu;
      const double *qq = Q.data() + k * nx * nx, *ss = S.data() + k * nx * nu, *rr = R.data() + k * nu * nu;
      double *Fx = d.Fx + (size_t)t * nx * nx, *Fu = d.Fu + (size_t)t * nx * nu, *Lxx = d.Lxx + (size_t)t * nx * nx,
             *Lxu = d.Lxu + (size_t)t * nx * nu, *Luu = d.Luu + (size_t)t * nu * nu;
      for (int i = 0; i < nx; ++i) {
        for (int j = 0; j < nx; ++j) { Fx[i + nx * j] = a[i * nx + j]; Lxx[i + nx * j] = qq[i * nx + j]; }
        for (int j = 0; j < nu; ++j) { Fu[i + nx * j] = bm[i * nu + j]; Lxu[i + nx * j] = ss[i * nu + j]; }
        double v = 0.0, w = 0.0;
        for (int j = 0; j < nx; ++j) v += qq[i * nx + j] * x[j];
        for (int j = 0; j < nu; ++j) w += ss[i * nu + j] * u[j];
        d.Lx[(size_t)t * nx + i] = v + w + q[k * nx + i];
      }
      for (int i = 0; i < nu; ++i) {
        for (int j = 0; j < nu; ++j) Luu[i + nu * j] = rr[i * nu + j];
        double v = 0.0, w = 0.0;
        for (int j = 0; j < nx; ++j) v += ss[j * nu + i] * x[j];
        for (int j = 0; j < nu; ++j) w += rr[i * nu + j] * u[j];
        d.Lu[(size_t)t * nu + i] = v + w + r[k * nu + i];
      }
      if (ng) {
        for (int i = 0; i < ng; ++i) {
          double v = 0.0, w = 0.0;
          for (int j = 0; j < nx; ++j) { d.Gx[(size_t)t * ng * nx + i + ng * j] = C[(k * ng + i) * nx + j]; v += C[(k * ng + i) * nx + j] * x[j]; }
          for (int j = 0; j < nu; ++j) { d.Gu[(size_t)t * ng * nu + i + ng * j] = D[(k * ng + i) * nu + j]; w += D[(k * ng + i) * nu + j] * u[j]; }
          d.g[(size_t)t * ng + i] = v + w + e[k * ng + i];
          d.g_lb[(size_t)t * ng + i] = g_lb[k * ng + i];
          d.g_ub[(size_t)t * ng + i] = g_ub[k * ng + i];
        }
      }
    }
    const double* xT = xs + (size_t)T * nx;
    const double* qq = QT.data() + (size_t)b * nx * nx;
    for (int i = 0; i < nx; ++i) {
      double v = 0.0;
      for (int j = 0; j < nx; ++j) { d.LxxT[i + nx * j] = qq[i * nx + j]; v += qq[i * nx + j] * xT[j]; }
      d.LxT[i] = v + qT[(size_t)b * nx + i];
    }
    for (int i = 0; i < ngT; ++i) {
      double v = 0.0;
      for (int j = 0; j < nx; ++j) { d.GxT[i + ngT * j] = CT[((size_t)b * ngT + i) * nx + j]; v += CT[((size_t)b * ngT + i) * nx + j] * xT[j]; }
      d.gT[i] = v + eT[(size_t)b * ngT + i];
      d.g_lbT[i] = g_lbT[(size_t)b * ngT + i];
      d.g_ubT[i] = g_ubT[(size_t)b * ngT + i];
    }
  }
};

/* LQ plus separable polynomial terms -- the analytic model family the DEVICE-side solve() evaluates
 * (include/mimgpu.h, mimgpu_poly_model):  x+ = A x + B u + f + c3 x^3,  l_t += w4/4 sum x_i^4,
 * l_T += w4/4 sum x_i^4.  c3 = w4 = 0 is LQBatchProblem.  Unconstrained. */
class PolyBatchProblem : public LQBatchProblem {
 public:
  double c3 = 0.0, w4 = 0.0;

  double calcKnot(int b, int t, const double* x, const double* u, double* xnext) override {
    double c = LQBatchProblem::calcKnot(b, t, x, u, xnext);
    for (int i = 0; i < sh.nx; ++i) { const double x2 = x[i] * x[i]; xnext[i] += c3 * x2 * x[i]; c += 0.25 * w4 * x2 * x2; }
    return c;
  }
  double calcTerminal(int b, const double* x) override {
    double c = LQBatchProblem::calcTerminal(b, x);
    for (int i = 0; i < sh.nx; ++i) { const double x2 = x[i] * x[i]; c += 0.25 * w4 * x2 * x2; }
    return c;
  }
  double calc(int b, const double* xs, const double* us, double* xnext, double* g, double* gT) override {
    const int T = sh.T, nx = sh.nx, nu = sh.nu;
    double c = 0.0;
    for (int t = 0; t < T; ++t) c += calcKnot(b, t, xs + (size_t)t * nx, us + (size_t)t * nu, xnext + (size_t)t * nx);
    (void)g; (void)gT;
    return c + calcTerminal(b, xs + (size_t)T * nx);
  }
  void calcDiff(int b, const double* xs, const double* us, const KnotDerivs& d) override {
    LQBatchProblem::calcDiff(b, xs, us, d);
    const int T = sh.T, nx = sh.nx;
    for (int t = 0; t <= T; ++t) {
      const double* x = xs + (size_t)t * nx;
      for (int i = 0; i < nx; ++i) {
        const double x2 = x[i] * x[i];
        if (t < T) {
          d.Fx[(size_t)t * nx * nx + i + nx * i] += 3.0 * c3 * x2;
          d.Lxx[(size_t)t * nx * nx + i + nx * i] += 3.0 * w4 * x2;
          d.Lx[(size_t)t * nx + i] += w4 * x2 * x[i];
        } else {
          d.LxxT[i + nx * i] += 3.0 * w4 * x2;
          d.LxT[i] += w4 * x2 * x[i];
        }
      }
    }
  }
  /* the model in the column-major block layouts of the C ABI */
  struct AbiArrays { Vec A, Bm, Q, S, R, QT; };
  AbiArrays abi() const {
    const size_t B = sh.B, T = sh.T, nx = sh.nx, nu = sh.nu;
    auto tr = [](const Vec& src, size_t nblk, size_t r, size_t c) { /* row-major r x c -> col-major */
      Vec o(src.size());
      for (size_t k = 0; k < nblk; ++k)
        for (size_t i = 0; i < r; ++i)
          for (size_t j = 0; j < c; ++j) o[k * r * c + i + r * j] = src[k * r * c + i * c + j];
      return o;
    };
    return AbiArrays{tr(A, B * T, nx, nx), tr(Bm, B * T, nx, nu), tr(Q, B * T, nx, nx), tr(S, B * T, nx, nu),
                     tr(R, B * T, nu, nu), tr(QT, B, nx, nx)};
  }
};

/* ------------------------------------------------------------------ RAII handle of one mimgpu_ctx */
class Engine {
 public:
  Engine(const mimgpu_shape& shape, int device = 0, void* cuda_stream = nullptr) : sh_(shape), stream_(cuda_stream) {
    check(mimgpu_create(&ctx_, device, &sh_), "mimgpu_create");
  }
  ~Engine() { if (ctx_) mimgpu_destroy(ctx_); }
  Engine(const Engine&) = delete;
  Engine& operator=(const Engine&) = delete;

  const mimgpu_shape& shape() const { return sh_; }
  mimgpu_ctx* ctx() const { return ctx_; }
  void model_attach(const mimgpu_poly_model& m) { check(mimgpu_model_attach(ctx_, &m, MIMGPU_MEM_HOST, stream_), "mimgpu_model_attach"); sync(); }
  int sqp_solve(const mimgpu_sqp_solve_params& p, const double* xs0, const double* us0) {
    int32_t n = 0;
    check(mimgpu_sqp_solve(ctx_, &p, xs0, us0, &n, stream_), "mimgpu_sqp_solve");
    return n;
  }
  void solve_info(std::vector<int32_t>& iters, std::vector<int32_t>& outcome) {
    iters.resize(sh_.B); outcome.resize(sh_.B);
    check(mimgpu_solve_info(ctx_, iters.data(), outcome.data(), stream_), "mimgpu_solve_info");
  }
  void sync() { check(mimgpu_stream_sync(ctx_, stream_), "mimgpu_stream_sync"); }
  void upload(const HostDerivs& d) {
    mimgpu_derivs v = d.c_view();
    check(mimgpu_upload(ctx_, &v, MIMGPU_MEM_HOST, stream_), "mimgpu_upload");
  }
  void set_reg(const Vec& preg, const Vec& dreg) {
    check(mimgpu_set_reg(ctx_, preg.data(), dreg.data(), stream_), "mimgpu_set_reg");
    sync();
  }
  /* back to the scalar preg / dreg of the params struct (remove_reg_: a ctx that earlier received per-problem
   * arrays must not keep using them) */
  void clear_reg() {
    check(mimgpu_set_reg(ctx_, nullptr, nullptr, stream_), "mimgpu_set_reg");
    sync();
  }
  void set_active(const std::vector<int32_t>* idx) {
    if (!idx) { check(mimgpu_set_active(ctx_, nullptr, -1, stream_), "mimgpu_set_active"); return; }
    check(mimgpu_set_active(ctx_, idx->data(), (int32_t)idx->size(), stream_), "mimgpu_set_active");
    sync();
  }
  void set_field(int field, const Vec& v) { check(mimgpu_set_field(ctx_, field, v.data(), stream_), "mimgpu_set_field"); sync(); }
  void sqp_direction(const mimgpu_sqp_params& p) { check(mimgpu_sqp_direction(ctx_, &p, stream_), "mimgpu_sqp_direction"); }
  void fddp_direction(const mimgpu_sqp_params& p) { check(mimgpu_fddp_direction(ctx_, &p, stream_), "mimgpu_fddp_direction"); }
  void csqp_calc_norms() { check(mimgpu_csqp_calc_norms(ctx_, stream_), "mimgpu_csqp_calc_norms"); }
  void csqp_reset_rho(const mimgpu_csqp_params& p) { check(mimgpu_csqp_reset_rho(ctx_, &p, stream_), "mimgpu_csqp_reset_rho"); }
  void csqp_direction(const mimgpu_csqp_params& p, bool with_kkt = true) {
    check(mimgpu_csqp_direction(ctx_, &p, with_kkt ? 1 : 0, stream_), "mimgpu_csqp_direction");
  }
  size_t field_size(int field) const {
    const size_t B = sh_.B, T = sh_.T, nx = sh_.nx, nu = sh_.nu, nc = T * sh_.ng + sh_.ngT;
    switch (field) {
      case MIMGPU_F_DX: case MIMGPU_F_LAG_MUL: case MIMGPU_F_DX_TILDE: case MIMGPU_F_VX: case MIMGPU_F_XS: return B * (T + 1) * nx;
      case MIMGPU_F_DU: case MIMGPU_F_DU_TILDE: case MIMGPU_F_KFF: case MIMGPU_F_QU: case MIMGPU_F_QUUK: case MIMGPU_F_US: return B * T * nu;
      case MIMGPU_F_Y: case MIMGPU_F_Z: case MIMGPU_F_RHO_VEC: return B * nc;
      case MIMGPU_F_VXX: return B * (T + 1) * nx * nx;
      case MIMGPU_F_K: return B * T * nu * nx;
      case MIMGPU_F_QUU_LLT: return B * T * nu * nu;
      default: throw std::invalid_argument("unknown mimgpu_field");
    }
  }
  Vec field(int f) {
    Vec out(field_size(f));
    if (!out.empty()) { check(mimgpu_download(ctx_, f, out.data(), stream_), "mimgpu_download"); sync(); }
    return out;
  }
  Vec scalar(int which) {
    Vec out((size_t)sh_.B);
    check(mimgpu_scalar(ctx_, which, out.data(), stream_), "mimgpu_scalar");
    sync();
    return out;
  }
  void iters_status(std::vector<int32_t>& qp_iters, std::vector<int32_t>& status) {
    qp_iters.resize(sh_.B); status.resize(sh_.B);
    check(mimgpu_iters_status(ctx_, qp_iters.data(), status.data(), stream_), "mimgpu_iters_status");
    sync();
  }
  mimgpu_summary summary(double termination_tol) {
    mimgpu_summary s;
    check(mimgpu_summary_reduce(ctx_, termination_tol, &s, stream_), "mimgpu_summary_reduce");
    return s;
  }
  int64_t launch_count() const { return mimgpu_launch_count(ctx_); }

 private:
  static void check(int rc, const char* what) {
    if (rc != MIMGPU_OK) {
      const char* m = mimgpu_last_error();
      throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " + (m ? m : ""));
    }
  }
  mimgpu_ctx* ctx_ = nullptr;
  mimgpu_shape sh_;
  void* stream_;
};

/* boost::circular_buffer<double> of capacity filter_size per problem, zero-filled by resize()
 * (src/sqp.cpp:32-33); recent(b, count) is the count-th most recent push_back. */
class FilterMemory {
 public:
  FilterMemory(int B, int size) : size_(size), buf_((size_t)B * size, 0.0) {}
  void push(int b, double v) {
    double* p = buf_.data() + (size_t)b * size_;
    for (int i = 0; i + 1 < size_; ++i) p[i] = p[i + 1];
    p[size_ - 1] = v;
  }
  double recent(int b, int count) const { return buf_[(size_t)b * size_ + size_ - 1 - count]; }
 private:
  int size_;
  Vec buf_;
};

/* ------------------------------------------------------------------ SolverDDP members (src/ddp.cpp:18-55) */
class BatchSolverBase {
 public:
  using Callback = std::function<void(BatchSolverBase&, const char*)>;

  explicit BatchSolverBase(std::shared_ptr<BatchShootingProblem> problem, int device = 0)
      : problem_(std::move(problem)), device_(device), sh_(problem_->shape()) {
    const size_t B = sh_.B, T = sh_.T, nx = sh_.nx, nu = sh_.nu;
    for (int n = 0; n < 10; ++n) alphas_.push_back(1.0 / std::pow(2.0, (double)n));
    if (th_stepinc_ < alphas_.back()) th_stepinc_ = alphas_.back();
    xs_.assign(B * (T + 1) * nx, 0.0); us_.assign(B * T * nu, 0.0);
    xs_try_ = xs_; us_try_ = us_;
    fs_.assign(B * (T + 1) * nx, 0.0); fs_try_ = fs_;
    preg_.assign(B, reg_min_); dreg_.assign(B, reg_min_);
    iter_.assign(B, 0);
    cost_.assign(B, 0.0); cost_try_ = merit_ = merit_try_ = gap_norm_ = gap_norm_try_ = x_grad_norm_ = u_grad_norm_ = cost_;
    KKT_.assign(B, std::numeric_limits<double>::infinity());
    steplength_.assign(B, 1.0);
    derivs_ = HostDerivs(sh_);
    xnext_.assign(B * T * nx, 0.0);
    g_.assign(B * T * sh_.ng, 0.0); gT_.assign(B * (size_t)sh_.ngT, 0.0);
  }
  virtual ~BatchSolverBase() = default;

  /* ---- crocoddyl::SolverAbstract plumbing */
  void setCandidate(const Vec& xs = Vec(), const Vec& us = Vec(), bool is_feasible = false) {
    const size_t B = sh_.B, T = sh_.T, nx = sh_.nx;
    if (xs.empty()) {
      for (size_t b = 0; b < B; ++b)
        for (size_t t = 0; t <= T; ++t) std::memcpy(&xs_[(b * (T + 1) + t) * nx], problem_->x0((int)b), nx * sizeof(double));
    } else {
      if (xs.size() != xs_.size()) throw std::invalid_argument("Invalid argument: xs has wrong dimension");
      xs_ = xs;
    }
    if (us.empty()) {
      std::fill(us_.begin(), us_.end(), 0.0);
    } else {
      if (us.size() != us_.size()) throw std::invalid_argument("Invalid argument: us has wrong dimension");
      us_ = us;
    }
    is_feasible_ = is_feasible;
  }
  void setCallbacks(std::vector<Callback> cbs) { callbacks_ = std::move(cbs); }
  const std::vector<Callback>& getCallbacks() const { return callbacks_; }

  /* src/ddp.cpp:431-445, per problem */
  void increaseRegularization(int b) { preg_[b] = std::min(preg_[b] * reg_incfactor_, reg_max_); dreg_[b] = preg_[b]; }
  void decreaseRegularization(int b) { preg_[b] = std::max(preg_[b] / reg_decfactor_, reg_min_); dreg_[b] = preg_[b]; }

  /* ---- accessors (names of bindings/ddp.cpp:63-238, bindings/sqp.cpp:41-117) */
  const mimgpu_shape& shape() const { return sh_; }
  std::shared_ptr<BatchShootingProblem> problem() const { return problem_; }
  const Vec& get_xs() const { return xs_; }
  const Vec& get_us() const { return us_; }
  const Vec& get_xs_try() const { return xs_try_; }
  const Vec& get_us_try() const { return us_try_; }
  const Vec& get_fs() const { return fs_; }
  const Vec& get_cost() const { return cost_; }
  const Vec& get_KKT() const { return KKT_; }
  const Vec& get_merit() const { return merit_; }
  const Vec& get_gap_norm() const { return gap_norm_; }
  const Vec& get_preg() const { return preg_; }
  const Vec& get_dreg() const { return dreg_; }
  const Vec& get_steplength() const { return steplength_; }
  const std::vector<int64_t>& get_iter() const { return iter_; }
  const std::vector<int32_t>& get_status() const { return status_; }
  Vec get_dx() { return engine().field(MIMGPU_F_DX); }
  Vec get_du() { return engine().field(MIMGPU_F_DU); }
  Vec get_lag_mul() { return engine().field(MIMGPU_F_LAG_MUL); }
  Vec get_Vxx() { return engine().field(MIMGPU_F_VXX); }
  Vec get_Vx() { return engine().field(MIMGPU_F_VX); }
  Vec get_K() { return engine().field(MIMGPU_F_K); }
  Vec get_k() { return engine().field(MIMGPU_F_KFF); }
  Vec get_Qu() { return engine().field(MIMGPU_F_QU); }
  double get_termination_tolerance() const { return termination_tol_; }
  void set_termination_tolerance(double v) { termination_tol_ = v; }
  bool get_use_filter_line_search() const { return use_filter_line_search_; }
  void set_use_filter_line_search(bool v) { use_filter_line_search_ = v; }
  int get_filter_size() const { return filter_size_; }
  void set_filter_size(int v) { filter_size_ = v; }
  bool get_extra_iteration_for_last_kkt() const { return extra_iteration_for_last_kkt_; }
  void set_extra_iteration_for_last_kkt(bool v) { extra_iteration_for_last_kkt_ = v; }
  double get_mu_dynamic() const { return mu_dynamic_; }
  void set_mu_dynamic(double v) { mu_dynamic_ = v; }
  double get_reg_min() const { return reg_min_; }
  void set_reg_min(double v) { reg_min_ = v; }
  double get_reg_max() const { return reg_max_; }
  void set_reg_max(double v) { reg_max_ = v; }
  double get_reg_incfactor() const { return reg_incfactor_; }
  void set_reg_incfactor(double v) { if (v <= 1.) throw std::invalid_argument("Invalid argument: reg_incfactor value is higher than 1."); reg_incfactor_ = v; }
  double get_reg_decfactor() const { return reg_decfactor_; }
  void set_reg_decfactor(double v) { if (v <= 1.) throw std::invalid_argument("Invalid argument: reg_decfactor value is higher than 1."); reg_decfactor_ = v; }
  double get_th_stepdec() const { return th_stepdec_; }
  void set_th_stepdec(double v) { if (v < 0. || v > 1.) throw std::invalid_argument("Invalid argument: th_stepdec value should between 0 and 1."); th_stepdec_ = v; }
  double get_th_stepinc() const { return th_stepinc_; }
  void set_th_stepinc(double v) { if (v < 0. || v > 1.) throw std::invalid_argument("Invalid argument: th_stepinc value should between 0 and 1."); th_stepinc_ = v; }
  double get_th_grad() const { return th_grad_; }
  void set_th_grad(double v) { if (v < 0.) throw std::invalid_argument("Invalid argument: th_grad value has to be positive."); th_grad_ = v; }
  const Vec& get_alphas() const { return alphas_; }
  void set_alphas(const Vec& a) {
    /* src/ddp.cpp:604-628 */
    double prev = a.empty() ? 0.0 : a[0];
    if (a.empty() || prev != 1.) throw std::invalid_argument("Invalid argument: alpha[0] should be 1");
    for (size_t i = 1; i < a.size(); ++i) {
      if (a[i] <= 0.) throw std::invalid_argument("Invalid argument: alpha values has to be positive.");
      if (a[i] >= prev) throw std::invalid_argument("Invalid argument: alpha values are monotonously decreasing.");
      prev = a[i];
    }
    alphas_ = a;
  }
  Engine& engine() {
    if (!engine_) throw std::logic_error("no derivatives uploaded yet: call solve() / computeDirection(true) first");
    return *engine_;
  }

 protected:
  size_t xoff(int b, int t = 0) const { return ((size_t)b * (sh_.T + 1) + t) * sh_.nx; }
  size_t uoff(int b, int t = 0) const { return ((size_t)b * sh_.T + t) * sh_.nu; }

  /* problem_->calc + calcDiff + the gaps fs_[0] = x0 (-) xs_[0], fs_[t+1] = xnext_t (-) xs_[t+1]
   * (src/ddp.cpp:253-263, src/csqp.cpp:411-416, 426-431), then the upload. */
  void calcDiffAndUpload() {
    const int B = sh_.B, T = sh_.T, nx = sh_.nx, ng = sh_.ng, ngT = sh_.ngT;
    for (int b = 0; b < B; ++b) {
      double* xn = xnext_.data() + (size_t)b * T * nx;
      cost_[b] = problem_->calc(b, &xs_[xoff(b)], &us_[uoff(b)], xn, ng ? &g_[(size_t)b * T * ng] : nullptr,
                                ngT ? &gT_[(size_t)b * ngT] : nullptr);
      problem_->calcDiff(b, &xs_[xoff(b)], &us_[uoff(b)], derivs_.problem(b));
      problem_->diff(&xs_[xoff(b)], problem_->x0(b), &fs_[xoff(b)]);
      for (int t = 0; t < T; ++t) problem_->diff(&xs_[xoff(b, t + 1)], xn + (size_t)t * nx, &fs_[xoff(b, t + 1)]);
    }
    derivs_.fs = fs_;
    derivs_.cost = cost_;
    if (!engine_) engine_ = std::make_unique<Engine>(sh_, device_);
    engine_->upload(derivs_);
  }

  void setActive(const Mask& m) {
    std::vector<int32_t> idx;
    for (int b = 0; b < sh_.B; ++b) if (m[b]) idx.push_back(b);
    if ((int)idx.size() == sh_.B) engine_->set_active(nullptr); else engine_->set_active(&idx);
  }

  /* `while (true) { try { computeDirection } catch { increaseRegularization; if (preg_ ==
   * reg_max_) return false; } }` (src/sqp.cpp:103-116, src/csqp.cpp:256-269) per problem: the
   * problems whose status word reports a failed LLT / non-finite value are retried alone at ten
   * times the regularisation.  Returns the problems that gave up (preg hit reg_max). */
  Mask directionWithRetry(const Mask& running, bool recalc, const std::function<void()>& direction) {
    const int B = sh_.B;
    Mask gave_up(B, 0), todo = running;
    if (recalc) calcDiffAndUpload();
    std::vector<int32_t> qp;
    while (any(todo)) {
      setActive(todo);
      direction();
      engine_->iters_status(qp, status_);
      Mask failed(B, 0);
      for (int b = 0; b < B; ++b) failed[b] = todo[b] && status_[b] != 0;
      if (!any(failed)) break;
      for (int b = 0; b < B; ++b) {
        todo[b] = 0;
        if (!failed[b]) continue;
        increaseRegularization(b);
        if (preg_[b] >= reg_max_) gave_up[b] = 1; else todo[b] = 1;
      }
    }
    engine_->set_active(nullptr);
    return gave_up;
  }

  /* filter line-search test (src/sqp.cpp:148-158, src/csqp.cpp:325-341): rejected iff some
   * remembered point is <= the trial in EVERY criterion; min(filter_size, iter+1) entries. */
  bool filterAccepts(int b, std::initializer_list<std::pair<const FilterMemory*, double>> crit) const {
    for (int count = 0; count < filter_size_; ++count) {
      if (count > iter_[b]) continue;
      bool dominated = true;
      for (const auto& c : crit) dominated = dominated && (c.first->recent(b, count) <= c.second);
      if (dominated) return false;
    }
    return true;
  }

  void pull(int which, Vec& dst, const Mask& m) {
    Vec v = engine_->scalar(which);
    for (int b = 0; b < sh_.B; ++b) if (m[b]) dst[b] = v[b];
  }
  void runCallbacks(const char* name) { for (auto& cb : callbacks_) cb(*this, name); }

  std::shared_ptr<BatchShootingProblem> problem_;
  int device_;
  mimgpu_shape sh_;
  std::unique_ptr<Engine> engine_;
  HostDerivs derivs_;
  Vec xs_, us_, xs_try_, us_try_, fs_, fs_try_, xnext_, g_, gT_;
  Vec preg_, dreg_, cost_, cost_try_, merit_, merit_try_, gap_norm_, gap_norm_try_, KKT_, x_grad_norm_, u_grad_norm_, steplength_;
  std::vector<int64_t> iter_;
  std::vector<int32_t> status_;
  Vec alphas_;
  std::vector<Callback> callbacks_;
  bool is_feasible_ = false;
  double reg_incfactor_ = 10., reg_decfactor_ = 10., reg_min_ = 1e-9, reg_max_ = 1e9;
  double th_stepdec_ = 0.5, th_stepinc_ = 0.01, th_grad_ = 1e-12, th_acceptstep_ = 0.1, th_stop_ = 1e-9;
  double termination_tol_ = 1e-6, mu_dynamic_ = 1.0;
  bool use_filter_line_search_ = true, extra_iteration_for_last_kkt_ = false;
  int filter_size_ = 1;
};

/* ------------------------------------------------------------------ SolverSQP (include/mim_solvers/sqp.hpp:39-191) */
class BatchSolverSQP : public BatchSolverBase {
 public:
  using BatchSolverBase::BatchSolverBase;

  /* src/sqp.cpp:224-242 for every problem of the batch: one CUDA launch */
  void computeDirection(bool recalcDiff = true) {
    if (recalcDiff || !engine_) calcDiffAndUpload();
    launchDirection();
    pullScalars(Mask(sh_.B, 1));
  }
  /* src/sqp.cpp:244-268: computed by the same launch as the direction */
  const Vec& checkKKTConditions() const { return KKT_; }

  /* src/sqp.cpp:307-374 for the problems of `mask`; bad[b] = the step raised "step_error" */
  void tryStep(double steplength, const Mask& mask, Mask& bad) {
    if (steplength > 1. || steplength < 0.) throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    const int B = sh_.B, T = sh_.T, nx = sh_.nx, nu = sh_.nu;
    Vec step(nx);
    bad.assign(B, 0);
    for (int b = 0; b < B; ++b) {
      if (!mask[b]) continue;
      for (int t = 0; t <= T; ++t) {
        for (int i = 0; i < nx; ++i) step[i] = steplength * dx_cache_[xoff(b, t) + i];
        problem_->integrate(&xs_[xoff(b, t)], step.data(), &xs_try_[xoff(b, t)]);
      }
      for (int i = 0; i < T * nu; ++i) us_try_[uoff(b) + i] = us_[uoff(b) + i] + steplength * du_cache_[uoff(b) + i];
      double* xn = xnext_.data() + (size_t)b * T * nx;
      const double c = problem_->calc(b, &xs_try_[xoff(b)], &us_try_[uoff(b)], xn, nullptr, nullptr);
      double gap = 0.0;
      for (int t = 0; t < T; ++t) {
        problem_->diff(&xs_try_[xoff(b, t + 1)], xn + (size_t)t * nx, &fs_try_[xoff(b, t + 1)]);
        for (int i = 0; i < nx; ++i) gap += std::fabs(fs_try_[xoff(b, t + 1) + i]);
      }
      cost_try_[b] = c;
      gap_norm_try_[b] = gap;
      merit_try_[b] = c + mu_dynamic_ * gap;
      bad[b] = raiseIfNaN(c);
    }
  }

  /* src/sqp.cpp:75-222 per problem.  Returns converged[b] = KKT_ <= termination_tol_. */
  std::vector<bool> solve(const Vec& init_xs = Vec(), const Vec& init_us = Vec(), size_t maxiter = 100,
                          bool /*is_feasible*/ = false, double reginit = NAN) {
    const int B = sh_.B, nx = sh_.nx;
    setCandidate(init_xs, init_us, false);
    for (int b = 0; b < B; ++b) {
      std::memcpy(&xs_[xoff(b)], problem_->x0(b), nx * sizeof(double));
      std::memcpy(&xs_try_[xoff(b)], problem_->x0(b), nx * sizeof(double));
    }
    const double reg0 = std::isnan(reginit) ? reg_min_ : reginit;
    std::fill(preg_.begin(), preg_.end(), reg0);
    std::fill(dreg_.begin(), dreg_.end(), reg0);
    std::vector<bool> converged(B, false);
    Mask running(B, 1);
    std::fill(iter_.begin(), iter_.end(), 0);
    FilterMemory gaps(B, filter_size_), costs(B, filter_size_);
    size_t it = 0;
    while (it < maxiter && any(running)) {
      iteration(running);
      for (int b = 0; b < B; ++b)
        if (running[b] && KKT_[b] <= termination_tol_) { converged[b] = true; running[b] = 0; }
      if (!any(running)) break;
      for (int b = 0; b < B; ++b) if (running[b]) { gaps.push(b, gap_norm_[b]); costs.push(b, cost_[b]); }
      dx_cache_ = engine_->field(MIMGPU_F_DX);
      du_cache_ = engine_->field(MIMGPU_F_DU);
      Mask searching = running, bad;
      for (double alpha : alphas_) {
        if (!any(searching)) break;
        for (int b = 0; b < B; ++b) if (searching[b]) steplength_[b] = alpha;
        tryStep(alpha, searching, bad);
        for (int b = 0; b < B; ++b) {
          if (!searching[b]) continue;
          const bool ok = use_filter_line_search_
                              ? filterAccepts(b, {{&costs, cost_try_[b]}, {&gaps, gap_norm_try_[b]}})
                              : merit_[b] > merit_try_[b];
          if (ok && !bad[b]) { accept(b); searching[b] = 0; }
        }
      }
      for (int b = 0; b < B; ++b) {
        if (!running[b]) continue;
        if (steplength_[b] > th_stepdec_) decreaseRegularization(b);
        if (steplength_[b] <= th_stepinc_) {
          increaseRegularization(b);
          if (preg_[b] >= reg_max_) running[b] = 0;
        }
      }
      ++it;
      for (int b = 0; b < B; ++b) if (running[b]) iter_[b] = (int64_t)it;
    }
    if (extra_iteration_for_last_kkt_ && any(running)) {
      iteration(running);
      for (int b = 0; b < B; ++b) if (running[b] && KKT_[b] <= termination_tol_) converged[b] = true;
    }
    return converged;
  }

  const Vec& get_xgrad_norm() const { return x_grad_norm_; }
  const Vec& get_ugrad_norm() const { return u_grad_norm_; }

  /* The same solve() with the WHOLE loop on the GPU (mimgpu_sqp_solve): model calc / calcDiff,
   * computeDirection, KKT test, line search, acceptance and regularisation schedule per problem; needs an
   * analytic model the device can evaluate (PolyBatchProblem).  Throws std::invalid_argument otherwise. */
  std::vector<bool> solveOnDevice(const Vec& init_xs = Vec(), const Vec& init_us = Vec(), size_t maxiter = 100,
                                  bool /*is_feasible*/ = false, double reginit = NAN) {
    auto* pm = dynamic_cast<PolyBatchProblem*>(problem_.get());
    if (!pm) throw std::invalid_argument("Invalid argument: solveOnDevice needs an analytic model (PolyBatchProblem)");
    if (sh_.ng || sh_.ngT) throw std::invalid_argument("Invalid argument: solveOnDevice handles unconstrained problems");
    const size_t B = sh_.B;
    if (!init_xs.empty() && init_xs.size() != xs_.size()) throw std::invalid_argument("Invalid argument: xs has wrong dimension");
    if (!init_us.empty() && init_us.size() != us_.size()) throw std::invalid_argument("Invalid argument: us has wrong dimension");
    if (!engine_) engine_ = std::make_unique<Engine>(sh_, device_);
    if (!model_attached_) {
      const PolyBatchProblem::AbiArrays a = pm->abi();
      mimgpu_poly_model m;
      m.A = a.A.data(); m.Bm = a.Bm.data(); m.f = pm->f.data(); m.Q = a.Q.data(); m.S = a.S.data(); m.R = a.R.data();
      m.q = pm->q.data(); m.r = pm->r.data(); m.QT = a.QT.data(); m.qT = pm->qT.data(); m.x0 = pm->X0.data();
      m.c3 = pm->c3; m.w4 = pm->w4;
      engine_->model_attach(m);
      model_attached_ = true;
    }
    mimgpu_sqp_solve_params sp;
    sp.maxiter = (int32_t)maxiter; sp.use_filter_line_search = use_filter_line_search_; sp.filter_size = filter_size_;
    sp.extra_iteration_for_last_kkt = extra_iteration_for_last_kkt_; sp.reg_init = reginit; sp.reg_min = reg_min_;
    sp.reg_max = reg_max_; sp.reg_incfactor = reg_incfactor_; sp.reg_decfactor = reg_decfactor_;
    sp.th_stepdec = th_stepdec_; sp.th_stepinc = th_stepinc_; sp.termination_tol = termination_tol_; sp.mu_dynamic = mu_dynamic_;
    /* crocoddyl's setCandidate: empty warm start = x0 at every knot here (as setCandidate above), zero controls */
    Vec xs0 = init_xs;
    if (xs0.empty()) {
      xs0.resize(xs_.size());
      for (size_t b = 0; b < B; ++b)
        for (int t = 0; t <= sh_.T; ++t) std::memcpy(&xs0[xoff((int)b, t)], pm->x0((int)b), sh_.nx * sizeof(double));
    }
    n_device_passes_ = engine_->sqp_solve(sp, xs0.data(), init_us.empty() ? nullptr : init_us.data());
    xs_ = engine_->field(MIMGPU_F_XS); us_ = engine_->field(MIMGPU_F_US);
    std::vector<int32_t> it, oc;
    engine_->solve_info(it, oc);
    const Mask all(B, 1);
    pull(MIMGPU_S_KKT, KKT_, all); pull(MIMGPU_S_COST, cost_, all); pull(MIMGPU_S_GAP_NORM, gap_norm_, all);
    pull(MIMGPU_S_MERIT, merit_, all); pull(MIMGPU_S_STEPLENGTH, steplength_, all); pull(MIMGPU_S_PREG, preg_, all);
    pull(MIMGPU_S_XGRAD_NORM, x_grad_norm_, all); pull(MIMGPU_S_UGRAD_NORM, u_grad_norm_, all);
    dreg_ = preg_;
    std::vector<bool> converged(B);
    for (size_t b = 0; b < B; ++b) { iter_[b] = it[b]; converged[b] = (oc[b] == 1); }
    return converged;
  }
  int get_device_passes() const { return n_device_passes_; }

 private:
  bool model_attached_ = false;
  int n_device_passes_ = 0;
  void launchDirection() {
    mimgpu_sqp_params p;
    p.preg = reg_min_; p.dreg = reg_min_; p.mu_dynamic = mu_dynamic_; p.is_feasible = 0; p.reserved = 0;
    engine_->set_reg(preg_, dreg_);
    engine_->sqp_direction(p);
  }
  void pullScalars(const Mask& m) {
    pull(MIMGPU_S_KKT, KKT_, m); pull(MIMGPU_S_GAP_NORM, gap_norm_, m); pull(MIMGPU_S_MERIT, merit_, m);
    pull(MIMGPU_S_XGRAD_NORM, x_grad_norm_, m); pull(MIMGPU_S_UGRAD_NORM, u_grad_norm_, m);
  }
  void iteration(Mask& running) {
    Mask gave_up = directionWithRetry(running, true, [this] { launchDirection(); });
    pullScalars(running);
    runCallbacks("SQP");
    for (int b = 0; b < sh_.B; ++b) if (gave_up[b]) running[b] = 0;
  }
  void accept(int b) {
    std::copy(&xs_try_[xoff(b)], &xs_try_[xoff(b)] + (size_t)(sh_.T + 1) * sh_.nx, &xs_[xoff(b)]);
    std::copy(&us_try_[uoff(b)], &us_try_[uoff(b)] + (size_t)sh_.T * sh_.nu, &us_[uoff(b)]);
  }
  Vec dx_cache_, du_cache_;
};

/* ------------------------------------------------------------------ SolverCSQP (include/mim_solvers/csqp.hpp:41-357) */
class BatchSolverCSQP : public BatchSolverBase {
 public:
  explicit BatchSolverCSQP(std::shared_ptr<BatchShootingProblem> problem, int device = 0)
      : BatchSolverBase(std::move(problem), device) {
    mu_dynamic_ = 1e1;  /* csqp.hpp:312 */
    const size_t B = sh_.B;
    constraint_norm_.assign(B, 0.0);
    constraint_norm_try_ = lag_mul_inf_norm_ = norm_primal_ = norm_dual_ = norm_primal_rel_ = norm_dual_rel_ = constraint_norm_;
    qp_iters_.assign(B, 0);
  }

  /* src/csqp.cpp:411-452: derivatives (upstream) + gap_norm_ / constraint_norm_ (device) */
  void calc(bool recalc = true) {
    if (recalc || !engine_) calcDiffAndUpload();
    engine_->csqp_calc_norms();
    gap_norm_ = engine_->scalar(MIMGPU_S_GAP_NORM);
    constraint_norm_ = engine_->scalar(MIMGPU_S_CONSTRAINT_NORM);
  }
  /* src/csqp.cpp:534-537 */
  void reset_rho_vec() { engine_->csqp_reset_rho(params()); }
  /* src/csqp.cpp:454-517: the whole ADMM loop of every problem in one CUDA launch */
  void computeDirection(bool recalcDiff = true) {
    if (recalcDiff && !engine_) calc(true);
    if (!remove_reg_) engine_->set_reg(preg_, dreg_);
    else engine_->clear_reg();
    engine_->csqp_direction(params(), true);
    pullScalars(Mask(sh_.B, 1));
  }
  /* src/csqp.cpp:584-628: computed at the end of the direction launch */
  const Vec& checkKKTConditions() const { return KKT_; }

  /* src/csqp.cpp:1342-1435 */
  void tryStep(double steplength, const Mask& mask, Mask& bad) {
    if (steplength > 1. || steplength < 0.) throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    const int B = sh_.B, T = sh_.T, nx = sh_.nx, nu = sh_.nu, ng = sh_.ng, ngT = sh_.ngT;
    Vec step(nx);
    bad.assign(B, 0);
    for (int b = 0; b < B; ++b) {
      if (!mask[b]) continue;
      for (int t = 0; t <= T; ++t) {
        for (int i = 0; i < nx; ++i) step[i] = steplength * dx_cache_[xoff(b, t) + i];
        problem_->integrate(&xs_[xoff(b, t)], step.data(), &xs_try_[xoff(b, t)]);
      }
      for (int i = 0; i < T * nu; ++i) us_try_[uoff(b) + i] = us_[uoff(b) + i] + steplength * du_cache_[uoff(b) + i];
      double* xn = xnext_.data() + (size_t)b * T * nx;
      double* gg = ng ? &g_[(size_t)b * T * ng] : nullptr;
      double* gt = ngT ? &gT_[(size_t)b * ngT] : nullptr;
      const double c = problem_->calc(b, &xs_try_[xoff(b)], &us_try_[uoff(b)], xn, gg, gt);
      double gap = 0.0, con = 0.0;
      for (int t = 0; t < T; ++t) {
        problem_->diff(&xs_try_[xoff(b, t + 1)], xn + (size_t)t * nx, &fs_try_[xoff(b, t + 1)]);
        for (int i = 0; i < nx; ++i) gap += std::fabs(fs_try_[xoff(b, t + 1) + i]);
      }
      for (int i = 0; i < T * ng; ++i) {
        const size_t k = (size_t)b * T * ng + i;
        con += std::max(derivs_.g_lb[k] - gg[i], 0.0) + std::max(gg[i] - derivs_.g_ub[k], 0.0);
      }
      for (int i = 0; i < ngT; ++i) {
        const size_t k = (size_t)b * ngT + i;
        con += std::max(derivs_.g_lbT[k] - gt[i], 0.0) + std::max(gt[i] - derivs_.g_ubT[k], 0.0);
      }
      cost_try_[b] = c;
      gap_norm_try_[b] = gap;
      constraint_norm_try_[b] = con;
      if (mu_dynamic_ < 0. || mu_constraint_ < 0.)
        merit_try_[b] = c + lag_mul_inf_norm_coef_ * lag_mul_inf_norm_[b] * (gap + con);
      else
        merit_try_[b] = c + mu_dynamic_ * gap + mu_constraint_ * con;
      bad[b] = raiseIfNaN(c);
    }
  }

  /* src/csqp.cpp:197-409 per problem */
  std::vector<bool> solve(const Vec& init_xs = Vec(), const Vec& init_us = Vec(), size_t maxiter = 100,
                          bool /*is_feasible*/ = false, double reginit = NAN) {
    const int B = sh_.B, nx = sh_.nx, T = sh_.T;
    setCandidate(init_xs, init_us, false);
    for (int b = 0; b < B; ++b) {
      std::memcpy(&xs_[xoff(b)], problem_->x0(b), nx * sizeof(double));
      std::memcpy(&xs_try_[xoff(b)], problem_->x0(b), nx * sizeof(double));
    }
    const double reg0 = remove_reg_ ? 0.0 : (std::isnan(reginit) ? reg_min_ : reginit);
    std::fill(preg_.begin(), preg_.end(), reg0);
    std::fill(dreg_.begin(), dreg_.end(), reg0);
    if (maxiter == 0) { calc(true); reset_rho_vec(); }
    max_solve_time_reached_ = false;
    std::vector<bool> converged(B, false);
    Mask running(B, 1);
    std::fill(iter_.begin(), iter_.end(), 0);
    FilterMemory gaps(B, filter_size_), costs(B, filter_size_), cons(B, filter_size_);
    size_t it = 0;
    while (it < maxiter && any(running)) {
      calc(true);
      if (it == 0 && !reset_rho_) reset_rho_vec();
      Mask gave_up = qp(running);
      for (int b = 0; b < B; ++b) if (gave_up[b]) running[b] = 0;
      pullScalars(running);
      for (int b = 0; b < B; ++b) if (qp_iters_[b] == 0) running[b] = 0;
      runCallbacks("CSQP");
      for (int b = 0; b < B; ++b)
        if (running[b] && KKT_[b] <= termination_tol_) { converged[b] = true; running[b] = 0; }
      if (!any(running)) break;
      for (int b = 0; b < B; ++b)
        if (running[b]) { cons.push(b, constraint_norm_[b]); gaps.push(b, gap_norm_[b]); costs.push(b, cost_[b]); }
      /* dx_/du_ hold dxtilde/dutilde after update_lagrangian_parameters (src/csqp.cpp:1241-1242) */
      dx_cache_ = engine_->field(MIMGPU_F_DX);
      du_cache_ = engine_->field(MIMGPU_F_DU);
      if (mu_dynamic_ < 0. || mu_constraint_ < 0.) {
        Vec lam = engine_->field(MIMGPU_F_LAG_MUL), yy = engine_->field(MIMGPU_F_Y);
        const size_t nl = (size_t)(T + 1) * nx, nc = (size_t)T * sh_.ng + sh_.ngT;
        for (int b = 0; b < B; ++b) {
          double m = 0.0;
          for (size_t i = 0; i < nl; ++i) m = std::max(m, std::fabs(lam[b * nl + i]));
          for (size_t i = 0; i < nc; ++i) m = std::max(m, std::fabs(yy[b * nc + i]));
          lag_mul_inf_norm_[b] = m;
          merit_[b] = cost_[b] + lag_mul_inf_norm_coef_ * m * (gap_norm_[b] + constraint_norm_[b]);
        }
      } else {
        for (int b = 0; b < B; ++b) merit_[b] = cost_[b] + mu_dynamic_ * gap_norm_[b] + mu_constraint_ * constraint_norm_[b];
      }
      Mask searching = running, bad;
      for (double alpha : alphas_) {
        if (!any(searching)) break;
        tryStep(alpha, searching, bad);
        for (int b = 0; b < B; ++b) {
          if (!searching[b]) continue;
          const bool ok = use_filter_line_search_
                              ? filterAccepts(b, {{&costs, cost_try_[b]}, {&gaps, gap_norm_try_[b]}, {&cons, constraint_norm_try_[b]}})
                              : merit_[b] > merit_try_[b];
          if (ok && !bad[b]) {
            std::copy(&xs_try_[xoff(b)], &xs_try_[xoff(b)] + (size_t)(T + 1) * nx, &xs_[xoff(b)]);
            std::copy(&us_try_[uoff(b)], &us_try_[uoff(b)] + (size_t)T * sh_.nu, &us_[uoff(b)]);
            searching[b] = 0;
          }
        }
      }
      for (int b = 0; b < B; ++b) if (searching[b]) running[b] = 0;   /* `if (!found) break;` */
      if (!remove_reg_) {
        /* The reference's loop variable shadows the member steplength_ (src/csqp.cpp:317), so
         * the schedule always sees the member (1.0): regularisation only decreases. */
        for (int b = 0; b < B; ++b) {
          if (!running[b]) continue;
          if (steplength_[b] > th_stepdec_) {
            decreaseRegularization(b);
          } else {
            increaseRegularization(b);
            if (preg_[b] >= reg_max_) running[b] = 0;
          }
        }
      }
      ++it;
      for (int b = 0; b < B; ++b) if (running[b]) iter_[b] = (int64_t)it;
    }
    if (extra_iteration_for_last_kkt_ && any(running) && !max_solve_time_reached_) {
      calc(true);
      Mask gave_up = qp(running);
      for (int b = 0; b < B; ++b) if (gave_up[b]) running[b] = 0;
      pullScalars(running);
      for (int b = 0; b < B; ++b) if (running[b] && KKT_[b] <= termination_tol_) converged[b] = true;
    }
    return converged;
  }

  /* ---- accessors (bindings/csqp.cpp:85-240) */
  Vec get_dx_tilde() { return engine().field(MIMGPU_F_DX_TILDE); }
  Vec get_du_tilde() { return engine().field(MIMGPU_F_DU_TILDE); }
  Vec get_y() { return engine().field(MIMGPU_F_Y); }
  Vec get_z() { return engine().field(MIMGPU_F_Z); }
  Vec get_rho_vec() { return engine().field(MIMGPU_F_RHO_VEC); }
  void set_y(const Vec& v) { engine().set_field(MIMGPU_F_Y, v); }
  void set_z(const Vec& v) { engine().set_field(MIMGPU_F_Z, v); }
  void set_rho_vec(const Vec& v) { engine().set_field(MIMGPU_F_RHO_VEC, v); }
  Vec get_rho_sparse() { return engine_ ? engine_->scalar(MIMGPU_S_RHO_SPARSE) : Vec((size_t)sh_.B, rho_sparse_base_); }
  void set_rho_sparse(double v) { rho_sparse_base_ = v; }
  const std::vector<int64_t>& get_qp_iters() const { return qp_iters_; }
  const Vec& get_constraint_norm() const { return constraint_norm_; }
  const Vec& get_norm_primal() const { return norm_primal_; }
  const Vec& get_norm_dual() const { return norm_dual_; }
  const Vec& get_norm_primal_rel() const { return norm_primal_rel_; }
  const Vec& get_norm_dual_rel() const { return norm_dual_rel_; }
  const Vec& get_xgrad_norm() const { return x_grad_norm_; }
  const Vec& get_ugrad_norm() const { return u_grad_norm_; }
  double get_sigma() const { return sigma_; }
  void set_sigma(double v) { sigma_ = v; }
  double get_alpha() const { return alpha_; }
  void set_alpha(double v) { alpha_ = v; }
  double get_eps_abs() const { return eps_abs_; }
  void set_eps_abs(double v) { eps_abs_ = v; }
  double get_eps_rel() const { return eps_rel_; }
  void set_eps_rel(double v) { eps_rel_ = v; }
  int get_max_qp_iters() const { return max_qp_iters_; }
  void set_max_qp_iters(int v) { max_qp_iters_ = v; }
  int get_rho_update_interval() const { return rho_update_interval_; }
  void set_rho_update_interval(int v) { rho_update_interval_ = v; }
  double get_adaptive_rho_tolerance() const { return adaptive_rho_tolerance_; }
  void set_adaptive_rho_tolerance(double v) { adaptive_rho_tolerance_ = v; }
  double get_rho_min() const { return rho_min_; }
  double get_rho_max() const { return rho_max_; }
  bool get_equality_qp_initial_guess() const { return equality_qp_initial_guess_; }
  void set_equality_qp_initial_guess(bool v) { equality_qp_initial_guess_ = v; }
  bool get_reset_y() const { return reset_y_; }
  void set_reset_y(bool v) { reset_y_ = v; }
  bool get_reset_rho() const { return reset_rho_; }
  void set_reset_rho(bool v) { reset_rho_ = v; }
  bool get_remove_reg() const { return remove_reg_; }
  void set_remove_reg(bool v) { remove_reg_ = v; }
  bool get_update_rho_with_heuristic() const { return update_rho_with_heuristic_; }
  void set_update_rho_with_heuristic(bool v) { update_rho_with_heuristic_ = v; }
  double get_mu_constraint() const { return mu_constraint_; }
  void set_mu_constraint(double v) { mu_constraint_ = v; }
  double get_lag_mul_inf_norm_coef() const { return lag_mul_inf_norm_coef_; }
  void set_lag_mul_inf_norm_coef(double v) { lag_mul_inf_norm_coef_ = v; }
  double get_max_solve_time() const { return max_solve_time_; }
  void set_max_solve_time(double v) { max_solve_time_ = v; }
  bool get_max_solve_time_reached() const { return max_solve_time_reached_; }

 private:
  mimgpu_csqp_params params() const {
    mimgpu_csqp_params p;
    std::memset(&p, 0, sizeof p);
    const double reg = remove_reg_ ? 0.0 : reg_min_;
    p.preg = reg; p.dreg = reg; p.sigma = sigma_; p.alpha = alpha_; p.rho_sparse_base = rho_sparse_base_;
    p.rho_min = rho_min_; p.rho_max = rho_max_; p.adaptive_rho_tolerance = adaptive_rho_tolerance_;
    p.eps_abs = eps_abs_; p.eps_rel = eps_rel_; p.max_qp_iters = max_qp_iters_;
    p.rho_update_interval = rho_update_interval_; p.equality_qp_initial_guess = equality_qp_initial_guess_;
    p.reset_y = reset_y_; p.reset_rho = reset_rho_; p.update_rho_with_heuristic = update_rho_with_heuristic_;
    return p;
  }
  void pullScalars(const Mask& m) {
    pull(MIMGPU_S_KKT, KKT_, m); pull(MIMGPU_S_XGRAD_NORM, x_grad_norm_, m); pull(MIMGPU_S_UGRAD_NORM, u_grad_norm_, m);
    pull(MIMGPU_S_NORM_PRIMAL, norm_primal_, m); pull(MIMGPU_S_NORM_DUAL, norm_dual_, m);
    pull(MIMGPU_S_NORM_PRIMAL_REL, norm_primal_rel_, m); pull(MIMGPU_S_NORM_DUAL_REL, norm_dual_rel_, m);
    std::vector<int32_t> qp;
    engine_->iters_status(qp, status_);
    for (int b = 0; b < sh_.B; ++b) if (m[b]) qp_iters_[b] = qp[b];
  }
  Mask qp(const Mask& running) {
    if (remove_reg_) {
      setActive(running);
      engine_->clear_reg();
      engine_->csqp_direction(params(), true);
      engine_->set_active(nullptr);
      return Mask(sh_.B, 0);
    }
    return directionWithRetry(running, false, [this] { engine_->set_reg(preg_, dreg_); engine_->csqp_direction(params(), true); });
  }

  /* defaults: include/mim_solvers/csqp.hpp:292-331 */
  double sigma_ = 1e-6, alpha_ = 1.6, rho_sparse_base_ = 1e-1, rho_min_ = 1e-6, rho_max_ = 1e3;
  int rho_update_interval_ = 25, max_qp_iters_ = 1000;
  double adaptive_rho_tolerance_ = 5., eps_abs_ = 1e-4, eps_rel_ = 1e-4;
  bool equality_qp_initial_guess_ = true, reset_y_ = false, reset_rho_ = false, remove_reg_ = false,
       update_rho_with_heuristic_ = false, max_solve_time_reached_ = false;
  double mu_constraint_ = 1e1, lag_mul_inf_norm_coef_ = 10., max_solve_time_ = std::numeric_limits<double>::infinity();
  Vec constraint_norm_, constraint_norm_try_, lag_mul_inf_norm_, norm_primal_, norm_dual_, norm_primal_rel_, norm_dual_rel_;
  std::vector<int64_t> qp_iters_;
  Vec dx_cache_, du_cache_;
};

/* ------------------------------------------------------------------ SolverFDDP (include/mim_solvers/fddp.hpp:42-166) */
class BatchSolverFDDP : public BatchSolverBase {
 public:
  explicit BatchSolverFDDP(std::shared_ptr<BatchShootingProblem> problem, int device = 0)
      : BatchSolverBase(std::move(problem), device) {
    const size_t B = sh_.B;
    dg_.assign(B, 0.0); dq_ = dv_ = dV_ = dVexp_ = dg_;
    lag_mul_.assign(B * (size_t)(sh_.T + 1) * sh_.nx, 0.0);
    dx_roll_ = lag_mul_;
  }

  /* src/fddp.cpp:189-204 + updateExpectedImprovement (:256-279) in one launch */
  void computeDirection(bool recalcDiff = true) {
    if (recalcDiff || !engine_) calcDiffAndUpload();
    launchDirection();
    pullScalars(Mask(sh_.B, 1));
  }

  /* src/fddp.cpp:229-254: d = [dg + dv, dq - 2 dv], dv = -sum fs_t . Vxx_t (xs_t (-) xs_try_t) */
  std::pair<double, double> expectedImprovement(int b) {
    const int T = sh_.T, nx = sh_.nx;
    dv_[b] = 0.0;
    if (!is_feasible_) {
      Vec d(nx);
      double s = 0.0;
      for (int t = 0; t <= T; ++t) {
        problem_->diff(&xs_try_[xoff(b, t)], &xs_[xoff(b, t)], d.data());
        const double* V = &Vxx_[xoff(b, t) * nx];
        for (int i = 0; i < nx; ++i) {
          double v = 0.0;
          for (int j = 0; j < nx; ++j) v += V[i + nx * j] * d[j];
          s += fs_[xoff(b, t) + i] * v;
        }
      }
      dv_[b] = -s;
    }
    return {dg_[b] + dv_[b], dq_[b] - 2.0 * dv_[b]};
  }

  /* src/fddp.cpp:206-227 (same residual as src/sqp.cpp:244-268), on the host from lag_mul_ */
  const Vec& checkKKTConditions() {
    const int B = sh_.B, T = sh_.T, nx = sh_.nx, nu = sh_.nu;
    for (int b = 0; b < B; ++b) {
      double k = 0.0;
      const KnotDerivs d = derivs_.problem(b);
      for (int t = 0; t < T; ++t) {
        const double *Fx = d.Fx + (size_t)t * nx * nx, *Fu = d.Fu + (size_t)t * nx * nu;
        const double *ln = &lag_mul_[xoff(b, t + 1)], *l = &lag_mul_[xoff(b, t)];
        for (int i = 0; i < nx; ++i) {
          double v = 0.0;
          for (int j = 0; j < nx; ++j) v += Fx[j + nx * i] * ln[j];
          k = std::max(k, std::fabs(d.Lx[(size_t)t * nx + i] + v - l[i]));
        }
        for (int i = 0; i < nu; ++i) {
          double v = 0.0;
          for (int j = 0; j < nx; ++j) v += Fu[j + nx * i] * ln[j];
          k = std::max(k, std::fabs(d.Lu[(size_t)t * nu + i] + v));
        }
      }
      for (int i = 0; i < nx; ++i) k = std::max(k, std::fabs(d.LxT[i] - lag_mul_[xoff(b, T) + i]));
      for (size_t i = 0; i < (size_t)(T + 1) * nx; ++i) k = std::max(k, std::fabs(fs_[xoff(b) + i]));
      KKT_[b] = k;
    }
    return KKT_;
  }

  /* src/fddp.cpp:281-383: nonlinear roll-out with the feedback gains, problem b */
  bool forwardPass(int b, double steplength, double& cost, double& gap) {
    if (steplength > 1. || steplength < 0.) throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    const int T = sh_.T, nx = sh_.nx, nu = sh_.nu;
    const bool closed = is_feasible_ || steplength == 1.;
    Vec xnext(problem_->x0(b), problem_->x0(b) + nx), shift(nx), xn2(nx), scaled(nx);
    cost = 0.0; gap = 0.0;
    bool bad = false;
    auto lagrange = [&](int t) {
      const double* V = &Vxx_[xoff(b, t) * nx];
      for (int i = 0; i < nx; ++i) shift[i] = closed ? dx_roll_[xoff(b, t) + i] : dx_roll_[xoff(b, t) + i] - fs_[xoff(b, t) + i];
      for (int i = 0; i < nx; ++i) {
        double v = 0.0;
        for (int j = 0; j < nx; ++j) v += V[i + nx * j] * shift[j];
        lag_mul_[xoff(b, t) + i] = v + Vx_[xoff(b, t) + i];
      }
    };
    auto place = [&](int t) {
      if (closed) {
        std::copy(xnext.begin(), xnext.end(), &xs_try_[xoff(b, t)]);
      } else {
        for (int i = 0; i < nx; ++i) scaled[i] = fs_[xoff(b, t) + i] * (steplength - 1);
        problem_->integrate(xnext.data(), scaled.data(), &xs_try_[xoff(b, t)]);
      }
    };
    for (int t = 0; t < T; ++t) {
      lagrange(t);
      place(t);
      problem_->diff(&xs_[xoff(b, t)], &xs_try_[xoff(b, t)], &dx_roll_[xoff(b, t)]);
      const double* K = &K_[uoff(b, t) * nx];
      for (int a = 0; a < nu; ++a) {
        double v = 0.0;
        for (int j = 0; j < nx; ++j) v += K[a * nx + j] * dx_roll_[xoff(b, t) + j];
        us_try_[uoff(b, t) + a] = us_[uoff(b, t) + a] - k_[uoff(b, t) + a] * steplength - v;
      }
      cost += problem_->calcKnot(b, t, &xs_try_[xoff(b, t)], &us_try_[uoff(b, t)], xn2.data());
      xnext = xn2;
      problem_->diff(&xs_try_[xoff(b, t)], xnext.data(), &fs_try_[xoff(b, t)]);
      double mx = 0.0;
      for (int i = 0; i < nx; ++i) { gap += std::fabs(fs_try_[xoff(b, t) + i]); mx = std::max(mx, std::fabs(xnext[i])); }
      bad = bad || raiseIfNaN(cost) || raiseIfNaN(mx);
    }
    lagrange(T);
    place(T);
    cost += problem_->calcTerminal(b, &xs_try_[xoff(b, T)]);
    bad = bad || raiseIfNaN(cost);
    return bad;
  }

  /* src/fddp.cpp:57-187 per problem (one shared is_feasible flag: all problems start from the
   * same kind of guess) */
  std::vector<bool> solve(const Vec& init_xs = Vec(), const Vec& init_us = Vec(), size_t maxiter = 100,
                          bool is_feasible = false, double reginit = NAN) {
    const int B = sh_.B, nx = sh_.nx, T = sh_.T;
    for (int b = 0; b < B; ++b) std::memcpy(&xs_try_[xoff(b)], problem_->x0(b), nx * sizeof(double));
    setCandidate(init_xs, init_us, is_feasible);
    const double reg0 = std::isnan(reginit) ? reg_min_ : reginit;
    std::fill(preg_.begin(), preg_.end(), reg0);
    std::fill(dreg_.begin(), dreg_.end(), reg0);
    std::vector<bool> converged(B, false);
    Mask running(B, 1);
    FilterMemory gaps(B, filter_size_), costs(B, filter_size_);
    std::fill(lag_mul_.begin(), lag_mul_.end(), 0.0);
    std::fill(dx_roll_.begin(), dx_roll_.end(), 0.0);
    std::fill(iter_.begin(), iter_.end(), 0);
    for (size_t it = 0; it < maxiter; ++it) {
      if (!any(running)) break;
      Mask gave_up = directionWithRetry(running, true, [this] { launchDirection(); });
      for (int b = 0; b < B; ++b) if (gave_up[b]) running[b] = 0;
      pullScalars(running);
      Vxx_ = engine_->field(MIMGPU_F_VXX); Vx_ = engine_->field(MIMGPU_F_VX);
      K_ = engine_->field(MIMGPU_F_K); k_ = engine_->field(MIMGPU_F_KFF);
      checkKKTConditions();
      for (int b = 0; b < B; ++b)
        if (running[b] && KKT_[b] <= termination_tol_) { converged[b] = true; running[b] = 0; }
      if (!any(running)) break;
      for (int b = 0; b < B; ++b) if (running[b]) { gaps.push(b, gap_norm_[b]); costs.push(b, cost_[b]); }
      Vec xs_new = xs_, us_new = us_;
      for (int b = 0; b < B; ++b) {
        if (!running[b]) continue;
        for (double alpha : alphas_) {
          steplength_[b] = alpha;
          double c, g;
          const bool bad = forwardPass(b, alpha, c, g);
          cost_try_[b] = c; gap_norm_try_[b] = g;
          dV_[b] = cost_[b] - c;
          bool ok;
          if (use_filter_line_search_) {
            ok = filterAccepts(b, {{&costs, c}, {&gaps, g}});
          } else {
            const auto d = expectedImprovement(b);
            dVexp_[b] = alpha * (d.first + 0.5 * alpha * d.second);
            ok = dVexp_[b] >= 0 ? (d.first < th_grad_ || dV_[b] > th_acceptstep_ * dVexp_[b])
                                : dV_[b] > th_acceptnegstep_ * dVexp_[b];
          }
          if (ok && !bad) {
            std::copy(&xs_try_[xoff(b)], &xs_try_[xoff(b)] + (size_t)(T + 1) * nx, &xs_new[xoff(b)]);
            std::copy(&us_try_[uoff(b)], &us_try_[uoff(b)] + (size_t)T * sh_.nu, &us_new[uoff(b)]);
            break;
          }
        }
      }
      xs_ = xs_new; us_ = us_new;
      is_feasible_ = false;
      for (int b = 0; b < B; ++b) {
        if (!running[b]) continue;
        if (steplength_[b] > th_stepdec_) decreaseRegularization(b);
        if (steplength_[b] <= th_stepinc_) {
          increaseRegularization(b);
          if (preg_[b] >= reg_max_) running[b] = 0;
        }
        if (running[b]) iter_[b] = (int64_t)it + 1;
      }
      runCallbacks("SQP");
    }
    return converged;
  }

  const Vec& get_dg() const { return dg_; }
  const Vec& get_dq() const { return dq_; }
  const Vec& get_dV() const { return dV_; }
  const Vec& get_dVexp() const { return dVexp_; }
  const Vec& get_lag_mul_host() const { return lag_mul_; }
  double get_th_acceptnegstep() const { return th_acceptnegstep_; }
  void set_th_acceptnegstep(double v) { if (v < 0.) throw std::invalid_argument("Invalid argument: th_acceptnegstep value has to be positive."); th_acceptnegstep_ = v; }

 private:
  void launchDirection() {
    mimgpu_sqp_params p;
    p.preg = reg_min_; p.dreg = reg_min_; p.mu_dynamic = mu_dynamic_; p.is_feasible = is_feasible_ ? 1 : 0; p.reserved = 0;
    engine_->set_reg(preg_, dreg_);
    engine_->fddp_direction(p);
  }
  void pullScalars(const Mask& m) { pull(MIMGPU_S_GAP_NORM, gap_norm_, m); pull(MIMGPU_S_DG, dg_, m); pull(MIMGPU_S_DQ, dq_, m); }

  double th_acceptnegstep_ = 2.;
  Vec dg_, dq_, dv_, dV_, dVexp_, lag_mul_, dx_roll_, Vxx_, Vx_, K_, k_;
};

}  // namespace mim_solvers_b200

#endif  // MIM_SOLVERS_B200_BATCH_SOLVERS_HPP_
