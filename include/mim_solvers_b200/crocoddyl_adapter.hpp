/*
 * crocoddyl_adapter.hpp -- B crocoddyl::ShootingProblem objects of one shape behind the
 * BatchShootingProblem interface of batch_solvers.hpp: the reference-side half of the drop-in.
 *
 * COMPILE-GUARDED: needs crocoddyl >= 3.0 and Eigen (CMake option MIM_SOLVERS_B200_WITH_CROCODDYL,
 * which defines the macro of the same name).  Neither is installed in the build container of this
 * repo, so this file is compiled only on a machine that has them; everything it relies on from
 * crocoddyl is the public API the reference itself uses at the upload point of the hot path:
 *   problem->calc / calcDiff                       src/ddp.cpp:253-258, src/csqp.cpp:413-414
 *   get_runningDatas()[t]->{Fx,Fu,Lxx,Lxu,Luu,Lx,Lu,xnext,cost,g,Gx,Gu}   src/ddp.cpp:283-310, src/csqp.cpp:727-780
 *   get_runningModels()[t]->get_g_lb() / get_g_ub()                      src/csqp.cpp:1237
 *   state->integrate / diff                        src/sqp.cpp:328,338
 * The blocks are dense column-major Eigen::MatrixXd, i.e. exactly the memory image the C ABI wants:
 * every block is ONE memcpy (KnotDerivs points into the SoA batch arrays, or -- with record_layout()
 * -- into knot records, so that no re-layout pass runs on the device).
 *
 * Usage (what a maintainer adds next to src/sqp.cpp):
 *   std::vector<std::shared_ptr<crocoddyl::ShootingProblem>> problems = ...;   // same T, ndx, nu, ng
 *   auto batch = std::make_shared<mim_solvers_b200::CrocoddylBatch>(problems);
 *   mim_solvers_b200::BatchSolverCSQP solver(batch);
 *   solver.solve(xs, us, 100);
 */
#ifndef MIM_SOLVERS_B200_CROCODDYL_ADAPTER_HPP_
#define MIM_SOLVERS_B200_CROCODDYL_ADAPTER_HPP_
#ifdef MIM_SOLVERS_B200_WITH_CROCODDYL

#include <Eigen/Core>
#include <crocoddyl/core/optctrl/shooting.hpp>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#include "batch_solvers.hpp"

namespace mim_solvers_b200 {

class CrocoddylBatch : public BatchShootingProblem {
 public:
  using Problem = crocoddyl::ShootingProblem;

  explicit CrocoddylBatch(std::vector<std::shared_ptr<Problem>> problems) : problems_(std::move(problems)) {
    if (problems_.empty()) throw std::invalid_argument("Invalid argument: empty batch");
    const auto& p0 = *problems_[0];
    sh_.B = (int32_t)problems_.size();
    sh_.T = (int32_t)p0.get_T();
    sh_.nx = (int32_t)p0.get_ndx();
    sh_.nu = (int32_t)p0.get_nu_max();
    sh_.ng = 0;
    for (const auto& m : p0.get_runningModels()) sh_.ng = std::max<int32_t>(sh_.ng, (int32_t)m->get_ng());
    sh_.ngT = (int32_t)p0.get_terminalModel()->get_ng();
    nq_state_ = (int)p0.get_nx();
    for (const auto& p : problems_) {
      if ((int)p->get_T() != sh_.T || (int)p->get_ndx() != sh_.nx || (int)p->get_nu_max() != sh_.nu)
        throw std::invalid_argument("Invalid argument: all problems of a batch need the same (T, ndx, nu); bucket by shape");
      x0_.push_back(p->get_x0());
    }
    xs_.resize(sh_.T + 1);
    us_.resize(sh_.T);
  }

  mimgpu_shape shape() const override { return sh_; }
  const double* x0(int b) const override { return x0_[b].data(); }

  /* ShootingProblem::calc: xnext_t, g_t (padded to ng rows with g = 0), total cost */
  double calc(int b, const double* xs, const double* us, double* xnext, double* g, double* gT) override {
    load(b, xs, us);
    auto& p = *problems_[b];
    const double cost = p.calc(xs_, us_);
    const int T = sh_.T, nx = nq_state_, ng = sh_.ng;
    for (int t = 0; t < T; ++t) {
      const auto& d = p.get_runningDatas()[t];
      std::memcpy(xnext + (size_t)t * nx, d->xnext.data(), nx * sizeof(double));
      if (ng && g) {
        std::fill(g + (size_t)t * ng, g + (size_t)(t + 1) * ng, 0.0);
        std::memcpy(g + (size_t)t * ng, d->g.data(), d->g.size() * sizeof(double));
      }
    }
    if (sh_.ngT && gT) std::memcpy(gT, p.get_terminalData()->g.data(), sh_.ngT * sizeof(double));
    return cost;
  }

  /* ShootingProblem::calcDiff: every Eigen block of every knot is one memcpy into the batch arrays */
  void calcDiff(int b, const double* xs, const double* us, const KnotDerivs& d) override {
    load(b, xs, us);
    auto& p = *problems_[b];
    p.calc(xs_, us_);
    p.calcDiff(xs_, us_);
    const size_t T = sh_.T, nx = sh_.nx, nu = sh_.nu, ng = sh_.ng, ngT = sh_.ngT;
    const double inf = std::numeric_limits<double>::infinity();
    for (size_t t = 0; t < T; ++t) {
      const auto& k = p.get_runningDatas()[t];
      const auto& m = p.get_runningModels()[t];
      std::memcpy(d.Fx + t * nx * nx, k->Fx.data(), nx * nx * sizeof(double));
      std::memcpy(d.Fu + t * nx * nu, k->Fu.data(), nx * nu * sizeof(double));
      std::memcpy(d.Lxx + t * nx * nx, k->Lxx.data(), nx * nx * sizeof(double));
      std::memcpy(d.Lxu + t * nx * nu, k->Lxu.data(), nx * nu * sizeof(double));
      std::memcpy(d.Luu + t * nu * nu, k->Luu.data(), nu * nu * sizeof(double));
      std::memcpy(d.Lx + t * nx, k->Lx.data(), nx * sizeof(double));
      std::memcpy(d.Lu + t * nu, k->Lu.data(), nu * sizeof(double));
      if (ng) {
        /* knots with fewer rows than ng are padded with unbounded all-zero rows (rho_min rows in
         * apply_rho_update, src/csqp.cpp:551-553: they never become active) */
        const size_t r = m->get_ng();
        for (size_t j = 0; j < nx; ++j)
          for (size_t i = 0; i < ng; ++i) d.Gx[t * ng * nx + i + ng * j] = i < r ? k->Gx(i, j) : 0.0;
        for (size_t j = 0; j < nu; ++j)
          for (size_t i = 0; i < ng; ++i) d.Gu[t * ng * nu + i + ng * j] = i < r ? k->Gu(i, j) : 0.0;
        for (size_t i = 0; i < ng; ++i) {
          d.g[t * ng + i] = i < r ? k->g[i] : 0.0;
          d.g_lb[t * ng + i] = i < r ? m->get_g_lb()[i] : -inf;
          d.g_ub[t * ng + i] = i < r ? m->get_g_ub()[i] : inf;
        }
      }
    }
    const auto& kT = p.get_terminalData();
    std::memcpy(d.LxxT, kT->Lxx.data(), nx * nx * sizeof(double));
    std::memcpy(d.LxT, kT->Lx.data(), nx * sizeof(double));
    if (ngT) {
      const auto& mT = p.get_terminalModel();
      std::memcpy(d.GxT, kT->Gx.data(), ngT * nx * sizeof(double));
      std::memcpy(d.gT, kT->g.data(), ngT * sizeof(double));
      std::memcpy(d.g_lbT, mT->get_g_lb().data(), ngT * sizeof(double));
      std::memcpy(d.g_ubT, mT->get_g_ub().data(), ngT * sizeof(double));
    }
  }

  double calcKnot(int b, int t, const double* x, const double* u, double* xnext) override {
    auto& p = *problems_[b];
    const auto& m = p.get_runningModels()[t];
    const auto& d = p.get_runningDatas()[t];
    m->calc(d, Eigen::Map<const Eigen::VectorXd>(x, nq_state_), Eigen::Map<const Eigen::VectorXd>(u, m->get_nu()));
    std::memcpy(xnext, d->xnext.data(), nq_state_ * sizeof(double));
    return d->cost;
  }
  double calcTerminal(int b, const double* x) override {
    auto& p = *problems_[b];
    p.get_terminalModel()->calc(p.get_terminalData(), Eigen::Map<const Eigen::VectorXd>(x, nq_state_));
    return p.get_terminalData()->cost;
  }
  /* manifold states (StateMultibody): dx lives in the tangent space */
  void integrate(const double* x, const double* dx, double* out) const override {
    const auto& st = problems_[0]->get_terminalModel()->get_state();
    Eigen::VectorXd o(nq_state_);
    st->integrate(Eigen::Map<const Eigen::VectorXd>(x, nq_state_), Eigen::Map<const Eigen::VectorXd>(dx, sh_.nx), o);
    std::memcpy(out, o.data(), nq_state_ * sizeof(double));
  }
  void diff(const double* xa, const double* xb, double* out) const override {
    const auto& st = problems_[0]->get_terminalModel()->get_state();
    Eigen::VectorXd o(sh_.nx);
    st->diff(Eigen::Map<const Eigen::VectorXd>(xa, nq_state_), Eigen::Map<const Eigen::VectorXd>(xb, nq_state_), o);
    std::memcpy(out, o.data(), sh_.nx * sizeof(double));
  }

 private:
  void load(int b, const double* xs, const double* us) {
    const auto& p = *problems_[b];
    for (int t = 0; t <= sh_.T; ++t) xs_[t] = Eigen::Map<const Eigen::VectorXd>(xs + (size_t)t * nq_state_, nq_state_);
    for (int t = 0; t < sh_.T; ++t)
      us_[t] = Eigen::Map<const Eigen::VectorXd>(us + (size_t)t * sh_.nu, p.get_runningModels()[t]->get_nu());
  }
  std::vector<std::shared_ptr<Problem>> problems_;
  std::vector<Eigen::VectorXd> x0_, xs_, us_;
  mimgpu_shape sh_{};
  int nq_state_ = 0; /* state->get_nx(); the solvers' `nx` is ndx.  StateVector: equal. */
};

}  // namespace mim_solvers_b200
#endif /* MIM_SOLVERS_B200_WITH_CROCODDYL */
#endif
