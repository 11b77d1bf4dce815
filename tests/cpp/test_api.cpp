/*
 * test_api.cpp -- defaults and argument checking of the C++ host layer, the invariants of the
 * reference's own API tests (tests/test_solvers.cpp:83-112 defaults; src/ddp.cpp:547-638,
 * src/sqp.cpp:308-311 invalid arguments throw to the caller).  Links the host-emulation shim
 * (CPU, test infrastructure); prints "OK" and returns 0 when every check holds.
 */
#include <cstdio>

#include "../../include/mim_solvers_b200/batch_solvers.hpp"

using namespace mim_solvers_b200;

static int fails = 0;
#define CHECK(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); ++fails; } } while (0)
template <class F>
static bool throws_invalid(F f) {
  try { f(); } catch (const std::invalid_argument&) { return true; } catch (...) { return false; }
  return false;
}

static std::shared_ptr<LQBatchProblem> tiny() {
  auto p = std::make_shared<LQBatchProblem>();
  const int B = 2, T = 3, nx = 2, nu = 1;
  p->sh = mimgpu_shape{B, T, nx, nu, 0, 0};
  p->X0.assign(B * nx, 1.0);
  p->A.assign(B * T * nx * nx, 0.0); p->Bm.assign(B * T * nx * nu, 0.1); p->f.assign(B * T * nx, 0.0);
  p->Q.assign(B * T * nx * nx, 0.0); p->S.assign(B * T * nx * nu, 0.0); p->R.assign(B * T * nu * nu, 1.0);
  p->q.assign(B * T * nx, 0.0); p->r.assign(B * T * nu, 0.0); p->QT.assign(B * nx * nx, 0.0); p->qT.assign(B * nx, 0.0);
  for (int k = 0; k < B * T; ++k) for (int i = 0; i < nx; ++i) { p->A[(k * nx + i) * nx + i] = 1.0; p->Q[(k * nx + i) * nx + i] = 1.0; }
  for (int b = 0; b < B; ++b) for (int i = 0; i < nx; ++i) p->QT[(b * nx + i) * nx + i] = 1.0;
  return p;
}

int main() {
  auto p = tiny();
  BatchSolverSQP sqp(p);
  CHECK(sqp.get_termination_tolerance() == 1e-6);
  CHECK(sqp.get_mu_dynamic() == 1.0 && sqp.get_reg_min() == 1e-9 && sqp.get_reg_max() == 1e9);
  CHECK(sqp.get_th_stepdec() == 0.5 && sqp.get_th_stepinc() == 0.01 && sqp.get_th_grad() == 1e-12);
  CHECK(sqp.get_use_filter_line_search() && sqp.get_filter_size() == 1 && !sqp.get_extra_iteration_for_last_kkt());
  CHECK(sqp.get_alphas().size() == 10 && sqp.get_alphas()[0] == 1.0 && sqp.get_alphas()[9] == 1.0 / 512.0);
  BatchSolverCSQP csqp(p);
  CHECK(csqp.get_max_qp_iters() == 1000 && csqp.get_eps_abs() == 1e-4 && csqp.get_eps_rel() == 1e-4);
  CHECK(csqp.get_rho_update_interval() == 25 && csqp.get_adaptive_rho_tolerance() == 5.0);
  CHECK(csqp.get_alpha() == 1.6 && csqp.get_sigma() == 1e-6 && csqp.get_rho_min() == 1e-6 && csqp.get_rho_max() == 1e3);
  CHECK(csqp.get_mu_dynamic() == 1e1 && csqp.get_mu_constraint() == 1e1 && csqp.get_lag_mul_inf_norm_coef() == 10.0);
  CHECK(csqp.get_equality_qp_initial_guess() && !csqp.get_reset_y() && !csqp.get_reset_rho() && !csqp.get_remove_reg());
  CHECK(csqp.get_rho_sparse().size() == 2 && csqp.get_rho_sparse()[0] == 1e-1);
  BatchSolverFDDP fddp(p);
  CHECK(fddp.get_th_acceptnegstep() == 2.0);

  /* invalid arguments throw to the caller */
  Mask all(2, 1), bad;
  CHECK(throws_invalid([&] { sqp.tryStep(1.5, all, bad); }));
  CHECK(throws_invalid([&] { csqp.tryStep(-0.1, all, bad); }));
  CHECK(throws_invalid([&] { sqp.set_reg_incfactor(0.5); }));
  CHECK(throws_invalid([&] { sqp.set_reg_decfactor(1.0); }));
  CHECK(throws_invalid([&] { sqp.set_th_stepdec(1.5); }));
  CHECK(throws_invalid([&] { sqp.set_th_grad(-1.0); }));
  CHECK(throws_invalid([&] { sqp.set_alphas({0.5, 0.25}); }));
  CHECK(throws_invalid([&] { sqp.set_alphas({1.0, 0.5, 0.5}); }));
  CHECK(throws_invalid([&] { sqp.set_alphas({1.0, -0.5}); }));
  CHECK(throws_invalid([&] { sqp.setCandidate(Vec(3, 0.0), Vec()); }));
  CHECK(throws_invalid([&] { sqp.setCandidate(Vec(), Vec(5, 0.0)); }));
  CHECK(throws_invalid([&] { fddp.set_th_acceptnegstep(-1.0); }));
  /* members are not readable before the first direction (no engine, no silent zeros) */
  bool logic = false;
  try { sqp.get_dx(); } catch (const std::logic_error&) { logic = true; }
  CHECK(logic);

  /* a solve on the tiny batch: LQ -> one Newton step, per-problem iteration counts */
  std::vector<bool> ok = sqp.solve(Vec(), Vec(), 5);
  CHECK(ok[0] && ok[1] && sqp.get_iter()[0] == 1 && sqp.get_iter()[1] == 1);
  CHECK(sqp.get_KKT()[0] <= 1e-6 && sqp.get_dx().size() == 2u * 4u * 2u);
  std::vector<bool> okc = csqp.solve(Vec(), Vec(), 5);
  CHECK(okc[0] && okc[1] && csqp.get_iter()[0] == 1 && csqp.get_qp_iters()[0] > 0);
  if (!fails) std::printf("OK\n");
  return fails ? 1 : 0;
}
