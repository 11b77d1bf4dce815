/*
 * pywrap.cpp -- pybind11 module `mim_solvers_b200._pywrap`: the batched C++ solvers of
 * include/mim_solvers_b200/batch_solvers.hpp with the reference's Python surface
 * (bindings/sqp.cpp:41-117, bindings/csqp.cpp:41-240, bindings/ddp.cpp:63-238,
 * bindings/fddp.cpp:48-100, bindings/utils/callbacks.cpp): classes SolverSQP / SolverCSQP /
 * SolverFDDP, methods solve / computeDirection / setCandidate / setCallbacks / getCallbacks,
 * the property list of the reference (xs, us, dx, du, lag_mul, KKT, merit, gap_norm, iter, cost,
 * termination_tolerance, max_qp_iters, eps_abs, eps_rel, rho_sparse, ... reg_incFactor, th_stepDec,
 * ...), CallbackAbstract / CallbackVerbose.  Every per-solver scalar of the reference is an array
 * with leading dimension B here; trajectories are numpy arrays [B, T+1, nx] / [B, T, nu].
 *
 * Built in-tree by __graft_entry__.build() (g++ + pybind11 headers, linked to libmimgpu.so).
 */
#include <pybind11/functional.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cmath>

#include "../include/mim_solvers_b200/batch_solvers.hpp"
#include "../include/mim_solvers_b200/callbacks.hpp"

namespace py = pybind11;
using namespace mim_solvers_b200;
using arr = py::array_t<double, py::array::c_style | py::array::forcecast>;

static Vec to_vec(const py::object& o) {
  if (o.is_none()) return Vec();
  arr a = arr::ensure(o);
  if (!a) throw std::invalid_argument("Invalid argument: expected a float64 array");
  return Vec(a.data(), a.data() + a.size());
}
static py::array_t<double> as_np(const Vec& v, std::vector<py::ssize_t> shape) {
  py::array_t<double> out(shape);
  if (!v.empty()) std::memcpy(out.mutable_data(), v.data(), v.size() * sizeof(double));
  return out;
}
template <class I>
static py::array_t<long long> ints_np(const std::vector<I>& v) {
  py::array_t<long long> out((py::ssize_t)v.size());
  for (size_t i = 0; i < v.size(); ++i) out.mutable_data()[i] = (long long)v[i];
  return out;
}
static py::array_t<bool> bools_np(const std::vector<bool>& v) {
  py::array_t<bool> out((py::ssize_t)v.size());
  for (size_t i = 0; i < v.size(); ++i) out.mutable_data()[i] = v[i];
  return out;
}

/* crocoddyl::CallbackAbstract analogue: Python subclasses implement __call__(solver, solver_type) */
struct CallbackAbstract {
  virtual ~CallbackAbstract() = default;
  virtual void call(py::object solver, const std::string& type) = 0;
};
struct PyCallbackAbstract : CallbackAbstract {
  using CallbackAbstract::CallbackAbstract;
  void call(py::object solver, const std::string& type) override {
    PYBIND11_OVERRIDE_PURE_NAME(void, CallbackAbstract, "__call__", call, solver, type);
  }
};
struct VerboseCb : CallbackAbstract {
  CallbackVerbose impl;
  explicit VerboseCb(int precision, int problem) : impl(precision, problem) {}
  void call(py::object solver, const std::string& type) override {
    impl(solver.cast<BatchSolverBase&>(), type.c_str());
  }
};

/* keeps the Python callback objects alive and forwards the C++ callback to them */
template <class S>
static void set_callbacks(S& self, py::object py_self, std::vector<py::object> cbs) {
  py_self.attr("_callbacks") = py::cast(cbs);
  std::vector<BatchSolverBase::Callback> v;
  py::weakref weak(py_self);
  for (auto& cb : cbs)
    v.push_back([cb, weak](BatchSolverBase&, const char* type) {
      py::gil_scoped_acquire gil;
      py::object s = weak();
      if (py::isinstance<CallbackAbstract>(cb)) cb.cast<CallbackAbstract&>().call(s, type);
      else cb(s, std::string(type));
    });
  self.setCallbacks(std::move(v));
}

template <class P>
static std::shared_ptr<P> make_lq(arr x0, arr A, arr Bm, arr f, arr Q, arr S, arr R, arr q, arr r, arr QT, arr qT,
                                  py::object C, py::object D, py::object e, py::object g_lb, py::object g_ub,
                                  py::object CT, py::object eT, py::object g_lbT, py::object g_ubT) {
  if (A.ndim() != 4 || Bm.ndim() != 4) throw std::invalid_argument("Invalid argument: A is [B,T,nx,nx], Bm is [B,T,nx,nu]");
  auto p = std::make_shared<P>();
  p->sh.B = (int)A.shape(0); p->sh.T = (int)A.shape(1); p->sh.nx = (int)A.shape(2); p->sh.nu = (int)Bm.shape(3);
  p->sh.ng = C.is_none() ? 0 : (int)arr::ensure(C).shape(2);
  p->sh.ngT = CT.is_none() ? 0 : (int)arr::ensure(CT).shape(1);
  auto v = [](const arr& a) { return Vec(a.data(), a.data() + a.size()); };
  p->X0 = v(x0); p->A = v(A); p->Bm = v(Bm); p->f = v(f); p->Q = v(Q); p->S = v(S); p->R = v(R); p->q = v(q); p->r = v(r);
  p->QT = v(QT); p->qT = v(qT);
  p->C = to_vec(C); p->D = to_vec(D); p->e = to_vec(e); p->g_lb = to_vec(g_lb); p->g_ub = to_vec(g_ub);
  p->CT = to_vec(CT); p->eT = to_vec(eT); p->g_lbT = to_vec(g_lbT); p->g_ubT = to_vec(g_ubT);
  return p;
}

#define RO_TRAJ(name, getter, d1, d2) \
  .def_property_readonly(name, [](S& s) { auto sh = s.shape(); return as_np(s.getter(), {sh.B, d1, d2}); })
#define RO_B(name, getter) .def_property_readonly(name, [](S& s) { return as_np(s.getter(), {s.shape().B}); })
#define RW(name, get, set) .def_property(name, &S::get, &S::set)

/* members every solver shares (bindings/ddp.cpp:63-238 + crocoddyl::SolverAbstract) */
template <class S, class PyC>
static void bind_common(PyC& c) {
  c.def("setCandidate",
        [](S& s, py::object xs, py::object us, bool feas) { s.setCandidate(to_vec(xs), to_vec(us), feas); },
        py::arg("xs") = py::none(), py::arg("us") = py::none(), py::arg("is_feasible") = false)
      .def("setCallbacks", [](py::object self, std::vector<py::object> cbs) { set_callbacks(self.cast<S&>(), self, cbs); })
      .def("getCallbacks", [](py::object self) -> py::object {
        if (py::hasattr(self, "_callbacks")) return self.attr("_callbacks");
        return py::list();
      })
      .def("computeDirection", &S::computeDirection, py::arg("recalc") = true)
      .def_property_readonly("problem", [](S& s) { return s.problem(); })
      RO_TRAJ("xs", get_xs, sh.T + 1, sh.nx) RO_TRAJ("us", get_us, sh.T, sh.nu)
      RO_TRAJ("xs_try", get_xs_try, sh.T + 1, sh.nx) RO_TRAJ("us_try", get_us_try, sh.T, sh.nu)
      RO_TRAJ("fs", get_fs, sh.T + 1, sh.nx)
      RO_TRAJ("dx", get_dx, sh.T + 1, sh.nx) RO_TRAJ("du", get_du, sh.T, sh.nu)
      RO_TRAJ("lag_mul", get_lag_mul, sh.T + 1, sh.nx)
      RO_TRAJ("Vx", get_Vx, sh.T + 1, sh.nx) RO_TRAJ("k", get_k, sh.T, sh.nu) RO_TRAJ("Qu", get_Qu, sh.T, sh.nu)
      .def_property_readonly("Vxx", [](S& s) { auto sh = s.shape(); return as_np(s.get_Vxx(), {sh.B, sh.T + 1, sh.nx, sh.nx}); })
      .def_property_readonly("K", [](S& s) { auto sh = s.shape(); return as_np(s.get_K(), {sh.B, sh.T, sh.nu, sh.nx}); })
      RO_B("cost", get_cost) RO_B("KKT", get_KKT) RO_B("merit", get_merit) RO_B("gap_norm", get_gap_norm)
      RO_B("preg", get_preg) RO_B("dreg", get_dreg) RO_B("stepLength", get_steplength)
      .def_property_readonly("iter", [](S& s) { return ints_np(s.get_iter()); })
      RW("termination_tolerance", get_termination_tolerance, set_termination_tolerance)
      RW("use_filter_line_search", get_use_filter_line_search, set_use_filter_line_search)
      RW("filter_size", get_filter_size, set_filter_size)
      RW("extra_iteration_for_last_kkt", get_extra_iteration_for_last_kkt, set_extra_iteration_for_last_kkt)
      RW("mu_dynamic", get_mu_dynamic, set_mu_dynamic)
      RW("reg_min", get_reg_min, set_reg_min) RW("reg_max", get_reg_max, set_reg_max)
      RW("reg_incFactor", get_reg_incfactor, set_reg_incfactor) RW("reg_decFactor", get_reg_decfactor, set_reg_decfactor)
      RW("th_stepDec", get_th_stepdec, set_th_stepdec) RW("th_stepInc", get_th_stepinc, set_th_stepinc)
      RW("th_grad", get_th_grad, set_th_grad)
      .def_property("alphas", [](S& s) { return s.get_alphas(); }, [](S& s, const Vec& a) { s.set_alphas(a); });
}

PYBIND11_MODULE(_pywrap, m) {
  m.doc() = "mim_solvers_b200: batched B200-native SolverSQP / SolverCSQP / SolverFDDP (C++ host layer over libmimgpu.so)";

  py::class_<BatchShootingProblem, std::shared_ptr<BatchShootingProblem>>(m, "BatchShootingProblem")
      .def_property_readonly("B", [](BatchShootingProblem& p) { return p.shape().B; })
      .def_property_readonly("T", [](BatchShootingProblem& p) { return p.shape().T; })
      .def_property_readonly("nx", [](BatchShootingProblem& p) { return p.shape().nx; })
      .def_property_readonly("nu", [](BatchShootingProblem& p) { return p.shape().nu; })
      .def_property_readonly("ng", [](BatchShootingProblem& p) { return p.shape().ng; })
      .def_property_readonly("ngT", [](BatchShootingProblem& p) { return p.shape().ngT; });
  py::class_<LQBatchProblem, BatchShootingProblem, std::shared_ptr<LQBatchProblem>>(m, "LQBatchProblem")
      .def(py::init(&make_lq<LQBatchProblem>), py::arg("x0"), py::arg("A"), py::arg("Bm"), py::arg("f"), py::arg("Q"),
           py::arg("S"), py::arg("R"), py::arg("q"), py::arg("r"), py::arg("QT"), py::arg("qT"),
           py::arg("C") = py::none(), py::arg("D") = py::none(), py::arg("e") = py::none(), py::arg("g_lb") = py::none(),
           py::arg("g_ub") = py::none(), py::arg("CT") = py::none(), py::arg("eT") = py::none(),
           py::arg("g_lbT") = py::none(), py::arg("g_ubT") = py::none());
  py::class_<PolyBatchProblem, LQBatchProblem, std::shared_ptr<PolyBatchProblem>>(m, "PolyBatchProblem")
      .def(py::init([](arr x0, arr A, arr Bm, arr f, arr Q, arr S, arr R, arr q, arr r, arr QT, arr qT, double c3, double w4) {
             auto p = make_lq<PolyBatchProblem>(x0, A, Bm, f, Q, S, R, q, r, QT, qT, py::none(), py::none(), py::none(),
                                                py::none(), py::none(), py::none(), py::none(), py::none(), py::none());
             p->c3 = c3; p->w4 = w4;
             return p;
           }),
           py::arg("x0"), py::arg("A"), py::arg("Bm"), py::arg("f"), py::arg("Q"), py::arg("S"), py::arg("R"), py::arg("q"),
           py::arg("r"), py::arg("QT"), py::arg("qT"), py::arg("c3") = 0.0, py::arg("w4") = 0.0)
      .def_readwrite("c3", &PolyBatchProblem::c3)
      .def_readwrite("w4", &PolyBatchProblem::w4);

  py::class_<CallbackAbstract, PyCallbackAbstract, std::shared_ptr<CallbackAbstract>>(m, "CallbackAbstract")
      .def(py::init<>())
      .def("__call__", &CallbackAbstract::call);
  py::class_<VerboseCb, CallbackAbstract, std::shared_ptr<VerboseCb>>(m, "CallbackVerbose")
      .def(py::init<int, int>(), py::arg("precision") = 3, py::arg("problem") = 0)
      .def_property("precision", [](VerboseCb& c) { return c.impl.get_precision(); },
                    [](VerboseCb& c, int p) { c.impl.set_precision(p); })
      .def_property("problem", [](VerboseCb& c) { return c.impl.get_problem(); },
                    [](VerboseCb& c, int b) { c.impl.set_problem(b); });

  py::class_<BatchSolverBase>(m, "SolverAbstract");

  {
    using S = BatchSolverSQP;
    py::class_<S, BatchSolverBase> c(m, "SolverSQP", py::dynamic_attr());
    c.def(py::init<std::shared_ptr<BatchShootingProblem>, int>(), py::arg("problem"), py::arg("device") = 0)
        .def("solve",
             [](S& s, py::object xs, py::object us, size_t maxiter, bool feas, double reg) {
               return bools_np(s.solve(to_vec(xs), to_vec(us), maxiter, feas, reg));
             },
             py::arg("init_xs") = py::none(), py::arg("init_us") = py::none(), py::arg("maxiter") = 100,
             py::arg("is_feasible") = false, py::arg("init_reg") = NAN)
        .def("solveOnDevice",
             [](S& s, py::object xs, py::object us, size_t maxiter, bool feas, double reg) {
               return bools_np(s.solveOnDevice(to_vec(xs), to_vec(us), maxiter, feas, reg));
             },
             py::arg("init_xs") = py::none(), py::arg("init_us") = py::none(), py::arg("maxiter") = 100,
             py::arg("is_feasible") = false, py::arg("init_reg") = NAN)
        .def("checkKKTConditions", [](S& s) { return as_np(s.checkKKTConditions(), {s.shape().B}); })
        .def_property_readonly("device_passes", &S::get_device_passes)
        RO_B("x_grad_norm", get_xgrad_norm) RO_B("u_grad_norm", get_ugrad_norm);
    bind_common<S>(c);
  }
  {
    using S = BatchSolverCSQP;
    py::class_<S, BatchSolverBase> c(m, "SolverCSQP", py::dynamic_attr());
    c.def(py::init<std::shared_ptr<BatchShootingProblem>, int>(), py::arg("problem"), py::arg("device") = 0)
        .def("solve",
             [](S& s, py::object xs, py::object us, size_t maxiter, bool feas, double reg) {
               return bools_np(s.solve(to_vec(xs), to_vec(us), maxiter, feas, reg));
             },
             py::arg("init_xs") = py::none(), py::arg("init_us") = py::none(), py::arg("maxiter") = 100,
             py::arg("is_feasible") = false, py::arg("init_reg") = NAN)
        .def("calc", &S::calc, py::arg("recalc") = true)
        .def("reset_rho_vec", &S::reset_rho_vec)
        .def("checkKKTConditions", [](S& s) { return as_np(s.checkKKTConditions(), {s.shape().B}); })
        RO_TRAJ("dx_tilde", get_dx_tilde, sh.T + 1, sh.nx) RO_TRAJ("du_tilde", get_du_tilde, sh.T, sh.nu)
        .def_property("y", [](S& s) { auto sh = s.shape(); return as_np(s.get_y(), {sh.B, sh.T * sh.ng + sh.ngT}); },
                      [](S& s, py::object v) { s.set_y(to_vec(v)); })
        .def_property("z", [](S& s) { auto sh = s.shape(); return as_np(s.get_z(), {sh.B, sh.T * sh.ng + sh.ngT}); },
                      [](S& s, py::object v) { s.set_z(to_vec(v)); })
        .def_property("rho_vec", [](S& s) { auto sh = s.shape(); return as_np(s.get_rho_vec(), {sh.B, sh.T * sh.ng + sh.ngT}); },
                      [](S& s, py::object v) { s.set_rho_vec(to_vec(v)); })
        .def_property("rho_sparse", [](S& s) { return as_np(s.get_rho_sparse(), {s.shape().B}); }, &S::set_rho_sparse)
        .def_property_readonly("qp_iters", [](S& s) { return ints_np(s.get_qp_iters()); })
        RO_B("constraint_norm", get_constraint_norm) RO_B("norm_primal", get_norm_primal) RO_B("norm_dual", get_norm_dual)
        RO_B("norm_primal_rel", get_norm_primal_rel) RO_B("norm_dual_rel", get_norm_dual_rel)
        RO_B("x_grad_norm", get_xgrad_norm) RO_B("u_grad_norm", get_ugrad_norm)
        RW("sigma", get_sigma, set_sigma) RW("alpha", get_alpha, set_alpha) RW("eps_abs", get_eps_abs, set_eps_abs)
        RW("eps_rel", get_eps_rel, set_eps_rel) RW("max_qp_iters", get_max_qp_iters, set_max_qp_iters)
        RW("rho_update_interval", get_rho_update_interval, set_rho_update_interval)
        RW("adaptive_rho_tolerance", get_adaptive_rho_tolerance, set_adaptive_rho_tolerance)
        .def_property_readonly("rho_min", &S::get_rho_min)
        .def_property_readonly("rho_max", &S::get_rho_max)
        RW("equality_qp_initial_guess", get_equality_qp_initial_guess, set_equality_qp_initial_guess)
        RW("reset_y", get_reset_y, set_reset_y) RW("reset_rho", get_reset_rho, set_reset_rho)
        RW("remove_reg", get_remove_reg, set_remove_reg)
        RW("update_rho_with_heuristic", get_update_rho_with_heuristic, set_update_rho_with_heuristic)
        RW("mu_constraint", get_mu_constraint, set_mu_constraint)
        RW("lag_mul_inf_norm_coef", get_lag_mul_inf_norm_coef, set_lag_mul_inf_norm_coef)
        RW("max_solve_time", get_max_solve_time, set_max_solve_time)
        .def_property_readonly("max_solve_time_reached", &S::get_max_solve_time_reached);
    bind_common<S>(c);
  }
  {
    using S = BatchSolverFDDP;
    py::class_<S, BatchSolverBase> c(m, "SolverFDDP", py::dynamic_attr());
    c.def(py::init<std::shared_ptr<BatchShootingProblem>, int>(), py::arg("problem"), py::arg("device") = 0)
        .def("solve",
             [](S& s, py::object xs, py::object us, size_t maxiter, bool feas, double reg) {
               return bools_np(s.solve(to_vec(xs), to_vec(us), maxiter, feas, reg));
             },
             py::arg("init_xs") = py::none(), py::arg("init_us") = py::none(), py::arg("maxiter") = 100,
             py::arg("is_feasible") = false, py::arg("init_reg") = NAN)
        RO_B("dg", get_dg) RO_B("dq", get_dq) RO_B("dV", get_dV) RO_B("dVexp", get_dVexp)
        RW("th_acceptNegStep", get_th_acceptnegstep, set_th_acceptnegstep);
    bind_common<S>(c);
  }
}
