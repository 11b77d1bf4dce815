/*
 * callbacks.hpp -- CallbackVerbose for the batched solvers (reference: src/utils/callbacks.cpp:19-167,
 * include/mim_solvers/utils/callbacks.hpp).  Same table as the reference prints (header every 10
 * iterations; columns iter, merit, cost, ||gaps||, [||Constraint||, ||(dx,du)||,] step, KKT criteria
 * [, QP iters]; "----" in the step columns on the converged line), for ONE problem of the batch
 * (`problem`, default 0) -- a batch has B independent iteration histories -- or, with problem < 0,
 * for the batch-wide worst values (max KKT, max gap, ...).
 */
#ifndef MIM_SOLVERS_B200_CALLBACKS_HPP_
#define MIM_SOLVERS_B200_CALLBACKS_HPP_

#include <algorithm>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>

#include "batch_solvers.hpp"

namespace mim_solvers_b200 {

class CallbackVerbose {
 public:
  explicit CallbackVerbose(int precision = 3, int problem = 0, std::ostream* os = nullptr)
      : precision_(precision), problem_(problem), os_(os ? os : &std::cout) {
    if (precision < 0) throw std::invalid_argument("Invalid argument: The precision needs to be at least 0.");
  }
  int get_precision() const { return precision_; }
  void set_precision(int p) {
    if (p < 0) throw std::invalid_argument("Invalid argument: The precision needs to be at least 0.");
    precision_ = p;
  }
  int get_problem() const { return problem_; }
  void set_problem(int b) { problem_ = b; }

  /* src/utils/callbacks.cpp:74-167 */
  void operator()(BatchSolverBase& solver, const char* solver_type_c) {
    const std::string type(solver_type_c);
    if (type != "SQP" && type != "CSQP" && type != "FDDP") throw std::invalid_argument("The solver type is not supported.");
    const int B = solver.shape().B;
    auto pick = [&](const Vec& v) {
      if (problem_ >= 0) return v[std::min(problem_, B - 1)];
      return *std::max_element(v.begin(), v.end());
    };
    const auto& iters = solver.get_iter();
    const long long it = problem_ >= 0 ? (long long)iters[std::min(problem_, B - 1)]
                                       : (long long)*std::max_element(iters.begin(), iters.end());
    std::ostream& o = *os_;
    if (it % 10 == 0) o << header(type) << std::endl;
    const double kkt = pick(solver.get_KKT());
    const bool done = kkt < solver.get_termination_tolerance();
    o << std::setw(4) << it << sep_;
    auto sci = [&](double v) { o << std::scientific << std::setprecision(precision_) << v; };
    sci(pick(solver.get_merit())); o << sep_;
    sci(pick(solver.get_cost())); o << sep_;
    sci(pick(solver.get_gap_norm())); o << sep_ << sep_;
    if (type == "CSQP") {
      auto& cs = static_cast<BatchSolverCSQP&>(solver);
      sci(pick(cs.get_constraint_norm())); o << sep_;
      if (done) {
        o << "       ---- " << sep_ << "     ---- " << sep_ << sep_;
      } else {
        o << sep_ << sep_;
        sci((pick(cs.get_xgrad_norm()) + pick(cs.get_ugrad_norm())) / 2); o << sep_ << sep_ << sep_;
        o << std::fixed << std::setprecision(precision_) << pick(solver.get_steplength()) << sep_ << sep_;
      }
      sci(kkt); o << sep_ << sep_short_;
      if (done) {
        o << "    -----" << sep_;
      } else {
        const auto& q = cs.get_qp_iters();
        o << std::fixed << std::setprecision(0) << sep_short_
          << (problem_ >= 0 ? q[std::min(problem_, B - 1)] : *std::max_element(q.begin(), q.end())) << sep_;
      }
    } else {
      if (done) { o << "     ---- " << sep_ << sep_; } else { sci(pick(solver.get_steplength())); o << sep_; }
      sci(kkt); o << sep_ << sep_;
    }
    o << std::endl << std::flush;
  }

 private:
  static std::string center(const std::string& s, int w) {
    const int pad = std::max(0, w - (int)s.size());
    return std::string(pad / 2, ' ') + s + std::string(pad - pad / 2, ' ');
  }
  /* src/utils/callbacks.cpp:35-72 */
  std::string header(const std::string& type) const {
    const int w = 6 + precision_;
    std::string h = "iter" + sep_;
    for (const char* c : {"merit", "cost", "||gaps||"}) h += center(c, w) + sep_;
    if (type == "CSQP") {
      h += center("||Constraint||", w) + sep_ + center("||(dx,du)||", w) + sep_;
      h += center("step", w) + sep_ + center("KKT criteria", w) + sep_ + center("QP iters", w);
    } else {
      h += center("step", w) + sep_ + center("KKT criteria", w) + sep_;
    }
    return h;
  }
  int precision_, problem_;
  std::ostream* os_;
  std::string sep_ = "  ", sep_short_ = " ";
};

}  // namespace mim_solvers_b200
#endif
