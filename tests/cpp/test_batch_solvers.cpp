/*
 * test_batch_solvers.cpp -- driver of the C++ host layer (include/mim_solvers_b200/batch_solvers.hpp).
 * Reads a batch of LQ problems written by tests/test_cpp_host.py, runs BatchSolverSQP / CSQP / FDDP
 * ::solve() and writes converged / iter / KKT / preg / qp_iters / xs / us for the Python side to
 * compare with the reference-mirroring checks (tests/test_solvers.cpp:255-348 invariants) and with
 * the Python host mirror on the same problems.  Linked either with libmimgpu.so (GPU) or with the
 * host-emulation shim (tests/cpp/mimgpu_emu_shim.cpp, CPU).
 *
 * usage: test_batch_solvers <sqp|csqp|fddp> problem.bin out.bin [key=value ...]
 */
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>

#include "../../include/mim_solvers_b200/batch_solvers.hpp"

using namespace mim_solvers_b200;

static void rd(FILE* f, Vec& v, size_t n) {
  v.resize(n);
  if (n && std::fread(v.data(), sizeof(double), n, f) != n) { std::fprintf(stderr, "short read\n"); std::exit(2); }
}
static void wr(FILE* f, const Vec& v) { if (!v.empty()) std::fwrite(v.data(), sizeof(double), v.size(), f); }

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: %s <sqp|csqp|fddp> problem.bin out.bin [key=value ...]\n", argv[0]); return 2; }
  const std::string solver = argv[1];
  std::map<std::string, double> opt;
  for (int i = 4; i < argc; ++i) {
    std::string a = argv[i];
    const size_t eq = a.find('=');
    if (eq == std::string::npos) { std::fprintf(stderr, "bad option %s\n", argv[i]); return 2; }
    opt[a.substr(0, eq)] = std::atof(a.c_str() + eq + 1);
  }
  auto get = [&](const char* k, double dflt) { return opt.count(k) ? opt[k] : dflt; };

  FILE* f = std::fopen(argv[2], "rb");
  if (!f) { std::perror(argv[2]); return 2; }
  int32_t h[6];
  if (std::fread(h, sizeof(int32_t), 6, f) != 6) return 2;
  auto p = std::make_shared<LQBatchProblem>();
  p->sh = mimgpu_shape{h[0], h[1], h[2], h[3], h[4], h[5]};
  const size_t B = h[0], T = h[1], nx = h[2], nu = h[3], ng = h[4], ngT = h[5];
  rd(f, p->X0, B * nx); rd(f, p->A, B * T * nx * nx); rd(f, p->Bm, B * T * nx * nu); rd(f, p->f, B * T * nx);
  rd(f, p->Q, B * T * nx * nx); rd(f, p->S, B * T * nx * nu); rd(f, p->R, B * T * nu * nu);
  rd(f, p->q, B * T * nx); rd(f, p->r, B * T * nu); rd(f, p->QT, B * nx * nx); rd(f, p->qT, B * nx);
  if (ng) { rd(f, p->C, B * T * ng * nx); rd(f, p->D, B * T * ng * nu); rd(f, p->e, B * T * ng); rd(f, p->g_lb, B * T * ng); rd(f, p->g_ub, B * T * ng); }
  if (ngT) { rd(f, p->CT, B * ngT * nx); rd(f, p->eT, B * ngT); rd(f, p->g_lbT, B * ngT); rd(f, p->g_ubT, B * ngT); }
  std::fclose(f);

  const size_t maxiter = (size_t)get("maxiter", 10);
  std::vector<bool> conv;
  Vec iters(B), qp(B, 0.0), kkt, preg, xs, us;
  try {
    if (solver == "sqp") {
      BatchSolverSQP s(p);
      s.set_termination_tolerance(get("termination_tolerance", 1e-6));
      s.set_use_filter_line_search(get("use_filter_line_search", 1) != 0);
      s.set_extra_iteration_for_last_kkt(get("extra_iteration_for_last_kkt", 0) != 0);
      conv = s.solve(Vec(), Vec(), maxiter);
      for (size_t b = 0; b < B; ++b) iters[b] = (double)s.get_iter()[b];
      kkt = s.get_KKT(); preg = s.get_preg(); xs = s.get_xs(); us = s.get_us();
    } else if (solver == "csqp") {
      BatchSolverCSQP s(p);
      s.set_termination_tolerance(get("termination_tolerance", 1e-6));
      s.set_eps_abs(get("eps_abs", 1e-4)); s.set_eps_rel(get("eps_rel", 1e-4));
      s.set_max_qp_iters((int)get("max_qp_iters", 1000));
      s.set_use_filter_line_search(get("use_filter_line_search", 1) != 0);
      if (opt.count("mu_dynamic")) s.set_mu_dynamic(opt["mu_dynamic"]);
      conv = s.solve(Vec(), Vec(), maxiter);
      for (size_t b = 0; b < B; ++b) { iters[b] = (double)s.get_iter()[b]; qp[b] = (double)s.get_qp_iters()[b]; }
      kkt = s.get_KKT(); preg = s.get_preg(); xs = s.get_xs(); us = s.get_us();
      const Vec& np_ = s.get_norm_primal(); const Vec& nd = s.get_norm_dual();
      for (size_t b = 0; b < B; ++b) std::printf("b=%zu norm_primal=%.3e norm_dual=%.3e\n", b, np_[b], nd[b]);
    } else if (solver == "fddp") {
      BatchSolverFDDP s(p);
      if (std::getenv("MIM_TRACE"))
        s.setCallbacks({[](BatchSolverBase& b, const char*) {
          for (int i = 0; i < b.shape().B; ++i)
            std::printf("trace b=%d iter=%d KKT=%.3e step=%g preg=%.1e cost=%.17g\n", i, (int)b.get_iter()[i], b.get_KKT()[i], b.get_steplength()[i], b.get_preg()[i], b.get_cost()[i]);
        }});
      s.set_termination_tolerance(get("termination_tolerance", 1e-6));
      s.set_use_filter_line_search(get("use_filter_line_search", 1) != 0);
      conv = s.solve(Vec(), Vec(), maxiter);
      for (size_t b = 0; b < B; ++b) iters[b] = (double)s.get_iter()[b];
      kkt = s.get_KKT(); preg = s.get_preg(); xs = s.get_xs(); us = s.get_us();
    } else {
      std::fprintf(stderr, "unknown solver %s\n", solver.c_str());
      return 2;
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "exception: %s\n", e.what());
    return 3;
  }
  Vec cv(B);
  for (size_t b = 0; b < B; ++b) cv[b] = conv[b] ? 1.0 : 0.0;
  FILE* o = std::fopen(argv[3], "wb");
  if (!o) { std::perror(argv[3]); return 2; }
  wr(o, cv); wr(o, iters); wr(o, kkt); wr(o, preg); wr(o, qp); wr(o, xs); wr(o, us);
  std::fclose(o);
  for (size_t b = 0; b < B; ++b)
    std::printf("b=%zu converged=%d iter=%d KKT=%.3e preg=%.1e qp_iters=%d\n", b, (int)cv[b], (int)iters[b], kkt[b], preg[b], (int)qp[b]);
  return 0;
}
