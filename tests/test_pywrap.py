"""The pybind11 module (mim_solvers_b200.pywrap): the reference's Python surface on the batched
C++ solvers.  CPU part: defaults table and argument checks of tests/test_solvers.cpp:31-162 restated
on the module (no compute call).  GPU part: the multi-iteration fixtures recorded from the
reference's python/sqp.py SQP.solve() driven through the C++ solve() loop and through
solveOnDevice(), the CSQP == SQP invariant (tests/test_solvers.cpp:346-347), callbacks."""
import numpy as np
import pytest

import solver_checks as sc


@pytest.fixture(scope="module")
def ms():
    import __graft_entry__ as g
    g.build_pywrap()
    from mim_solvers_b200 import pywrap
    return pywrap


def _lq(ms, **kw):
    lq = sc.make_lq(**kw)
    extra = {k: getattr(lq, k) for k in ("C", "D", "e", "g_lb", "g_ub", "CT", "eT", "g_lbT", "g_ubT") if getattr(lq, k) is not None}
    return ms.LQBatchProblem(lq.x0, lq.A, lq.Bm, lq.f, lq.Q, lq.S, lq.R, lq.q, lq.r, lq.QT, lq.qT, **extra), lq


def test_defaults_and_argument_checks(ms):
    p, _ = _lq(ms, B=2, T=4)
    assert (p.B, p.T, p.nx, p.nu, p.ng) == (2, 4, 6, 3, 0)
    s = ms.SolverSQP(p)
    assert (s.termination_tolerance, s.mu_dynamic, s.reg_min, s.reg_max) == (1e-6, 1.0, 1e-9, 1e9)
    assert (s.th_stepDec, s.th_stepInc, s.th_grad, s.reg_incFactor, s.reg_decFactor) == (0.5, 0.01, 1e-12, 10.0, 10.0)
    assert s.use_filter_line_search and s.filter_size == 1 and not s.extra_iteration_for_last_kkt
    assert s.alphas == [2.0 ** -n for n in range(10)]
    assert s.xs.shape == (2, 5, 6) and s.us.shape == (2, 4, 3) and np.isinf(s.KKT).all() and (s.iter == 0).all()
    c = ms.SolverCSQP(p)
    assert (c.max_qp_iters, c.eps_abs, c.eps_rel, c.rho_update_interval, c.adaptive_rho_tolerance) == (1000, 1e-4, 1e-4, 25, 5.0)
    assert (c.alpha, c.sigma, c.rho_min, c.rho_max, c.mu_dynamic, c.mu_constraint) == (1.6, 1e-6, 1e-6, 1e3, 1e1, 1e1)
    assert c.lag_mul_inf_norm_coef == 10.0 and c.equality_qp_initial_guess
    assert not (c.reset_y or c.reset_rho or c.remove_reg or c.update_rho_with_heuristic or c.max_solve_time_reached)
    assert np.isinf(c.max_solve_time) and np.all(c.rho_sparse == 1e-1)
    f = ms.SolverFDDP(p)
    assert f.th_acceptNegStep == 2.0
    for bad in (lambda: setattr(s, "reg_incFactor", 0.5), lambda: setattr(s, "reg_decFactor", 1.0),
                lambda: setattr(s, "th_stepDec", 1.5), lambda: setattr(s, "th_grad", -1.0),
                lambda: setattr(s, "alphas", [0.5, 0.25]), lambda: s.setCandidate(np.zeros(3), None),
                lambda: setattr(f, "th_acceptNegStep", -1.0)):
        with pytest.raises(ValueError):
            bad()
    with pytest.raises(RuntimeError):   # members are not readable before the first direction
        s.dx
    cb = [ms.CallbackVerbose(precision=4), ms.CallbackLogger()]
    s.setCallbacks(cb)
    assert s.getCallbacks() == cb and cb[0].precision == 4
    with pytest.raises(ValueError):
        ms.CallbackVerbose(precision=-1)


def _poly(ms, z):
    return ms.PolyBatchProblem(z["x0"], z["A"], z["Bm"], z["f"], z["Q"], z["S"], z["R"], z["q"], z["r"], z["QT"], z["qT"],
                               c3=float(z["c3"]), w4=float(z["w4"]))


@pytest.mark.gpu
@pytest.mark.parametrize("device", [False, True], ids=["host-loop", "device-loop"])
@pytest.mark.parametrize("name", sc.GOLDEN_SOLVE)
def test_cpp_solve_against_reference_python_solve(ms, name, device):
    _, z = sc.load_solve_golden(name)
    p = _poly(ms, z)
    s = ms.SolverSQP(p)
    s.use_filter_line_search, s.filter_size, s.reg_min = bool(z["use_filter"]), int(z["filter_size"]), 0.0
    log = ms.CallbackLogger()
    s.setCallbacks([log])
    xs0 = np.repeat(z["x0"][:, None, :], p.T + 1, axis=1)
    us0 = np.zeros((p.B, p.T, p.nu))
    fn = s.solveOnDevice if device else s.solve
    ok = fn(xs0, us0, int(z["maxiter"]), False, 0.0)
    assert np.array_equal(ok.astype(int), z["out_converged"]) and np.array_equal(s.iter, z["out_iter"])
    for got, ref in ((s.xs, z["out_xs"]), (s.us, z["out_us"])):
        err = np.abs(got - ref).reshape(p.B, -1).max(axis=1) / np.abs(ref).reshape(p.B, -1).max(axis=1)
        assert err.max() <= 1e-8
    if not device:   # the logger saw every iteration of the slowest problem (+ the converged pass)
        assert len(log.convergence_data["KKT"]) >= int(z["out_iter"].max())
        assert log.convergence_data["xs"][0].shape == (p.B, p.T + 1, p.nx)


@pytest.mark.gpu
def test_cpp_csqp_equals_sqp_without_constraints_and_verbose(ms, capfd):
    """tests/test_solvers.cpp:346-347 (CSQP == SQP on an unconstrained problem to 1e-6) + CallbackVerbose's table."""
    p, _ = _lq(ms, B=3, T=10)
    a, b = ms.SolverSQP(p), ms.SolverCSQP(p)
    b.setCallbacks([ms.CallbackVerbose(problem=1)])
    assert a.solve(None, None, 10).all() and b.solve(None, None, 10).all()
    assert np.abs(a.xs - b.xs).max() < 1e-6 and np.abs(a.us - b.us).max() < 1e-6
    assert (a.iter == 1).all() and (b.qp_iters > 0).all()
    out = capfd.readouterr().out
    assert "KKT criteria" in out and "QP iters" in out and "||Constraint||" in out
    pc, _ = _lq(ms, B=3, T=10, constrained=True)
    c = ms.SolverCSQP(pc)
    c.max_qp_iters, c.eps_abs, c.eps_rel = 2000, 1e-8, 1e-8
    ok = c.solve(None, None, 20)
    assert ok.all() and (np.abs(c.us).max(axis=(1, 2)) <= 0.3 + 1e-6).all()
    assert c.y.shape == (3, pc.T * pc.ng + pc.ngT)
