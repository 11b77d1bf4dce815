"""GPU: SolverSQP::solve() with the whole loop on the device (mimgpu_sqp_solve)
against the host-loop solver (mim_solvers_b200/solvers.py: same direction kernels, model / line search
/ filter / regularisation schedule in numpy, itself a restatement of src/sqp.cpp:75-222,307-374).
Problems are LQ plus separable polynomial terms (PolyBatchProblem): several SQP iterations with
data-dependent counts, step lengths below 1 under the filter, and regularisation increases up to
reg_max under the merit rule."""
import numpy as np
import pytest

import solver_checks as sc
from mim_solvers_b200.solvers import PolyBatchProblem, SolverSQP

pytestmark = pytest.mark.gpu


def make_poly(c3, w4, scale=1.0, B=6, T=15, nx=6, nu=3, seed=3):
    lq = sc.make_lq(B=B, T=T, nx=nx, nu=nu, seed=seed)
    return PolyBatchProblem(lq.x0 * scale, lq.A, lq.Bm, lq.f, lq.Q, lq.S, lq.R, lq.q, lq.r, lq.QT, lq.qT, c3=c3, w4=w4)


def run_both(p, maxiter, **attrs):
    host, dev = SolverSQP(p), SolverSQP(p)
    steps = []
    for s in (host, dev):
        for k, v in attrs.items():
            setattr(s, k, v)
    host.setCallbacks([lambda sol, name: steps.append(sol.steplength.copy())])
    xs0 = np.repeat(p.x0[:, None, :], p.T + 1, axis=1)   # the same explicit warm start for both
    us0 = np.zeros((p.B, p.T, p.nu))
    ok_h = host.solve(xs0, us0, maxiter=maxiter)
    ok_d = dev.solve_on_device(xs0, us0, maxiter=maxiter)
    return host, dev, ok_h, ok_d, steps


@pytest.mark.parametrize("c3,w4,scale,filt,fsize,maxiter", [
    (0.0, 0.0, 1.0, True, 1, 10),      # LQ: one Newton step (tests/test_solvers.cpp:267)
    (0.05, 0.5, 1.0, True, 1, 60),     # 10..32 iterations, all alpha = 1
    (0.02, 0.1, 1.0, True, 3, 60),     # filter memory of 3
    (0.0, 1.0, 2.0, False, 1, 60),     # merit rule, converges
    (-0.05, 1.0, 1.5, False, 1, 60),
])
def test_device_solve_matches_host_loop(c3, w4, scale, filt, fsize, maxiter):
    p = make_poly(c3, w4, scale)
    host, dev, ok_h, ok_d, _ = run_both(p, maxiter, use_filter_line_search=filt, filter_size=fsize)
    assert ok_h.all() and np.array_equal(ok_h, ok_d)
    assert np.array_equal(host.iter, dev.iter), (host.iter, dev.iter)
    if c3 == 0.0 and w4 == 0.0:
        assert (dev.iter == 1).all()
    assert len(set(dev.iter.tolist())) > 1 or (c3 == 0.0 and w4 == 0.0)
    for a, b, name in ((host.xs, dev.xs, "xs"), (host.us, dev.us, "us")):
        err = np.abs(a - b).reshape(p.B, -1).max(axis=1) / np.maximum(1e-300, np.abs(a).reshape(p.B, -1).max(axis=1))
        assert err.max() <= 1e-9, (name, err)
    assert np.all(np.abs(host.KKT - dev.KKT) <= 1e-9 * np.maximum(1.0, host.KKT))
    assert np.all(dev.KKT <= dev.termination_tolerance)
    assert np.allclose(host.cost, dev.cost, rtol=1e-12, atol=0)
    assert np.array_equal(host.preg, dev.preg)
    # passes of the device loop: the slowest problem's iterations plus the pass that finds it converged
    assert dev.n_device_iters == dev.iter.max() + 1


def test_device_solve_short_steps_and_reg_limit():
    """Step lengths below 1 (filter) and the give-up at reg_max (merit rule): per-problem outcome,
    iteration count, regularisation and accepted step of the first iterations agree with the host loop."""
    p = make_poly(0.1, 0.3, 1.0)
    host, dev, ok_h, ok_d, steps = run_both(p, 12, use_filter_line_search=True)
    assert np.array_equal(ok_h, ok_d) and np.array_equal(host.iter, dev.iter)
    assert min(s.min() for s in steps) < 1.0, "the case must exercise alpha < 1"
    assert np.array_equal(host.steplength, dev.steplength)
    assert np.abs(host.xs - dev.xs).max() <= 1e-8 * np.abs(host.xs).max()
    p = make_poly(0.02, 0.1, 1.0)
    host, dev, ok_h, ok_d, steps = run_both(p, 60, use_filter_line_search=False)
    assert not ok_d.any() and np.array_equal(ok_h, ok_d)
    assert (dev.preg == dev.reg_max).all() and np.array_equal(host.preg, dev.preg)   # stopped by the regularisation limit
    assert np.array_equal(host.iter, dev.iter), (host.iter, dev.iter)
    it, outcome = dev._dev_engine.solve_info()
    assert (outcome == 2).all()


def test_device_solve_large_batch_lq():
    """B = 2048 UR5-shaped LQ problems (the knot-record SQP kernel underneath): every problem converges
    in one iteration with the gaps closed, as the reference asserts for LQ problems."""
    lq = sc.make_lq(B=2048, T=40, nx=12, nu=6, seed=7)
    p = PolyBatchProblem(lq.x0, lq.A, lq.Bm, lq.f, lq.Q, lq.S, lq.R, lq.q, lq.r, lq.QT, lq.qT)
    s = SolverSQP(p)
    ok = s.solve_on_device(maxiter=5)
    assert ok.all() and (s.iter == 1).all()
    c = p.calc(s.xs, s.us)
    assert np.abs(c["xnext"] - s.xs[:, 1:]).max() < 1e-9


@pytest.mark.parametrize("device", [False, True], ids=["host-loop", "device-loop"])
@pytest.mark.parametrize("name", sc.GOLDEN_SOLVE)
def test_solve_against_reference_python_solve(name, device):
    """Multi-iteration fixtures recorded from the reference's python/sqp.py SQP.solve(): iteration counts,
    outcome, accepted step lengths, final xs / us / KKT, through the host loop and through the device loop."""
    sc.check_solve_golden(name, None, device=device)
