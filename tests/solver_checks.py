"""Host-logic checks of the batched solver classes (mim_solvers_b200/solvers.py), shared by
the CPU suite (kernel bodies through the host emulator, tests/emu) and the GPU suite
(libmimgpu.so).  The invariants are the reference's own (tests/test_solvers.cpp:267,290:
LQ problems converge with iter == 1; :346-347: CSQP == SQP without constraints to 1e-6;
:294-299: QP residuals <= tolerance and qp_iters < max; :83-112: defaults)."""
import numpy as np

from mim_solvers_b200.solvers import LQBatchProblem, SolverCSQP, SolverFDDP, SolverSQP


def make_lq(B=3, T=12, nx=6, nu=3, seed=0, constrained=False):
    r = np.random.default_rng(seed)
    dt = 0.05
    A = np.eye(nx)[None, None] + dt * r.uniform(-1, 1, (B, T, nx, nx))
    Bm = dt * r.uniform(-1, 1, (B, T, nx, nu))
    f = dt * r.normal(size=(B, T, nx))
    W = r.normal(size=(B, T, nx + nu, nx + nu))
    H = dt * (W @ np.swapaxes(W, -1, -2) / (nx + nu) + np.eye(nx + nu) * 0.1)
    Q, S, R = H[..., :nx, :nx], H[..., :nx, nx:], H[..., nx:, nx:]
    q, rr = dt * r.normal(size=(B, T, nx)), dt * r.normal(size=(B, T, nu))
    WT = r.normal(size=(B, nx, nx))
    QT = WT @ np.swapaxes(WT, -1, -2) / nx + np.eye(nx)
    qT = r.normal(size=(B, nx))
    x0 = r.normal(size=(B, nx))
    kw = {}
    if constrained:  # control box (Gu = I) + a terminal state box
        ng = nu
        kw = dict(C=np.zeros((B, T, ng, nx)), D=np.tile(np.eye(nu), (B, T, 1, 1)), e=np.zeros((B, T, ng)),
                  g_lb=np.full((B, T, ng), -0.3), g_ub=np.full((B, T, ng), 0.3),
                  CT=np.tile(np.eye(nx), (B, 1, 1)), eT=np.zeros((B, nx)),
                  g_lbT=np.full((B, nx), -50.0), g_ubT=np.full((B, nx), 50.0))
    return LQBatchProblem(x0, A, Bm, f, Q, S, R, q, rr, QT, qT, **kw)


def check_defaults(factory):
    p = make_lq(B=1, T=3)
    s = SolverCSQP(p, factory)
    assert (s.termination_tolerance, s.max_qp_iters, s.eps_abs, s.eps_rel) == (1e-6, 1000, 1e-4, 1e-4)
    assert (s.rho_update_interval, s.adaptive_rho_tolerance, s.alpha, s.sigma) == (25, 5.0, 1.6, 1e-6)
    assert (s.rho_min, s.rho_max, s.mu_dynamic, s.mu_constraint) == (1e-6, 1e3, 1e1, 1e1)
    assert s.equality_qp_initial_guess and s.use_filter_line_search and s.filter_size == 1
    q = SolverSQP(p, factory)
    assert (q.mu_dynamic, q.reg_min, q.reg_max, q.th_stepDec, q.th_stepInc) == (1.0, 1e-9, 1e9, 0.5, 0.01)
    assert q.alphas == [2.0 ** -n for n in range(10)]


def check_sqp_lq(factory):
    """tests/test_solvers.cpp:255-270: an LQ problem is solved by one full Newton step."""
    p = make_lq()
    s = SolverSQP(p, factory)
    ok = s.solve(maxiter=10)
    assert ok.all(), s.KKT
    assert (s.iter == 1).all(), s.iter
    assert (s.KKT <= s.termination_tolerance).all()
    # the accepted trajectory closes every gap
    c = p.calc(s.xs, s.us)
    assert np.abs(c["xnext"] - s.xs[:, 1:]).max() < 1e-9
    return s


def check_csqp_equals_sqp_without_constraints(factory):
    """tests/test_solvers.cpp:330-348."""
    p = make_lq(seed=3)
    a, b = SolverSQP(p, factory), SolverCSQP(p, factory)
    oa, ob = a.solve(maxiter=10), b.solve(maxiter=10)
    assert oa.all() and ob.all()
    assert (b.iter == 1).all() and (b.qp_iters == 25).all(), (b.iter, b.qp_iters)
    assert np.abs(a.xs - b.xs).max() < 1e-6 and np.abs(a.us - b.us).max() < 1e-6


def check_csqp_constrained(factory):
    """tests/test_solvers.cpp:285-300: residuals below tolerance, qp_iters < max, and the box
    is respected by the solution."""
    p = make_lq(seed=5, constrained=True)
    s = SolverCSQP(p, factory)
    s.termination_tolerance = 1e-4
    s.eps_abs, s.eps_rel, s.max_qp_iters = 1e-8, 0.0, 4000
    ok = s.solve(maxiter=20)
    assert ok.all(), (s.KKT, s.iter, s.qp_iters)
    assert (s.qp_iters < s.max_qp_iters).all()
    assert (s.norm_primal <= 1e-8 + 1e-12).all() and (s.norm_dual <= 1e-8 + 1e-12).all()
    assert s.us.max() <= 0.3 + 1e-5 and s.us.min() >= -0.3 - 1e-5
    assert (np.abs(s.us) > 0.3 - 1e-5).any(), "box never active: weak test"
    return s


def check_regularisation_retry(factory):
    """src/sqp.cpp:103-116: an indefinite Quu makes computeDirection fail ("backward_error");
    solve() raises preg/dreg tenfold for that problem only and retries."""
    p = make_lq(B=3, seed=7)
    p.R[1] = p.R[1] - 0.02 * np.eye(p.nu)          # problem 1: Quu not PD at reg_min
    s = SolverSQP(p, factory)
    s.solve(maxiter=1)
    assert s.preg[1] > 1e-9 * 10 and s.preg[0] == 1e-9 and s.preg[2] == 1e-9, s.preg


def check_remove_reg_on_reuse(factory):
    """src/csqp.cpp:216-219: with remove_reg_ the QP runs with preg = dreg = 0.  A solver (and its ctx) that
    earlier received per-problem regularisation arrays must not keep using them: its converged QP solution
    must be the one of a fresh remove_reg solver (1e-6; a stale reg of 1e-2 moves it by 1e-3 and more)."""
    p = make_lq(seed=11, constrained=True)

    def tight(s):
        s.eps_abs, s.eps_rel, s.max_qp_iters = 1e-10, 0.0, 8000
        return s
    a = tight(SolverCSQP(p, factory))
    a.reg_min = 1e-2                       # visible regularisation: the first solve uploads per-problem arrays
    a.solve(maxiter=1)
    xs_reg = a.xs.copy()
    a.remove_reg = True
    a.solve(maxiter=1)
    b = tight(SolverCSQP(p, factory))
    b.remove_reg = True
    b.solve(maxiter=1)
    assert (a.qp_iters < a.max_qp_iters).all() and (b.qp_iters < b.max_qp_iters).all()
    assert np.abs(xs_reg - b.xs).max() > 1e-4, "the regularised solve must differ visibly: weak test"
    assert np.abs(a.xs - b.xs).max() < 1e-6 and np.abs(a.us - b.us).max() < 1e-6, np.abs(a.xs - b.xs).max()


def check_fddp_lq(factory):
    p = make_lq(seed=9)
    s = SolverFDDP(p, factory)
    ok = s.solve(maxiter=10)
    assert ok.all(), (s.KKT, s.iter)
    # one Newton step reaches the optimum; lag_mul_ (hence KKT_) is refreshed by the NEXT roll-out
    # (src/fddp.cpp:302,343), so the reference needs up to two more iterations to see it
    assert (s.iter <= 3).all(), s.iter
    ref = SolverSQP(p, factory)
    ref.solve(maxiter=10)
    assert np.abs(ref.us - s.us).max() < 1e-5


# ------------------------------------------------------------------ multi-iteration goldens
GOLDEN_SOLVE = ("solve_lq", "solve_poly_filter", "solve_poly_filter3", "solve_poly_short_steps", "solve_poly_merit",
                "solve_poly_merit2")


def load_solve_golden(name):
    """Fixtures recorded from the reference's own python/sqp.py SQP.solve() (tests/golden/make_golden_solve.py)."""
    from pathlib import Path
    from mim_solvers_b200.solvers import PolyBatchProblem
    z = np.load(Path(__file__).resolve().parent / "golden" / f"{name}.npz")
    p = PolyBatchProblem(z["x0"], z["A"], z["Bm"], z["f"], z["Q"], z["S"], z["R"], z["q"], z["r"], z["QT"], z["qT"],
                         c3=float(z["c3"]), w4=float(z["w4"]))
    return p, z


def check_solve_golden(name, factory=None, device=False, tol=1e-8):
    """Replay one multi-iteration fixture through the batched SolverSQP (host loop, or the loop on the
    device): identical iteration counts and outcome, the accepted step length of every iteration, final
    xs / us / KKT.  The NumPy reference has no regularisation: reg_min = reginit = 0."""
    p, z = load_solve_golden(name)
    s = SolverSQP(p, factory)
    s.use_filter_line_search = bool(z["use_filter"])
    s.filter_size = int(z["filter_size"])
    s.reg_min = 0.0
    maxiter = int(z["maxiter"])
    xs0 = np.repeat(p.x0[:, None, :], p.T + 1, axis=1)
    us0 = np.zeros((p.B, p.T, p.nu))
    steps = []
    if device:
        ok = s.solve_on_device(xs0, us0, maxiter=maxiter, reginit=0.0)
    else:
        prev = {"xs": xs0.copy()}

        def cb(sol, _name):   # called after every computeDirection: which step did the previous search take?
            steps.append(sol.steplength.copy())
        s.setCallbacks([cb])
        ok = s.solve(xs0, us0, maxiter=maxiter, reginit=0.0)
        del prev
    assert np.array_equal(ok.astype(int), z["out_converged"]), (name, ok, z["out_converged"])
    assert np.array_equal(s.iter, z["out_iter"]), (name, s.iter, z["out_iter"])
    for got, ref, what in ((s.xs, z["out_xs"], "xs"), (s.us, z["out_us"], "us")):
        err = np.abs(got - ref).reshape(p.B, -1).max(axis=1) / np.abs(ref).reshape(p.B, -1).max(axis=1)
        assert err.max() <= tol, (name, what, err)
    conv = z["out_converged"].astype(bool)
    assert np.all(np.abs(s.KKT - z["out_KKT"])[conv] <= tol * np.maximum(1.0, z["out_KKT"][conv]))
    if not device and len(steps) > 1:
        # steps[k] (k >= 1) holds, for the problems still running, the step accepted by line search k - 1
        ref_a = 2.0 ** np.round(np.log2(z["out_alphas"]))   # recorded as a ratio: snap to the 2^-n grid
        for b in range(p.B):
            n = int(np.isfinite(ref_a[b]).sum())
            mine = [steps[k][b] for k in range(1, min(len(steps), n + 1))]
            assert np.array_equal(np.array(mine), ref_a[b, :len(mine)]), (name, b, mine, ref_a[b, :n])
    return s
