"""Batched host-side mirror of the reference solver classes (SolverSQP, SolverCSQP, SolverFDDP).

Same names, argument meaning and error behaviour as the reference's Python bindings
(bindings/sqp.cpp:41-117, bindings/csqp.cpp:41-240, bindings/fddp.cpp:48-100,
bindings/ddp.cpp:63-238), for a batch of B independent shooting problems of one shape.
`computeDirection` -- the hot path -- is ONE call into the CUDA library through the C ABI
(include/mimgpu.h); everything else here is the thin host logic of `solve()`:
candidate bookkeeping, merit / filter line search, regularisation schedule
(src/sqp.cpp:75-222, 307-374; src/csqp.cpp:197-409, 411-452, 1342-1435;
src/fddp.cpp:57-187, 229-383; src/ddp.cpp:431-445), vectorised over the batch with a
per-problem "still running" mask (each problem keeps its own iter / preg / filter memory,
so iteration counts are the ones B separate reference solvers would report).

What differs from the reference by construction: every per-solver scalar (KKT, merit,
iter, preg, ...) is an array of B values, `solve()` returns a bool array, and the model
evaluations (`problem.calc / calcDiff`, crocoddyl's job) go through the small
`BatchShootingProblem` interface below.  There is NO CPU fallback: the default engine is
`BatchEngine` (libmimgpu.so on a CUDA device) and constructing it without one raises.
"""
from __future__ import annotations

import numpy as np

from . import _capi
from .problems import OCPBatch

_RAISE = 1e30  # crocoddyl::raiseIfNaN: NaN, +-inf or >= 1e30


def _raise_if_nan(v):
    return ~np.isfinite(v) | (v >= _RAISE)


# ------------------------------------------------------------------------------ problems
class BatchShootingProblem:
    """What the solvers need from B `crocoddyl.ShootingProblem`s of one shape.

    Sub-classes provide (all arrays float64, leading dimension B):
      x0 [B, nx];  T, nx (tangent dimension), nu, ng, ngT
      calc(xs, us)      -> dict(cost [B], xnext [B, T, nx], g [B, T, ng], gT [B, ngT])
      calcDiff(xs, us)  -> OCPBatch holding Fx Fu Lxx Lxu Luu Lx Lu LxxT LxT (Gx Gu g g_lb
                           g_ub GxT gT g_lbT g_ubT); `fs` and `cost` are filled by the solver
      integrate(x, dx), diff(x0, x1) (= x1 (-) x0), calc_knot(t, x, u) -> (xnext, cost),
      calc_terminal(x) -> cost      (FDDP's nonlinear roll-out only)
    """
    B = T = nx = nu = 0
    ng = ngT = 0
    x0: np.ndarray

    def integrate(self, x, dx):
        return x + dx

    def diff(self, x0, x1):
        return x1 - x0


class LQBatchProblem(BatchShootingProblem):
    """B discrete-time LQ(-constrained) problems; the batched analogue of the LQR fixtures the
    reference tests use (tests/python/problems.py, crocoddyl ActionModelLQR):
      x+ = A_t x + B_t u + f_t
      l_t = 1/2 [x;u]^T [[Q_t, S_t],[S_t^T, R_t]] [x;u] + q_t.x + r_t.u,   l_T = 1/2 x^T Q_T x + q_T.x
      g_lb <= C_t x + D_t u + e_t <= g_ub   (optional)
    All inputs are mathematical (row, col) arrays with leading [B, T]."""

    def __init__(self, x0, A, Bm, f, Q, S, R, q, r, QT, qT, C=None, D=None, e=None, g_lb=None, g_ub=None,
                 CT=None, eT=None, g_lbT=None, g_ubT=None):
        self.x0 = np.asarray(x0, dtype=np.float64)
        self.A, self.Bm, self.f = A, Bm, f
        self.Q, self.S, self.R, self.q, self.r, self.QT, self.qT = Q, S, R, q, r, QT, qT
        self.B, self.T, self.nx, self.nu = A.shape[0], A.shape[1], A.shape[2], Bm.shape[3]
        self.C, self.D, self.e, self.g_lb, self.g_ub = C, D, e, g_lb, g_ub
        self.CT, self.eT, self.g_lbT, self.g_ubT = CT, eT, g_lbT, g_ubT
        self.ng = 0 if C is None else C.shape[2]
        self.ngT = 0 if CT is None else CT.shape[1]

    def calc(self, xs, us):
        x, u = xs[:, :-1], us
        xnext = np.einsum("btij,btj->bti", self.A, x) + np.einsum("btij,btj->bti", self.Bm, u) + self.f
        cost = 0.5 * np.einsum("bti,btij,btj->b", x, self.Q, x) + np.einsum("bti,btij,btj->b", x, self.S, u) \
            + 0.5 * np.einsum("bti,btij,btj->b", u, self.R, u) + np.einsum("bti,bti->b", self.q, x) \
            + np.einsum("bti,bti->b", self.r, u)
        xT = xs[:, -1]
        cost = cost + 0.5 * np.einsum("bi,bij,bj->b", xT, self.QT, xT) + np.einsum("bi,bi->b", self.qT, xT)
        out = dict(cost=cost, xnext=xnext, g=np.zeros((self.B, self.T, 0)), gT=np.zeros((self.B, 0)))
        if self.ng:
            out["g"] = np.einsum("btij,btj->bti", self.C, x) + np.einsum("btij,btj->bti", self.D, u) + self.e
        if self.ngT:
            out["gT"] = np.einsum("bij,bj->bi", self.CT, xT) + self.eT
        return out

    def calc_knot(self, t, x, u):
        xn = np.einsum("bij,bj->bi", self.A[:, t], x) + np.einsum("bij,bj->bi", self.Bm[:, t], u) + self.f[:, t]
        c = 0.5 * np.einsum("bi,bij,bj->b", x, self.Q[:, t], x) + np.einsum("bi,bij,bj->b", x, self.S[:, t], u) \
            + 0.5 * np.einsum("bi,bij,bj->b", u, self.R[:, t], u) + np.einsum("bi,bi->b", self.q[:, t], x) \
            + np.einsum("bi,bi->b", self.r[:, t], u)
        return xn, c

    def calc_terminal(self, x):
        return 0.5 * np.einsum("bi,bij,bj->b", x, self.QT, x) + np.einsum("bi,bi->b", self.qT, x)

    def calcDiff(self, xs, us):
        x, u = xs[:, :-1], us
        b = OCPBatch.zeros(self.B, self.T, self.nx, self.nu, self.ng, self.ngT)
        b.set("Fx", self.A); b.set("Fu", self.Bm)
        b.set("Lxx", self.Q); b.set("Lxu", self.S); b.set("Luu", self.R)
        b.set("Lx", np.einsum("btij,btj->bti", self.Q, x) + np.einsum("btij,btj->bti", self.S, u) + self.q)
        b.set("Lu", np.einsum("btji,btj->bti", self.S, x) + np.einsum("btij,btj->bti", self.R, u) + self.r)
        b.set("LxxT", self.QT)
        b.set("LxT", np.einsum("bij,bj->bi", self.QT, xs[:, -1]) + self.qT)
        if self.ng:
            c = self.calc(xs, us)
            b.set("Gx", self.C); b.set("Gu", self.D); b.set("g", c["g"])
            b.set("g_lb", self.g_lb); b.set("g_ub", self.g_ub)
            if self.ngT:
                b.set("gT", c["gT"])
        if self.ngT:
            b.set("GxT", self.CT); b.set("g_lbT", self.g_lbT); b.set("g_ubT", self.g_ubT)
            if not self.ng:
                b.set("gT", np.einsum("bij,bj->bi", self.CT, xs[:, -1]) + self.eT)
        return b


class PolyBatchProblem(LQBatchProblem):
    """LQ plus separable polynomial terms -- the analytic model family the device-side solve()
    evaluates (include/mimgpu.h, mimgpu_poly_model):
      x+ = A x + B u + f + c3 x^3,   l_t += w4/4 sum x_i^4,   l_T += w4/4 sum x_i^4.
    c3 = w4 = 0 is LQBatchProblem.  Unconstrained (ng = ngT = 0)."""

    def __init__(self, x0, A, Bm, f, Q, S, R, q, r, QT, qT, c3=0.0, w4=0.0):
        super().__init__(x0, A, Bm, f, Q, S, R, q, r, QT, qT)
        self.c3, self.w4 = float(c3), float(w4)

    def calc(self, xs, us):
        out = super().calc(xs, us)
        x = xs[:, :-1]
        out["xnext"] = out["xnext"] + self.c3 * x**3
        out["cost"] = out["cost"] + 0.25 * self.w4 * ((x**4).sum(axis=(1, 2)) + (xs[:, -1]**4).sum(axis=1))
        return out

    def calc_knot(self, t, x, u):
        xn, c = super().calc_knot(t, x, u)
        return xn + self.c3 * x**3, c + 0.25 * self.w4 * (x**4).sum(axis=1)

    def calc_terminal(self, x):
        return super().calc_terminal(x) + 0.25 * self.w4 * (x**4).sum(axis=1)

    def calcDiff(self, xs, us):
        b = super().calcDiff(xs, us)
        x, xT = xs[:, :-1], xs[:, -1]
        eye = np.eye(self.nx)
        b.set("Fx", self.A + 3.0 * self.c3 * (x**2)[..., None] * eye)
        b.set("Lxx", self.Q + 3.0 * self.w4 * (x**2)[..., None] * eye)
        b.set("Lx", b.arrays["Lx"] + self.w4 * x**3)
        b.set("LxxT", self.QT + 3.0 * self.w4 * (xT**2)[..., None] * eye)
        b.set("LxT", b.arrays["LxT"] + self.w4 * xT**3)
        return b

    def model_arrays(self) -> dict:
        """The model in the block layouts of the C ABI (column-major blocks)."""
        from .problems import to_colmajor
        return dict(A=to_colmajor(self.A), Bm=to_colmajor(self.Bm), f=np.ascontiguousarray(self.f),
                    Q=to_colmajor(self.Q), S=to_colmajor(self.S), R=to_colmajor(self.R),
                    q=np.ascontiguousarray(self.q), r=np.ascontiguousarray(self.r), QT=to_colmajor(self.QT),
                    qT=np.ascontiguousarray(self.qT), x0=np.ascontiguousarray(self.x0))


def _default_engine_factory(batch: OCPBatch):
    from .engine import BatchEngine  # CUDA only; raises without libmimgpu.so / a device
    return BatchEngine(batch)


class _Filter:
    """boost::circular_buffer<double> of capacity filter_size per problem, pre-filled with
    zeros by resize() (src/sqp.cpp:32-33); index filter_size-1-count is the count-th most
    recent push_back."""

    def __init__(self, B, size):
        self.size = size
        self.buf = np.zeros((B, size))

    def push(self, mask, values):
        self.buf[mask] = np.roll(self.buf[mask], -1, axis=1)
        self.buf[mask, -1] = values[mask]

    def recent(self, count):
        return self.buf[:, self.size - 1 - count]


# ------------------------------------------------------------------------------ SolverDDP base
class _SolverBase:
    """Members shared through SolverDDP (src/ddp.cpp:18-55): regularisation schedule, alphas,
    thresholds, candidate trajectories."""
    _solver = "ddp"

    def __init__(self, problem: BatchShootingProblem, engine_factory=None):
        self.problem = problem
        self._mk_engine = engine_factory or _default_engine_factory
        B, T, nx, nu = problem.B, problem.T, problem.nx, problem.nu
        self.B, self.T = B, T
        self.reg_incFactor = 10.0
        self.reg_decFactor = 10.0
        self.reg_min = 1e-9
        self.reg_max = 1e9
        self.th_stepDec = 0.5
        self.th_stepInc = 0.01
        self.th_grad = 1e-12
        self.th_acceptStep = 0.1      # SolverAbstract::th_acceptstep_
        self.th_stop = 1e-9           # SolverAbstract::th_stop_
        self.alphas = [1.0 / 2.0 ** n for n in range(10)]
        if self.th_stepInc < self.alphas[-1]:
            self.th_stepInc = self.alphas[-1]
        self.termination_tolerance = 1e-6
        self.use_filter_line_search = True
        self.filter_size = 1
        self.extra_iteration_for_last_kkt = False
        self.mu_dynamic = 1.0
        self.xs = np.zeros((B, T + 1, nx))
        self.us = np.zeros((B, T, nu))
        self.xs_try = np.zeros((B, T + 1, nx))
        self.us_try = np.zeros((B, T, nu))
        self.fs = np.zeros((B, T + 1, nx))
        self.fs_try = np.zeros((B, T + 1, nx))
        self.is_feasible = False
        self.preg = np.full(B, self.reg_min)
        self.dreg = np.full(B, self.reg_min)
        self.iter = np.zeros(B, dtype=np.int64)
        self.cost = np.zeros(B)
        self.cost_try = np.zeros(B)
        self.merit = np.zeros(B)
        self.merit_try = np.zeros(B)
        self.gap_norm = np.zeros(B)
        self.gap_norm_try = np.zeros(B)
        self.KKT = np.full(B, np.inf)
        self.x_grad_norm = np.zeros(B)
        self.u_grad_norm = np.zeros(B)
        self.steplength = np.ones(B)  # SolverAbstract::steplength_ (1.)
        self.callbacks = []
        self._engine = None
        self._batch = None

    # ---- crocoddyl::SolverAbstract plumbing ------------------------------------------------
    def setCandidate(self, xs=None, us=None, is_feasible=False):
        """SolverAbstract::setCandidate: copies (or zero / x0 defaults) and sets is_feasible_."""
        p = self.problem
        if xs is None or len(xs) == 0:
            self.xs = np.repeat(p.x0[:, None, :], self.T + 1, axis=1).copy()
        else:
            xs = np.asarray(xs, dtype=np.float64)
            if xs.shape != self.xs.shape:
                raise ValueError(f"Invalid argument: xs has wrong dimension (it should be {self.xs.shape})")
            self.xs = xs.copy()
        if us is None or len(us) == 0:
            self.us = np.zeros_like(self.us)
        else:
            us = np.asarray(us, dtype=np.float64)
            if us.shape != self.us.shape:
                raise ValueError(f"Invalid argument: us has wrong dimension (it should be {self.us.shape})")
            self.us = us.copy()
        self.is_feasible = bool(is_feasible)

    def setCallbacks(self, callbacks):
        self.callbacks = list(callbacks)

    def getCallbacks(self):
        return self.callbacks

    def increaseRegularization(self, mask=None):
        """src/ddp.cpp:431-437 (per problem)."""
        m = slice(None) if mask is None else mask
        self.preg[m] = np.minimum(self.preg[m] * self.reg_incFactor, self.reg_max)
        self.dreg[m] = self.preg[m]

    def decreaseRegularization(self, mask=None):
        """src/ddp.cpp:439-445."""
        m = slice(None) if mask is None else mask
        self.preg[m] = np.maximum(self.preg[m] / self.reg_decFactor, self.reg_min)
        self.dreg[m] = self.preg[m]

    # ---- derivatives -> device -------------------------------------------------------------
    def _calc_diff(self):
        """problem_->calc + calcDiff + computeDynamicFeasibility (src/ddp.cpp:253-263): fills the
        gaps fs_[t+1] = xnext_t (-) xs_[t+1], fs_[0] = x0 (-) xs_[0], uploads the batch."""
        p = self.problem
        c = p.calc(self.xs, self.us)
        batch = p.calcDiff(self.xs, self.us)
        self.cost = np.asarray(c["cost"], dtype=np.float64).copy()
        self.fs[:, 0] = p.diff(self.xs[:, 0], p.x0)
        self.fs[:, 1:] = p.diff(self.xs[:, 1:], c["xnext"])
        batch.arrays["fs"][...] = self.fs
        batch.arrays["cost"][...] = self.cost
        self._batch = batch
        if self._engine is None:
            self._engine = self._mk_engine(batch)
        else:
            self._engine.upload(batch)
        self._calc = c
        return batch

    def _field(self, name):
        return self._engine.field(name)

    # solver members the reference exposes (bindings/ddp.cpp:149-193)
    Vxx = property(lambda s: s._field("Vxx"))
    Vx = property(lambda s: s._field("Vx"))
    K = property(lambda s: s._field("K"))
    k = property(lambda s: s._field("k"))
    Qu = property(lambda s: s._field("Qu"))
    dx = property(lambda s: s._field("dx"))
    du = property(lambda s: s._field("du"))
    lag_mul = property(lambda s: s._field("lag_mul"))

    @property
    def status(self):
        return np.asarray(self._engine.status)

    def _set_active(self, mask):
        idx = np.flatnonzero(mask)
        if len(idx) == self.B:
            self._engine.set_active(None)
        else:
            self._engine.set_active(idx)
        return idx

    def _direction_with_retry(self, running, recalc, direction_call):
        """`while (true) { try { computeDirection } catch { increaseRegularization; if (preg_ ==
        reg_max_) return false; } }` (src/sqp.cpp:103-116) per problem: problems whose status word
        reports a failed LLT / non-finite value are retried alone at 10x the regularisation.
        Returns the mask of problems that gave up (preg hit reg_max)."""
        gave_up = np.zeros(self.B, dtype=bool)
        todo = running.copy()
        if recalc:
            self._calc_diff()
        while todo.any():
            self._set_active(todo)
            direction_call()
            st = self.status
            failed = todo & (st != 0)
            if not failed.any():
                break
            self.increaseRegularization(failed)
            hit = failed & (self.preg >= self.reg_max)
            gave_up |= hit
            todo = failed & ~hit
        self._engine.set_active(None)
        return gave_up

    def _filter_accepts(self, lists, tries, iters):
        """Filter line-search test (src/sqp.cpp:148-158): rejected iff some remembered point is
        <= the trial in EVERY criterion; looks at min(filter_size, iter+1) entries."""
        worse = np.zeros(self.B, dtype=bool)
        for count in range(self.filter_size):
            use = (count <= iters) & ~worse
            dom = np.ones(self.B, dtype=bool)
            for lst, tr in zip(lists, tries):
                dom &= lst.recent(count) <= tr
            worse |= use & dom
        return ~worse


# ------------------------------------------------------------------------------ SolverSQP
class SolverSQP(_SolverBase):
    """Batched SolverSQP (include/mim_solvers/sqp.hpp:39-191)."""
    _solver = "sqp"

    def _params(self):
        return _capi.default_sqp_params(preg=self.reg_min, dreg=self.reg_min, mu_dynamic=self.mu_dynamic,
                                        is_feasible=False)

    def computeDirection(self, recalcDiff=True):
        """src/sqp.cpp:224-242 for every problem of the batch: one CUDA launch."""
        if recalcDiff or self._engine is None:
            self._calc_diff()
        self._engine.set_reg(self.preg, self.dreg)
        self._engine.sqp_direction(self._params(), with_kkt=True)
        self._pull_scalars(np.ones(self.B, dtype=bool))

    def _pull_scalars(self, mask):
        e = self._engine
        for name in ("KKT", "gap_norm", "merit", "x_grad_norm", "u_grad_norm"):
            getattr(self, name)[mask] = e.scalar(name)[mask]

    def checkKKTConditions(self):
        """src/sqp.cpp:244-268: computed by the same launch as the direction."""
        return self.KKT

    def tryStep(self, steplength, mask=None):
        """src/sqp.cpp:307-374, for the problems of `mask`; returns merit_try (NaN where the step
        raised "step_error")."""
        if np.any(np.asarray(steplength) > 1.0) or np.any(np.asarray(steplength) < 0.0):
            raise ValueError("Invalid argument: invalid step length, value is between 0. to 1.")
        p = self.problem
        m = np.ones(self.B, dtype=bool) if mask is None else mask
        a = np.broadcast_to(np.asarray(steplength, dtype=np.float64), (self.B,))[:, None, None]
        dx, du = self._dx_cache, self._du_cache
        xs_try = p.integrate(self.xs, a * dx)
        us_try = self.us + a * du
        c = p.calc(xs_try, us_try)
        fs_try = p.diff(xs_try[:, 1:], c["xnext"])
        gap = np.abs(fs_try).sum(axis=(1, 2))
        bad = _raise_if_nan(c["cost"])
        self.xs_try[m], self.us_try[m] = xs_try[m], us_try[m]
        self.fs_try[m, 1:] = fs_try[m]
        self.cost_try[m] = c["cost"][m]
        self.gap_norm_try[m] = gap[m]
        self.merit_try[m] = (c["cost"] + self.mu_dynamic * gap)[m]
        return np.where(bad, np.nan, self.merit_try), bad

    def _iteration(self, running):
        gave_up = self._direction_with_retry(
            running, True, lambda: (self._engine.set_reg(self.preg, self.dreg),
                                    self._engine.sqp_direction(self._params(), with_kkt=True)))
        self._pull_scalars(running)
        for cb in self.callbacks:
            cb(self, "SQP")
        return gave_up

    def solve(self, init_xs=None, init_us=None, maxiter=100, is_feasible=False, reginit=float("nan")):
        """src/sqp.cpp:75-222 per problem.  Returns converged [B] (KKT_ <= termination_tol_)."""
        p, B = self.problem, self.B
        self.setCandidate(init_xs, init_us, False)
        self.xs[:, 0] = p.x0
        self.xs_try[:, 0] = p.x0
        reg0 = self.reg_min if np.isnan(reginit) else float(reginit)
        self.preg[:] = reg0
        self.dreg[:] = reg0
        converged = np.zeros(B, dtype=bool)
        running = np.ones(B, dtype=bool)
        self.iter[:] = 0
        gaps, costs = _Filter(B, self.filter_size), _Filter(B, self.filter_size)
        it = 0
        while it < maxiter and running.any():
            gave_up = self._iteration(running)
            running &= ~gave_up
            done = running & (self.KKT <= self.termination_tolerance)
            converged |= done
            running &= ~done
            if not running.any():
                break
            gaps.push(running, self.gap_norm)
            costs.push(running, self.cost)
            self._dx_cache, self._du_cache = self._engine.field("dx"), self._engine.field("du")
            searching = running.copy()
            for alpha in self.alphas:
                if not searching.any():
                    break
                self.steplength[searching] = alpha
                merit_try, bad = self.tryStep(alpha, searching)
                if self.use_filter_line_search:
                    ok = self._filter_accepts((costs, gaps), (self.cost_try, self.gap_norm_try), self.iter)
                else:
                    ok = self.merit > self.merit_try
                acc = searching & ok & ~bad
                self.xs[acc], self.us[acc] = self.xs_try[acc], self.us_try[acc]
                searching &= ~acc
            dec = running & (self.steplength > self.th_stepDec)
            self.decreaseRegularization(dec)
            inc = running & (self.steplength <= self.th_stepInc)
            self.increaseRegularization(inc)
            running &= ~(inc & (self.preg >= self.reg_max))
            it += 1
            self.iter[running] = it
        if self.extra_iteration_for_last_kkt and running.any():
            gave_up = self._iteration(running)
            running &= ~gave_up
            converged |= running & (self.KKT <= self.termination_tolerance)
        return converged


    def solve_on_device(self, init_xs=None, init_us=None, maxiter=100, is_feasible=False, reginit=float("nan"),
                        device=0):
        """The same solve(), with the whole loop on the GPU (mimgpu_sqp_solve): model calc / calcDiff,
        computeDirection, KKT test, filter / merit line search and the regularisation schedule run there for
        every problem; nothing but two counters per iteration crosses PCIe.  Needs a PolyBatchProblem (an
        analytic model the device can evaluate).  Returns converged [B]."""
        from .engine import BatchEngine
        p = self.problem
        if not isinstance(p, PolyBatchProblem):
            raise TypeError("solve_on_device needs an analytic model (PolyBatchProblem)")
        if getattr(self, "_dev_engine", None) is None:
            shape = _capi.Shape(p.B, p.T, p.nx, p.nu, 0, 0)
            self._dev_engine = BatchEngine(shape=shape, device=device)
            self._dev_engine.attach_model(p.model_arrays(), p.c3, p.w4)
        e = self._dev_engine
        sp = _capi.default_sqp_solve_params(
            maxiter=maxiter, use_filter_line_search=int(self.use_filter_line_search), filter_size=self.filter_size,
            extra_iteration_for_last_kkt=int(self.extra_iteration_for_last_kkt), reg_init=reginit,
            reg_min=self.reg_min, reg_max=self.reg_max, reg_incfactor=self.reg_incFactor,
            reg_decfactor=self.reg_decFactor, th_stepdec=self.th_stepDec, th_stepinc=self.th_stepInc,
            termination_tol=self.termination_tolerance, mu_dynamic=self.mu_dynamic)
        self.n_device_iters = e.sqp_solve(sp, init_xs, init_us)
        self.xs, self.us = e.field("xs"), e.field("us")
        it, outcome = e.solve_info()
        self.iter[:] = it
        for name, attr in (("KKT", "KKT"), ("cost", "cost"), ("gap_norm", "gap_norm"), ("merit", "merit"),
                           ("steplength", "steplength"), ("preg", "preg"), ("preg", "dreg"),
                           ("x_grad_norm", "x_grad_norm"), ("u_grad_norm", "u_grad_norm")):
            getattr(self, attr)[:] = e.scalar(name)
        self._engine = e
        return outcome == 1


# ------------------------------------------------------------------------------ SolverFDDP
class SolverFDDP(_SolverBase):
    """Batched SolverFDDP direction + expected improvement (src/fddp.cpp:189-279) and the
    solve() loop with the nonlinear roll-out (src/fddp.cpp:57-187, 281-383)."""
    _solver = "fddp"

    def __init__(self, problem, engine_factory=None):
        super().__init__(problem, engine_factory)
        self.th_acceptNegStep = 2.0
        self.dg = np.zeros(self.B)
        self.dq = np.zeros(self.B)
        self.dv = np.zeros(self.B)
        self.dV = np.zeros(self.B)
        self.dVexp = np.zeros(self.B)

    def _params(self):
        return _capi.default_sqp_params(preg=self.reg_min, dreg=self.reg_min, mu_dynamic=self.mu_dynamic,
                                        is_feasible=self.is_feasible)

    def computeDirection(self, recalcDiff=True):
        """src/fddp.cpp:189-204 + updateExpectedImprovement (:256-279) in one launch."""
        if recalcDiff or self._engine is None:
            self._calc_diff()
        self._engine.set_reg(self.preg, self.dreg)
        self._engine.fddp_direction(self._params())
        self._pull_scalars(np.ones(self.B, dtype=bool))

    def _pull_scalars(self, mask):
        e = self._engine
        for name in ("gap_norm", "dg", "dq"):
            getattr(self, name)[mask] = e.scalar(name)[mask]

    def updateExpectedImprovement(self):
        return self.dg, self.dq

    def expectedImprovement(self):
        """src/fddp.cpp:229-254: d = [dg + dv, dq - 2 dv], dv = -sum fs_t . Vxx_t (xs_t (-) xs_try_t)."""
        self.dv[:] = 0.0
        if not self.is_feasible:
            d = self.problem.diff(self.xs_try, self.xs)
            Vxx = np.swapaxes(self._Vxx_cache, -1, -2)
            self.dv = -np.einsum("bti,btij,btj->b", self.fs, Vxx, d)
        return np.stack([self.dg + self.dv, self.dq - 2.0 * self.dv], axis=1)

    def checkKKTConditions(self):
        """src/fddp.cpp:206-227 (same residual as src/sqp.cpp:244-268) on the host from lag_mul_."""
        b = self._batch
        lam = self._lag_mul
        Fx, Fu = b.mat("Fx"), b.mat("Fu")
        rx = b.arrays["Lx"] + np.einsum("btji,btj->bti", Fx, lam[:, 1:]) - lam[:, :-1]
        ru = b.arrays["Lu"] + np.einsum("btji,btj->bti", Fu, lam[:, 1:])
        k = np.maximum(np.abs(rx).max(axis=(1, 2)), np.abs(ru).max(axis=(1, 2)) if ru.size else 0.0)
        k = np.maximum(k, np.abs(b.arrays["LxT"] - lam[:, -1]).max(axis=1))
        self.KKT = np.maximum(k, np.abs(self.fs).max(axis=(1, 2)))
        return self.KKT

    def forwardPass(self, steplength, mask=None):
        """src/fddp.cpp:281-383: nonlinear roll-out with the feedback gains.  Only the problems of
        `mask` commit their members (lag_mul_, dx_, xs_try_, us_try_, fs_try_): a problem that has
        already accepted a step must not be rolled again by its neighbours' smaller alphas."""
        p, T = self.problem, self.T
        m = np.ones(self.B, dtype=bool) if mask is None else mask
        K, kff = self._K_cache, self._k_cache
        Vxx = np.swapaxes(self._Vxx_cache, -1, -2)
        closed = self.is_feasible or steplength == 1
        xnext = p.x0.copy()
        cost = np.zeros(self.B)
        gap = np.zeros(self.B)
        dxs, lag = self._dx_roll.copy(), self._lag_mul.copy()
        xs_try, us_try, fs_try_all = self.xs_try.copy(), self.us_try.copy(), self.fs_try.copy()
        bad = np.zeros(self.B, dtype=bool)
        for t in range(T):
            shift = dxs[:, t] if closed else dxs[:, t] - self.fs[:, t]
            lag[:, t] = np.einsum("bij,bj->bi", Vxx[:, t], shift) + self._Vx_cache[:, t]
            xs_try[:, t] = xnext if closed else p.integrate(xnext, self.fs[:, t] * (steplength - 1))
            dxs[:, t] = p.diff(self.xs[:, t], xs_try[:, t])
            us_try[:, t] = self.us[:, t] - kff[:, t] * steplength - np.einsum("bij,bj->bi", K[:, t], dxs[:, t])
            xnext, c = p.calc_knot(t, xs_try[:, t], us_try[:, t])
            cost += c
            fs_try = p.diff(xs_try[:, t], xnext)
            fs_try_all[:, t] = fs_try
            gap += np.abs(fs_try).sum(axis=1)
            bad |= _raise_if_nan(cost) | _raise_if_nan(np.abs(xnext).max(axis=1))
        shift = dxs[:, T] if closed else dxs[:, T] - self.fs[:, T]
        lag[:, T] = np.einsum("bij,bj->bi", Vxx[:, T], shift) + self._Vx_cache[:, T]
        xs_try[:, T] = xnext if closed else p.integrate(xnext, self.fs[:, T] * (steplength - 1))
        cost += p.calc_terminal(xs_try[:, T])
        bad |= _raise_if_nan(cost)
        self._dx_roll[m], self._lag_mul[m] = dxs[m], lag[m]
        self.xs_try[m], self.us_try[m], self.fs_try[m] = xs_try[m], us_try[m], fs_try_all[m]
        return cost, gap, bad

    def solve(self, init_xs=None, init_us=None, maxiter=100, is_feasible=False, reginit=float("nan")):
        """src/fddp.cpp:57-187 with the batch stepping in lock-step (one shared is_feasible flag,
        as all problems start from the same kind of guess)."""
        p, B = self.problem, self.B
        self.xs_try[:, 0] = p.x0
        self.setCandidate(init_xs, init_us, is_feasible)
        reg0 = self.reg_min if np.isnan(reginit) else float(reginit)
        self.preg[:] = reg0
        self.dreg[:] = reg0
        converged = np.zeros(B, dtype=bool)
        running = np.ones(B, dtype=bool)
        gaps, costs = _Filter(B, self.filter_size), _Filter(B, self.filter_size)
        self._lag_mul = np.zeros((B, self.T + 1, p.nx))
        self._dx_roll = np.zeros((B, self.T + 1, p.nx))
        self.iter[:] = 0
        recalc = True
        for it in range(maxiter):
            if not running.any():
                break
            gave_up = self._direction_with_retry(
                running, recalc, lambda: (self._engine.set_reg(self.preg, self.dreg),
                                          self._engine.fddp_direction(self._params())))
            running &= ~gave_up
            self._pull_scalars(running)
            self._Vxx_cache, self._Vx_cache = self._field("Vxx"), self._field("Vx")
            self._K_cache, self._k_cache = self._field("K"), self._field("k")
            self.checkKKTConditions()
            done = running & (self.KKT <= self.termination_tolerance)
            converged |= done
            running &= ~done
            if not running.any():
                break
            gaps.push(running, self.gap_norm)
            costs.push(running, self.cost)
            searching = running.copy()
            xs_new, us_new = self.xs.copy(), self.us.copy()
            for alpha in self.alphas:
                if not searching.any():
                    break
                self.steplength[searching] = alpha
                cost_try, gap_try, bad = self.forwardPass(alpha, searching)
                self.cost_try[searching], self.gap_norm_try[searching] = cost_try[searching], gap_try[searching]
                self.dV = self.cost - cost_try
                if self.use_filter_line_search:
                    ok = self._filter_accepts((costs, gaps), (cost_try, gap_try), self.iter)
                else:
                    d = self.expectedImprovement()
                    self.dVexp = alpha * (d[:, 0] + 0.5 * alpha * d[:, 1])
                    ok = np.where(self.dVexp >= 0,
                                  (d[:, 0] < self.th_grad) | (self.dV > self.th_acceptStep * self.dVexp),
                                  self.dV > self.th_acceptNegStep * self.dVexp)
                acc = searching & ok & ~bad
                xs_new[acc], us_new[acc] = self.xs_try[acc], self.us_try[acc]
                searching &= ~acc
            self.xs, self.us = xs_new, us_new
            self.is_feasible = False
            recalc = True
            dec = running & (self.steplength > self.th_stepDec)
            self.decreaseRegularization(dec)
            inc = running & (self.steplength <= self.th_stepInc)
            self.increaseRegularization(inc)
            running &= ~(inc & (self.preg >= self.reg_max))
            self.iter[running] = it + 1
            for cb in self.callbacks:
                cb(self, "SQP")
        return converged


# ------------------------------------------------------------------------------ SolverCSQP
class SolverCSQP(_SolverBase):
    """Batched SolverCSQP (include/mim_solvers/csqp.hpp:41-357)."""
    _solver = "csqp"

    def __init__(self, problem, engine_factory=None):
        super().__init__(problem, engine_factory)
        B = self.B
        # defaults: include/mim_solvers/csqp.hpp:292-331
        self.sigma = 1e-6
        self.alpha = 1.6
        self.rho_sparse_base = 1e-1      # property "rho_sparse" of the bindings sets the base value
        self.rho_min = 1e-6
        self.rho_max = 1e3
        self.rho_update_interval = 25
        self.adaptive_rho_tolerance = 5.0
        self.eps_abs = 1e-4
        self.eps_rel = 1e-4
        self.max_qp_iters = 1000
        self.equality_qp_initial_guess = True
        self.reset_y = False
        self.reset_rho = False
        self.remove_reg = False
        self.update_rho_with_heuristic = False
        self.with_qp_callbacks = False
        self.mu_constraint = 1e1
        self.mu_dynamic = 1e1
        self.lag_mul_inf_norm_coef = 10.0
        self.max_solve_time = float("inf")
        self.max_solve_time_reached = False
        self.constraint_norm = np.zeros(B)
        self.constraint_norm_try = np.zeros(B)
        self.lag_mul_inf_norm = np.zeros(B)
        self.qp_iters = np.zeros(B, dtype=np.int64)
        self.norm_primal = np.zeros(B)
        self.norm_dual = np.zeros(B)
        self.norm_primal_rel = np.zeros(B)
        self.norm_dual_rel = np.zeros(B)

    rho_sparse = property(lambda s: s._engine.scalar("rho_sparse") if s._engine else np.full(s.B, s.rho_sparse_base))
    dx_tilde = property(lambda s: s._field("dx_tilde"))
    du_tilde = property(lambda s: s._field("du_tilde"))
    y = property(lambda s: s._field("y"))
    z = property(lambda s: s._field("z"))
    rho_vec = property(lambda s: s._field("rho_vec"))

    def _params(self):
        reg = 0.0 if self.remove_reg else self.reg_min
        return _capi.default_csqp_params(
            preg=reg, dreg=reg, sigma=self.sigma, alpha=self.alpha, rho_sparse_base=self.rho_sparse_base,
            rho_min=self.rho_min, rho_max=self.rho_max, adaptive_rho_tolerance=self.adaptive_rho_tolerance,
            eps_abs=self.eps_abs, eps_rel=self.eps_rel, max_qp_iters=self.max_qp_iters,
            rho_update_interval=self.rho_update_interval,
            equality_qp_initial_guess=int(self.equality_qp_initial_guess), reset_y=int(self.reset_y),
            reset_rho=int(self.reset_rho), update_rho_with_heuristic=int(self.update_rho_with_heuristic))

    def calc(self, recalc=True):
        """src/csqp.cpp:411-452: derivatives (upstream) + gap_norm_ / constraint_norm_ (device)."""
        if recalc or self._engine is None:
            self._calc_diff()
        self._engine.csqp_calc_norms()
        self.gap_norm = self._engine.scalar("gap_norm")
        self.constraint_norm = self._engine.scalar("constraint_norm")

    def reset_rho_vec(self):
        """src/csqp.cpp:534-537."""
        self._engine.csqp_reset_rho(self._params())

    def computeDirection(self, recalcDiff=True):
        """src/csqp.cpp:454-517: the whole ADMM loop of every problem in one CUDA launch."""
        if recalcDiff and self._engine is None:
            self.calc(True)
        if not self.remove_reg:
            self._engine.set_reg(self.preg, self.dreg)
        else:
            self._engine.set_reg(None, None)
        self._engine.csqp_direction(self._params(), with_kkt=True)
        self._pull_scalars(np.ones(self.B, dtype=bool))

    def _pull_scalars(self, mask):
        e = self._engine
        for name in ("KKT", "x_grad_norm", "u_grad_norm", "norm_primal", "norm_dual", "norm_primal_rel",
                     "norm_dual_rel"):
            getattr(self, name)[mask] = e.scalar(name)[mask]
        self.qp_iters[mask] = np.asarray(e.qp_iters)[mask]

    def checkKKTConditions(self):
        """src/csqp.cpp:584-628: computed at the end of the direction launch (with_kkt)."""
        return self.KKT

    def tryStep(self, steplength, mask=None):
        """src/csqp.cpp:1342-1435."""
        if np.any(np.asarray(steplength) > 1.0) or np.any(np.asarray(steplength) < 0.0):
            raise ValueError("Invalid argument: invalid step length, value is between 0. to 1.")
        p = self.problem
        m = np.ones(self.B, dtype=bool) if mask is None else mask
        a = float(steplength)
        xs_try = p.integrate(self.xs, a * self._dx_cache)
        us_try = self.us + a * self._du_cache
        c = p.calc(xs_try, us_try)
        fs_try = p.diff(xs_try[:, 1:], c["xnext"])
        gap = np.abs(fs_try).sum(axis=(1, 2))
        con = self._constraint_norm(c)
        bad = _raise_if_nan(c["cost"])
        if self.mu_dynamic < 0.0 or self.mu_constraint < 0.0:
            merit = c["cost"] + self.lag_mul_inf_norm_coef * self.lag_mul_inf_norm * (gap + con)
        else:
            merit = c["cost"] + self.mu_dynamic * gap + self.mu_constraint * con
        self.xs_try[m], self.us_try[m] = xs_try[m], us_try[m]
        self.fs_try[m, 1:] = fs_try[m]
        self.cost_try[m], self.gap_norm_try[m] = c["cost"][m], gap[m]
        self.constraint_norm_try[m], self.merit_try[m] = con[m], merit[m]
        return np.where(bad, np.nan, self.merit_try), bad

    def _constraint_norm(self, c):
        con = np.zeros(self.B)
        b = self._batch
        if self.problem.ng:
            con += np.maximum(b.arrays["g_lb"] - c["g"], 0.0).sum(axis=(1, 2))
            con += np.maximum(c["g"] - b.arrays["g_ub"], 0.0).sum(axis=(1, 2))
        if self.problem.ngT:
            con += np.maximum(b.arrays["g_lbT"] - c["gT"], 0.0).sum(axis=1)
            con += np.maximum(c["gT"] - b.arrays["g_ubT"], 0.0).sum(axis=1)
        return con

    def _qp(self, running):
        if self.remove_reg:
            self._set_active(running)
            self._engine.set_reg(None, None)   # a ctx that earlier got per-problem reg arrays must not keep them
            self._engine.csqp_direction(self._params(), with_kkt=True)
            self._engine.set_active(None)
            return np.zeros(self.B, dtype=bool)
        return self._direction_with_retry(
            running, False, lambda: (self._engine.set_reg(self.preg, self.dreg),
                                     self._engine.csqp_direction(self._params(), with_kkt=True)))

    def solve(self, init_xs=None, init_us=None, maxiter=100, is_feasible=False, reginit=float("nan")):
        """src/csqp.cpp:197-409 per problem."""
        p, B = self.problem, self.B
        self.setCandidate(init_xs, init_us, False)
        self.xs[:, 0] = p.x0
        self.xs_try[:, 0] = p.x0
        if self.remove_reg:
            self.preg[:] = 0.0
            self.dreg[:] = 0.0
        else:
            reg0 = self.reg_min if np.isnan(reginit) else float(reginit)
            self.preg[:] = reg0
            self.dreg[:] = reg0
        if maxiter == 0:
            self.calc(True)
            self.reset_rho_vec()
        self.max_solve_time_reached = False
        converged = np.zeros(B, dtype=bool)
        running = np.ones(B, dtype=bool)
        self.iter[:] = 0
        gaps, costs, cons = (_Filter(B, self.filter_size) for _ in range(3))
        it = 0
        while it < maxiter and running.any():
            self.calc(True)
            if it == 0 and not self.reset_rho:
                self.reset_rho_vec()
            gave_up = self._qp(running)
            running &= ~gave_up
            self._pull_scalars(running)
            running &= ~(self.qp_iters == 0)
            for cb in self.callbacks:
                cb(self, "CSQP")
            done = running & (self.KKT <= self.termination_tolerance)
            converged |= done
            running &= ~done
            if not running.any():
                break
            cons.push(running, self.constraint_norm)
            gaps.push(running, self.gap_norm)
            costs.push(running, self.cost)
            # dx_/du_ hold dxtilde/dutilde after update_lagrangian_parameters (src/csqp.cpp:1241-1242)
            self._dx_cache, self._du_cache = self._engine.field("dx"), self._engine.field("du")
            if self.mu_dynamic < 0.0 or self.mu_constraint < 0.0:
                lam = np.abs(self._engine.field("lag_mul")).max(axis=(1, 2))
                yy = self._engine.field("y")
                self.lag_mul_inf_norm = np.maximum(lam, np.abs(yy).max(axis=1) if yy.size else 0.0)
                self.merit = self.cost + self.lag_mul_inf_norm_coef * self.lag_mul_inf_norm * (
                    self.gap_norm + self.constraint_norm)
            else:
                self.merit = self.cost + self.mu_dynamic * self.gap_norm + self.mu_constraint * self.constraint_norm
            searching = running.copy()
            for alpha in self.alphas:
                if not searching.any():
                    break
                merit_try, bad = self.tryStep(alpha, searching)
                if self.use_filter_line_search:
                    ok = self._filter_accepts((costs, gaps, cons),
                                              (self.cost_try, self.gap_norm_try, self.constraint_norm_try), self.iter)
                else:
                    ok = self.merit > self.merit_try
                acc = searching & ok & ~bad
                self.xs[acc], self.us[acc] = self.xs_try[acc], self.us_try[acc]
                searching &= ~acc
            running &= ~searching                      # `if (!found) break;`
            if not self.remove_reg:
                # The reference's loop variable shadows the member steplength_ (src/csqp.cpp:317), so
                # the schedule below always sees the member (1.0): regularisation only decreases.
                dec = running & (self.steplength > self.th_stepDec)
                self.decreaseRegularization(dec)
                inc = running & ~dec
                self.increaseRegularization(inc)
                running &= ~(inc & (self.preg >= self.reg_max))
            it += 1
            self.iter[running] = it
        if self.extra_iteration_for_last_kkt and running.any() and not self.max_solve_time_reached:
            self.calc(True)
            gave_up = self._qp(running)
            running &= ~gave_up
            self._pull_scalars(running)
            converged |= running & (self.KKT <= self.termination_tolerance)
        return converged


__all__ = ["BatchShootingProblem", "LQBatchProblem", "SolverSQP", "SolverCSQP", "SolverFDDP"]
