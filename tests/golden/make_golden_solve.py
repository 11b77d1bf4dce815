#!/usr/bin/env python
"""Multi-iteration golden vectors: the reference's OWN NumPy SolverSQP.solve() loop
(/root/reference/python/sqp.py:254-340, unmodified: computeDirection, KKT_check, the filter /
merit line search over alpha = 2^-k, acceptStep) run on nonlinear shooting problems, recording
per-iteration accepted step lengths, the iteration count, the final xs / us and KKT.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_solve.py
The models are OUR analytic family (LQ plus separable polynomial terms, include/mimgpu.h
mimgpu_poly_model) written as crocoddyl-style action models on the refshim stand-in, so that the
same problems can be replayed through the batched solvers (host loop and device loop).

Differences between the NumPy prototype and the C++ solver that the replay accounts for
(SURVEY.md 8c.2): no preg / dreg in Python (replay with reg_min = reginit = 0), KKT and filter
comparisons are strict in Python (`<`) and non-strict in C++ (`<=`) -- identical except on exact
ties.
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
REF = Path("/root/reference")
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))
sys.path.insert(0, str(REPO / "oracle" / "refshim"))
sys.path.insert(1, str(REF / "python"))

import crocoddyl  # noqa: E402  (the stand-in)
from sqp import SQP  # noqa: E402  (reference)

import solver_checks as sc  # noqa: E402


class PolyKnot(crocoddyl.ActionModelAbstract):
    """x+ = A x + B u + f + c3 x^3;  l = 1/2 [x;u]^T H [x;u] + q.x + r.u + w4/4 sum x^4."""

    def __init__(self, A, Bm, f, Q, S, R, q, r, c3, w4):
        super().__init__(crocoddyl.StateVector(A.shape[0]), Bm.shape[1])
        self.A, self.Bm, self.f, self.Q, self.S, self.R, self.q, self.r = A, Bm, f, Q, S, R, q, r
        self.c3, self.w4 = c3, w4

    def calc(self, data, x, u=None):
        data.xnext = self.A @ x + self.Bm @ u + self.f + self.c3 * x**3
        data.cost = float(0.5 * x @ self.Q @ x + x @ self.S @ u + 0.5 * u @ self.R @ u + self.q @ x + self.r @ u
                          + 0.25 * self.w4 * np.sum(x**4))

    def calcDiff(self, data, x, u=None):
        data.Fx = self.A + np.diag(3.0 * self.c3 * x**2)
        data.Fu = self.Bm.copy()
        data.Lx = self.Q @ x + self.S @ u + self.q + self.w4 * x**3
        data.Lu = self.S.T @ x + self.R @ u + self.r
        data.Lxx = self.Q + np.diag(3.0 * self.w4 * x**2)
        data.Lxu = self.S.copy()
        data.Luu = self.R.copy()


class PolyTerminal(crocoddyl.ActionModelAbstract):
    def __init__(self, QT, qT, w4, nu):
        super().__init__(crocoddyl.StateVector(QT.shape[0]), nu)
        self.QT, self.qT, self.w4 = QT, qT, w4

    def calc(self, data, x, u=None):
        data.xnext = np.array(x, dtype=float)
        data.cost = float(0.5 * x @ self.QT @ x + self.qT @ x + 0.25 * self.w4 * np.sum(x**4))

    def calcDiff(self, data, x, u=None):
        data.Lx = self.QT @ x + self.qT + self.w4 * x**3
        data.Lxx = self.QT + np.diag(3.0 * self.w4 * x**2)
        data.Fx = np.eye(len(x))


CASES = {  # name -> (c3, w4, x0 scale, use_filter, filter_size, maxiter, seed)
    "solve_lq": (0.0, 0.0, 1.0, True, 1, 10, 3),
    "solve_poly_filter": (0.05, 0.5, 1.0, True, 1, 60, 3),
    "solve_poly_filter3": (0.02, 0.1, 1.0, True, 3, 60, 3),
    "solve_poly_short_steps": (0.1, 0.3, 1.0, True, 1, 12, 3),
    "solve_poly_merit": (0.0, 1.0, 2.0, False, 1, 60, 3),
    "solve_poly_merit2": (-0.05, 1.0, 1.5, False, 1, 60, 5),
}


def main():
    B, T, nx, nu = 4, 15, 6, 3
    for name, (c3, w4, scale, filt, fsize, maxiter, seed) in CASES.items():
        lq = sc.make_lq(B=B, T=T, nx=nx, nu=nu, seed=seed)
        out = dict(c3=c3, w4=w4, use_filter=int(filt), filter_size=fsize, maxiter=maxiter,
                   x0=lq.x0 * scale, A=lq.A, Bm=lq.Bm, f=lq.f, Q=lq.Q, S=lq.S, R=lq.R, q=lq.q, r=lq.r, QT=lq.QT, qT=lq.qT)
        iters, conv, xs_all, us_all, kkt, alphas = [], [], [], [], [], []
        for b in range(B):
            running = [PolyKnot(lq.A[b, t], lq.Bm[b, t], lq.f[b, t], lq.Q[b, t], lq.S[b, t], lq.R[b, t], lq.q[b, t],
                                lq.r[b, t], c3, w4) for t in range(T)]
            prob = crocoddyl.ShootingProblem(lq.x0[b] * scale, running, PolyTerminal(lq.QT[b], lq.qT[b], w4, nu))
            s = SQP(prob, use_filter_line_search=filt, with_callbacks=False)
            s.filter_size = fsize
            taken = []
            accept = s.acceptStep

            def spy(_accept=accept, _s=s, _taken=taken):   # record the step length of every accepted step
                _taken.append(float(np.max(np.abs(np.array(_s.xs_try[1]) - np.array(_s.xs[1]))) /
                                    max(1e-300, float(np.max(np.abs(_s.dx[1]))))))
                _accept()
            s.acceptStep = spy
            xs0 = [prob.x0.copy() for _ in range(T + 1)]
            us0 = [np.zeros(nu) for _ in range(T)]
            n_dir = [0]
            cd = s.computeDirection

            def count(recalc=True, _cd=cd, _n=n_dir):
                _n[0] += 1
                return _cd(recalc=recalc)
            s.computeDirection = count
            ok = bool(s.solve(xs0, us0, maxiter))
            # iter_ of the C++ solver = line searches completed = computeDirection calls - 1 when converged
            iters.append(n_dir[0] - 1 if ok else n_dir[0])
            conv.append(int(ok))
            xs_all.append(np.array(s.xs)); us_all.append(np.array(s.us)); kkt.append(float(s.KKT))
            a = np.full(maxiter, np.nan); a[:len(taken)] = taken
            alphas.append(a)
        out.update(out_iter=np.array(iters), out_converged=np.array(conv), out_xs=np.array(xs_all),
                   out_us=np.array(us_all), out_KKT=np.array(kkt), out_alphas=np.array(alphas))
        np.savez_compressed(HERE / f"{name}.npz", **out)
        print(name, "iter", iters, "converged", conv, "min alpha", np.nanmin(np.array(alphas)) if np.isfinite(np.array(alphas)).any() else None)


if __name__ == "__main__":
    main()
