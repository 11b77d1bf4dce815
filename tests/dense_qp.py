"""Algorithm-independent cross-check of CONVERGED CSQP solutions: the constrained QP of
one shooting problem as ONE dense QP in the layout of the reference's ProxQP back-end
(src/csqp_proxqp.cpp:296-345):

  z = [dx_1 .. dx_T | du_0 .. du_{T-1}]           (dx_0 = 0 is eliminated)
  min 1/2 z^T P z + q^T z   s.t.  A z = b,  l <= C z <= u
  A row block t:  dx_{t+1} - Fx_t dx_t - Fu_t du_t = fs_{t+1}
  C row block t:  Gx_t dx_t + Gu_t du_t  (no Gx at t = 0),  l = g_lb - g, u = g_ub - g;  terminal: GxT dx_T

The check does not use ADMM: from the solver's (z, y) it reads an active set, solves the equality-constrained
KKT system on that set with numpy, and accepts the result only with a full optimality CERTIFICATE of the
convex QP (primal feasibility of every row, multiplier signs, stationarity).  The certified optimum is then
compared with the solver's dx~, du~, y and dynamics multipliers."""
import numpy as np


def build_proxqp_layout(batch, b):
    T, nx, nu, ng, ngT = batch.T, batch.nx, batch.nu, batch.ng, batch.ngT
    M = lambda n: batch.mat(n)[b]
    V = lambda n: batch.arrays[n][b]
    nX, nU = T * nx, T * nu
    n = nX + nU
    xs = lambda t: slice((t - 1) * nx, t * nx)            # dx_t, t = 1..T
    us = lambda t: slice(nX + t * nu, nX + (t + 1) * nu)  # du_t, t = 0..T-1
    P = np.zeros((n, n)); q = np.zeros(n)
    A = np.zeros((nX, n)); rhs = np.zeros(nX)
    nin = T * ng + ngT
    Cm = np.zeros((nin, n)); lo = np.full(nin, -np.inf); hi = np.full(nin, np.inf)
    Lxx, Lxu, Luu, Fx, Fu = M("Lxx"), M("Lxu"), M("Luu"), M("Fx"), M("Fu")
    for t in range(T):
        if t > 0:
            P[xs(t), xs(t)] = Lxx[t]; P[xs(t), us(t)] = Lxu[t]; P[us(t), xs(t)] = Lxu[t].T
            q[xs(t)] = V("Lx")[t]
            A[t * nx:(t + 1) * nx, xs(t)] = -Fx[t]
        P[us(t), us(t)] = Luu[t]; q[us(t)] = V("Lu")[t]
        A[t * nx:(t + 1) * nx, us(t)] = -Fu[t]
        A[t * nx:(t + 1) * nx, xs(t + 1)] = np.eye(nx)
        rhs[t * nx:(t + 1) * nx] = V("fs")[t + 1]
        if ng:
            r = slice(t * ng, (t + 1) * ng)
            if t > 0:
                Cm[r, xs(t)] = M("Gx")[t]
            Cm[r, us(t)] = M("Gu")[t]
            lo[r] = V("g_lb")[t] - V("g")[t]; hi[r] = V("g_ub")[t] - V("g")[t]
    P[xs(T), xs(T)] = M("LxxT"); q[xs(T)] = V("LxT")
    if ngT:
        r = slice(T * ng, T * ng + ngT)
        Cm[r, xs(T)] = M("GxT")
        lo[r] = V("g_lbT") - V("gT"); hi[r] = V("g_ubT") - V("gT")
    return P, q, A, rhs, Cm, lo, hi


def certified_optimum(P, q, A, rhs, Cm, lo, hi, z0, y0, act_tol=1e-6, feas_tol=1e-8):
    """Active set read off (z0, y0) -> equality-constrained KKT solve -> optimality certificate.
    Returns (z, nu, y) of the dense QP, raises AssertionError when the certificate fails."""
    n, ne = P.shape[0], A.shape[0]
    c0 = Cm @ z0
    eq = np.isfinite(lo) & np.isfinite(hi) & (hi - lo <= 1e-12)
    at_hi = ((np.abs(c0 - hi) <= act_tol) & (y0 > act_tol)) | eq
    at_lo = ((np.abs(c0 - lo) <= act_tol) & (y0 < -act_tol)) & ~eq
    act = at_hi | at_lo
    bound = np.where(at_hi, hi, lo)[act]
    Ca = Cm[act]
    na = Ca.shape[0]
    K = np.block([[P, A.T, Ca.T], [A, np.zeros((ne, ne)), np.zeros((ne, na))],
                  [Ca, np.zeros((na, ne)), np.zeros((na, na))]])
    sol = np.linalg.lstsq(K, np.concatenate([-q, rhs, bound]), rcond=None)[0]
    z, nu, ya = sol[:n], sol[n:n + ne], sol[n + ne:]
    y = np.zeros(Cm.shape[0]); y[act] = ya
    c = Cm @ z
    scale = max(1.0, np.abs(z).max())
    assert np.abs(A @ z - rhs).max() <= feas_tol * scale, "dynamics rows"
    assert np.all(c >= lo - feas_tol * scale) and np.all(c <= hi + feas_tol * scale), "inequality rows violated"
    assert np.all(y[at_hi & ~eq] >= -1e-9) and np.all(y[at_lo] <= 1e-9), "multiplier signs"
    assert np.abs(P @ z + q + A.T @ nu + Cm.T @ y).max() <= 1e-8 * max(1.0, np.abs(q).max()), "stationarity"
    return z, nu, y


def check_against_dense_qp(make_engine, shapes=None, tol=2e-6):
    """Converged ADMM solutions (tight eps, preg = dreg = 0) against the certified dense optimum."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    # feasible problems (violate_frac = 0: the generator's "currently violated" rows can make a QP infeasible, and
    # ADMM then never meets a tight eps): UR5-like, Solo12-like (torque boxes + dense rows), equality rows
    shapes = shapes or [(3, 6, 4, 2, 3, 2, dict(box_rows=1)), (2, 8, 12, 6, 3, 3, dict(box_rows=0)),
                        (2, 5, 8, 4, 4, 3, dict(box_rows=4)), (2, 10, 36, 12, 16, 0, dict(box_rows=12)),
                        (1, 6, 16, 8, 8, 4, dict(box_rows=8)), (2, 12, 12, 6, 6, 0, dict(box_rows=6, eq_frac=0.3))]
    checked = 0
    for (B, T, nx, nu, ng, ngT, kw) in shapes:
        batch = make_random_batch(B, T, nx, nu, ng, ngT, violate_frac=0.0, **kw)
        e = make_engine(batch)
        p = _capi.default_csqp_params(preg=0.0, dreg=0.0, eps_abs=1e-9, eps_rel=0.0, max_qp_iters=50000)
        e.csqp_calc_norms(); e.csqp_reset_rho(p)
        e.csqp_direction(p, with_kkt=True)
        conv = np.asarray(e.converged)
        dxt, dut, ys, lam = e.field("dx_tilde"), e.field("du_tilde"), e.field("y"), e.field("lag_mul")
        for b in range(B):
            if not conv[b]:
                continue
            P, q, A, rhs, Cm, lo, hi = build_proxqp_layout(batch, b)
            z0 = np.concatenate([dxt[b, 1:].ravel(), dut[b].ravel()])
            z, nu_, y = certified_optimum(P, q, A, rhs, Cm, lo, hi, z0, ys[b])
            s = max(1.0, np.abs(z).max())
            assert np.abs(z0 - z).max() <= tol * s, ("primal", (B, T, nx, nu, ng, ngT), b, np.abs(z0 - z).max())
            assert np.abs(ys[b] - y).max() <= 50 * tol * max(1.0, np.abs(y).max()), ("y", b, np.abs(ys[b] - y).max())
            # the solver's lag_mul_[t+1] is minus the multiplier of dynamics row t in this layout (A has +I on dx_{t+1})
            lam_d = -nu_.reshape(T, nx)
            err = np.abs(lam[b, 1:] - lam_d).max()
            assert err <= 50 * tol * max(1.0, np.abs(lam_d).max()), ("lag_mul", b, err)
            checked += 1
    assert checked >= 12, f"only {checked} problems converged at eps 1e-9"
    return checked
