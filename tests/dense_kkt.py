"""Algorithm-independent cross-check: the SQP sub-problem as one dense KKT
system, in the layout of the reference's dense solver (src/kkt.cpp:235-293: primal block
[dx_0..dx_T, du_0..du_{T-1}], one dynamics row block per knot), solved with numpy.

  min  sum_t 1/2 [dx;du]^T [[Lxx,Lxu],[Lxu^T,Luu]] [dx;du] + Lx.dx + Lu.du  + 1/2 dx_T^T Lxx_T dx_T + Lx_T.dx_T
  s.t. dx_0 = 0,  dx_{t+1} = Fx_t dx_t + Fu_t du_t + fs_{t+1}

With preg = dreg = 0 the Riccati recursion of SolverDDP::backwardPass / SolverSQP::forwardPass
solves exactly this QP, and lag_mul_ are its dynamics multipliers."""
import numpy as np


def solve_dense(batch, b):
    T, nx, nu = batch.T, batch.nx, batch.nu
    Fx, Fu = batch.mat("Fx")[b], batch.mat("Fu")[b]
    Lxx, Lxu, Luu = batch.mat("Lxx")[b], batch.mat("Lxu")[b], batch.mat("Luu")[b]
    Lx, Lu, fs = batch.arrays["Lx"][b], batch.arrays["Lu"][b], batch.arrays["fs"][b]
    nX, nU = (T + 1) * nx, T * nu
    n = nX + nU
    H = np.zeros((n, n)); g = np.zeros(n)
    A = np.zeros((nX, n)); rhs = np.zeros(nX)
    xs = lambda t: slice(t * nx, (t + 1) * nx)
    us = lambda t: slice(nX + t * nu, nX + (t + 1) * nu)
    for t in range(T):
        H[xs(t), xs(t)] = Lxx[t]; H[xs(t), us(t)] = Lxu[t]; H[us(t), xs(t)] = Lxu[t].T; H[us(t), us(t)] = Luu[t]
        g[xs(t)] = Lx[t]; g[us(t)] = Lu[t]
        # row block t+1:  Fx dx_t + Fu du_t - dx_{t+1} = -fs_{t+1}
        A[xs(t + 1), xs(t)] = Fx[t]; A[xs(t + 1), us(t)] = Fu[t]; A[xs(t + 1), xs(t + 1)] = -np.eye(nx)
        rhs[xs(t + 1)] = -fs[t + 1]
    H[xs(T), xs(T)] = batch.mat("LxxT")[b]; g[xs(T)] = batch.arrays["LxT"][b]
    A[xs(0), xs(0)] = -np.eye(nx); rhs[xs(0)] = -fs[0]      # dx_0 = fs_0 (= 0)
    K = np.block([[H, A.T], [A, np.zeros((nX, nX))]])
    sol = np.linalg.solve(K, np.concatenate([-g, rhs]))
    dx = sol[:nX].reshape(T + 1, nx); du = sol[nX:n].reshape(T, nu); lam = sol[n:].reshape(T + 1, nx)
    return dx, du, lam


def check_against_dense(make_engine, tol=1e-8):
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    for (B, T, nx, nu) in [(3, 6, 4, 2), (2, 9, 12, 6), (2, 4, 7, 3), (2, 10, 36, 12), (1, 6, 16, 8), (2, 5, 32, 16),
                           (1, 4, 24, 12), (1, 3, 64, 32), (1, 12, 8, 4)]:
        batch = make_random_batch(B, T, nx, nu)
        e = make_engine(batch)
        e.sqp_direction(_capi.default_sqp_params(preg=0.0, dreg=0.0))
        dx, du, lam = e.field("dx"), e.field("du"), e.field("lag_mul")
        for b in range(B):
            dxd, dud, lamd = solve_dense(batch, b)
            for got, ref, what in ((dx[b], dxd, "dx"), (du[b], dud, "du"), (lam[b, 1:], lamd[1:], "lag_mul")):
                err = np.abs(got - ref).max() / max(1.0, np.abs(ref).max())
                assert err <= tol, (what, (B, T, nx, nu), b, err)
