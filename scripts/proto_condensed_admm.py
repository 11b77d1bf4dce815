#!/usr/bin/env python
"""Numerical prototype (numpy, CPU) of the CONDENSED light ADMM iteration used by the CUDA
CSQP kernel (kernels_admm.cuh), checked against the oracle.

Reference iteration (src/csqp.cpp:1025-1108 light backward, :630-647 forward, :1196-1340
update) is restated with the closed-loop matrix Acl_t = Fx_t - Fu_t K_t:

  backward chain   Vx_t = r_t + Acl_t^T tmp_t,  tmp_{t-1} = Vx_t + Vxx_fs_{t-1}
                   r_t  = (Lx - sigma dx + Gx^T d) - K^T (Lu - sigma du + Gu^T d)
  gains            Qu_t = (Lu - sigma du + Gu^T d) + Fu^T tmp_t,  k_t = Quu^-1 Qu_t
  forward chain    x_{t+1} = (fs_{t+1} - Fu k_t) + Acl_t x_t,  u_t = -k_t - K_t x_t

Test infrastructure only: runs the oracle for 1 iteration (one full sweep), continues with 24
condensed iterations in numpy and compares with the oracle's own 25 iterations.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from mim_solvers_b200 import _capi  # noqa: E402
from mim_solvers_b200.problems import make_config  # noqa: E402
from oracle import oracle  # noqa: E402


def run(name, n_it=25, b=0):
    batch = make_config(name, b + 1).subset([b])
    T, nx, nu, ng, ngT = batch.T, batch.nx, batch.nu, batch.ng, batch.ngT
    kw = dict(eps_abs=0.0, eps_rel=0.0, equality_qp_initial_guess=0)
    p1 = _capi.default_csqp_params(max_qp_iters=1, **kw)
    pn = _capi.default_csqp_params(max_qp_iters=n_it, **kw)
    o1, on = oracle.Oracle(batch), oracle.Oracle(batch)
    for o, p in ((o1, p1), (on, pn)):
        o.csqp_calc_norms(); o.csqp_reset_rho(p); o.csqp_direction(p, with_kkt=False)
    A = {k: batch.mat(k)[0] if k in ("Fx", "Fu", "Gx", "Gu", "GxT") else batch.arrays[k][0] for k in batch.arrays}
    Fx, Fu, Lx, Lu, fs = A["Fx"], A["Fu"], A["Lx"], A["Lu"], A["fs"]
    Gx, Gu = A["Gx"], A["Gu"]
    lo, hi = A["g_lb"] - A["g"], A["g_ub"] - A["g"]
    K = o1.field("K")[0].copy()                      # (T, nu, nx) row-major
    Lf = np.swapaxes(o1.field("Quu_llt")[0], -1, -2).copy()  # (T, nu, nu) lower
    Lf = np.tril(Lf)
    Vf = np.ctypeslib.as_array(o1._st.contents.Vxx_fs, shape=(T * nx,)).reshape(T, nx).copy()
    x = o1.field("dx_tilde")[0].copy(); u = o1.field("du_tilde")[0].copy()
    y = o1.field("y")[0].copy(); z = o1.field("z")[0].copy(); rho = o1.field("rho_vec")[0].copy()
    yr, zr, rr = (v[:T * ng].reshape(T, ng) for v in (y, z, rho))
    yT, zT, rT = y[T * ng:], z[T * ng:], rho[T * ng:]
    sigma, alpha = pn.sigma, pn.alpha
    Acl = Fx - np.einsum("tia,taj->tij", Fu, K)
    import scipy.linalg as sl
    for it in range(2, n_it + 1):
        d = yr - rr * zr
        qx = Lx - sigma * x[:-1] + np.einsum("tri,tr->ti", Gx, d) * (np.arange(T)[:, None] > 0)
        qu = Lu - sigma * u + np.einsum("tra,tr->ta", Gu, d)
        r = qx - np.einsum("tai,ta->ti", K, qu)
        VxT = A["LxT"] - sigma * x[T]
        if ngT:
            VxT = VxT + A["GxT"].T @ (yT - rT * zT)
        tmp = np.zeros((T, nx))
        Vx = VxT
        for t in range(T - 1, -1, -1):
            tmp[t] = Vf[t] + Vx
            Vx = r[t] + Acl[t].T @ tmp[t]
        Qu = qu + np.einsum("tia,ti->ta", Fu, tmp)
        k = np.stack([sl.cho_solve((Lf[t], True), Qu[t]) for t in range(T)])
        bvec = fs[1:] - np.einsum("tia,ta->ti", Fu, k)
        xn = np.zeros_like(x)
        for t in range(T):
            xn[t + 1] = bvec[t] + Acl[t] @ xn[t]
        un = -k - np.einsum("tai,ti->ta", K, xn[:-1])
        c = np.einsum("tri,ti->tr", Gx, xn[:-1]) + np.einsum("tra,ta->tr", Gu, un)
        zrel = alpha * c + (1 - alpha) * zr
        znew = np.minimum(np.maximum(zrel + yr / rr, lo), hi)
        yr = yr + rr * (zrel - znew)
        zr = znew
        if ngT:
            cT = A["GxT"] @ xn[T]
            zrel = alpha * cT + (1 - alpha) * zT
            zn = np.minimum(np.maximum(zrel + yT / rT, A["g_lbT"] - A["gT"]), A["g_ubT"] - A["gT"])
            yT = yT + rT * (zrel - zn); zT = zn
        x, u = xn, un

    def rel(a, bb):
        return np.abs(a - bb).max() / max(np.abs(bb).max(), 1e-300)
    ex = rel(x, on.field("dx_tilde")[0]); eu = rel(u, on.field("du_tilde")[0])
    ey = rel(np.concatenate([yr.ravel(), yT]), on.field("y")[0]); ez = rel(np.concatenate([zr.ravel(), zT]), on.field("z")[0])
    print(f"{name}: after {n_it} its  dx~ {ex:.2e}  du~ {eu:.2e}  y {ey:.2e}  z {ez:.2e}")
    return max(ex, eu, ey, ez)


if __name__ == "__main__":
    worst = 0.0
    for name in ("C1_ur5_csqp", "C3_solo12_csqp", "C3b_solo12_bench", "C4_clq_nx32", "C4_clq_nx64", "C5_talos_csqp"):
        for b in (0, 1):
            worst = max(worst, run(name, 25, b))
    print("worst", worst)
    assert worst < 1e-10
