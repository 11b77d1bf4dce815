"""Engine-vs-engine parity (CUDA or emulated kernels against the CPU oracle) on the same
seeded inputs.  Tolerance: BASELINE.json north_star -- 1e-9 relative on dx/du/lag_mul and
the KKT residuals, identical iteration counts.  Every array is normalised PER PROBLEM (axis 0 = batch):
the worst problem must meet the bar, not the batch-wide maximum."""
import numpy as np

from helpers import assert_close
from mim_solvers_b200 import _capi

TOL = 1e-9


def _lower(L):
    return np.tril(np.swapaxes(L, -1, -2))  # storage is col-major: (.., col, row)


def compare_engines_sqp(make_a, make_b, batch, params=None, tol=TOL):
    a, b = make_a(batch), make_b(batch)
    p = params or _capi.default_sqp_params()
    a.sqp_direction(p, with_kkt=True)
    b.sqp_direction(p, with_kkt=True)
    assert np.array_equal(np.asarray(a.status), np.asarray(b.status))
    ok = np.asarray(b.status) == 0
    for f in ("dx", "du", "lag_mul", "Vxx", "Vx", "K", "k", "Qu", "Quuk"):
        assert_close(a.field(f)[ok], b.field(f)[ok], tol, f, per_problem=True)
    assert_close(_lower(a.field("Quu_llt")[ok]), _lower(b.field("Quu_llt")[ok]), tol, "Quu_llt", per_problem=True)
    for s in ("KKT", "gap_norm", "merit", "x_grad_norm", "u_grad_norm"):
        assert_close(a.scalar(s)[ok], b.scalar(s)[ok], tol, s, per_problem=True)
    return a, b


def compare_engines_fddp(make_a, make_b, batch, params=None, tol=TOL):
    a, b = make_a(batch), make_b(batch)
    p = params or _capi.default_sqp_params()
    a.fddp_direction(p)
    b.fddp_direction(p)
    assert np.array_equal(np.asarray(a.status), np.asarray(b.status))
    for f in ("Vxx", "Vx", "K", "k", "Qu", "Quuk"):
        assert_close(a.field(f), b.field(f), tol, f, per_problem=True)
    for s in ("gap_norm", "dg", "dq"):
        assert_close(a.scalar(s), b.scalar(s), tol, s, per_problem=True)
    return a, b


def compare_engines_csqp(make_a, make_b, batch, params=None, tol=TOL, n_calls=2, reset_rho_first=True):
    a, b = make_a(batch), make_b(batch)
    p = params or _capi.default_csqp_params()
    for e in (a, b):
        e.csqp_calc_norms()
        if reset_rho_first:
            e.csqp_reset_rho(p)
    for s in ("gap_norm", "constraint_norm"):
        assert_close(a.scalar(s), b.scalar(s), 1e-12, s)
    for c in range(n_calls):
        a.csqp_direction(p, with_kkt=True)
        b.csqp_direction(p, with_kkt=True)
        w = f"[call {c}] "
        assert np.array_equal(np.asarray(a.status), np.asarray(b.status)), w + "status"
        assert np.array_equal(np.asarray(a.qp_iters), np.asarray(b.qp_iters)), \
            w + f"qp_iters {np.asarray(a.qp_iters)} vs {np.asarray(b.qp_iters)}"
        assert np.array_equal(np.asarray(a.converged), np.asarray(b.converged)), w + "converged"
        for f in ("dx_tilde", "du_tilde", "dx", "du", "lag_mul", "y", "z", "rho_vec", "Vxx", "Vx", "K", "k"):
            assert_close(a.field(f), b.field(f), tol, w + f, per_problem=True)
        for s in ("KKT", "rho_sparse", "x_grad_norm", "u_grad_norm"):
            assert_close(a.scalar(s), b.scalar(s), tol, w + s, per_problem=True)
        for s in ("norm_primal", "norm_dual", "norm_primal_rel", "norm_dual_rel"):
            x, y = a.scalar(s), b.scalar(s)
            assert np.array_equal(np.isfinite(x), np.isfinite(y)), w + s
            fin = np.isfinite(y)
            assert np.all(np.abs(x[fin] - y[fin]) <= tol * np.maximum(1.0, np.abs(y[fin]))), (w + s, x, y)
    return a, b
