"""Shared test helpers: golden-file loading and engine-agnostic comparisons.

An "engine" is either oracle.oracle.Oracle (CPU restatement) or
mim_solvers_b200.engine.BatchEngine (CUDA, through the C ABI); both expose
sqp_direction / fddp_direction / csqp_calc_norms / csqp_reset_rho / csqp_direction,
field(name), scalar(name), qp_iters, status.
"""
from pathlib import Path

import numpy as np

from mim_solvers_b200 import _capi
from mim_solvers_b200.problems import OCPBatch

GOLDEN = Path(__file__).resolve().parent / "golden"


def golden_names(prefix):
    return sorted(p.stem for p in GOLDEN.glob(f"{prefix}*.npz"))


def load_golden(name):
    z = np.load(GOLDEN / f"{name}.npz")
    B, T, nx, nu, ng, ngT = (int(v) for v in z["shape"])
    batch = OCPBatch(B, T, nx, nu, ng, ngT)
    batch.arrays = {k[3:]: np.ascontiguousarray(z[k]) for k in z.files if k.startswith("in_")}
    params = {k[2:]: z[k][()] for k in z.files if k.startswith("p_")}
    out = {k[4:]: z[k] for k in z.files if k.startswith("out_")}
    return batch, params, out


def relerr(a, b, per_problem=False):
    """max |a - b| / max |b|.  per_problem=True: axis 0 is the batch; every problem is normalised by
    ITS OWN max |b| (north_star: 1e-9 relative per solve) and the worst problem is returned -- a
    problem whose values are 1e-3 of the batch maximum cannot hide behind its neighbours."""
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    fin = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), fin), "non-finite pattern differs"
    if not fin.all():
        assert np.array_equal(a[~fin], b[~fin])
    if not fin.any():
        return 0.0
    if per_problem and a.ndim >= 1:
        B = a.shape[0]
        a2, b2, f2 = a.reshape(B, -1), b.reshape(B, -1), fin.reshape(B, -1)
        err = np.where(f2, np.abs(np.where(f2, a2, 0.0) - np.where(f2, b2, 0.0)), 0.0).max(axis=1)
        scale = np.maximum(np.where(f2, np.abs(np.where(f2, b2, 0.0)), 0.0).max(axis=1), 1e-300)
        return float((err / scale).max())
    scale = max(np.abs(b[fin]).max(), 1e-300)
    return float(np.abs(a[fin] - b[fin]).max() / scale)


def assert_close(a, b, tol, what="", per_problem=False):
    e = relerr(a, b, per_problem)
    assert e <= tol, f"{what}: relative error {e:.3e} > {tol:.1e}" + (" (worst problem)" if per_problem else "")


def csqp_params_from_golden(P, **over):
    d = {k: P[k] for k in ("preg", "dreg", "sigma", "alpha", "rho_sparse_base", "rho_min", "rho_max",
                           "adaptive_rho_tolerance", "eps_abs", "eps_rel", "max_qp_iters",
                           "rho_update_interval", "equality_qp_initial_guess", "reset_y", "reset_rho")}
    d.update(over)
    return _capi.default_csqp_params(**d)


def real_row_mask(batch, nrows):
    """Mask over the [T*ng+ngT] constraint vector selecting rows that exist in the
    reference problem (knots with fewer rows were padded with unbounded zero rows)."""
    T, ng, ngT = batch.T, batch.ng, batch.ngT
    m = np.zeros(T * ng + ngT, dtype=bool)
    for t in range(T):
        m[t * ng: t * ng + int(nrows[t])] = True
    m[T * ng: T * ng + int(nrows[T])] = True
    return m
