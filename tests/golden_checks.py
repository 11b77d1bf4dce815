"""Engine-agnostic replays of the golden files (reference NumPy outputs).

`make_engine(batch)` returns an Oracle or a CUDA BatchEngine.  Tolerances are relative
to the max-abs of the reference array.  The reference's own C++-vs-NumPy tests use
1e-6 / 1e-8 absolute (tests/python/test_clqr_stagewise_admm.py:64,
tests/python/test_ur5_csqp.py:147); the NumPy prototypes order their arithmetic
differently from the C++ (closed-loop forward pass, S - L^T H L Riccati form), so the
bar here is 1e-8 relative on vectors and EXACT equality on iteration counts.
"""
import numpy as np

from helpers import (assert_close, csqp_params_from_golden, load_golden, real_row_mask)
from mim_solvers_b200 import _capi
from mim_solvers_b200.problems import from_colmajor

TOL_SQP = 1e-9
TOL_CSQP = 1e-8


def check_sqp_golden(make_engine, name, tol=TOL_SQP):
    batch, P, out = load_golden(name)
    eng = make_engine(batch)
    p = _capi.default_sqp_params(preg=float(P["preg"]), dreg=float(P["dreg"]),
                                 mu_dynamic=float(P["mu_dynamic"]))
    eng.sqp_direction(p, with_kkt=True)
    assert int(eng.status[0]) == 0
    assert_close(eng.field("dx")[0], out["dx"], tol, "dx")
    assert_close(eng.field("du")[0], out["du"], tol, "du")
    assert_close(eng.field("lag_mul")[0], out["lag_mul"], tol, "lag_mul")
    Vxx = from_colmajor(eng.field("Vxx")[0])
    assert_close(Vxx, out["S"], tol, "Vxx")
    # C++ Vx_[t] carries the gap term Vxx_[t] fs_[t] (src/ddp.cpp:336-338); python s[t] does not
    fs = batch.arrays["fs"][0]
    s_eq = eng.field("Vx")[0] - np.einsum("tij,tj->ti", Vxx, fs)
    assert_close(s_eq, out["s"], tol, "Vx (gap-free)")
    assert_close(eng.field("K")[0], -out["L"], tol, "K")
    assert_close(eng.field("k")[0], -out["l"], tol, "k")
    assert_close(eng.scalar("KKT")[0], out["KKT"], tol, "KKT")
    assert_close(eng.scalar("gap_norm")[0], out["gap_norm"], 1e-12, "gap_norm")
    assert_close(eng.scalar("merit")[0], out["merit"], 1e-12, "merit")
    assert_close(eng.scalar("x_grad_norm")[0], out["x_grad_norm"], tol, "x_grad_norm")
    assert_close(eng.scalar("u_grad_norm")[0], out["u_grad_norm"], tol, "u_grad_norm")


def _python_kkt(batch, eng):
    """KKT as python/csqp.py:212-240 defines it (no state-stationarity row at t = 0),
    evaluated on the engine's outputs -- lets the golden KKT be compared although the
    C++ definition (src/csqp.cpp:598-606) also includes the t = 0 row."""
    T, nx, nu, ng, ngT = batch.T, batch.nx, batch.nu, batch.ng, batch.ngT
    lag = eng.field("lag_mul")[0]
    y = eng.field("y")[0]
    Fx, Fu = batch.mat("Fx")[0], batch.mat("Fu")[0]
    kkt = 0.0
    for t in range(T):
        yt = y[t * ng:(t + 1) * ng]
        lu = batch.arrays["Lu"][0, t] + Fu[t].T @ lag[t + 1]
        if ng:
            lu = lu + batch.mat("Gu")[0, t].T @ yt
        kkt = max(kkt, np.abs(lu).max())
        if t == 0:
            continue
        lx = batch.arrays["Lx"][0, t] + Fx[t].T @ lag[t + 1] - lag[t]
        if ng:
            lx = lx + batch.mat("Gx")[0, t].T @ yt
        kkt = max(kkt, np.abs(lx).max())
    lx = batch.arrays["LxT"][0] - lag[T]
    if ngT:
        lx = lx + batch.mat("GxT")[0].T @ y[T * ng:]
    kkt = max(kkt, np.abs(lx).max(), np.abs(batch.arrays["fs"][0]).max(),
              float(eng.scalar("constraint_norm")[0]))
    return kkt


def check_csqp_golden(make_engine, name, tol=TOL_CSQP):
    batch, P, out = load_golden(name)
    eng = make_engine(batch)
    p = csqp_params_from_golden(P)
    n_calls = int(P["n_calls"])
    eng.csqp_calc_norms()
    assert_close(eng.scalar("gap_norm")[0], out["gap_norm"], 1e-12, "gap_norm")
    assert_close(eng.scalar("constraint_norm")[0], out["constraint_norm"], 1e-12, "constraint_norm")
    if not int(P["reset_rho"]):
        eng.csqp_reset_rho(p)
    mask = real_row_mask(batch, out["nrows"])
    for c in range(n_calls):
        o = {k: out[k if c == 0 else f"call{c}_{k}"] for k in (
            "dx_tilde", "du_tilde", "dx", "du", "y", "z", "rho_vec", "rho_sparse", "norm_primal",
            "norm_dual", "norm_primal_rel", "norm_dual_rel", "qp_iters", "lag_mul", "KKT",
            "Vxx", "Vx", "K", "k")}
        eng.csqp_direction(p, with_kkt=True)
        assert int(eng.status[0]) == 0
        # iteration count: identical; python reports max_qp_iters on exhaustion, C++ max+1
        # (src/csqp.cpp:470 vs python/qp_solvers/stagewise_qp.py:154,206)
        qp_py = int(o["qp_iters"])
        exhausted = not bool(eng.converged[0]) if hasattr(eng, "converged") else None
        qp = int(eng.qp_iters[0])
        if qp == int(P["max_qp_iters"]) + 1:
            assert qp_py == int(P["max_qp_iters"]), (qp, qp_py)
            if exhausted is not None:
                assert exhausted
        else:
            assert qp == qp_py, f"qp_iters {qp} != reference {qp_py} (call {c})"
        w = f"[call {c}] "
        assert_close(eng.field("dx_tilde")[0], o["dx_tilde"], tol, w + "dx_tilde")
        assert_close(eng.field("du_tilde")[0], o["du_tilde"], tol, w + "du_tilde")
        assert_close(eng.field("dx")[0], o["dx"], tol, w + "dx")
        assert_close(eng.field("du")[0], o["du"], tol, w + "du")
        assert_close(eng.field("lag_mul")[0], o["lag_mul"], tol, w + "lag_mul")
        assert_close(eng.scalar("rho_sparse")[0], o["rho_sparse"], tol, w + "rho_sparse")
        if mask.any():
            assert_close(eng.field("y")[0][mask], o["y"], tol, w + "y")
            assert_close(eng.field("z")[0][mask], o["z"], tol, w + "z")
            assert_close(eng.field("rho_vec")[0][mask], o["rho_vec"], tol, w + "rho_vec")
        for nm in ("norm_primal", "norm_dual", "norm_primal_rel", "norm_dual_rel"):
            ref = float(o[nm])
            got = float(eng.scalar(nm)[0])
            if np.isfinite(ref) and abs(ref) < 1e-13:
                assert abs(got) < 1e-10, (nm, got, ref)   # residual at round-off level
            elif np.isneginf(ref) and mask.sum() < mask.size and got == 0.0:
                pass  # only padded rows exist: 0 instead of -inf (see helpers.real_row_mask)
            else:
                # residuals that are differences of O(1) numbers: tolerance relative to 1
                assert abs(got - ref) <= tol * max(1.0, abs(ref)) or got == ref, (nm, got, ref)
        assert_close(from_colmajor(eng.field("Vxx")[0]), o["Vxx"], tol, w + "Vxx")
        assert_close(eng.field("Vx")[0], o["Vx"], tol, w + "Vx")
        assert_close(eng.field("K")[0], o["K"], tol, w + "K")
        assert_close(eng.field("k")[0], o["k"], tol, w + "k")
        kkt_py_def = _python_kkt(batch, eng)
        assert abs(kkt_py_def - float(o["KKT"])) <= tol * max(1.0, abs(float(o["KKT"]))), \
            (kkt_py_def, float(o["KKT"]))
        assert float(eng.scalar("KKT")[0]) >= kkt_py_def - 1e-12  # C++ adds the t = 0 row
