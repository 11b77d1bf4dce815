"""GPU: the BASELINE.json configurations at (or near) their full sizes.

The oracle is too slow for whole batches of the large configs, so each test checks
  (1) the fixed-seed parity subset (the first problems of the batch) against the
      oracle at 1e-9, with identical status / qp_iters, and
  (2) size-independent properties on EVERY problem of the batch, recomputed with numpy from
      the downloaded members: the direction satisfies the linearised dynamics
      dx_{t+1} = Fx dx_t + Fu du_t + fs_{t+1} (src/sqp.cpp:287-289), du_t = -K_t dx_t - k_t
      (:285-286), and the reported KKT_ equals the residual of src/sqp.cpp:244-268 evaluated
      from lag_mul_.
"""
import numpy as np
import pytest

from helpers import assert_close
from mim_solvers_b200 import _capi
from mim_solvers_b200.problems import make_config, tile_batch

pytestmark = pytest.mark.gpu
TOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    from mim_solvers_b200.engine import BatchEngine
    _capi.load_library()
    return BatchEngine


def _rel(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


def _assert_norms(g, o, nsub, name):
    """ADMM residual norms: |c - z| and |G^T rho dz| are differences of O(1) quantities (they reach
    1e-11 on converged problems), so they are compared at 1e-9 of max(1, |value|) per problem --
    the scale of c, z, y they are formed from -- not relative to themselves."""
    for s in ("norm_primal", "norm_dual", "norm_primal_rel", "norm_dual_rel"):
        x, y = g.scalar(s)[:nsub], o.scalar(s)
        assert np.array_equal(np.isfinite(x), np.isfinite(y)), f"{name} {s}"
        fin = np.isfinite(y)
        assert np.all(np.abs(x[fin] - y[fin]) <= TOL * np.maximum(1.0, np.abs(y[fin]))), (name, s, x, y)


def _sqp_properties(batch, g):
    dx, du, lam = g.field("dx"), g.field("du"), g.field("lag_mul")
    Fx, Fu, fs = batch.mat("Fx"), batch.mat("Fu"), batch.arrays["fs"]
    nxt = np.einsum("btij,btj->bti", Fx, dx[:, :-1]) + np.einsum("btij,btj->bti", Fu, du) + fs[:, 1:]
    assert _rel(dx[:, 1:], nxt) <= TOL, "linearised dynamics"
    assert np.abs(dx[:, 0]).max() == 0.0
    K, k = g.field("K"), g.field("k")
    assert _rel(du, -np.einsum("btai,bti->bta", K, dx[:, :-1]) - k) <= TOL, "feedback law"
    rx = batch.arrays["Lx"] + np.einsum("btji,btj->bti", Fx, lam[:, 1:]) - lam[:, :-1]
    ru = batch.arrays["Lu"] + np.einsum("btji,btj->bti", Fu, lam[:, 1:])
    kkt = np.maximum(np.abs(rx).max(axis=(1, 2)), np.abs(ru).max(axis=(1, 2)))
    kkt = np.maximum(kkt, np.abs(batch.arrays["LxT"] - lam[:, -1]).max(axis=1))
    kkt = np.maximum(kkt, np.abs(fs).max(axis=(1, 2)))
    got = g.scalar("KKT")
    assert np.all(np.abs(got - kkt) <= 1e-9 * np.maximum(1.0, kkt)), "KKT residual"
    assert np.all(np.abs(g.scalar("gap_norm") - np.abs(fs).sum(axis=(1, 2))) <= 1e-9 * np.abs(fs).sum(axis=(1, 2)))


@pytest.mark.parametrize("name,B,nsub", [("C2_ur5_sqp", 4096, 64), ("C1_ur5_sqp", 4096, 64), ("C4_lq_nx16", 2048, 32),
                                         ("C4_lq_nx64", 96, 4), ("C5_talos_sqp", 32, 2)])
def test_fullsize_sqp(eng, oracle_lib, name, B, nsub):
    batch = make_config(name, B)
    g = eng(batch)
    g.sqp_direction()
    assert not g.status.any()
    sub = batch.subset(np.arange(nsub))
    o = oracle_lib.Oracle(sub)
    o.sqp_direction()
    for f in ("dx", "du", "lag_mul", "Vxx", "Vx", "K", "k"):
        assert_close(g.field(f)[:nsub], o.field(f), TOL, f"{name} {f}", per_problem=True)
    for s in ("KKT", "gap_norm", "merit", "x_grad_norm", "u_grad_norm"):
        assert_close(g.scalar(s)[:nsub], o.scalar(s), TOL, f"{name} {s}", per_problem=True)
    _sqp_properties(batch, g)


@pytest.mark.parametrize("name,B,nsub", [("C1_ur5_csqp", 512, 16), ("C3_solo12_csqp", 64, 4), ("C3b_solo12_bench", 64, 4),
                                         ("C5_talos_csqp", 32, 2), ("C4_clq_nx32", 64, 4), ("C4_clq_nx64", 32, 2),
                                         ("C4_clq_nx16", 128, 8)])
def test_fullsize_csqp_protocol(eng, oracle_lib, name, B, nsub):
    """The benchmark protocol (benchmarks/solo12.cpp:213-216: 50 ADMM iterations, eps = 0, no
    equality-QP guess) on the bench's own generator; every problem runs the 50 iterations, the
    projected z respects the bounds, y and z are complementary in sign, and the forward roll-out
    satisfies the linearised dynamics."""
    batch = make_config(name, B)
    p = _capi.default_csqp_params(eps_abs=0.0, eps_rel=0.0, max_qp_iters=50, equality_qp_initial_guess=0)
    g = eng(batch)
    g.csqp_calc_norms(); g.csqp_reset_rho(p)
    g.csqp_direction(p, with_kkt=True)
    assert not g.status.any() and np.all(g.qp_iters == 51)
    sub = batch.subset(np.arange(nsub))
    o = oracle_lib.Oracle(sub)
    o.csqp_calc_norms(); o.csqp_reset_rho(p)
    o.csqp_direction(p, with_kkt=True)
    assert np.array_equal(np.asarray(o.qp_iters), g.qp_iters[:nsub])
    for f in ("dx_tilde", "du_tilde", "lag_mul", "y", "z", "rho_vec"):
        assert_close(g.field(f)[:nsub], o.field(f), TOL, f"{name} {f}", per_problem=True)
    for s in ("KKT", "rho_sparse"):
        assert_close(g.scalar(s)[:nsub], o.scalar(s), TOL, f"{name} {s}", per_problem=True)
    _assert_norms(g, o, nsub, name)
    dxt, dut = g.field("dx_tilde"), g.field("du_tilde")
    nxt = np.einsum("btij,btj->bti", batch.mat("Fx"), dxt[:, :-1]) + np.einsum("btij,btj->bti", batch.mat("Fu"), dut) \
        + batch.arrays["fs"][:, 1:]
    assert _rel(dxt[:, 1:], nxt) <= TOL
    T, ng = batch.T, batch.ng
    z = g.field("z")[:, :T * ng].reshape(B, T, ng)
    lo, hi = batch.arrays["g_lb"] - batch.arrays["g"], batch.arrays["g_ub"] - batch.arrays["g"]
    assert np.all(z >= lo - 1e-12 * np.maximum(1.0, np.abs(lo))) and np.all(z <= hi + 1e-12 * np.maximum(1.0, np.abs(hi)))


@pytest.mark.parametrize("name,B,nsub,heur", [("C3_solo12_tol", 64, 16, 0), ("C1_ur5_csqp", 256, 64, 0),
                                              ("C3_solo12_csqp", 32, 4, 0), ("C1_ur5_csqp", 64, 32, 1),
                                              ("C3_solo12_tol", 32, 8, 1)])
def test_fullsize_csqp_default_tolerances(eng, oracle_lib, name, B, nsub, heur):
    """Threshold-driven ADMM (src/csqp.cpp:497-511, 519-532): default eps_abs = eps_rel = 1e-4,
    max_qp_iters = 1000, equality-QP guess, adaptive rho; heur = 1 also exercises
    update_rho_with_heuristic (src/csqp.cpp:1245-1257).  qp_iters / converged are compared
    problem by problem and every mismatch is listed."""
    batch = make_config(name, B)
    p = _capi.default_csqp_params(update_rho_with_heuristic=heur)
    g = eng(batch)
    g.csqp_calc_norms(); g.csqp_reset_rho(p)
    g.csqp_direction(p, with_kkt=True)
    assert not g.status.any()
    sub = batch.subset(np.arange(nsub))
    o = oracle_lib.Oracle(sub)
    o.csqp_calc_norms(); o.csqp_reset_rho(p)
    o.csqp_direction(p, with_kkt=True)
    gi, oi = g.qp_iters[:nsub], np.asarray(o.qp_iters)
    bad = [(int(b), int(gi[b]), int(oi[b])) for b in np.nonzero(gi != oi)[0]]
    assert not bad, f"{name}: qp_iters mismatches (problem, gpu, oracle): {bad}"
    assert np.array_equal(g.converged[:nsub], np.asarray(o.converged))
    if name != "C3_solo12_csqp":   # (its violated rows keep every problem at max_qp_iters + 1)
        assert len(set(oi.tolist())) > 1, "iteration counts should be data-dependent"
    for f in ("dx_tilde", "du_tilde", "lag_mul", "y", "z", "rho_vec"):
        assert_close(g.field(f)[:nsub], o.field(f), TOL, f"{name} {f}", per_problem=True)
    for s in ("KKT", "rho_sparse"):
        assert_close(g.scalar(s)[:nsub], o.scalar(s), TOL, f"{name} {s}", per_problem=True)
    _assert_norms(g, o, nsub, name)
    # second call on the same objects: warm-started y / rho (solver members persist, csqp.hpp:250-357)
    g.csqp_direction(p, with_kkt=True); o.csqp_direction(p, with_kkt=True)
    assert np.array_equal(g.qp_iters[:nsub], np.asarray(o.qp_iters))
    for f in ("dx_tilde", "du_tilde", "y", "z"):
        assert_close(g.field(f)[:nsub], o.field(f), TOL, f"{name} warm {f}", per_problem=True)


def test_fullsize_tiled_batches_are_identical(eng):
    """Problems are independent: a batch made of 8 copies of 512 problems returns 8 identical
    blocks bit for bit (no cross-problem state, no dependence on the slot a problem lands in)."""
    base = make_config("C2_ur5_sqp", 512)
    g = eng(tile_batch(base, 4096))
    g.sqp_direction()
    dx = g.field("dx").reshape(8, 512, -1)
    assert all(np.array_equal(dx[0], dx[i]) for i in range(1, 8))
