#!/usr/bin/env python
"""Generate golden input/output vectors by RUNNING THE REFERENCE'S OWN NumPy solvers.

Run in the build container only (needs /root/reference, which does not exist on the
GPU box):      python tests/golden/make_golden.py

What runs: the unmodified reference files
  /root/reference/python/sqp.py                       (class SQP)
  /root/reference/python/csqp.py                      (class CSQP)
  /root/reference/python/qp_solvers/stagewise_qp.py   (class StagewiseADMM)
  /root/reference/tests/python/problems.py            (create_lqr_problem, create_clqr_problem)
on top of the crocoddyl/eigenpy stand-ins of oracle/refshim (those only provide
model/problem plumbing, i.e. the INPUT blocks; see oracle/refshim/README.md).

Each golden file holds, for ONE problem (B = 1): the inputs in the C-ABI storage
layout of include/mimgpu.h (`in_<field>`), the solver parameters (`p_<name>`) and the
reference's outputs (`out_<name>`).  tests/test_oracle_golden.py replays the inputs
through the C oracle, tests/test_gpu_golden.py through the CUDA library.

Known, documented differences between the NumPy prototypes and the C++ solvers
 that the tests account for:
  * no preg/dreg in Python            -> oracle/GPU run with preg = dreg = 0
  * Python always does the full backward pass (same maths as the C++ light sweep)
  * qp_iters on exhaustion: Python max_qp_iters, C++ max_qp_iters + 1
  * Python's SQP keeps Vx gap-free (s[t]); C++ folds Vxx_t fs_t into Vx_[t]
  * Python KKT_check skips the t = 0 state-stationarity row, C++ keeps it
  * gap[t] (Python) == fs_[t+1] (C++)
"""
import os
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
REF = Path("/root/reference")
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "oracle" / "refshim"))
sys.path.insert(1, str(REF / "python"))
sys.path.insert(1, str(REF / "tests" / "python"))

import crocoddyl  # noqa: E402  (the stand-in)
from csqp import CSQP  # noqa: E402  (reference)
from problems import create_clqr_problem, create_lqr_problem  # noqa: E402 (reference)
from sqp import SQP  # noqa: E402  (reference)

from mim_solvers_b200.problems import OCPBatch, make_random_batch  # noqa: E402


# ------------------------------------------------------------------ helpers
def batch_from_problem(problem, solver_gap) -> OCPBatch:
    """Read the derivative blocks that problem.calcDiff left in the datas (the upload
    point of the hot path) into a B=1 OCPBatch; knots with fewer constraint rows than
    the maximum are padded with unbounded all-zero rows."""
    T = problem.T
    nx = problem.terminalModel.state.ndx
    nu = problem.runningModels[0].nu
    ng = max(m.ng for m in problem.runningModels)
    ngT = problem.terminalModel.ng
    b = OCPBatch.zeros(1, T, nx, nu, ng, ngT)
    for t, (m, d) in enumerate(zip(problem.runningModels, problem.runningDatas)):
        for k in ("Fx", "Fu", "Lxx", "Lxu", "Luu"):
            b.mat(k)[0, t] = np.asarray(getattr(d, k)).reshape(b.mat(k)[0, t].shape)
        b.arrays["Lx"][0, t] = d.Lx
        b.arrays["Lu"][0, t] = d.Lu
        b.arrays["fs"][0, t + 1] = solver_gap[t]
        if ng:
            b.arrays["g_lb"][0, t] = -np.inf
            b.arrays["g_ub"][0, t] = np.inf
            if m.ng:
                b.mat("Gx")[0, t, : m.ng] = d.Gx
                b.mat("Gu")[0, t, : m.ng] = d.Gu
                b.arrays["g"][0, t, : m.ng] = d.g
                b.arrays["g_lb"][0, t, : m.ng] = m.g_lb
                b.arrays["g_ub"][0, t, : m.ng] = m.g_ub
    dT, mT = problem.terminalData, problem.terminalModel
    b.mat("LxxT")[0] = dT.Lxx
    b.arrays["LxT"][0] = dT.Lx
    if ngT:
        b.mat("GxT")[0] = dT.Gx
        b.arrays["gT"][0] = dT.g
        b.arrays["g_lbT"][0] = mT.g_lb
        b.arrays["g_ubT"][0] = mT.g_ub
    return b


def frozen_problem(batch: OCPBatch, p: int = 0):
    """Build a (stand-in) ShootingProblem whose knots return the blocks of problem `p`
    of `batch`; xs = 0 and xnext_t = fs[t+1] so that the solver's gap[t] == fs[t+1]."""
    T, nx, nu, ng, ngT = batch.T, batch.nx, batch.nu, batch.ng, batch.ngT
    A = batch.arrays
    running = []
    for t in range(T):
        blk = {k: batch.mat(k)[p, t].copy() for k in ("Fx", "Fu", "Lxx", "Lxu", "Luu")}
        blk["Lx"], blk["Lu"] = A["Lx"][p, t].copy(), A["Lu"][p, t].copy()
        blk["xnext"] = A["fs"][p, t + 1].copy()
        blk["cost"] = A["cost"][p] / (T + 1)
        if ng:
            blk["Gx"], blk["Gu"] = batch.mat("Gx")[p, t].copy(), batch.mat("Gu")[p, t].copy()
            blk["g"] = A["g"][p, t].copy()
        running.append(crocoddyl.FrozenActionModel(
            nx, nu, ng, blk, A["g_lb"][p, t] if ng else None, A["g_ub"][p, t] if ng else None))
    blk = {"Lxx": batch.mat("LxxT")[p].copy(), "Lx": A["LxT"][p].copy(), "cost": A["cost"][p] / (T + 1),
           "Fx": np.eye(nx), "Fu": np.zeros((nx, nu))}
    if ngT:
        blk["Gx"] = batch.mat("GxT")[p].copy()
        blk["g"] = A["gT"][p].copy()
    terminal = crocoddyl.FrozenActionModel(nx, nu, ngT, blk, A["g_lbT"][p] if ngT else None,
                                           A["g_ubT"][p] if ngT else None)
    problem = crocoddyl.ShootingProblem(np.zeros(nx), running, terminal)
    xs = [np.zeros(nx) for _ in range(T + 1)]
    us = [np.zeros(nu) for _ in range(T)]
    return problem, xs, us


def save(name, batch: OCPBatch, params: dict, outputs: dict):
    d = {f"in_{k}": v for k, v in batch.arrays.items()}
    d["shape"] = np.array([batch.B, batch.T, batch.nx, batch.nu, batch.ng, batch.ngT])
    d.update({f"p_{k}": np.asarray(v) for k, v in params.items()})
    d.update({f"out_{k}": np.asarray(v) for k, v in outputs.items()})
    path = HERE / f"{name}.npz"
    np.savez_compressed(path, **d)
    print(f"  wrote {path.name}  ({path.stat().st_size / 1024:.1f} KB)")


# ------------------------------------------------------------------ SQP cases
def run_sqp(name, problem, xs, us):
    """reference python/sqp.py: SQP.computeDirection (calc + backwardPass + computeUpdates)
    and KKT_check, at the initial guess."""
    s = SQP(problem)
    xs = [np.array(x, dtype=float) for x in xs]
    xs[0] = problem.x0.copy()
    s.setCandidate(xs, us, False)
    s.computeDirection()
    s.KKT_check()
    batch = batch_from_problem(problem, s.gap)
    batch.arrays["cost"][0] = s.cost
    out = dict(
        dx=np.array(s.dx), du=np.array(s.du), lag_mul=np.array(s.lag_mul),
        S=np.array(s.S),            # == C++ Vxx_ (math layout, symmetric)
        s=np.array(s.s),            # == C++ Vx_[t] - Vxx_[t] fs_[t]
        L=np.array(s.L), l=np.array(s.l),   # == -K_, -k_
        KKT=s.KKT, gap_norm=s.gap_norm, merit=s.merit, cost=s.cost,
        x_grad_norm=s.x_grad_norm, u_grad_norm=s.u_grad_norm,
    )
    save(name, batch, dict(preg=0.0, dreg=0.0, mu_dynamic=s.mu), out)


# ------------------------------------------------------------------ CSQP cases
def csqp_outputs(s):
    return dict(
        dx_tilde=np.array(s.dx_tilde), du_tilde=np.array(s.du_tilde),
        dx=np.array(s.dx), du=np.array(s.du),
        y=np.concatenate([np.ravel(v) for v in s.y]) if sum(len(v) for v in s.y) else np.zeros(0),
        z=np.concatenate([np.ravel(v) for v in s.z]) if sum(len(v) for v in s.z) else np.zeros(0),
        rho_vec=np.concatenate([np.ravel(v) for v in s.rho_vec]) if sum(len(v) for v in s.rho_vec) else np.zeros(0),
        nrows=np.array([len(v) for v in s.y]),
        rho_sparse=s.rho_sparse, norm_primal=s.norm_primal, norm_dual=s.norm_dual,
        norm_primal_rel=s.norm_primal_rel, norm_dual_rel=s.norm_dual_rel,
        qp_iters=s.qp_iters, lag_mul=np.array(s.lag_mul), KKT=s.KKT,
        Vxx=np.array(s.Vxx), Vx=np.array(s.Vx), K=np.array(s.K), k=np.array(s.k),
        gap_norm=s.gap_norm, constraint_norm=s.constraint_norm, cost=s.cost,
    )


def run_csqp(name, problem, xs, us, n_calls=1, **params):
    """reference python/csqp.py + qp_solvers/stagewise_qp.py: the QP part of one SQP
    iteration of CSQP.solve (python/csqp.py:284-296): calc(True); reset_rho_vec() at
    iter 0; computeDirection(); KKT_check().  n_calls > 1 repeats computeDirection on the
    same derivatives (warm-started y / rho, as in the reference's QP benchmark loop,
    benchmarks/solo12.cpp:221-225) and records the outputs of every call."""
    s = CSQP(problem, "StagewiseQP")
    for k, v in params.items():
        assert hasattr(s, k), k
        setattr(s, k, v)
    xs = [np.array(x, dtype=float) for x in xs]
    xs[0] = problem.x0.copy()
    s.setCandidate(xs, us, False)
    s.calc(True)
    if not s.reset_rho:
        s.reset_rho_vec()
    batch = batch_from_problem(problem, s.gap)
    batch.arrays["cost"][0] = s.cost
    outs = {}
    for c in range(n_calls):
        s.computeDirection()
        s.KKT_check()
        o = csqp_outputs(s)
        outs.update({(f"{k}" if c == 0 else f"call{c}_{k}"): v for k, v in o.items()})
    P = dict(preg=0.0, dreg=0.0, sigma=s.sigma_sparse, alpha=s.alpha,
             rho_sparse_base=s.rho_sparse_base, rho_min=s.rho_min, rho_max=s.rho_max,
             adaptive_rho_tolerance=s.adaptive_rho_tolerance, eps_abs=s.eps_abs,
             eps_rel=s.eps_rel, max_qp_iters=s.max_qp_iters,
             rho_update_interval=s.rho_update_interval,
             equality_qp_initial_guess=int(s.equality_qp_initial_guess),
             reset_y=int(s.reset_y), reset_rho=int(s.reset_rho), n_calls=n_calls)
    save(name, batch, P, outs)


def main():
    np.seterr(all="ignore")
    print("SQP (reference python/sqp.py)")
    p, xs, us = create_lqr_problem()
    run_sqp("sqp_lqr", p, xs, us)
    for i, (T, nx, nu) in enumerate([(10, 6, 3), (15, 12, 6), (5, 4, 2), (7, 8, 4), (4, 16, 8)]):
        b = make_random_batch(1, T, nx, nu, first=100 + i)
        p, xs, us = frozen_problem(b)
        run_sqp(f"sqp_rand{i}_T{T}_nx{nx}_nu{nu}", p, xs, us)

    print("CSQP (reference python/csqp.py + qp_solvers/stagewise_qp.py)")
    # the reference's own C++-vs-NumPy test settings, tests/python/test_clqr_stagewise_admm.py:32-63
    p, xs, us = create_clqr_problem()
    run_csqp("csqp_clqr_tight", p, xs, us, eps_abs=1e-8, eps_rel=0.0, max_qp_iters=1000,
             reset_rho=True, reset_y=True)
    p, xs, us = create_clqr_problem()
    run_csqp("csqp_clqr_default", p, xs, us)
    p, xs, us = create_clqr_problem()
    run_csqp("csqp_clqr_noguess_50", p, xs, us, eps_abs=0.0, eps_rel=0.0, max_qp_iters=50,
             equality_qp_initial_guess=False)
    p, xs, us = create_clqr_problem()
    run_csqp("csqp_clqr_warm3", p, xs, us, n_calls=3, eps_abs=1e-6, eps_rel=0.0)
    p, xs, us = create_clqr_problem()
    run_csqp("csqp_clqr_warm3_resets", p, xs, us, n_calls=3, eps_abs=1e-6, eps_rel=0.0,
             reset_rho=True, reset_y=True)
    shapes = [  # (T, nx, nu, ng, ngT, kwargs)
        (8, 6, 3, 4, 2, {}),
        (12, 12, 6, 3, 3, dict(box_rows=0)),                   # UR5-like dense rows
        (6, 8, 4, 12, 8, dict(box_rows=-1, eq_frac=0.15)),     # box + equalities
        (10, 4, 2, 2, 0, dict(box_rows=2)),                    # control box only, no terminal rows
        (5, 10, 5, 0, 0, {}),                                  # unconstrained: exits at 25, norms -inf
    ]
    for i, (T, nx, nu, ng, ngT, kw) in enumerate(shapes):
        b = make_random_batch(1, T, nx, nu, ng, ngT, first=200 + i, **kw)
        p, xs, us = frozen_problem(b)
        run_csqp(f"csqp_rand{i}_T{T}_nx{nx}_nu{nu}_ng{ng}", p, xs, us)
        if ng == 0 and ngT == 0:
            # eps_rel = 0 with no constraint rows: the NumPy prototype evaluates
            # -inf <= eps_abs + 0 * -inf = NaN (never converges) while the C++ has an explicit
            # eps_rel == 0 branch (src/csqp.cpp:496-503) and exits at 25.  C++ wins; no golden.
            continue
        p, xs, us = frozen_problem(b)
        run_csqp(f"csqp_rand{i}_tight", p, xs, us, eps_abs=1e-7, eps_rel=0.0, n_calls=2)
    b = make_random_batch(1, 9, 6, 3, 5, 3, first=300)
    p, xs, us = frozen_problem(b)
    run_csqp("csqp_rand_exhaust", p, xs, us, eps_abs=0.0, eps_rel=0.0, max_qp_iters=60)


if __name__ == "__main__":
    os.chdir(HERE)
    main()
