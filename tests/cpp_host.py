"""Helpers shared by the CPU and GPU tests of the C++ host layer
(include/mim_solvers_b200/batch_solvers.hpp): build the driver tests/cpp/test_batch_solvers.cpp
against (a) the host-emulation shim of the C ABI (CPU, test infrastructure) or (b) libmimgpu.so
(GPU), hand it LQ problems, read back its results."""
from __future__ import annotations

import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent / "cpp"
_ROOT = _DIR.parent.parent
_BUILD = _DIR / "_build"
EMU_BIN = _BUILD / "test_batch_solvers_emu"
GPU_BIN = _BUILD / "test_batch_solvers_gpu"
_HDRS = [_ROOT / "include" / "mim_solvers_b200" / "batch_solvers.hpp", _ROOT / "include" / "mimgpu.h",
         _DIR / "test_batch_solvers.cpp"]


def _stale(out: Path, deps) -> bool:
    return not out.exists() or out.stat().st_mtime < max(d.stat().st_mtime for d in deps)


def build_emu(force=False) -> Path:
    deps = _HDRS + [_DIR / "mimgpu_emu_shim.cpp", _ROOT / "mim_solvers_b200" / "csrc" / "kernels_generic.cuh"]
    if force or _stale(EMU_BIN, deps):
        _BUILD.mkdir(exist_ok=True)
        subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-ffp-contract=off", "-o", str(EMU_BIN),
                        str(_DIR / "test_batch_solvers.cpp"), str(_DIR / "mimgpu_emu_shim.cpp")], check=True)
    return EMU_BIN


def build_api_emu(force=False) -> Path:
    out = _BUILD / "test_api_emu"
    deps = [_HDRS[0], _HDRS[1], _DIR / "test_api.cpp", _DIR / "mimgpu_emu_shim.cpp",
            _ROOT / "mim_solvers_b200" / "csrc" / "kernels_generic.cuh"]
    if force or _stale(out, deps):
        _BUILD.mkdir(exist_ok=True)
        subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-ffp-contract=off", "-o", str(out),
                        str(_DIR / "test_api.cpp"), str(_DIR / "mimgpu_emu_shim.cpp")], check=True)
    return out


def build_gpu(force=False) -> Path:
    """g++ only: the driver links the prebuilt libmimgpu.so (rpath relative to the binary)."""
    lib = _ROOT / "mim_solvers_b200" / "csrc" / "libmimgpu.so"
    if force or _stale(GPU_BIN, _HDRS + [lib]):
        _BUILD.mkdir(exist_ok=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", str(GPU_BIN), str(_DIR / "test_batch_solvers.cpp"),
                        "-L" + str(lib.parent), "-lmimgpu", "-Wl,-rpath,$ORIGIN/../../../mim_solvers_b200/csrc"],
                       check=True)
    return GPU_BIN


def write_problem(path: Path, p) -> None:
    """LQBatchProblem (mim_solvers_b200/solvers.py) -> the driver's binary format (row-major)."""
    with open(path, "wb") as f:
        np.array([p.B, p.T, p.nx, p.nu, p.ng, p.ngT], dtype=np.int32).tofile(f)
        arrs = [p.x0, p.A, p.Bm, p.f, p.Q, p.S, p.R, p.q, p.r, p.QT, p.qT]
        if p.ng:
            arrs += [p.C, p.D, p.e, p.g_lb, p.g_ub]
        if p.ngT:
            arrs += [p.CT, p.eT, p.g_lbT, p.g_ubT]
        for a in arrs:
            np.ascontiguousarray(a, dtype=np.float64).tofile(f)


def run(binary: Path, solver: str, p, tmp: Path, **opts) -> dict:
    prob, out = tmp / f"{solver}_problem.bin", tmp / f"{solver}_out.bin"
    write_problem(prob, p)
    cmd = [str(binary), solver, str(prob), str(out)] + [f"{k}={v}" for k, v in opts.items()]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout[-2000:], r.stderr[-2000:])
    v = np.fromfile(out, dtype=np.float64)
    B, T, nx, nu = p.B, p.T, p.nx, p.nu
    o, res = 0, {}
    for name, n in (("converged", B), ("iter", B), ("KKT", B), ("preg", B), ("qp_iters", B),
                    ("xs", B * (T + 1) * nx), ("us", B * T * nu)):
        res[name] = v[o:o + n]
        o += n
    assert o == v.size
    res["xs"] = res["xs"].reshape(B, T + 1, nx)
    res["us"] = res["us"].reshape(B, T, nu)
    res["stdout"] = r.stdout
    return res


# ---- the checks (reference invariants of tests/test_solvers.cpp + agreement with the Python mirror)
def check_sqp(binary, tmp, factory):
    import solver_checks as sc
    from mim_solvers_b200.solvers import SolverSQP
    p = sc.make_lq()
    r = run(binary, "sqp", p, tmp, maxiter=10)
    assert r["converged"].all() and (r["iter"] == 1).all() and (r["KKT"] <= 1e-6).all(), r["stdout"]
    c = p.calc(r["xs"], r["us"])
    assert np.abs(c["xnext"] - r["xs"][:, 1:]).max() < 1e-9
    s = SolverSQP(p, factory)
    s.solve(maxiter=10)
    assert np.abs(s.xs - r["xs"]).max() < 1e-9 and np.abs(s.us - r["us"]).max() < 1e-9
    assert (s.iter == r["iter"]).all()


def check_csqp_unconstrained(binary, tmp, factory):
    import solver_checks as sc
    p = sc.make_lq(seed=3)
    a = run(binary, "sqp", p, tmp, maxiter=10)
    b = run(binary, "csqp", p, tmp, maxiter=10)
    assert a["converged"].all() and b["converged"].all()
    assert (b["iter"] == 1).all() and (b["qp_iters"] == 25).all(), b["stdout"]
    assert np.abs(a["xs"] - b["xs"]).max() < 1e-6 and np.abs(a["us"] - b["us"]).max() < 1e-6


def check_csqp_constrained(binary, tmp, factory):
    import solver_checks as sc
    from mim_solvers_b200.solvers import SolverCSQP
    p = sc.make_lq(seed=5, constrained=True)
    r = run(binary, "csqp", p, tmp, maxiter=20, termination_tolerance=1e-4, eps_abs=1e-8, eps_rel=0.0,
            max_qp_iters=4000)
    assert r["converged"].all(), r["stdout"]
    assert (r["qp_iters"] < 4000).all()
    assert r["us"].max() <= 0.3 + 1e-5 and r["us"].min() >= -0.3 - 1e-5
    assert (np.abs(r["us"]) > 0.3 - 1e-5).any()
    s = SolverCSQP(p, factory)
    s.termination_tolerance = 1e-4
    s.eps_abs, s.eps_rel, s.max_qp_iters = 1e-8, 0.0, 4000
    s.solve(maxiter=20)
    assert (s.iter == r["iter"]).all() and (s.qp_iters == r["qp_iters"]).all(), \
        (s.iter, r["iter"], s.qp_iters, r["qp_iters"])
    assert np.abs(s.xs - r["xs"]).max() < 1e-9 and np.abs(s.us - r["us"]).max() < 1e-9


def check_retry(binary, tmp, factory):
    import solver_checks as sc
    p = sc.make_lq(B=3, seed=7)
    p.R[1] = p.R[1] - 0.02 * np.eye(p.nu)
    r = run(binary, "sqp", p, tmp, maxiter=1)
    assert r["preg"][1] > 1e-9 * 10 and r["preg"][0] == 1e-9 and r["preg"][2] == 1e-9, r["preg"]


def check_fddp(binary, tmp, factory):
    import solver_checks as sc
    from mim_solvers_b200.solvers import SolverFDDP
    p = sc.make_lq(seed=9)
    r = run(binary, "fddp", p, tmp, maxiter=10)
    assert r["converged"].all() and (r["iter"] <= 3).all(), r["stdout"]
    a = run(binary, "sqp", p, tmp, maxiter=10)
    assert np.abs(a["us"] - r["us"]).max() < 1e-5
    # at the optimum cost_try == cost to rounding, so whether the filter takes alpha = 1 (and with
    # it the iteration at which KKT_ sees the fresh lag_mul_) depends on summation order: compare
    # the solutions (both inside termination_tolerance of the optimum), not the counts
    s = SolverFDDP(p, factory)
    s.solve(maxiter=10)
    assert (np.abs(s.iter - r["iter"]) <= 1).all(), (s.iter, r["iter"])
    assert np.abs(s.xs - r["xs"]).max() < 1e-5 and np.abs(s.us - r["us"]).max() < 1e-5


CHECKS = [check_sqp, check_csqp_unconstrained, check_csqp_constrained, check_retry, check_fddp]
