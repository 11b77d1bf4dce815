"""Batch container for per-knot OCP derivatives and the synthetic generator.

`OCPBatch` holds what `problem_->calcDiff` leaves in the crocoddyl
ActionDataAbstract objects of B shooting problems of one shape
(Fx Fu Lxx Lxu Luu Lx Lu, gaps fs, Gx Gu g g_lb g_ub; reference: src/ddp.cpp:253-263,
src/csqp.cpp:411-452), in the C-ABI layout of include/mimgpu.h: one array per field,
[B][T][block], dense COLUMN-major blocks (the memory image of Eigen::MatrixXd).

numpy view convention: storage arrays have shape (B, T, ncols, nrows) and are
C-contiguous; `batch.mat("Fx")` returns the mathematical (B, T, nrows, ncols) view.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import _capi

# (rows, cols) of each matrix field as functions of the shape
_MAT = {
    "Fx": ("nx", "nx"), "Fu": ("nx", "nu"), "Lxx": ("nx", "nx"), "Lxu": ("nx", "nu"),
    "Luu": ("nu", "nu"), "Gx": ("ng", "nx"), "Gu": ("ng", "nu"),
    "LxxT": ("nx", "nx"), "GxT": ("ngT", "nx"),
}
_VEC = {"Lx": "nx", "Lu": "nu", "g": "ng", "g_lb": "ng", "g_ub": "ng",
        "LxT": "nx", "gT": "ngT", "g_lbT": "ngT", "g_ubT": "ngT"}
_TERMINAL = {"LxxT", "LxT", "GxT", "gT", "g_lbT", "g_ubT"}


def to_colmajor(a: np.ndarray) -> np.ndarray:
    """math (..., r, c) -> C-contiguous storage (..., c, r) holding col-major blocks."""
    return np.ascontiguousarray(np.swapaxes(a, -1, -2))


def from_colmajor(s: np.ndarray) -> np.ndarray:
    """storage (..., c, r) -> math view (..., r, c)."""
    return np.swapaxes(s, -1, -2)


@dataclass
class OCPBatch:
    B: int
    T: int
    nx: int
    nu: int
    ng: int = 0
    ngT: int = 0
    arrays: dict = field(default_factory=dict)  # name -> storage ndarray

    # ---- construction ---------------------------------------------------------
    @classmethod
    def zeros(cls, B, T, nx, nu, ng=0, ngT=0) -> "OCPBatch":
        self = cls(B, T, nx, nu, ng, ngT)
        dims = dict(nx=nx, nu=nu, ng=ng, ngT=ngT)
        for name, (r, c) in _MAT.items():
            lead = (B,) if name in _TERMINAL else (B, T)
            self.arrays[name] = np.zeros(lead + (dims[c], dims[r]))
        for name, n in _VEC.items():
            lead = (B,) if name in _TERMINAL else (B, T)
            self.arrays[name] = np.zeros(lead + (dims[n],))
        self.arrays["fs"] = np.zeros((B, T + 1, nx))
        self.arrays["cost"] = np.zeros((B,))
        return self

    def shape(self) -> _capi.Shape:
        return _capi.Shape(self.B, self.T, self.nx, self.nu, self.ng, self.ngT)

    def mat(self, name: str) -> np.ndarray:
        """Mathematical (rows x cols) view of a matrix field; vectors as stored."""
        a = self.arrays[name]
        return from_colmajor(a) if name in _MAT else a

    def set(self, name: str, value: np.ndarray):
        if name in _MAT:
            self.mat(name)[...] = value
        else:
            self.arrays[name][...] = value

    def derivs(self) -> _capi.Derivs:
        """ctypes struct of host pointers (keeps no references: keep `self` alive)."""
        d = _capi.Derivs()
        for n in _capi.DERIV_FIELDS:
            a = self.arrays[n]
            assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
            setattr(d, n, a.ctypes.data if a.size else None)
        return d

    def nbytes(self) -> int:
        return int(sum(a.nbytes for a in self.arrays.values()))

    def subset(self, idx) -> "OCPBatch":
        idx = np.asarray(idx)
        o = OCPBatch(len(idx), self.T, self.nx, self.nu, self.ng, self.ngT)
        o.arrays = {k: np.ascontiguousarray(v[idx]) for k, v in self.arrays.items()}
        return o

    def algorithmic_bytes_per_qp(self) -> int:
        """SURVEY 8(d) bytes_QP: inputs read once + dx/du/lag_mul(+y,z) written once."""
        T, nx, nu, ng, ngT = self.T, self.nx, self.nu, self.ng, self.ngT
        ind = T * (2 * nx * nx + 2 * nx * nu + nu * nu + 2 * nx + nu
                   + (ng * (nx + nu) + 2 * ng if ng else 0)) \
            + nx * nx + nx + (ngT * nx + 2 * ngT if ngT else 0) + nx
        outd = 2 * (T + 1) * nx + T * nu + (2 * (T * ng + ngT) if (ng or ngT) else 0) + 8
        return 8 * (ind + outd)

    def flops_sqp_qp(self) -> float:
        """SURVEY 8(d) flops_SQP_QP = T*(full + fwd_sqp + kkt)."""
        T, nx, nu = self.T, self.nx, self.nu
        full = 4 * nx**3 + 6 * nx * nx * nu + 4 * nx * nu * nu + nu**3 / 3 \
            + 5 * nx * nx + 4 * nx * nu + 2 * nu * nu
        fwd = 4 * nx * nx + 4 * nx * nu
        kkt = 2 * nx * nx + 2 * nx * nu
        return T * (full + fwd + kkt)

    def flops_csqp_qp(self, n_it: int, eq_guess: bool, interval: int = 25) -> float:
        """SURVEY 8(d) flops_CSQP_QP."""
        T, nx, nu, ng = self.T, self.nx, self.nu, self.ng
        full = 4 * nx**3 + 6 * nx * nx * nu + 4 * nx * nu * nu + nu**3 / 3 \
            + 5 * nx * nx + 4 * nx * nu + 2 * nu * nu
        full_c = full + 2 * ng * (nx * nx + nu * nu + nx * nu) + 3 * ng * (nx + nu) + 2 * nx * nx
        light = 2 * nx * nx + 4 * nx * nu + 2 * nu * nu + 2 * ng * (nx + nu)
        fwd_c = 2 * nx * nx + 4 * nx * nu
        upd = 2 * ng * (nx + nu) + 8 * ng
        kkt = 2 * nx * nx + 2 * nx * nu + 2 * ng * (nx + nu)
        n_full = -(-n_it // interval)
        n_light = n_it - n_full
        eq = (full + fwd_c) if eq_guess else 0.0
        return T * (eq + n_full * full_c + n_light * light + n_it * (fwd_c + upd) + kkt)


# ------------------------------------------------------------------ generator
SEED0 = 5489  # default-constructed std::mt19937 seed (tests/random_generator.hpp:20)


def _gen_one(rng: np.random.Generator, T, nx, nu, ng, ngT, dt, box_rows, eq_frac,
             onesided_frac, violate_frac):
    """One synthetic problem, recipe of SURVEY 8(d) (Euler-discretised random
    second-order dynamics, SPD cost Hessian, mixed box/dense constraint rows)."""
    out = {}
    nq = nx // 2
    # dynamics: Fx = I + [[dt^2 Aq, dt I + dt^2 Av],[dt Aq, dt Av]], Fu = [dt^2 Bm; dt Bm]
    Fx = np.tile(np.eye(nx), (T, 1, 1))
    if nx % 2 == 0 and nq > 0:
        Aq = rng.uniform(-1, 1, (T, nq, nq))
        Av = rng.uniform(-1, 1, (T, nq, nq))
        Bm = rng.uniform(-1, 1, (T, nq, nu))
        Fx[:, :nq, :nq] += dt * dt * Aq
        Fx[:, :nq, nq:] += dt * np.eye(nq) + dt * dt * Av
        Fx[:, nq:, :nq] += dt * Aq
        Fx[:, nq:, nq:] += dt * Av
        Fu = np.concatenate([dt * dt * Bm, dt * Bm], axis=1)
    else:  # odd nx: generic stable-ish dynamics
        Fx += dt * rng.uniform(-1, 1, (T, nx, nx))
        Fu = dt * rng.uniform(-1, 1, (T, nx, nu))
    out["Fx"], out["Fu"] = Fx, Fu
    # cost Hessian: dt * (W W^T/(nx+nu) + diag(w))
    n = nx + nu
    W = rng.standard_normal((T, n, n))
    w = np.concatenate([rng.uniform(1e-2, 1, (T, nx)), rng.uniform(1e-3, 1e-1, (T, nu))], axis=1)
    H = dt * (W @ np.swapaxes(W, 1, 2) / n + np.einsum("ti,ij->tij", w, np.eye(n)))
    H = 0.5 * (H + np.swapaxes(H, 1, 2))
    out["Lxx"], out["Lxu"], out["Luu"] = H[:, :nx, :nx], H[:, :nx, nx:], H[:, nx:, nx:]
    WT = rng.standard_normal((nx, nx))
    HT = WT @ WT.T / nx + 10 * np.eye(nx)
    out["LxxT"] = 0.5 * (HT + HT.T)
    out["Lx"] = dt * rng.standard_normal((T, nx))
    out["Lu"] = dt * rng.standard_normal((T, nu))
    out["LxT"] = rng.standard_normal(nx)
    fs = 0.1 * rng.standard_normal((T + 1, nx))
    fs[0] = 0.0
    out["fs"] = fs
    out["cost"] = float(rng.uniform(0, 10))

    def rows(nk, lead, with_u):
        """constraint rows: first `box_rows` are signed unit vectors over [x;u]
        (or x only at the terminal knot), the rest dense N(0,1)/sqrt(nx+nu)."""
        ncol = nx + (nu if with_u else 0)
        G = rng.standard_normal(lead + (nk, ncol)) / np.sqrt(nx + nu)
        nb = min(box_rows if box_rows >= 0 else nk, nk)
        if nb:
            G[..., :nb, :] = 0.0
            # box rows cycle over the u columns first (torque limits) then x
            order = (list(range(nx, ncol)) + list(range(nx))) if with_u else list(range(nx))
            sign = rng.choice([-1.0, 1.0], lead + (nb,))
            for r in range(nb):
                G[..., r, order[r % ncol]] = sign[..., r]
        g = np.sqrt(0.1) * rng.standard_normal(lead + (nk,))
        lb = g - rng.uniform(0.05, 0.5, lead + (nk,))
        ub = g + rng.uniform(0.05, 0.5, lead + (nk,))
        u = rng.uniform(0, 1, lead + (nk,))
        lb = np.where(u < onesided_frac / 2, -np.inf, lb)
        ub = np.where((u >= onesided_frac / 2) & (u < onesided_frac), np.inf, ub)
        both = rng.uniform(0, 1, lead + (nk,)) < 0.02 * (onesided_frac > 0)
        lb = np.where(both, -np.inf, lb)
        ub = np.where(both, np.inf, ub)
        e = rng.uniform(0, 1, lead + (nk,)) < eq_frac
        fin = np.isfinite(lb) & np.isfinite(ub)
        ub = np.where(e & fin, lb, ub)
        # a fraction of rows currently violated (g outside the bounds)
        v = rng.uniform(0, 1, lead + (nk,)) < violate_frac
        shift = rng.uniform(0.6, 1.0, lead + (nk,))
        g = np.where(v & np.isfinite(ub), g + shift, g)
        return G, g, lb, ub

    if ng:
        G, g, lb, ub = rows(ng, (T,), True)
        out["Gx"], out["Gu"] = G[..., :nx], G[..., nx:]
        out["g"], out["g_lb"], out["g_ub"] = g, lb, ub
    if ngT:
        G, g, lb, ub = rows(ngT, (), False)
        out["GxT"], out["gT"], out["g_lbT"], out["g_ubT"] = G, g, lb, ub
    return out


def make_random_batch(B, T, nx, nu, ng=0, ngT=0, *, first=0, dt=0.05, box_rows=-1,
                      eq_frac=0.0, onesided_frac=0.25, violate_frac=0.1,
                      seed0=SEED0) -> OCPBatch:
    """B synthetic problems; problem `first + b` is drawn from MT19937(seed0 + first + b),
    so a rank's shard does not depend on the world size."""
    batch = OCPBatch.zeros(B, T, nx, nu, ng, ngT)
    for b in range(B):
        rng = np.random.Generator(np.random.MT19937(seed0 + first + b))
        one = _gen_one(rng, T, nx, nu, ng, ngT, dt, box_rows, eq_frac, onesided_frac,
                       violate_frac)
        for k, v in one.items():
            if k in _MAT:
                batch.mat(k)[b] = v
            else:
                batch.arrays[k][b] = v
    return batch


def tile_batch(base: OCPBatch, B: int) -> OCPBatch:
    """Repeat a smaller batch up to B problems (fast fill for large benchmarks)."""
    reps = -(-B // base.B)
    o = OCPBatch(B, base.T, base.nx, base.nu, base.ng, base.ngT)
    o.arrays = {k: np.ascontiguousarray(np.concatenate([v] * reps, axis=0)[:B])
                for k, v in base.arrays.items()}
    return o


# named workloads of BASELINE.json / SURVEY 8(d)
CONFIGS = {
    "C1_ur5_sqp": dict(T=40, nx=12, nu=6),
    "C1_ur5_csqp": dict(T=40, nx=12, nu=6, ng=3, ngT=3, box_rows=0),
    "C2_ur5_sqp": dict(T=100, nx=12, nu=6),
    "C3_solo12_csqp": dict(T=100, nx=36, nu=12, ng=16, ngT=0, box_rows=12),
    # C3 without currently-violated rows: the ADMM loop converges at the default tolerances with
    # data-dependent iteration counts (200..650) -- the threshold-driven parity / bench case
    "C3_solo12_tol": dict(T=100, nx=36, nu=12, ng=16, ngT=0, box_rows=12, violate_frac=0.0),
    "C3b_solo12_bench": dict(T=50, nx=36, nu=12, ng=36, ngT=36, box_rows=36),
    "C5_talos_csqp": dict(T=100, nx=76, nu=32, ng=32, ngT=0, box_rows=32),
    "C5_talos_sqp": dict(T=120, nx=76, nu=32),
}
for _nx in (8, 16, 32, 64):
    CONFIGS[f"C4_lq_nx{_nx}"] = dict(T=50, nx=_nx, nu=_nx // 2)
    CONFIGS[f"C4_clq_nx{_nx}"] = dict(T=50, nx=_nx, nu=_nx // 2, ng=_nx + _nx // 2,
                                      ngT=_nx, box_rows=-1, eq_frac=0.1)


def make_config(name: str, B: int, first: int = 0) -> OCPBatch:
    return make_random_batch(B, first=first, **CONFIGS[name])
