"""ctypes mirror of include/mimgpu.h and loader of the CUDA library.

The product path has NO CPU fallback: if `libmimgpu.so` is missing or no CUDA
device is present every compute entry point raises (`MimGpuError`).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = _PKG_DIR / "csrc" / "libmimgpu.so"

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)

# ---- return codes / enums (include/mimgpu.h) -------------------------------
OK, ERR_INVALID, ERR_CUDA, ERR_NODEVICE, ERR_STATE, ERR_UNSUPPORTED = range(6)
STATUS_OK, STATUS_LLT, STATUS_NONFINITE = 0, 1, 2
MEM_HOST, MEM_DEVICE = 0, 1

FIELDS = dict(
    dx=0, du=1, lag_mul=2, dx_tilde=3, du_tilde=4, y=5, z=6, rho_vec=7,
    Vxx=8, Vx=9, K=10, k=11, Qu=12, Quuk=13, Quu_llt=14, xs=15, us=16,
)
SCALARS = dict(
    KKT=0, gap_norm=1, merit=2, x_grad_norm=3, u_grad_norm=4, norm_primal=5,
    norm_dual=6, norm_primal_rel=7, norm_dual_rel=8, rho_sparse=9,
    constraint_norm=10, dg=11, dq=12, cost=13, steplength=14, preg=15,
)
N_SCALARS = 16
NOALLOC_SCRATCH_BYTES = 32 * 1024 * 1024   # MIMGPU_NOALLOC_SCRATCH_BYTES


class Shape(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("B", "T", "nx", "nu", "ng", "ngT")]

    def field_shape(self, name: str):
        """numpy STORAGE shape of a solver member (C ABI layout)."""
        B, T, nx, nu, ng, ngT = self.B, self.T, self.nx, self.nu, self.ng, self.ngT
        nc = T * ng + ngT
        return {
            "dx": (B, T + 1, nx), "du": (B, T, nu), "lag_mul": (B, T + 1, nx),
            "dx_tilde": (B, T + 1, nx), "du_tilde": (B, T, nu),
            "y": (B, nc), "z": (B, nc), "rho_vec": (B, nc),
            "Vxx": (B, T + 1, nx, nx), "Vx": (B, T + 1, nx),
            "K": (B, T, nu, nx), "k": (B, T, nu), "Qu": (B, T, nu),
            "Quuk": (B, T, nu), "Quu_llt": (B, T, nu, nu),
            "xs": (B, T + 1, nx), "us": (B, T, nu),
        }[name]


DERIV_FIELDS = ("Fx", "Fu", "Lxx", "Lxu", "Luu", "Lx", "Lu", "fs", "LxxT", "LxT",
                "Gx", "Gu", "g", "g_lb", "g_ub", "GxT", "gT", "g_lbT", "g_ubT", "cost")


class Derivs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in DERIV_FIELDS]


class SqpParams(C.Structure):
    _fields_ = [("preg", C.c_double), ("dreg", C.c_double), ("mu_dynamic", C.c_double),
                ("is_feasible", C.c_int32), ("reserved", C.c_int32)]


class CsqpParams(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "preg", "dreg", "sigma", "alpha", "rho_sparse_base", "rho_min", "rho_max",
        "adaptive_rho_tolerance", "eps_abs", "eps_rel")] + [(n, C.c_int32) for n in (
        "max_qp_iters", "rho_update_interval", "equality_qp_initial_guess", "reset_y",
        "reset_rho", "update_rho_with_heuristic", "warm_start_y", "reserved")]


class RecordLayout(C.Structure):
    """mimgpu_record_layout: block offsets (doubles) of the knot-record input layout."""
    _fields_ = [(n, C.c_int32) for n in (
        "rec_doubles", "term_doubles", "problem_doubles", "off_fs", "off_Lxx", "off_Lxu", "off_Luu",
        "off_Lx", "off_Lu", "off_Fx", "off_Fu", "toff_fs", "toff_Lxx", "toff_Lx", "sym_packed", "reserved")]


class PolyModel(C.Structure):
    """mimgpu_poly_model: analytic model of the device-side solve()."""
    _fields_ = [(n, C.c_void_p) for n in ("A", "Bm", "f", "Q", "S", "R", "q", "r", "QT", "qT", "x0")] + [
        ("c3", C.c_double), ("w4", C.c_double)]


class SqpSolveParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("maxiter", "use_filter_line_search", "filter_size",
                                         "extra_iteration_for_last_kkt")] + [(n, C.c_double) for n in (
        "reg_init", "reg_min", "reg_max", "reg_incfactor", "reg_decfactor", "th_stepdec", "th_stepinc",
        "termination_tol", "mu_dynamic")]


def default_sqp_solve_params(**kw) -> SqpSolveParams:
    """Defaults of src/ddp.cpp:25-32 and include/mim_solvers/sqp.hpp:164-181."""
    d = dict(maxiter=100, use_filter_line_search=1, filter_size=1, extra_iteration_for_last_kkt=0,
             reg_init=float("nan"), reg_min=1e-9, reg_max=1e9, reg_incfactor=10.0, reg_decfactor=10.0,
             th_stepdec=0.5, th_stepinc=0.01, termination_tol=1e-6, mu_dynamic=1.0)
    for k, v in kw.items():
        if k not in d:
            raise TypeError(f"unknown solve parameter {k!r}")
        d[k] = type(d[k])(v)
    return SqpSolveParams(**d)


class Summary(C.Structure):
    _fields_ = [("max_kkt", C.c_double), ("sum_qp_iters", C.c_int64),
                ("max_qp_iters", C.c_int64), ("n_converged", C.c_int64),
                ("n_failed", C.c_int64)]


def default_sqp_params(preg=1e-9, dreg=1e-9, mu_dynamic=1.0, is_feasible=False) -> SqpParams:
    """Defaults: reg_min_ (src/ddp.cpp:27), mu_dynamic_ (include/mim_solvers/sqp.hpp:164)."""
    return SqpParams(preg, dreg, mu_dynamic, int(is_feasible), 0)


def default_csqp_params(**kw) -> CsqpParams:
    """Defaults of include/mim_solvers/csqp.hpp:292-331 and src/ddp.cpp:27."""
    d = dict(preg=1e-9, dreg=1e-9, sigma=1e-6, alpha=1.6, rho_sparse_base=1e-1,
             rho_min=1e-6, rho_max=1e3, adaptive_rho_tolerance=5.0, eps_abs=1e-4,
             eps_rel=1e-4, max_qp_iters=1000, rho_update_interval=25,
             equality_qp_initial_guess=1, reset_y=0, reset_rho=0,
             update_rho_with_heuristic=0, warm_start_y=0, reserved=0)
    for k, v in kw.items():
        if k not in d:
            raise TypeError(f"unknown CSQP parameter {k!r}")
        d[k] = type(d[k])(v)
    return CsqpParams(**d)


class MimGpuError(RuntimeError):
    pass


# every symbol include/mimgpu.h declares (tests check the .so exports them all)
SYMBOLS = (
    "mimgpu_abi_version", "mimgpu_device_count", "mimgpu_last_error", "mimgpu_create",
    "mimgpu_destroy", "mimgpu_get_shape", "mimgpu_upload", "mimgpu_set_reg",
    "mimgpu_set_active", "mimgpu_set_field", "mimgpu_sqp_direction",
    "mimgpu_sqp_direction_host", "mimgpu_sqp_direction_host_records", "mimgpu_get_record_layout",
    "mimgpu_upload_records", "mimgpu_model_attach", "mimgpu_sqp_solve", "mimgpu_solve_info", "mimgpu_fddp_direction", "mimgpu_csqp_calc_norms", "mimgpu_csqp_reset_rho",
    "mimgpu_csqp_direction", "mimgpu_csqp_direction_host", "mimgpu_download", "mimgpu_field_ptr", "mimgpu_scalar",
    "mimgpu_iters_status", "mimgpu_converged", "mimgpu_summary_reduce", "mimgpu_launch_count", "mimgpu_stream_sync", "mimgpu_measure_peaks",
    "mimgpu_sqp_direction_records_noalloc", "mimgpu_noalloc_scratch_needed",
)

_lib = None


def _declare(lib):
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.mimgpu_abi_version.restype = C.c_int
    lib.mimgpu_device_count.argtypes = [C.POINTER(C.c_int)]
    lib.mimgpu_last_error.restype = C.c_char_p
    lib.mimgpu_create.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(Shape)]
    lib.mimgpu_destroy.argtypes = [vp]
    lib.mimgpu_destroy.restype = None
    lib.mimgpu_get_shape.argtypes = [vp, C.POINTER(Shape)]
    lib.mimgpu_upload.argtypes = [vp, C.POINTER(Derivs), C.c_int, vp]
    lib.mimgpu_set_reg.argtypes = [vp, vp, vp, vp]
    lib.mimgpu_set_active.argtypes = [vp, vp, i32, vp]
    lib.mimgpu_set_field.argtypes = [vp, C.c_int, vp, vp]
    lib.mimgpu_sqp_direction.argtypes = [vp, C.POINTER(SqpParams), vp]
    lib.mimgpu_sqp_direction_host.argtypes = [vp, C.POINTER(Derivs), C.POINTER(SqpParams), vp, vp, vp, vp, i32]
    lib.mimgpu_sqp_direction_host_records.argtypes = [vp, vp, vp, C.POINTER(SqpParams), vp, vp, vp, vp, i32]
    lib.mimgpu_get_record_layout.argtypes = [vp, C.POINTER(RecordLayout)]
    lib.mimgpu_upload_records.argtypes = [vp, vp, vp, C.c_int, vp]
    lib.mimgpu_model_attach.argtypes = [vp, C.POINTER(PolyModel), C.c_int, vp]
    lib.mimgpu_sqp_solve.argtypes = [vp, C.POINTER(SqpSolveParams), vp, vp, C.POINTER(i32), vp]
    lib.mimgpu_solve_info.argtypes = [vp, vp, vp, vp]
    lib.mimgpu_fddp_direction.argtypes = [vp, C.POINTER(SqpParams), vp]
    lib.mimgpu_csqp_calc_norms.argtypes = [vp, vp]
    lib.mimgpu_csqp_reset_rho.argtypes = [vp, C.POINTER(CsqpParams), vp]
    lib.mimgpu_csqp_direction.argtypes = [vp, C.POINTER(CsqpParams), C.c_int, vp]
    lib.mimgpu_csqp_direction_host.argtypes = [vp, C.POINTER(Derivs), C.POINTER(CsqpParams), C.c_int, C.c_int, C.c_int,
                                               vp, vp, vp, vp, vp, vp, vp, i32]
    lib.mimgpu_download.argtypes = [vp, C.c_int, vp, vp]
    lib.mimgpu_field_ptr.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(i64), vp]
    lib.mimgpu_scalar.argtypes = [vp, C.c_int, vp, vp]
    lib.mimgpu_iters_status.argtypes = [vp, vp, vp, vp]
    lib.mimgpu_converged.argtypes = [vp, vp, vp]
    lib.mimgpu_summary_reduce.argtypes = [vp, C.c_double, C.POINTER(Summary), vp]
    lib.mimgpu_stream_sync.argtypes = [vp, vp]
    lib.mimgpu_measure_peaks.argtypes = [C.c_int, c_double_p, c_double_p, c_double_p]
    lib.mimgpu_launch_count.argtypes = [vp]
    lib.mimgpu_launch_count.restype = i64
    lib.mimgpu_sqp_direction_records_noalloc.argtypes = [C.POINTER(Shape), vp, vp, C.POINTER(SqpParams), vp, vp, vp, vp,
                                                         vp, vp]
    lib.mimgpu_noalloc_scratch_needed.argtypes = [C.POINTER(Shape)]
    lib.mimgpu_noalloc_scratch_needed.restype = i64
    for name in SYMBOLS:
        f = getattr(lib, name)
        if f.restype is C.c_int and name not in ("mimgpu_abi_version",):
            pass
    return lib


def load_library(path: os.PathLike | None = None):
    """dlopen libmimgpu.so (built in-tree by __graft_entry__.build()).  Raises
    MimGpuError if it is absent -- there is no CPU fallback."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    # MIMGPU_LIB: another build of the same library (tuning variants, scripts/build_variant.sh)
    p = Path(path) if path else Path(os.environ.get("MIMGPU_LIB") or LIB_PATH)
    if not p.exists():
        raise MimGpuError(
            f"{p} not found: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
    lib = _declare(C.CDLL(str(p)))
    if path is None:
        _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != OK:
        lib = load_library()
        msg = lib.mimgpu_last_error()
        raise MimGpuError(f"{what} failed with code {rc}: {msg.decode() if msg else ''}")


def as_ptr(a: np.ndarray | None):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], "need C-contiguous float64"
    return a.ctypes.data
