"""ctypes front-end of the CPU ORACLE (test infrastructure, NOT product code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It loads oracle/_build/libmim_oracle.so (built by
oracle/Makefile from oracle/mim_oracle.c, the plain-C restatement of
/root/reference/src/{ddp,sqp,csqp,fddp}.cpp).
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from mim_solvers_b200 import _capi
from mim_solvers_b200.problems import OCPBatch

_DIR = Path(__file__).resolve().parent
_LIB_PATH = _DIR / "_build" / "libmim_oracle.so"
_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


class _OraState(C.Structure):
    _fields_ = [("shape", _capi.Shape)] + [(n, _dp) for n in (
        "dx", "du", "lag_mul", "dxt", "dut", "y", "z", "rho_vec", "Vxx", "Vx", "K", "kff",
        "Qu", "Quuk", "L", "Vxx_fs", "z_prev", "z_relaxed", "inv_rho_vec", "scal")] + [
        (n, _ip) for n in ("qp_iters", "status", "converged")]


def build(force: bool = False) -> Path:
    src_m = max((_DIR / f).stat().st_mtime for f in ("mim_oracle.c", "mim_oracle.h"))
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src_m:
        subprocess.run(["make", "-C", str(_DIR)], check=True, capture_output=True)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not _LIB_PATH.exists():
            build()
        L = C.CDLL(str(_LIB_PATH))
        L.ora_state_create.restype = C.POINTER(_OraState)
        L.ora_state_create.argtypes = [C.POINTER(_capi.Shape)]
        L.ora_state_destroy.argtypes = [C.POINTER(_OraState)]
        L.ora_state_destroy.restype = None
        P = C.POINTER
        L.ora_sqp_direction.argtypes = [P(_OraState), P(_capi.Derivs), P(_capi.SqpParams), C.c_int, C.c_int]
        L.ora_sqp_direction_reg.argtypes = [P(_OraState), P(_capi.Derivs), P(_capi.SqpParams),
                                            C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ora_fddp_direction.argtypes = [P(_OraState), P(_capi.Derivs), P(_capi.SqpParams), C.c_int]
        L.ora_csqp_calc_norms.argtypes = [P(_OraState), P(_capi.Derivs), C.c_int]
        L.ora_csqp_reset_rho.argtypes = [P(_OraState), P(_capi.Derivs), P(_capi.CsqpParams), C.c_int]
        L.ora_csqp_direction.argtypes = [P(_OraState), P(_capi.Derivs), P(_capi.CsqpParams), C.c_int, C.c_int]
        _lib = L
    return _lib


_FIELD_TO_MEMBER = {"dx": "dx", "du": "du", "lag_mul": "lag_mul", "dx_tilde": "dxt",
                    "du_tilde": "dut", "y": "y", "z": "z", "rho_vec": "rho_vec", "Vxx": "Vxx",
                    "Vx": "Vx", "K": "K", "k": "kff", "Qu": "Qu", "Quuk": "Quuk",
                    "Quu_llt": "L"}


class Oracle:
    """One solver object per batch: mirrors the members of SolverSQP / SolverCSQP /
    SolverFDDP for B problems (state persists between calls like the C++ objects)."""

    def __init__(self, batch: OCPBatch, nthreads: int = 0):
        self.batch = batch
        self.nthreads = nthreads
        self._shape = batch.shape()
        self._st = lib().ora_state_create(C.byref(self._shape))
        if not self._st:
            raise ValueError("bad shape")
        self._derivs = batch.derivs()

    def __del__(self):
        if getattr(self, "_st", None):
            lib().ora_state_destroy(self._st)
            self._st = None

    # ---- state views -----------------------------------------------------------
    def field(self, name: str) -> np.ndarray:
        """numpy view (no copy) in C-ABI storage layout."""
        shp = self._shape.field_shape(name)
        n = int(np.prod(shp))
        if n == 0:
            return np.zeros(shp)
        ptr = getattr(self._st.contents, _FIELD_TO_MEMBER[name])
        return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(shp)

    def scalar(self, name: str) -> np.ndarray:
        B = self._shape.B
        a = np.ctypeslib.as_array(self._st.contents.scal, shape=(_capi.N_SCALARS * B,))
        return a.reshape(_capi.N_SCALARS, B)[_capi.SCALARS[name]]

    @property
    def qp_iters(self):
        return np.ctypeslib.as_array(self._st.contents.qp_iters, shape=(self._shape.B,))

    @property
    def status(self):
        return np.ctypeslib.as_array(self._st.contents.status, shape=(self._shape.B,))

    @property
    def converged(self):
        return np.ctypeslib.as_array(self._st.contents.converged, shape=(self._shape.B,))

    # ---- reference entry points --------------------------------------------------
    def sqp_direction(self, params=None, with_kkt=True, preg=None, dreg=None):
        p = params or _capi.default_sqp_params()
        if preg is None and dreg is None:
            rc = lib().ora_sqp_direction(self._st, C.byref(self._derivs), C.byref(p),
                                         int(with_kkt), self.nthreads)
        else:
            preg = None if preg is None else np.ascontiguousarray(preg, dtype=np.float64)
            dreg = None if dreg is None else np.ascontiguousarray(dreg, dtype=np.float64)
            rc = lib().ora_sqp_direction_reg(
                self._st, C.byref(self._derivs), C.byref(p),
                None if preg is None else preg.ctypes.data,
                None if dreg is None else dreg.ctypes.data, int(with_kkt), self.nthreads)
        assert rc == 0, rc

    def fddp_direction(self, params=None):
        p = params or _capi.default_sqp_params()
        rc = lib().ora_fddp_direction(self._st, C.byref(self._derivs), C.byref(p), self.nthreads)
        assert rc == 0, rc

    def csqp_calc_norms(self):
        rc = lib().ora_csqp_calc_norms(self._st, C.byref(self._derivs), self.nthreads)
        assert rc == 0, rc

    def csqp_reset_rho(self, params=None):
        p = params or _capi.default_csqp_params()
        rc = lib().ora_csqp_reset_rho(self._st, C.byref(self._derivs), C.byref(p), self.nthreads)
        assert rc == 0, rc

    def csqp_direction(self, params=None, with_kkt=True):
        p = params or _capi.default_csqp_params()
        rc = lib().ora_csqp_direction(self._st, C.byref(self._derivs), C.byref(p),
                                      int(with_kkt), self.nthreads)
        assert rc == 0, rc


def max_threads() -> int:
    return int(lib().ora_max_threads())
