"""TEST INFRASTRUCTURE ONLY: host emulation of the CUDA kernel bodies (tests/emu/emu.cpp).

Lets the CPU test-suite replay golden vectors through the *kernel source* (serial CTA
emulation) in the GPU-less build container.  The product never imports this module.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from mim_solvers_b200 import _capi
from mim_solvers_b200.problems import OCPBatch

_DIR = Path(__file__).resolve().parent
_ROOT = _DIR.parent.parent
_LIB = _DIR / "_build" / "libmimgpu_emu.so"
_lib = None

_PTRS = ("Fx", "Fu", "Lxx", "Lxu", "Luu", "Lx", "Lu", "fs", "LxxT", "LxT",
         "Gx", "Gu", "g", "glb", "gub", "GxT", "gT", "glbT", "gubT", "cost",
         "dx", "du", "lag", "dxt", "dut", "y", "z", "rho",
         "Vxx", "Vx", "K", "kff", "Qu", "Quuk", "L", "Vxx_fs", "scal",
         "qp_iters", "status", "converged", "preg_arr", "dreg_arr", "active")


class KArgs(C.Structure):
    _fields_ = [("sh", _capi.Shape)] + [(n, C.c_void_p) for n in _PTRS] + [
        ("n_active", C.c_int32), ("stage_fx", C.c_int32), ("pack", C.c_void_p), ("spill", C.c_void_p),
        ("first", C.c_int32), ("pad_", C.c_int32)]


def build(force=False):
    srcs = [_DIR / "emu.cpp", _ROOT / "mim_solvers_b200" / "csrc" / "kernels_generic.cuh",
            _ROOT / "include" / "mimgpu.h"]
    m = max(s.stat().st_mtime for s in srcs)
    if force or not _LIB.exists() or _LIB.stat().st_mtime < m:
        _LIB.parent.mkdir(exist_ok=True)
        subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                        "-o", str(_LIB), str(_DIR / "emu.cpp")], check=True)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(_LIB))
        assert _lib.emu_sizeof_kargs() == C.sizeof(KArgs), (_lib.emu_sizeof_kargs(), C.sizeof(KArgs))
    return _lib


_DERIV_MAP = dict(g_lb="glb", g_ub="gub", g_lbT="glbT", g_ubT="gubT")
_FIELD_MAP = dict(dx="dx", du="du", lag_mul="lag", dx_tilde="dxt", du_tilde="dut", y="y", z="z",
                  rho_vec="rho", Vxx="Vxx", Vx="Vx", K="K", k="kff", Qu="Qu", Quuk="Quuk", Quu_llt="L")


class EmuEngine:
    def __init__(self, batch: OCPBatch, stage_fx: int = 1):
        self.batch = batch
        self._shape = batch.shape()
        sh = self._shape
        self._f = {n: np.zeros(sh.field_shape(n)) for n in _FIELD_MAP}
        self._vxx_fs = np.zeros((sh.B, sh.T, sh.nx))
        self._scal = np.zeros((_capi.N_SCALARS, sh.B))
        self._scal[_capi.SCALARS["KKT"]] = np.inf
        self._scal[_capi.SCALARS["rho_sparse"]] = 1e-1
        self.qp_iters = np.zeros(sh.B, dtype=np.int32)
        self.status = np.zeros(sh.B, dtype=np.int32)
        self.converged = np.zeros(sh.B, dtype=np.int32)
        k = KArgs()
        k.sh = sh
        for n in _capi.DERIV_FIELDS:
            a = batch.arrays[n]
            setattr(k, _DERIV_MAP.get(n, n), a.ctypes.data if a.size else None)
        for n, m in _FIELD_MAP.items():
            setattr(k, m, self._f[n].ctypes.data)
        k.Vxx_fs = self._vxx_fs.ctypes.data
        k.scal = self._scal.ctypes.data
        k.qp_iters = self.qp_iters.ctypes.data
        k.status = self.status.ctypes.data
        k.converged = self.converged.ctypes.data
        k.n_active = sh.B
        k.stage_fx = stage_fx
        self._k = k

    def field(self, name):
        return self._f[name].copy()

    def upload(self, batch):
        self.batch = batch
        for n in _capi.DERIV_FIELDS:
            a = batch.arrays[n]
            setattr(self._k, _DERIV_MAP.get(n, n), a.ctypes.data if a.size else None)

    def set_reg(self, preg=None, dreg=None):
        self._preg = None if preg is None else np.ascontiguousarray(preg, dtype=np.float64).copy()
        self._dreg = None if dreg is None else np.ascontiguousarray(dreg, dtype=np.float64).copy()
        self._k.preg_arr = None if preg is None else self._preg.ctypes.data
        self._k.dreg_arr = None if dreg is None else self._dreg.ctypes.data

    def set_active(self, idx=None):
        if idx is None:
            self._k.active, self._k.n_active = None, self._shape.B
            return
        self._active = np.ascontiguousarray(idx, dtype=np.int32).copy()
        self._k.active, self._k.n_active = self._active.ctypes.data, len(self._active)

    def scalar(self, name):
        return self._scal[_capi.SCALARS[name]].copy()

    def sqp_direction(self, params=None, with_kkt=True):
        p = params or _capi.default_sqp_params()
        assert lib().emu_sqp_direction(C.byref(self._k), C.byref(p)) == 0

    def fddp_direction(self, params=None):
        p = params or _capi.default_sqp_params()
        assert lib().emu_fddp_direction(C.byref(self._k), C.byref(p)) == 0

    def csqp_calc_norms(self):
        assert lib().emu_csqp_calc_norms(C.byref(self._k)) == 0

    def csqp_reset_rho(self, params=None):
        p = params or _capi.default_csqp_params()
        assert lib().emu_csqp_reset_rho(C.byref(self._k), C.byref(p)) == 0

    def csqp_direction(self, params=None, with_kkt=True):
        p = params or _capi.default_csqp_params()
        assert lib().emu_csqp_direction(C.byref(self._k), C.byref(p), int(with_kkt)) == 0
