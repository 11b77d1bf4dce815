"""BatchEngine: Python handle of one `mimgpu_ctx` (include/mimgpu.h) -- the CUDA
implementation of the computeDirection hot path for a batch of B problems of one shape.

Every compute call goes through the C ABI of libmimgpu.so.  There is NO CPU fallback:
constructing an engine without the library or without a CUDA device raises MimGpuError.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import MimGpuError, check
from .problems import OCPBatch


class BatchEngine:
    """Same method names as the reference solver members they replace
    (`computeDirection` pieces of SolverSQP / SolverCSQP / SolverFDDP)."""

    def __init__(self, batch: OCPBatch | None = None, *, shape: _capi.Shape | None = None,
                 device: int = 0, stream: int | None = None):
        self._lib = _capi.load_library()
        self._ctx = C.c_void_p()
        self._shape = batch.shape() if batch is not None else shape
        if self._shape is None:
            raise ValueError("need a batch or a shape")
        self.stream = C.c_void_p(stream) if stream else None
        self.device = device
        check(self._lib.mimgpu_create(C.byref(self._ctx), device, C.byref(self._shape)), "mimgpu_create")
        self._keep = None
        if batch is not None:
            self.upload(batch)

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._lib.mimgpu_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ------------------------------------------------------------------
    @property
    def shape(self) -> _capi.Shape:
        return self._shape

    def sync(self):
        check(self._lib.mimgpu_stream_sync(self._ctx, self.stream), "mimgpu_stream_sync")

    def upload(self, batch: OCPBatch):
        """Host arrays -> device (async H2D on the engine's stream)."""
        d = batch.derivs()
        self._keep = batch
        check(self._lib.mimgpu_upload(self._ctx, C.byref(d), _capi.MEM_HOST, self.stream), "mimgpu_upload")

    def attach_device(self, tensors: dict):
        """Device-resident inputs (torch CUDA float64 tensors in the C-ABI layout)."""
        d = _capi.Derivs()
        for n in _capi.DERIV_FIELDS:
            t = tensors.get(n)
            if t is not None and t.numel():
                assert t.is_cuda and t.is_contiguous() and str(t.dtype) == "torch.float64", n
                setattr(d, n, t.data_ptr())
        self._keep = tensors
        check(self._lib.mimgpu_upload(self._ctx, C.byref(d), _capi.MEM_DEVICE, self.stream), "mimgpu_upload")

    # ---- knot-record inputs (mimgpu_upload_records) -------------------------------
    def record_layout(self) -> _capi.RecordLayout:
        """Block offsets of the knot-record layout; raises MimGpuError for shapes without it."""
        L = _capi.RecordLayout()
        check(self._lib.mimgpu_get_record_layout(self._ctx, C.byref(L)), "mimgpu_get_record_layout")
        return L

    def pack_records(self, batch: OCPBatch, out: np.ndarray | None = None) -> np.ndarray:
        """What a producer-side packer does: every block of every knot copied to its place in the
        record layout (host side, numpy).  Returns a [B, problem_doubles] float64 array."""
        L = self.record_layout()
        B, T, nx, nu = batch.B, batch.T, batch.nx, batch.nu
        rec = np.zeros((B, L.problem_doubles)) if out is None else out
        run = rec[:, :T * L.rec_doubles].reshape(B, T, L.rec_doubles)
        term = rec[:, T * L.rec_doubles:]
        A = batch.arrays

        def put_sym(dst, off, M, n):
            """symmetric n x n block (storage [..., col, row]): packed lower triangle, column j = rows j..n-1"""
            if not L.sym_packed:
                dst[..., off:off + n * n] = M.reshape(M.shape[:-2] + (n * n,))
                return
            for j in range(n):
                o = off + j * n - j * (j - 1) // 2
                dst[..., o:o + n - j] = M[..., j, j:]
        put_sym(run, L.off_Lxx, A["Lxx"].reshape(B, T, nx, nx), nx)
        put_sym(run, L.off_Luu, A["Luu"].reshape(B, T, nu, nu), nu)
        for name, off, n in (("Lxu", L.off_Lxu, nx * nu), ("Lx", L.off_Lx, nx), ("Lu", L.off_Lu, nu),
                             ("Fx", L.off_Fx, nx * nx), ("Fu", L.off_Fu, nx * nu)):
            run[:, :, off:off + n] = A[name].reshape(B, T, n)
        run[:, :, L.off_fs:L.off_fs + nx] = A["fs"][:, :T]
        term[:, L.toff_fs:L.toff_fs + nx] = A["fs"][:, T]
        put_sym(term, L.toff_Lxx, A["LxxT"].reshape(B, nx, nx), nx)
        term[:, L.toff_Lx:L.toff_Lx + nx] = A["LxT"]
        return rec

    def upload_records(self, records: np.ndarray, cost: np.ndarray | None = None):
        """Host knot records -> device (one async H2D copy)."""
        assert records.dtype == np.float64 and records.flags["C_CONTIGUOUS"]
        self._keep = (records, cost)
        check(self._lib.mimgpu_upload_records(self._ctx, records.ctypes.data, _capi.as_ptr(cost), _capi.MEM_HOST,
                                              self.stream), "mimgpu_upload_records")

    def attach_records(self, records, cost=None):
        """Device-resident knot records (torch CUDA float64 tensors): zero-copy, no re-layout pass."""
        assert records.is_cuda and records.is_contiguous() and str(records.dtype) == "torch.float64"
        self._keep = (records, cost)
        check(self._lib.mimgpu_upload_records(self._ctx, records.data_ptr(), None if cost is None else cost.data_ptr(),
                                              _capi.MEM_DEVICE, self.stream), "mimgpu_upload_records")

    def sqp_direction_host_records(self, records: np.ndarray, cost=None, params=None, *, dx=None, du=None,
                                   lag_mul=None, kkt=None, n_chunks: int = 8):
        """mimgpu_sqp_direction_host_records: host records in, dx/du/lag_mul/KKT out, overlapped slices."""
        p = params or _capi.default_sqp_params()
        self._keep = (records, cost)
        ptr = lambda a: None if a is None else a.ctypes.data
        check(self._lib.mimgpu_sqp_direction_host_records(self._ctx, records.ctypes.data, ptr(cost), C.byref(p), ptr(dx),
                                                          ptr(du), ptr(lag_mul), ptr(kkt), int(n_chunks)),
              "mimgpu_sqp_direction_host_records")

    # ---- solve() on the device (analytic models) -----------------------------------
    def attach_model(self, arrays: dict, c3: float = 0.0, w4: float = 0.0):
        """mimgpu_model_attach: `arrays` maps A Bm f Q S R q r QT qT x0 to float64 arrays in the C-ABI block
        layouts (numpy: copied to the device; torch CUDA tensors: attached)."""
        m = _capi.PolyModel()
        m.c3, m.w4 = float(c3), float(w4)
        dev = all(hasattr(v, "data_ptr") for v in arrays.values())
        keep = {}
        for n in ("A", "Bm", "f", "Q", "S", "R", "q", "r", "QT", "qT", "x0"):
            v = arrays[n]
            if dev:
                assert v.is_cuda and v.is_contiguous()
                setattr(m, n, v.data_ptr() if v.numel() else None)
            else:
                v = np.ascontiguousarray(v, dtype=np.float64)
                setattr(m, n, v.ctypes.data if v.size else None)
            keep[n] = v
        self._keep_model = keep
        check(self._lib.mimgpu_model_attach(self._ctx, C.byref(m), _capi.MEM_DEVICE if dev else _capi.MEM_HOST,
                                            self.stream), "mimgpu_model_attach")
        self.sync()

    def sqp_solve(self, params=None, xs0=None, us0=None) -> int:
        """mimgpu_sqp_solve: SolverSQP::solve for all problems on the device; returns the number of SQP
        iterations of the slowest problem."""
        p = params or _capi.default_sqp_solve_params()
        x = None if xs0 is None else np.ascontiguousarray(xs0, dtype=np.float64)
        u = None if us0 is None else np.ascontiguousarray(us0, dtype=np.float64)
        n = C.c_int32(0)
        check(self._lib.mimgpu_sqp_solve(self._ctx, C.byref(p), _capi.as_ptr(x), _capi.as_ptr(u), C.byref(n),
                                         self.stream), "mimgpu_sqp_solve")
        return int(n.value)

    def solve_info(self):
        """(iter_ [B], outcome [B]: 1 converged, 2 stopped at the regularisation limit, 0 maxiter)."""
        B = self._shape.B
        it, oc = np.empty(B, dtype=np.int32), np.empty(B, dtype=np.int32)
        check(self._lib.mimgpu_solve_info(self._ctx, it.ctypes.data, oc.ctypes.data, self.stream), "mimgpu_solve_info")
        return it, oc

    def set_reg(self, preg=None, dreg=None):
        p = None if preg is None else np.ascontiguousarray(preg, dtype=np.float64)
        d = None if dreg is None else np.ascontiguousarray(dreg, dtype=np.float64)
        check(self._lib.mimgpu_set_reg(self._ctx, _capi.as_ptr(p), _capi.as_ptr(d), self.stream), "mimgpu_set_reg")
        self.sync()

    def set_active(self, idx=None):
        if idx is None:
            check(self._lib.mimgpu_set_active(self._ctx, None, -1, self.stream), "mimgpu_set_active")
            return
        a = np.ascontiguousarray(idx, dtype=np.int32)
        check(self._lib.mimgpu_set_active(self._ctx, a.ctypes.data, len(a), self.stream), "mimgpu_set_active")
        self.sync()

    def set_field(self, name: str, value):
        a = np.ascontiguousarray(value, dtype=np.float64).reshape(self._shape.field_shape(name))
        check(self._lib.mimgpu_set_field(self._ctx, _capi.FIELDS[name], a.ctypes.data, self.stream),
              "mimgpu_set_field")
        self.sync()

    # ---- the hot path --------------------------------------------------------------
    def sqp_direction(self, params=None, with_kkt=True, preg=None, dreg=None):
        p = params or _capi.default_sqp_params()
        if preg is not None or dreg is not None:
            self.set_reg(preg, dreg)
        check(self._lib.mimgpu_sqp_direction(self._ctx, C.byref(p), self.stream), "mimgpu_sqp_direction")

    def sqp_direction_host(self, batch: OCPBatch, params=None, *, dx=None, du=None, lag_mul=None, kkt=None,
                           n_chunks: int = 8):
        """Upload + computeDirection + download in overlapped slices (mimgpu_sqp_direction_host).
        `batch` arrays and the output arrays should be pinned host memory; synchronous."""
        p = params or _capi.default_sqp_params()
        d = batch.derivs()
        self._keep = batch
        ptr = lambda a: None if a is None else a.ctypes.data
        check(self._lib.mimgpu_sqp_direction_host(self._ctx, C.byref(d), C.byref(p), ptr(dx), ptr(du), ptr(lag_mul),
                                                  ptr(kkt), int(n_chunks)), "mimgpu_sqp_direction_host")

    def fddp_direction(self, params=None):
        p = params or _capi.default_sqp_params()
        check(self._lib.mimgpu_fddp_direction(self._ctx, C.byref(p), self.stream), "mimgpu_fddp_direction")

    def csqp_calc_norms(self):
        check(self._lib.mimgpu_csqp_calc_norms(self._ctx, self.stream), "mimgpu_csqp_calc_norms")

    def csqp_reset_rho(self, params=None):
        p = params or _capi.default_csqp_params()
        check(self._lib.mimgpu_csqp_reset_rho(self._ctx, C.byref(p), self.stream), "mimgpu_csqp_reset_rho")

    def csqp_direction(self, params=None, with_kkt=True):
        p = params or _capi.default_csqp_params()
        check(self._lib.mimgpu_csqp_direction(self._ctx, C.byref(p), int(with_kkt), self.stream),
              "mimgpu_csqp_direction")

    def csqp_direction_host(self, batch: OCPBatch, params=None, *, with_kkt=True, calc_norms=True, reset_rho=True,
                            dx_tilde=None, du_tilde=None, lag_mul=None, y=None, z=None, kkt=None, qp_iters=None,
                            n_chunks: int = 8):
        """mimgpu_csqp_direction_host: host derivatives in, ADMM results out, overlapped slices; synchronous."""
        p = params or _capi.default_csqp_params()
        d = batch.derivs()
        self._keep = batch
        ptr = lambda a: None if a is None else a.ctypes.data
        check(self._lib.mimgpu_csqp_direction_host(self._ctx, C.byref(d), C.byref(p), int(with_kkt), int(calc_norms),
                                                   int(reset_rho), ptr(dx_tilde), ptr(du_tilde), ptr(lag_mul), ptr(y),
                                                   ptr(z), ptr(kkt), ptr(qp_iters), int(n_chunks)),
              "mimgpu_csqp_direction_host")

    # ---- data out ------------------------------------------------------------------
    def field(self, name: str) -> np.ndarray:
        out = np.empty(self._shape.field_shape(name))
        if out.size:
            check(self._lib.mimgpu_download(self._ctx, _capi.FIELDS[name], out.ctypes.data, self.stream),
                  "mimgpu_download")
            self.sync()
        return out

    def download_into(self, name: str, out: np.ndarray):
        """Async D2H into a caller-provided (pinned) array; caller synchronises."""
        check(self._lib.mimgpu_download(self._ctx, _capi.FIELDS[name], out.ctypes.data, self.stream),
              "mimgpu_download")

    def scalar(self, name: str) -> np.ndarray:
        out = np.empty(self._shape.B)
        check(self._lib.mimgpu_scalar(self._ctx, _capi.SCALARS[name], out.ctypes.data, self.stream), "mimgpu_scalar")
        self.sync()
        return out

    def _ints(self):
        B = self._shape.B
        it = np.empty(B, dtype=np.int32)
        st = np.empty(B, dtype=np.int32)
        cv = np.empty(B, dtype=np.int32)
        check(self._lib.mimgpu_iters_status(self._ctx, it.ctypes.data, st.ctypes.data, self.stream),
              "mimgpu_iters_status")
        check(self._lib.mimgpu_converged(self._ctx, cv.ctypes.data, self.stream), "mimgpu_converged")
        self.sync()
        return it, st, cv

    @property
    def qp_iters(self):
        return self._ints()[0]

    @property
    def status(self):
        return self._ints()[1]

    @property
    def converged(self):
        return self._ints()[2]

    def summary(self, termination_tol: float = 1e-6) -> dict:
        s = _capi.Summary()
        check(self._lib.mimgpu_summary_reduce(self._ctx, termination_tol, C.byref(s), self.stream),
              "mimgpu_summary_reduce")
        return dict(max_kkt=s.max_kkt, sum_qp_iters=s.sum_qp_iters, max_qp_iters=s.max_qp_iters,
                    n_converged=s.n_converged, n_failed=s.n_failed)

    @property
    def launch_count(self) -> int:
        return int(self._lib.mimgpu_launch_count(self._ctx))


__all__ = ["BatchEngine", "MimGpuError"]
