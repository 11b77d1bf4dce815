"""mimgpu_sqp_direction_records_noalloc (include/mimgpu.h): the ctx-less, allocation-free variant of
the knot-record SQP direction (scratch = a static __device__ array).  No counterpart in the reference;
checked against the existing entry point
(mimgpu_upload_records + mimgpu_sqp_direction + downloads), which the other suites check against the
oracle.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from mim_solvers_b200 import _capi

OUT_SCALARS = ("KKT", "gap_norm", "merit", "x_grad_norm", "u_grad_norm")


def _existing(batch, params):
    """The existing entry points on a ctx: records in, dx/du/lag_mul/scalars/status out."""
    from mim_solvers_b200.engine import BatchEngine
    e = BatchEngine(shape=batch.shape(), device=0)
    rec = e.pack_records(batch)
    cost = np.ascontiguousarray(batch.arrays["cost"])
    e.upload_records(rec, cost)
    e.sqp_direction(params)
    out = {n: e.field(n) for n in ("dx", "du", "lag_mul")}
    out.update({n: e.scalar(n) for n in OUT_SCALARS})
    out["status"] = e.status.copy()
    e.close()
    return rec, cost, out


def _noalloc(lib, torch, batch, rec, cost, params, bufs):
    """One call of the new entry point through the C ABI; every buffer is caller-owned device memory
    that exists BEFORE the call (bufs), so the call itself has nothing to allocate."""
    sh = batch.shape()
    B, T, nx, nu = sh.B, sh.T, sh.nx, sh.nu
    bufs["records"][:rec.size].copy_(torch.from_numpy(rec.ravel()))
    bufs["cost"][:B].copy_(torch.from_numpy(cost))
    for n in ("dx", "du", "lag_mul", "scalars"):
        bufs[n].fill_(float("nan"))
    bufs["status"].fill_(-1)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.mimgpu_sqp_direction_records_noalloc(
        C.byref(sh), bufs["records"].data_ptr(), bufs["cost"].data_ptr(), C.byref(params), bufs["dx"].data_ptr(),
        bufs["du"].data_ptr(), bufs["lag_mul"].data_ptr(), bufs["scalars"].data_ptr(), bufs["status"].data_ptr(), None)
    assert rc == _capi.OK, lib.mimgpu_last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    scal = bufs["scalars"][:_capi.N_SCALARS * B].cpu().numpy().reshape(_capi.N_SCALARS, B)
    out = {"dx": bufs["dx"][:B * (T + 1) * nx].cpu().numpy().reshape(B, T + 1, nx),
           "du": bufs["du"][:B * T * nu].cpu().numpy().reshape(B, T, nu),
           "lag_mul": bufs["lag_mul"][:B * (T + 1) * nx].cpu().numpy().reshape(B, T + 1, nx),
           "status": bufs["status"][:B].cpu().numpy()}
    out.update({n: scal[_capi.SCALARS[n]] for n in OUT_SCALARS})
    return out


@pytest.mark.gpu
def test_no_alloc_sqp_direction_matches_existing_entry_point():
    import torch
    from helpers import assert_close
    from mim_solvers_b200.problems import make_random_batch
    lib = _capi.load_library()
    torch.cuda.set_device(0)
    params = _capi.default_sqp_params()
    # caller-owned device buffers, sized for the larger of the two calls
    cap = dict(B=48, T=40, nx=12, nu=6)
    nrec = cap["B"] * (cap["T"] * 600 + 200)
    bufs = {"records": torch.empty(nrec, dtype=torch.float64, device="cuda"),
            "cost": torch.empty(cap["B"], dtype=torch.float64, device="cuda"),
            "dx": torch.empty(cap["B"] * (cap["T"] + 1) * cap["nx"], dtype=torch.float64, device="cuda"),
            "du": torch.empty(cap["B"] * cap["T"] * cap["nu"], dtype=torch.float64, device="cuda"),
            "lag_mul": torch.empty(cap["B"] * (cap["T"] + 1) * cap["nx"], dtype=torch.float64, device="cuda"),
            "scalars": torch.empty(_capi.N_SCALARS * cap["B"], dtype=torch.float64, device="cuda"),
            "status": torch.empty(cap["B"], dtype=torch.int32, device="cuda")}
    # two calls in a row on the shared scratch: a larger problem set first, then a smaller one of another
    # shape whose spill records land on top of the first call's leftovers
    for (B, T, nx, nu, first) in ((48, 40, 12, 6, 0), (29, 17, 8, 4, 500)):
        batch = make_random_batch(B, T, nx, nu, first=first)
        rec, cost, ref = _existing(batch, params)
        got = _noalloc(lib, torch, batch, rec, cost, params, bufs)
        assert np.array_equal(got["status"], ref["status"]) and not ref["status"].any()
        for n in ("dx", "du", "lag_mul"):
            assert np.isfinite(got[n]).all(), n
            assert_close(got[n], ref[n], 1e-9, n, per_problem=True)
        for n in OUT_SCALARS:
            assert_close(got[n], ref[n], 1e-9, n, per_problem=True)
    # the same inputs again (third call): identical bits, the scratch's old contents do not leak into the results
    again = _noalloc(lib, torch, batch, rec, cost, params, bufs)
    for n in ("dx", "du", "lag_mul", "KKT"):
        assert np.array_equal(again[n], got[n]), n


@pytest.mark.gpu
def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    lib = _capi.load_library()
    torch.cuda.set_device(0)
    params = _capi.default_sqp_params()
    ok = _capi.Shape(177, 100, 12, 6, 0, 0)
    need = lib.mimgpu_noalloc_scratch_needed(C.byref(ok))
    assert 0 < need <= _capi.NOALLOC_SCRATCH_BYTES
    d = torch.zeros(16, dtype=torch.float64, device="cuda")   # never touched: the call must fail before launching
    for sh, code in ((_capi.Shape(4096, 100, 12, 6, 0, 0), _capi.ERR_INVALID),      # beyond the maximum
                     (_capi.Shape(8, 10, 36, 12, 0, 0), _capi.ERR_UNSUPPORTED),     # no record kernel for the shape
                     (_capi.Shape(8, 10, 12, 6, 3, 0), _capi.ERR_UNSUPPORTED)):     # constrained
        rc = lib.mimgpu_sqp_direction_records_noalloc(C.byref(sh), d.data_ptr(), None, C.byref(params), d.data_ptr(),
                                                      d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), None)
        assert rc == code, (rc, lib.mimgpu_last_error())
    big = _capi.Shape(4096, 100, 12, 6, 0, 0)
    assert lib.mimgpu_noalloc_scratch_needed(C.byref(big)) > _capi.NOALLOC_SCRATCH_BYTES
    rc = lib.mimgpu_sqp_direction_records_noalloc(C.byref(big), d.data_ptr(), None, C.byref(params), d.data_ptr(),
                                                  d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), None)
    assert rc == _capi.ERR_INVALID and b"MIMGPU_NOALLOC_SCRATCH_BYTES" in lib.mimgpu_last_error()
    torch.cuda.synchronize()


def test_no_alloc_scratch_accounting_on_cpu():
    """No GPU needed: the size check is host arithmetic (spill records of all problems + qp_iters)."""
    lib = _capi.load_library()
    sh = _capi.Shape(10, 20, 12, 6, 0, 0)
    srec = 12 * 12 + 12 + 12 * 6 + 6          # [Vxx | Vx | K | k] of one knot (kernels_small_gj.cuh::GjCfg, dense rows at nx = 12)
    spill = -(-10 * 21 * srec * 8 // 256) * 256
    assert lib.mimgpu_noalloc_scratch_needed(C.byref(sh)) == spill + 256
    assert lib.mimgpu_noalloc_scratch_needed(C.byref(_capi.Shape(10, 20, 36, 12, 0, 0))) == -1
