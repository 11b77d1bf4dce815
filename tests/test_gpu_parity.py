"""GPU parity tests proper: the CUDA path (through the C ABI of libmimgpu.so) against the
golden vectors (reference NumPy solvers) and against the CPU oracle on seeded batches.
Bar: 1e-9 relative on dx/du/lag_mul/KKT (1e-8 against the NumPy goldens, see
golden_checks.py), identical qp_iters."""
import numpy as np
import pytest

import golden_checks as gc
from helpers import golden_names
from parity_checks import compare_engines_csqp, compare_engines_fddp, compare_engines_sqp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.engine import BatchEngine
    lib = _capi.load_library()           # raises if the extension is missing: no fallback
    n = __import__("ctypes").c_int(0)
    assert lib.mimgpu_device_count(__import__("ctypes").byref(n)) == 0 and n.value > 0, "no CUDA device"
    return BatchEngine


@pytest.mark.parametrize("name", golden_names("sqp_"))
def test_gpu_sqp_golden(eng, name):
    gc.check_sqp_golden(lambda b: eng(b), name)


@pytest.mark.parametrize("name", golden_names("csqp_"))
def test_gpu_csqp_golden(eng, name):
    gc.check_csqp_golden(lambda b: eng(b), name)


SQP_SHAPES = [(16, 100, 12, 6), (5, 40, 12, 6), (4, 13, 7, 3), (3, 10, 36, 12), (2, 6, 64, 32),
              (2, 5, 76, 32), (3, 9, 5, 1), (2, 4, 3, 4)]


@pytest.mark.parametrize("B,T,nx,nu", SQP_SHAPES)
def test_gpu_sqp_vs_oracle(eng, oracle_lib, B, T, nx, nu):
    from mim_solvers_b200.problems import make_random_batch
    compare_engines_sqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), make_random_batch(B, T, nx, nu))


@pytest.mark.parametrize("B,T,nx,nu", [(4, 30, 12, 6), (2, 8, 36, 12)])
def test_gpu_fddp_vs_oracle(eng, oracle_lib, B, T, nx, nu):
    from mim_solvers_b200.problems import make_random_batch
    compare_engines_fddp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), make_random_batch(B, T, nx, nu))


CSQP_SHAPES = [  # B, T, nx, nu, ng, ngT, kwargs
    (6, 40, 12, 6, 3, 3, dict(box_rows=0)),
    (4, 20, 12, 6, 0, 0, {}),
    (3, 12, 36, 12, 16, 0, dict(box_rows=12)),
    (2, 10, 36, 12, 36, 36, dict(box_rows=36)),
    (3, 10, 16, 8, 24, 16, dict(eq_frac=0.1)),
    (2, 5, 64, 32, 96, 64, dict(eq_frac=0.1)),
    (2, 4, 76, 32, 32, 0, dict(box_rows=32)),
    # the other shapes of the record / streaming CSQP kernels: 8/4 and 4/2 (DMMA full sweeps), 6/2 and 10/4
    # (nx % 4 != 0: shape-generic full sweeps inside the specialised kernel), odd row counts
    (5, 9, 8, 4, 12, 8, dict(eq_frac=0.1)),
    (4, 30, 4, 2, 2, 2, {}),
    (3, 8, 6, 2, 4, 2, {}),
    (3, 11, 10, 4, 5, 3, {}),
]


@pytest.mark.parametrize("B,T,nx,nu,ng,ngT,kw", CSQP_SHAPES)
def test_gpu_csqp_vs_oracle(eng, oracle_lib, B, T, nx, nu, ng, ngT, kw):
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, **kw)
    p = _capi.default_csqp_params(max_qp_iters=200)
    compare_engines_csqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch, params=p, n_calls=2)


@pytest.mark.parametrize("B,T,nx,nu,ng,ngT,kw", [(6, 40, 12, 6, 3, 3, dict(box_rows=0)), (3, 12, 36, 12, 16, 0, dict(box_rows=12)),
                                                 (2, 6, 64, 32, 40, 16, {}), (3, 9, 7, 3, 5, 2, {})])
def test_gpu_csqp_rho_heuristic(eng, oracle_lib, B, T, nx, nu, ng, ngT, kw):
    """update_rho_with_heuristic = 1 (src/csqp.cpp:1245-1257) on the streaming, direct and generic kernels."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, violate_frac=0.0, **kw)
    p = _capi.default_csqp_params(max_qp_iters=300, update_rho_with_heuristic=1)
    compare_engines_csqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch, params=p, n_calls=2)


@pytest.mark.parametrize("B,T,nx,nu,ng,ngT", [(16, 12, 76, 32, 32, 0), (8, 9, 48, 24, 30, 20), (16, 12, 36, 12, 16, 0)])
def test_gpu_csqp_device_attached_separate_allocations(oracle_lib, B, T, nx, nu, ng, ngT):
    """Each derivative array in its own cudaMalloc block (the caller-owned layout of the drop-in): an index one
    element outside an array faults here instead of landing in a neighbouring array of the engine's single upload
    buffer.  Covers the streaming kernel in staged and in direct (L2-prefetch) mode."""
    import torch
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.engine import BatchEngine
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, box_rows=min(ng, nx))
    p = _capi.default_csqp_params(max_qp_iters=40)

    def mk(b):
        e = BatchEngine(shape=b.shape())
        e._keepalive = {k: torch.from_numpy(v).cuda() for k, v in b.arrays.items()}
        e.attach_device(e._keepalive)
        return e
    compare_engines_csqp(mk, lambda b: oracle_lib.Oracle(b), batch, params=p, n_calls=2)


def test_gpu_csqp_benchmark_protocol(eng, oracle_lib):
    """benchmarks/solo12.cpp:213-216: eps_abs = eps_rel = 0, 50 iterations, no eq. guess."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(4, 15, 12, 6, ng=6, ngT=6)
    p = _capi.default_csqp_params(max_qp_iters=50, eps_abs=0.0, eps_rel=0.0, equality_qp_initial_guess=0)
    g, o = compare_engines_csqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch, params=p, n_calls=1)
    assert np.all(g.qp_iters == 51)


def test_gpu_llt_failure_status(eng, oracle_lib):
    """Fault injection: indefinite Luu -> status LLT at the failing knot (src/ddp.cpp:415-419)."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(4, 8, 6, 3)
    batch.mat("Luu")[1, 5] = -10.0 * np.eye(3)
    g, o = eng(batch), oracle_lib.Oracle(batch)
    g.sqp_direction(); o.sqp_direction()
    assert np.array_equal(g.status, np.asarray(o.status))
    assert g.status[1] == (_capi.STATUS_LLT | (5 << 8)) and g.status[0] == 0
    # retry only the failed problem with a larger regularisation (src/sqp.cpp:103-116)
    g.set_active([1])
    preg = np.full(4, 1e-9); preg[1] = 20.0
    g.sqp_direction(preg=preg, dreg=preg)
    o.sqp_direction(preg=preg, dreg=preg)
    assert g.status[1] == 0
    np.testing.assert_allclose(g.field("dx")[1], o.field("dx")[1], rtol=0, atol=1e-9 * np.abs(o.field("dx")[1]).max())


def test_gpu_summary_and_device_inputs(eng, oracle_lib):
    import torch
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(7, 12, 12, 6)
    g = eng(shape=batch.shape())
    dev = {k: torch.from_numpy(v).cuda() for k, v in batch.arrays.items()}
    g.attach_device(dev)
    g.sqp_direction()
    o = oracle_lib.Oracle(batch); o.sqp_direction()
    s = g.summary(1e-6)
    assert abs(s["max_kkt"] - o.scalar("KKT").max()) <= 1e-9 * o.scalar("KKT").max()
    assert s["n_failed"] == 0 and s["sum_qp_iters"] == 0
    assert g.launch_count >= 2


@pytest.mark.parametrize("B,T,nx,nu", [(37, 25, 12, 6), (9, 7, 8, 4), (6, 5, 16, 8), (5, 30, 4, 2), (5, 9, 6, 2),
                                       (7, 11, 10, 4)])
def test_gpu_small_kernel_vs_oracle(eng, oracle_lib, B, T, nx, nu):
    """Shapes served by the warp-per-problem kernel (kernels_small.cuh); ragged B (not a
    multiple of the warps per CTA), feasible and infeasible variants."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu)
    compare_engines_sqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch)
    compare_engines_sqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch,
                        params=_capi.default_sqp_params(preg=0.0, dreg=0.0, is_feasible=True))


def test_gpu_small_kernel_failure_and_subset(eng, oracle_lib):
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(6, 9, 12, 6)
    batch.mat("Luu")[4, 3] = -10.0 * np.eye(6)
    batch.mat("Lxx")[2, 6][0, 0] = np.nan
    g, o = eng(batch), oracle_lib.Oracle(batch)
    g.sqp_direction(); o.sqp_direction()
    assert np.array_equal(g.status, np.asarray(o.status)), (g.status, np.asarray(o.status))
    assert g.status[4] == (_capi.STATUS_LLT | (3 << 8)) and (g.status[2] & 0xff) == _capi.STATUS_NONFINITE
    ok = g.status == 0
    np.testing.assert_allclose(g.field("dx")[ok], o.field("dx")[ok], rtol=0, atol=1e-9 * np.abs(o.field("dx")[ok]).max())
    g.set_active([4, 1])
    preg = np.full(6, 1e-9); preg[4] = 30.0
    g.sqp_direction(preg=preg, dreg=preg)
    o.sqp_direction(preg=preg, dreg=preg)
    assert g.status[4] == 0 and (g.status[2] & 0xff) == _capi.STATUS_NONFINITE   # problem 2 untouched
    np.testing.assert_allclose(g.field("dx")[4], o.field("dx")[4], rtol=0, atol=1e-9 * np.abs(o.field("dx")[4]).max())


@pytest.mark.parametrize("n_chunks", [1, 3, 8])
def test_gpu_sqp_direction_host_matches_stepwise(eng, n_chunks):
    """mimgpu_sqp_direction_host (overlapped slices) == upload + direction + download, bit for bit;
    ragged last slice; the members stay downloadable afterwards."""
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(37, 20, 12, 6)
    a = eng(batch)
    a.sqp_direction()
    b = eng(shape=batch.shape())
    sh = batch.shape()
    out = {n: np.empty(sh.field_shape(n)) for n in ("dx", "du", "lag_mul")}
    kkt = np.empty(sh.B)
    b.sqp_direction_host(batch, None, dx=out["dx"], du=out["du"], lag_mul=out["lag_mul"], kkt=kkt, n_chunks=n_chunks)
    for n in out:
        assert np.array_equal(out[n], a.field(n)), n
    assert np.array_equal(kkt, a.scalar("KKT"))
    for n in ("K", "Vxx", "Quu_llt"):
        assert np.array_equal(b.field(n), a.field(n)), n


def test_gpu_knot_records_layout(eng, oracle_lib):
    """Second input layout of the ABI (mimgpu_upload_records): the producer writes knot records, no
    k_pack_knots pass.  Host records, device-attached records (zero copy, re-read at every call) and
    the chunked host path are bit-identical to the field-major path; entry points the record kernel
    does not serve rebuild the field-major arrays on the device."""
    import torch
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(37, 20, 12, 6)
    a = eng(batch)
    a.sqp_direction()
    kkt, st0 = a.scalar("KKT"), a.status.copy()
    ref = {n: a.field(n) for n in ("dx", "du", "lag_mul", "K", "Vxx", "Qu", "Quu_llt")}
    # asking for Qu_ / Quu_llt_ reruns the full-spill kernel for the members only: the direction, the
    # scalars and the status of the fast launch are not rewritten
    assert np.array_equal(a.scalar("KKT"), kkt) and np.array_equal(a.status, st0)
    assert np.array_equal(a.field("dx"), ref["dx"])
    b = eng(shape=batch.shape())
    L = b.record_layout()
    assert L.problem_doubles == batch.T * L.rec_doubles + L.term_doubles
    rec = b.pack_records(batch)
    cost = np.ascontiguousarray(batch.arrays["cost"])
    l0 = b.launch_count
    b.upload_records(rec, cost)
    b.sqp_direction()
    assert b.launch_count - l0 == 1, "records path must not launch a re-layout kernel"
    for n in ("dx", "du", "lag_mul", "K", "Vxx"):
        assert np.array_equal(b.field(n), ref[n]), n
    assert np.array_equal(b.scalar("KKT"), kkt), np.abs(b.scalar("KKT") - kkt).max()
    assert np.array_equal(b.scalar("merit"), a.scalar("merit")), (b.scalar("merit")[:4], a.scalar("merit")[:4], cost[:4])
    for n in ("Qu", "Quu_llt"):   # full-spill recomputation needs the field-major arrays: rebuilt on demand
        assert np.array_equal(b.field(n), ref[n]), n
    # device-attached records, mutated in place between two calls
    c = eng(shape=batch.shape())
    drec = torch.from_numpy(rec).cuda()
    c.attach_records(drec, torch.from_numpy(cost).cuda())
    c.sqp_direction()
    assert np.array_equal(c.field("dx"), ref["dx"])
    batch2 = make_random_batch(37, 20, 12, 6, first=1000)
    drec.copy_(torch.from_numpy(c.pack_records(batch2)))
    c.sqp_direction()
    o = oracle_lib.Oracle(batch2); o.sqp_direction()
    from helpers import assert_close
    for n in ("dx", "du", "lag_mul"):
        assert_close(c.field(n), o.field(n), 1e-9, n, per_problem=True)
    assert_close(c.field("Qu"), o.field("Qu"), 1e-9, "Qu after in-place change", per_problem=True)
    # chunked host path on records
    d = eng(shape=batch.shape())
    sh = batch.shape()
    out = {n: np.empty(sh.field_shape(n)) for n in ("dx", "du", "lag_mul")}
    k2 = np.empty(sh.B)
    d.sqp_direction_host_records(rec, cost, None, dx=out["dx"], du=out["du"], lag_mul=out["lag_mul"], kkt=k2, n_chunks=5)
    for n in out:
        assert np.array_equal(out[n], ref[n]), n
    assert np.array_equal(k2, kkt)
    # FDDP on a records ctx (needs the field-major arrays)
    d.fddp_direction()
    o1 = oracle_lib.Oracle(batch); o1.fddp_direction()
    assert_close(d.scalar("dg"), o1.scalar("dg"), 1e-9, "dg", per_problem=True)


@pytest.mark.parametrize("B,T,nx,nu", [(9, 15, 12, 6), (4, 8, 36, 12)])
def test_gpu_attached_inputs_are_reread(eng, oracle_lib, B, T, nx, nu):
    """MIMGPU_MEM_DEVICE inputs are the caller's memory: updating them in place and calling the
    direction again must use the NEW values (the small shapes re-layout them at every call)."""
    import torch
    from helpers import assert_close
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu)
    g = eng(shape=batch.shape())
    dev = {k: torch.from_numpy(v).cuda() for k, v in batch.arrays.items()}
    g.attach_device(dev)
    g.sqp_direction()
    o = oracle_lib.Oracle(batch); o.sqp_direction()
    assert_close(g.field("dx"), o.field("dx"), 1e-9, "dx", per_problem=True)
    batch2 = make_random_batch(B, T, nx, nu, first=77)
    for k, v in batch2.arrays.items():
        dev[k].copy_(torch.from_numpy(v))
    g.sqp_direction()
    o2 = oracle_lib.Oracle(batch2); o2.sqp_direction()
    for n in ("dx", "du", "lag_mul", "K", "Qu"):
        assert_close(g.field(n), o2.field(n), 1e-9, n, per_problem=True)


def test_gpu_member_flags_do_not_go_stale(eng, oracle_lib):
    """After a fast SQP launch the members Vxx_ Vx_ K_ k_ live in spill records until asked for; a
    CSQP direction on the same ctx, a new upload or mimgpu_set_field must drop that pending state."""
    from helpers import assert_close
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(5, 12, 12, 6)
    g = eng(batch)
    g.sqp_direction()
    p = _capi.default_csqp_params(max_qp_iters=30)
    g.csqp_calc_norms(); g.csqp_reset_rho(p); g.csqp_direction(p)
    o = oracle_lib.Oracle(batch)
    o.csqp_calc_norms(); o.csqp_reset_rho(p); o.csqp_direction(p)
    for n in ("K", "k", "Vx", "Vxx", "Qu"):
        assert_close(g.field(n), o.field(n), 1e-9, "after CSQP: " + n, per_problem=True)
    s = g.summary(1e-6)
    assert s["sum_qp_iters"] == int(np.asarray(o.qp_iters).sum())
    g.sqp_direction()
    g.set_field("K", np.zeros(batch.shape().field_shape("K")))
    assert not g.field("K").any()


@pytest.mark.parametrize("switch", ["MIMGPU_CSQP_ADMM", "MIMGPU_CSQP_STREAM", "MIMGPU_CSQP_REC", "MIMGPU_REC_W1=0",
                                    "MIMGPU_REC_W1=1"])
@pytest.mark.parametrize("B,T,nx,nu,ng,ngT,kw", [(6, 40, 12, 6, 3, 3, dict(box_rows=0)), (3, 12, 36, 12, 16, 0, dict(box_rows=12)),
                                                 (3, 10, 16, 8, 24, 16, dict(eq_frac=0.1)), (3, 9, 32, 16, 20, 8, {}),
                                                 (2, 7, 8, 4, 0, 0, {}), (3, 8, 6, 2, 4, 2, {}), (3, 11, 10, 4, 5, 3, {})])
def test_gpu_csqp_both_kernels(eng, oracle_lib, monkeypatch, switch, B, T, nx, nu, ng, ngT, kw):
    """The CSQP kernels (condensed-chain kernels_admm.cuh, sweep-by-sweep kernels_stream.cuh, record-based
    kernels_rec.cuh for nx <= 16) are dispatched by shape; each is forced here on shapes another one serves
    by default (MIMGPU_CSQP_REC only acts on the shapes the record kernel is compiled for)."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    name, _, val = switch.partition("=")
    monkeypatch.setenv(name, val or "1")     # read once, at mimgpu_create
    if name == "MIMGPU_REC_W1":
        monkeypatch.setenv("MIMGPU_CSQP_REC", "1")  # both forms of the record kernel: multi-warp CTA, one warp per problem
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, **kw)
    p = _capi.default_csqp_params(max_qp_iters=120)
    compare_engines_csqp(lambda b: eng(b), lambda b: oracle_lib.Oracle(b), batch, params=p, n_calls=2)


@pytest.mark.parametrize("B,T,nx,nu,ng,ngT,n_chunks", [(37, 12, 12, 6, 3, 3, 5), (9, 10, 36, 12, 16, 0, 4), (5, 6, 7, 3, 4, 2, 2)])
def test_gpu_csqp_direction_host_matches_stepwise(eng, B, T, nx, nu, ng, ngT, n_chunks):
    """mimgpu_csqp_direction_host (overlapped slices, row n3) == upload + calc norms + rho reset + direction +
    download, bit for bit; ragged last slice; a second call continues from the persistent y / rho state."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT)
    p = _capi.default_csqp_params(max_qp_iters=60)
    a = eng(batch)
    b = eng(shape=batch.shape())
    sh = batch.shape()
    for call in range(2):
        a.csqp_calc_norms(); a.csqp_reset_rho(p); a.csqp_direction(p, with_kkt=True)
        out = {n: np.empty(sh.field_shape(n)) for n in ("dx_tilde", "du_tilde", "lag_mul", "y", "z")}
        kkt, it = np.empty(sh.B), np.empty(sh.B, dtype=np.int32)
        b.csqp_direction_host(batch, p, kkt=kkt, qp_iters=it, n_chunks=n_chunks, **out)
        for n in out:
            assert np.array_equal(out[n], a.field(n)), (call, n)
        assert np.array_equal(kkt, a.scalar("KKT")) and np.array_equal(it, a.qp_iters)
        assert np.array_equal(b.field("rho_vec"), a.field("rho_vec"))
