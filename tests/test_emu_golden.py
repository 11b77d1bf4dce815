"""CPU: the CUDA kernel SOURCE (kernels_generic.cuh), compiled for the host with the
serial CTA emulator of tests/emu, replays the golden vectors and matches the oracle on
random batches.  Debug aid for the GPU-less container; the GPU parity tests proper are
tests/test_gpu_*.py."""
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "emu"))
import golden_checks as gc  # noqa: E402
from helpers import golden_names  # noqa: E402
from parity_checks import compare_engines_csqp, compare_engines_fddp, compare_engines_sqp  # noqa: E402


@pytest.fixture(scope="module")
def emu():
    import emu_engine
    emu_engine.build()
    return emu_engine


@pytest.mark.parametrize("name", golden_names("sqp_"))
def test_emu_sqp_golden(emu, name):
    gc.check_sqp_golden(lambda b: emu.EmuEngine(b), name)


@pytest.mark.parametrize("name", golden_names("csqp_"))
def test_emu_csqp_golden(emu, name):
    gc.check_csqp_golden(lambda b: emu.EmuEngine(b), name)


@pytest.mark.parametrize("stage_fx", [0, 1])
def test_emu_vs_oracle_random(emu, oracle_lib, stage_fx):
    from mim_solvers_b200.problems import make_random_batch
    mk_e = lambda b: emu.EmuEngine(b, stage_fx=stage_fx)
    mk_o = lambda b: oracle_lib.Oracle(b, nthreads=1)
    compare_engines_sqp(mk_e, mk_o, make_random_batch(3, 9, 6, 3))
    compare_engines_fddp(mk_e, mk_o, make_random_batch(2, 7, 5, 2))
    compare_engines_csqp(mk_e, mk_o, make_random_batch(3, 8, 6, 3, ng=5, ngT=4, eq_frac=0.1))


def test_emu_vs_oracle_rho_heuristic(emu, oracle_lib):
    """update_rho_with_heuristic branch of update_lagrangian_parameters (src/csqp.cpp:1245-1257):
    dual norms without the G^T products; kernel source (host emulation) against the oracle."""
    from mim_solvers_b200 import _capi
    from mim_solvers_b200.problems import make_random_batch
    p = _capi.default_csqp_params(update_rho_with_heuristic=1, max_qp_iters=300)
    batch = make_random_batch(4, 10, 6, 3, ng=4, ngT=3, violate_frac=0.0)
    a, b = compare_engines_csqp(lambda x: emu.EmuEngine(x), lambda x: oracle_lib.Oracle(x, nthreads=1), batch, params=p)
    p0 = _capi.default_csqp_params(max_qp_iters=300)
    c = oracle_lib.Oracle(batch, nthreads=1)
    c.csqp_calc_norms(); c.csqp_reset_rho(p0); c.csqp_direction(p0)
    # the branch is really taken: the heuristic norms differ from the G^T-product norms
    assert not np.allclose(b.scalar("norm_dual"), c.scalar("norm_dual"))
