"""CPU: host logic of the batched SolverSQP / SolverCSQP / SolverFDDP (solve loop, line
search, filter, regularisation retry) driven through the kernel SOURCE compiled for the host
(tests/emu, test infrastructure).  The same checks run on the CUDA library in
tests/test_gpu_solvers.py."""
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "emu"))
import solver_checks as sc  # noqa: E402


@pytest.fixture(scope="module")
def factory():
    import emu_engine
    emu_engine.build()
    return lambda batch: emu_engine.EmuEngine(batch)


@pytest.mark.parametrize("check", [sc.check_defaults, sc.check_sqp_lq, sc.check_csqp_equals_sqp_without_constraints,
                                   sc.check_csqp_constrained, sc.check_regularisation_retry, sc.check_remove_reg_on_reuse,
                                   sc.check_fddp_lq],
                         ids=lambda f: f.__name__)
def test_solver_host(factory, check):
    check(factory)


@pytest.mark.parametrize("name", sc.GOLDEN_SOLVE)
def test_solve_loop_against_reference_python_solve(factory, name):
    """Row a15: the solve() loop (filter / merit acceptance, alpha < 1, iteration counts) against fixtures
    recorded from the reference's own NumPy SQP.solve() -- not against itself."""
    sc.check_solve_golden(name, factory)


def test_default_engine_is_cuda_only():
    """No CPU fallback: without a device the default engine factory raises."""
    import ctypes
    from mim_solvers_b200 import MimGpuError, _capi
    from mim_solvers_b200.solvers import SolverSQP
    n = ctypes.c_int(0)
    _capi.load_library().mimgpu_device_count(ctypes.byref(n))
    if n.value > 0:
        pytest.skip("CUDA device present")
    s = SolverSQP(sc.make_lq(B=1, T=3))
    with pytest.raises(MimGpuError):
        s.solve(maxiter=1)
