"""CPU: the C++ host layer (BatchSolverSQP / CSQP / FDDP of include/mim_solvers_b200/
batch_solvers.hpp) driven through the kernel SOURCE compiled for the host (tests/cpp/
mimgpu_emu_shim.cpp, test infrastructure).  Same checks on libmimgpu.so: test_gpu_cpp_host.py."""
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "emu"))
import cpp_host  # noqa: E402


@pytest.fixture(scope="module")
def binary():
    return cpp_host.build_emu()


@pytest.fixture(scope="module")
def factory():
    import emu_engine
    emu_engine.build()
    return lambda batch: emu_engine.EmuEngine(batch)


@pytest.mark.parametrize("check", cpp_host.CHECKS, ids=lambda f: f.__name__)
def test_cpp_host_emu(binary, factory, tmp_path, check):
    check(binary, tmp_path, factory)


def test_cpp_host_has_no_cpu_fallback(tmp_path):
    """The driver linked with libmimgpu.so must fail loudly (Engine throws on mimgpu_create)
    when there is no CUDA device."""
    import ctypes
    import subprocess
    import solver_checks as sc
    from mim_solvers_b200 import _capi
    n = ctypes.c_int(0)
    _capi.load_library().mimgpu_device_count(ctypes.byref(n))
    if n.value > 0:
        pytest.skip("CUDA device present")
    b = cpp_host.build_gpu()
    cpp_host.write_problem(tmp_path / "p.bin", sc.make_lq(B=1, T=3))
    r = subprocess.run([str(b), "sqp", str(tmp_path / "p.bin"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 3 and "mimgpu_create failed" in r.stderr, (r.returncode, r.stderr)


def test_cpp_api_defaults_and_argument_checks():
    """tests/cpp/test_api.cpp: reference defaults (tests/test_solvers.cpp:83-112), invalid arguments
    throw std::invalid_argument (src/ddp.cpp:547-638, src/sqp.cpp:308-311), tiny solves."""
    import subprocess
    b = cpp_host.build_api_emu()
    r = subprocess.run([str(b)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), (r.stdout, r.stderr)
