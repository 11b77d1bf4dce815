"""CPU: the C-ABI library loads and exports every symbol include/mimgpu.h declares; the
product path fails loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from mim_solvers_b200 import _capi

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    return _capi.load_library()


def test_header_symbols_exported(lib):
    hdr = (ROOT / "include" / "mimgpu.h").read_text()
    declared = set(re.findall(r"\b(mimgpu_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_capi.SYMBOLS), declared ^ set(_capi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.mimgpu_abi_version() == 3


def test_struct_sizes_match_header():
    assert C.sizeof(_capi.Shape) == 24
    assert C.sizeof(_capi.SqpParams) == 32
    assert C.sizeof(_capi.CsqpParams) == 10 * 8 + 8 * 4
    assert C.sizeof(_capi.Derivs) == 20 * 8


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from mim_solvers_b200.engine import BatchEngine
    from mim_solvers_b200.problems import make_random_batch
    with pytest.raises(_capi.MimGpuError):
        BatchEngine(make_random_batch(1, 3, 4, 2))


def test_defaults_match_reference():
    """tests/test_solvers.cpp:83-112 (CSQP defaults)."""
    p = _capi.default_csqp_params()
    assert (p.sigma, p.alpha, p.rho_sparse_base, p.rho_min, p.rho_max) == (1e-6, 1.6, 1e-1, 1e-6, 1e3)
    assert (p.eps_abs, p.eps_rel, p.max_qp_iters, p.rho_update_interval) == (1e-4, 1e-4, 1000, 25)
    assert p.adaptive_rho_tolerance == 5 and p.equality_qp_initial_guess == 1
    assert (p.reset_y, p.reset_rho, p.update_rho_with_heuristic) == (0, 0, 0)
