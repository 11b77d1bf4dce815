"""CPU: the C oracle replays the golden vectors produced by the reference's own NumPy
solvers (tests/golden/make_golden.py).  This is what pins the oracle."""
import pytest

import golden_checks as gc
from helpers import golden_names


def _make(oracle_lib):
    return lambda batch: oracle_lib.Oracle(batch, nthreads=1)


@pytest.mark.parametrize("name", golden_names("sqp_"))
def test_oracle_sqp_golden(oracle_lib, name):
    gc.check_sqp_golden(_make(oracle_lib), name)


@pytest.mark.parametrize("name", golden_names("csqp_"))
def test_oracle_csqp_golden(oracle_lib, name):
    gc.check_csqp_golden(_make(oracle_lib), name)
