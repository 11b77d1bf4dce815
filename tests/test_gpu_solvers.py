"""GPU: the batched solver classes on the CUDA library (default engine factory)."""
import pytest

import solver_checks as sc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("check", [sc.check_sqp_lq, sc.check_csqp_equals_sqp_without_constraints,
                                   sc.check_csqp_constrained, sc.check_regularisation_retry, sc.check_remove_reg_on_reuse,
                                   sc.check_fddp_lq],
                         ids=lambda f: f.__name__)
def test_solver_gpu(check):
    check(None)
