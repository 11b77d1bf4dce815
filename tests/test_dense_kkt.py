"""The oracle (CPU) and the CUDA path (GPU) against dense solves of the same QP: the unconstrained SQP sub-problem
as one KKT system (src/kkt.cpp:235-293 layout) and converged CSQP solutions against the certified optimum of the
inequality QP in the ProxQP layout (src/csqp_proxqp.cpp:296-345)."""
import pytest

import dense_kkt
import dense_qp


def test_oracle_matches_dense_kkt(oracle_lib):
    dense_kkt.check_against_dense(lambda b: oracle_lib.Oracle(b, nthreads=1))


@pytest.mark.gpu
def test_gpu_matches_dense_kkt():
    from mim_solvers_b200.engine import BatchEngine
    dense_kkt.check_against_dense(lambda b: BatchEngine(b))


def test_oracle_csqp_matches_certified_dense_qp(oracle_lib):
    assert dense_qp.check_against_dense_qp(lambda b: oracle_lib.Oracle(b, nthreads=0)) >= 12


@pytest.mark.gpu
def test_gpu_csqp_matches_certified_dense_qp():
    from mim_solvers_b200.engine import BatchEngine
    assert dense_qp.check_against_dense_qp(lambda b: BatchEngine(b)) >= 12
