"""The oracle (CPU) and the CUDA path (GPU) against the dense KKT solve of the same QP."""
import pytest

import dense_kkt


def test_oracle_matches_dense_kkt(oracle_lib):
    dense_kkt.check_against_dense(lambda b: oracle_lib.Oracle(b, nthreads=1))


@pytest.mark.gpu
def test_gpu_matches_dense_kkt():
    from mim_solvers_b200.engine import BatchEngine
    dense_kkt.check_against_dense(lambda b: BatchEngine(b))
