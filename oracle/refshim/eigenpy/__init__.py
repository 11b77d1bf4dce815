"""Stand-in for eigenpy: only LLT, as used by python/qp_solvers/stagewise_qp.py:504-506."""
import numpy as np
import scipy.linalg as scl


class LLT:
    def __init__(self, A):
        self._c = scl.cho_factor(np.asarray(A, dtype=float), lower=True)

    def solve(self, b):
        return scl.cho_solve(self._c, np.asarray(b, dtype=float))

    def matrixL(self):
        return np.tril(self._c[0])
