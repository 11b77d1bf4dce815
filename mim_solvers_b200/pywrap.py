"""mim_solvers_b200.pywrap -- the drop-in Python surface: the batched C++ solvers (pybind11 module
`_pywrap` over include/mim_solvers_b200/batch_solvers.hpp) under the reference's names
(bindings/mim_solvers/__init__.py): SolverSQP, SolverCSQP, SolverFDDP, CallbackAbstract,
CallbackVerbose, CallbackLogger.  Where the reference returns one float per solver these return an
array with one entry per problem of the batch.

    from mim_solvers_b200 import pywrap as mim_solvers
    solver = mim_solvers.SolverSQP(problem)          # problem: LQBatchProblem / PolyBatchProblem / adapter
    solver.setCallbacks([mim_solvers.CallbackVerbose()])
    converged = solver.solve(xs0, us0, maxiter=100)
"""
import copy

try:
    from ._pywrap import (BatchShootingProblem, CallbackAbstract, CallbackVerbose, LQBatchProblem,  # noqa: F401
                          PolyBatchProblem, SolverAbstract, SolverCSQP, SolverFDDP, SolverSQP)
except ImportError as e:  # pragma: no cover
    raise ImportError("mim_solvers_b200._pywrap is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(g++ + pybind11; links libmimgpu.so)") from e


class CallbackLogger(CallbackAbstract):
    """bindings/mim_solvers/__init__.py:14-51: appends the solver's state to `convergence_data` at
    every iteration (arrays with one entry per problem of the batch)."""

    def __init__(self, verbose=False):
        super().__init__()
        self.convergence_data = {}

    def __call__(self, solver, solver_type):
        def safe_append(key, value):
            self.convergence_data.setdefault(key, []).append(value)

        if solver_type not in ("CSQP", "SQP"):
            raise NotImplementedError("CallbackLogger is implemented for CSQP and SQP.")
        safe_append("us", copy.copy(solver.us))
        safe_append("xs", copy.copy(solver.xs))
        safe_append("fs", copy.copy(solver.fs))
        safe_append("iter", solver.iter)
        safe_append("cost", solver.cost)
        safe_append("merit", solver.merit)
        safe_append("preg", solver.preg)
        safe_append("dreg", solver.dreg)
        safe_append("step", solver.stepLength)
        if solver_type == "CSQP":
            safe_append("gap_norm", solver.gap_norm)
            safe_append("constraint_norm", solver.constraint_norm)
            safe_append("qp_iter", solver.qp_iters)
        safe_append("KKT", solver.KKT)


__all__ = ["SolverSQP", "SolverCSQP", "SolverFDDP", "SolverAbstract", "CallbackAbstract", "CallbackVerbose",
           "CallbackLogger", "LQBatchProblem", "PolyBatchProblem", "BatchShootingProblem"]
