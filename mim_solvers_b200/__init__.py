"""B200-native batched OCP-QP engine: drop-in for the computeDirection hot path of
mim_solvers' SolverSQP / SolverCSQP / SolverFDDP (include/mimgpu.h)."""
from . import _capi, problems  # noqa: F401
from ._capi import MimGpuError  # noqa: F401
from .problems import OCPBatch, make_random_batch, make_config  # noqa: F401

__version__ = "0.1.0"
