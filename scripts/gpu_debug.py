"""Run one CSQP scenario on the GPU against the oracle (debug aid).
usage: gpu_debug.py B T nx nu ng ngT eq_guess max_iters eps"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
from mim_solvers_b200 import _capi
from mim_solvers_b200.engine import BatchEngine
from mim_solvers_b200.problems import make_random_batch
from oracle import oracle
from parity_checks import compare_engines_csqp
B, T, nx, nu, ng, ngT, eq, it = (int(v) for v in sys.argv[1:9])
eps = float(sys.argv[9])
batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, box_rows=min(12, ng))
p = _capi.default_csqp_params(max_qp_iters=it, equality_qp_initial_guess=eq, eps_abs=eps, eps_rel=eps)
try:
    g, o = compare_engines_csqp(lambda b: BatchEngine(b), lambda b: oracle.Oracle(b), batch, params=p, n_calls=1)
    print("OK", sys.argv[1:], g.qp_iters.tolist())
except Exception as e:
    print("FAIL", sys.argv[1:], str(e)[:300])
