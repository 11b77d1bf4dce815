import sys, os, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from mim_solvers_b200 import _capi
from mim_solvers_b200.engine import BatchEngine
from mim_solvers_b200.problems import make_random_batch
from oracle import oracle
from parity_checks import compare_engines_csqp
import subprocess
cases = [(64, 100, 76, 32, 32, 0, 3), (64, 30, 76, 32, 32, 0, 3), (64, 100, 76, 32, 32, 0, 1), (64, 100, 36, 12, 16, 0, 3)]
which = int(sys.argv[1]) if len(sys.argv) > 1 else -1
if which < 0:
    for i in range(len(cases)):
        r = subprocess.run([sys.executable, __file__, str(i)], capture_output=True, text=True)
        print(cases[i], "rc", r.returncode, (r.stdout.strip().splitlines() or [""])[-1][:200], (r.stderr.strip().splitlines() or [""])[-1][:200], flush=True)
else:
    B, T, nx, nu, ng, ngT, iters = cases[which]
    batch = make_random_batch(B, T, nx, nu, ng=ng, ngT=ngT, box_rows=ng)
    p = _capi.default_csqp_params(max_qp_iters=iters, eps_abs=0.0, eps_rel=0.0, equality_qp_initial_guess=0)
    import torch
    def mk(b):
        e = BatchEngine(shape=b.shape())
        e.attach_device({k: torch.from_numpy(v).cuda() for k, v in b.arrays.items()})
        return e
    steps = os.environ.get("STEPS", "full")
    if steps == "full":
        compare_engines_csqp(mk, lambda b: oracle.Oracle(b), batch, params=p, n_calls=1)
    else:
        e = mk(batch); e.csqp_calc_norms(); e.csqp_reset_rho(p); e.sync(); print("pre ok", flush=True)
        e.csqp_direction(p, with_kkt=(steps != "nokkt")); e.sync()
    print("OK")
