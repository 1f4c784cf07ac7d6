"""DEVELOPMENT TOOL: compare library builds (occupancy variants) on the Th2 population."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import _capi
lib = sys.argv[1]
_capi._LIB_PATH = os.path.join(REPO, "lassim_b200", lib)
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
for method in ("dop853", "dp45"):
    eng = Engine(0)
    eng.set_method(method)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    eng.eval(X)
    eng.eval(X); ms1 = eng.last_kernel_ms()
    Xs = np.tile(X, (16, 1))
    eng.eval(Xs); r = eng.eval(Xs, details=True); ms16 = eng.last_kernel_ms()
    att = int(r["counters"][:, :2].sum())
    print(lib, method, "N=1024: %.2f ms | N=16384: %.2f ms -> %.0f evals/s, %.2f TFLOP/s algorithmic"
          % (ms1, ms16, 16384 / ms16 * 1e3, att * eng.flops_per_attempt / (ms16 * 1e-3) / 1e12))
    eng.close()
