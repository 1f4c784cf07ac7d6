"""DEVELOPMENT TOOL: a tiny evaluation of every kernel family for compute-sanitizer."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200 import Engine
from workloads import synthetic_problem
th2 = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
g = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_late.npz")))
for method in ("dop853", "dp45"):
    eng = Engine(0); eng.set_method(method)
    eng.set_tolerances(1e-5, 1e-5)          # few steps: the sanitizer is slow
    eng.set_problem(th2["data"], th2["sigma"], th2["times"], th2["y0"], th2["mask"], th2["pert"], 0.5)
    r = eng.eval(g["X"][:5], details=True); print(method, "th2", r["cost"][:2], r["status"])
    eng.trajectories(g["X"][:3])
    eng.set_peripheral(g["X"][0], np.random.default_rng(0).uniform(0.1, 1, (4, 6)), 0.2, np.full(4, 0.5))
    print(method, "periph", eng.eval_peripheral(np.arange(4), np.random.default_rng(1).uniform(0, 1, (4, 14))))
    for n in (50, 200):
        p = synthetic_problem(n, seed=n)
        eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
        X = np.random.default_rng(2).uniform(p["lb"], p["ub"], (3, p["lb"].size)) * 0.1
        print(method, n, eng.eval(X))
    eng.close()
