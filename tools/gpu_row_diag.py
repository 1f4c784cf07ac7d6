"""DEVELOPMENT TOOL: time-series cost of selected rows of the seed-0 Th2 population under every
time-series kernel and a few tolerances, against LSODA at 1e-13 (tests/golden/th2_uniform_full.npz)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
f = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_uniform_full.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
rows = [int(a) for a in sys.argv[1:]] or [373, 970, 647, 714, 415, 97]
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
tight = f["tight_cost_main"][rows]
print("rows", rows)
print("reference vs tight", np.abs(f["ref_cost_main"][rows] - tight) / tight)
for method in ("dop853", "dp45"):
    eng.set_method(method)
    for mode in ("off", "on", "stages", "stages3", "handover"):
        for tol in (1.49012e-8, 1e-9, 1e-10):
            eng.set_speculation(mode); eng.set_tolerances(tol, tol)
            r = eng.eval(X[rows], details=True)
            st = eng.last_traj_steps(len(rows))[:, 0]
            print("%-7s %-9s tol %.0e  err" % (method, mode, tol),
                  " ".join("%.1e" % e for e in np.abs(r["cost_main"] - tight) / tight), " steps", st.tolist(), flush=True)
