import os, sys
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
from lassim_b200._capi import Engine
from workloads import uniform_population
p = dict(np.load("/root/repo/tests/golden/th2_problem.npz"))
X = uniform_population(p["lb"], p["ub"], 1024, seed=0)
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 1.0)
res = {}
for mode in ("off", "on", "stages"):
    eng.set_speculation(mode); res[mode] = eng.eval(X, details=True)
eng.set_tolerances(1e-12, 1e-12); eng.set_speculation("off"); t = eng.eval(X, details=True)
for mode in res:
    d = np.abs(res[mode]["cost_main"] - t["cost_main"]) / np.abs(t["cost_main"])
    order = np.argsort(-np.nan_to_num(d))[:5]
    print(mode, "cost_main vs tight(1e-12): worst", [(int(i), "%.1e" % d[i]) for i in order], "pct99 %.1e" % np.nanpercentile(d, 99))
