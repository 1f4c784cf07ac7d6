"""DEVELOPMENT TOOL: configs 3 and 4 (reduced) for a library variant (LASSIM_LIB=...)."""
import os, sys, time
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200 import _capi
if os.environ.get("LASSIM_LIB"):
    _capi._LIB_PATH = os.path.join(REPO, "lassim_b200", os.environ["LASSIM_LIB"])
from lassim_b200 import Engine
from workloads import stiff_population, synthetic_problem, uniform_population
eng = Engine(0)
for name, n, N in (("config3", 50, 1024), ("config4/1024", 200, 1024), ("n=100", 100, 512), ("n=400", 400, 128)):
    p = synthetic_problem(n, seed=n)
    X = stiff_population(p, N, seed=200) if n == 200 else uniform_population(p["lb"], p["ub"], N, seed=n)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    eng.eval(X[:64])
    r = eng.eval(X, details=True)
    print(os.environ.get("LASSIM_LIB", "default"), name, eng.kernel_name, "ms %.1f" % eng.last_kernel_ms(),
          "attempts %d" % int(r["counters"][:, :2].sum()), "bad", int((r["status"] != 0).sum()), flush=True)
