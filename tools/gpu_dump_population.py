"""DEVELOPMENT TOOL: costs, parts and knock-down end states of the seed-0 Th2 population (1024
rows) at the default tolerances and at 1e-12, for offline comparison with
tests/golden/th2_uniform_full.npz.  Writes gpurun_out/th2_population_gpu.npz."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
out = {}
for tag, tol in (("", 1.49012e-8), ("tight_", 1e-12)):
    eng.set_tolerances(tol, tol)
    r = eng.eval(X, details=True)
    for k in ("cost", "cost_main", "pert_cost", "status"):
        out[tag + k] = r[k]
    out[tag + "E"] = eng.last_end_states(1024)
    out[tag + "steps"] = eng.last_traj_steps(1024)
os.makedirs(os.path.join(REPO, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(REPO, "gpurun_out", "th2_population_gpu.npz"), **out)
print("ok", np.isnan(out["cost"]).sum(), np.isnan(out["tight_cost"]).sum())
