"""DEVELOPMENT TOOL: the knock-down end states of one row of the seed-0 Th2 population (why a
row's knock-down cost is not finite): GPU end states next to the reference's and LSODA at 1e-13."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
f = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_uniform_full.npz")))
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
row = int(sys.argv[1]) if len(sys.argv) > 1 else 798
np.set_printoptions(linewidth=220, precision=3)
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for N in (1, 1024):
    r = eng.eval(X[row:row + 1] if N == 1 else X, details=True)
    i = 0 if N == 1 else row
    E = eng.last_end_states(N)[i]
    print("batch of", N, "cost", r["cost"][i], "cost_main", r["cost_main"][i], "pert", r["pert_cost"][i], "status", r["status"][i])
    print("steps", eng.last_traj_steps(N)[i])
    print("ctrl ends gpu\n", E[1:3])
    print("ctrl ends ref\n", f["ref_ctrl_end"][row])
    print("knock ends gpu: zeros per solve", (E[3:] == 0).sum(axis=1), "nan", np.isnan(E[3:]).sum(axis=1), "inf", np.isinf(E[3:]).sum(axis=1))
    print("knock ends gpu\n", E[3:])
    print("knock ends tight\n", f["tight_knock_end"][row])
