"""DEVELOPMENT TOOL: one lone trajectory (the longest of the Th2 population) for ncu."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
X = np.random.default_rng(0).uniform(p["lb"], p["ub"], size=(1024, 87))
worst = 330 if len(sys.argv) < 2 else int(sys.argv[1])
eng.eval(X[:8])
Y = eng.trajectories(X[worst])
print("steps", eng.last_traj_steps(1)[0, 0], "ms", eng.last_kernel_ms())
