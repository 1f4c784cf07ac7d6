"""DEVELOPMENT TOOL: one lone time series with a chosen time-series kernel (for ncu)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import uniform_population
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
X = uniform_population(p["lb"], p["ub"], 1024, seed=0)
eng.set_speculation(sys.argv[2] if len(sys.argv) > 2 else "stages")
row = int(sys.argv[1]) if len(sys.argv) > 1 else 820
Y = eng.trajectories(X[row])
print("attempts", int(eng.last_traj_steps(1)[0, 0]), "ms", eng.last_kernel_ms())
