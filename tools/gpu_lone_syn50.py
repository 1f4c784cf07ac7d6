"""DEVELOPMENT TOOL: the longest time series of the 50-gene golden set alone (for ncu)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import synthetic_problem
p = synthetic_problem(50, seed=50)
g = np.load(os.path.join(REPO, "tests", "golden", "syn50.npz"))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.0)
eng.trajectories(g["X"])
st = eng.last_traj_steps(len(g["X"]))[:, 0]
row = int(st.argmax())
eng.trajectories(g["X"][row]); eng.trajectories(g["X"][row])
n = int(eng.last_traj_steps(1)[0, 0])
print("row", row, "attempts", n, "ms", eng.last_kernel_ms(), "us/attempt", 1e3 * eng.last_kernel_ms() / n, eng.kernel_name)
