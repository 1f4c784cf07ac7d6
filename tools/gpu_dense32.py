"""DEVELOPMENT TOOL: 17..32-gene problems (one warp per trajectory, dense rows in registers):
evaluation time of a batch, for A/B builds of that shape (LASSIM_VARIANT_LIB = another .so)."""
import os, sys, ctypes
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
import lassim_b200._capi as capi
if os.environ.get("LASSIM_VARIANT_LIB"):
    capi._LIB_PATH = os.path.join(REPO, "lassim_b200", os.environ["LASSIM_VARIANT_LIB"])
from lassim_b200._capi import Engine
from workloads import synthetic_problem, uniform_population
for n in (20, 24, 32):
    p = synthetic_problem(n, seed=n, in_degree=6)
    X = uniform_population(p["lb"], p["ub"], 2048, seed=1)
    eng = Engine(0)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    eng.eval(X); ms = []
    for _ in range(3):
        a = eng.eval(X, details=True); ms.append(eng.last_kernel_ms())
    print(f"n={n} {eng.kernel_name}: {min(ms):.1f} ms for 2048 x {eng.trajectories_per_candidate} trajectories, attempts {int(a['counters'][:, :2].sum())}, status ok {(a['status'] == 0).all()}", flush=True)
    eng.close()
