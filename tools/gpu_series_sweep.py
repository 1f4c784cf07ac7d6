"""DEVELOPMENT TOOL: evaluation time against batch size for the three time-series
kernels (where does 'auto' have to switch back to the standard kernel?)."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import uniform_population
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
eng = Engine(0)
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
for N in (64, 256, 512, 1024, 2048, 4096, 8192):
    X = uniform_population(p["lb"], p["ub"], N, seed=0)
    line = f"N={N:5d}"
    for mode in ("off", "on", "stages", "stages2", "handover"):
        eng.set_speculation(mode)
        eng.eval(X)
        ms = []
        for _ in range(3):
            eng.eval(X); ms.append(eng.last_kernel_ms())
        line += f"  {mode} {min(ms):7.2f} ms"
    print(line, flush=True)
