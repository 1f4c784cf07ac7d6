"""DEVELOPMENT TOOL: batches of 70..400-gene problems (CTA-wide groups): evaluation time."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO); sys.path.insert(0, os.path.join(REPO, "tools"))
from lassim_b200._capi import Engine
from workloads import synthetic_problem, uniform_population
for n, N in [(int(a), int(b)) for a, b in (x.split(':') for x in sys.argv[1:])] or ((70, 512), (150, 256), (200, 128), (300, 64), (600, 16)):
    p = synthetic_problem(n, seed=n)
    X = uniform_population(p["lb"], p["ub"], N, seed=1)
    eng = Engine(0)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    eng.eval(X); ms = []
    for _ in range(2):
        a = eng.eval(X, details=True); ms.append(eng.last_kernel_ms())
    print(f"n={n} {eng.kernel_name}: {min(ms):.1f} ms for {N} x {eng.trajectories_per_candidate} trajectories, attempts {int(a['counters'][:, :2].sum())}, status ok {(a['status'] == 0).all()}", flush=True)
    eng.close()
