import os, sys
import numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
from lassim_b200._capi import Engine
from workloads import uniform_population
p = dict(np.load("/root/repo/tests/golden/th2_problem.npz"))
X = uniform_population(p["lb"], p["ub"], 1024, seed=0)
eng = Engine(0)
for pert in (True, False):
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"] if pert else None, 0.5)
    for mode in ("stages", "stages2"):
        eng.set_speculation(mode)
        eng.eval(X); ms=[]
        for _ in range(3):
            eng.eval(X); ms.append(eng.last_kernel_ms())
        st = eng.last_traj_steps(1024)[:,0]
        print("knock-downs", pert, mode, "%.2f ms"%min(ms), "attempt percentiles 50/90/99/99.9:", np.percentile(st,[50,90,99,99.9]).astype(int), "count>3000", int((st>3000).sum()), ">5000", int((st>5000).sum()), ">8000", int((st>8000).sum()))
eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], None, 0.5)
eng.set_speculation("stages2")
for rows in ([415,820], list(range(400,432))+[820]):
    eng.eval(X[rows]); eng.eval(X[rows]); print("rows", len(rows), "%.2f ms"%eng.last_kernel_ms())
