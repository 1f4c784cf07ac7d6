"""DEVELOPMENT TOOL: how many time series of an EVOLVING population of 1024 pass the hand-over cap
(the stage-parallel kernel holds 296 at a time) and what an evaluation of it takes."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import _capi
if os.environ.get("LASSIM_LIB"):
    _capi._LIB_PATH = os.path.join(REPO, "lassim_b200", os.environ["LASSIM_LIB"])
from lassim_b200._capi import Engine
p = dict(np.load(os.path.join(REPO, "tests", "golden", "th2_problem.npz")))
for seed in (1000, 1003):
    eng = Engine(0)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    X0 = np.random.default_rng(seed).uniform(p["lb"], p["ub"], size=(1, 1024, 87))
    eng.de_init(X0, p["lb"], p["ub"], variant=3, f=0.85, cr=1.0, seed=42)
    for gen in range(0, 41):
        if gen % 5 == 0:
            X, F = eng.de_get()
            eng.eval(X[0]); ms = []
            for _ in range(3):
                eng.eval(X[0]); ms.append(eng.last_kernel_ms())
            st = eng.last_traj_steps(1024)[:, 0]
            print("seed %d gen %2d: best %.2f  eval %.2f ms (stage-parallel %.2f)  over cap %d  longest %d  median %d"
                  % (seed, gen, np.nanmin(F), np.median(ms), eng.last_stage_parallel_ms(), int((st > 768).sum()),
                     int(st.max()), int(np.median(st))), flush=True)
        eng.de_evolve(1)
    eng.close()
